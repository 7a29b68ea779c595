// cnf_solver.cu — on-device Tsit5 step driver: RK combination, embedded-error RMS norm,
// PI controller / accept-reject, Hairer initial step, final-state epilogue.
// Replaces the host-driven loop of SciMLBase.solve (call site reference src/base_icnf.jl:141)
// and inference_sol (src/base_icnf.jl:167-189).  Semantics: SURVEY.md Appendix A.
#include "cnf_internal.h"

void cnf_upload_constants() {
    static const float A[7][6] = {
        {0},
        {(float)CNF_A21},
        {(float)CNF_A31, (float)CNF_A32},
        {(float)CNF_A41, (float)CNF_A42, (float)CNF_A43},
        {(float)CNF_A51, (float)CNF_A52, (float)CNF_A53, (float)CNF_A54},
        {(float)CNF_A61, (float)CNF_A62, (float)CNF_A63, (float)CNF_A64, (float)CNF_A65},
        {(float)CNF_A71, (float)CNF_A72, (float)CNF_A73, (float)CNF_A74, (float)CNF_A75, (float)CNF_A76}};
    static const float BT[7] = {(float)CNF_BT1, (float)CNF_BT2, (float)CNF_BT3, (float)CNF_BT4,
                                (float)CNF_BT5, (float)CNF_BT6, (float)CNF_BT7};
    cudaMemcpyToSymbol(c_A, A, sizeof(A));
    cudaMemcpyToSymbol(c_BT, BT, sizeof(BT));
}

// ---------------------------------------------------------------- layout conversion
// caller layout: Julia column-major rows x B (sample-major) ; internal: SoA rows x Bs.
__global__ void k_pack_rows(const float* __restrict__ src, int rows, long long B,
                            float* __restrict__ dst, long long Bs) {
    long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bs) return;
    for (int r = 0; r < rows; ++r) dst[(long long)r * Bs + b] = (b < B) ? src[b * rows + r] : 0.0f;
}
__global__ void k_unpack_rows(const float* __restrict__ src, int rows, long long B, long long Bs,
                              float* __restrict__ dst) {
    long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    for (int r = 0; r < rows; ++r) dst[b * rows + r] = src[(long long)r * Bs + b];
}
void launch_pack_rows(cnf_handle* h, const float* src, int rows, long long B, float* dst, long long Bs) {
    k_pack_rows<<<(unsigned)((Bs + CNF_NT - 1) / CNF_NT), CNF_NT, 0, h->stream>>>(src, rows, B, dst, Bs);
    h->launches++;
}
void launch_unpack_rows(cnf_handle* h, const float* src, int rows, long long B, long long Bs, float* dst) {
    k_unpack_rows<<<(unsigned)((B + CNF_NT - 1) / CNF_NT), CNF_NT, 0, h->stream>>>(src, rows, B, Bs, dst);
    h->launches++;
}

// ---------------------------------------------------------------- RK combination (generic path)
// stage s in 1..5: S = u + dt*sum_{j<s+... } a[s][j] k_j on the D z rows (f reads only those);
// stage 6: unew = u + dt*sum a7j k_j on all R rows; stage -1: S = u + tdir*dt0init*k1 (init probe).
// (device bodies d_*: the same code runs as its own launch and, for batches of one tile, inside the
// single-CTA whole-solve kernels of cnf_tile.cu, where the control block and the buffers change while
// the kernel runs — hence no __restrict__/read-only loads on them)
__device__ __forceinline__ void d_combine(const CnfCtrl* c, const CnfBufs& bf, int stage, int also_S) {
    if (c->done) return;
    const int nrows = (stage == 6) ? c->R : c->D;
    const long long n = (long long)nrows * bf.Bs;
    const float* u = bf.slots[cnf_slot(c, c->naccept)];
    float* out = (stage == 6) ? bf.slots[cnf_slot(c, c->naccept + 1)] : bf.S;
    const float* k1 = bf.K[c->k1idx];
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        if (stage < 0) {
            out[e] = fmaf(c->tdir * c->dt0init, k1[e], u[e]);
        } else {
            float acc = c_A[stage][0] * k1[e];
            for (int j = 1; j < stage; ++j) acc = fmaf(c_A[stage][j], bf.K[j][e], acc);
            const float v = fmaf(c->dt, acc, u[e]);
            out[e] = v;
            if (also_S && e < (long long)c->D * bf.Bs) bf.S[e] = v;   // z rows of the new state
        }
    }
}
__global__ void k_combine(const CnfCtrl* c, CnfBufs bf, int stage, int also_S) { d_combine(c, bf, stage, also_S); }
static inline unsigned grid_for(long long n) {
    long long g = (n + CNF_NT - 1) / CNF_NT;
    if (g > 148LL * 16) g = 148LL * 16;
    return (unsigned)(g < 1 ? 1 : g);
}
// The same combination four samples at a time (rows are 1 KB-aligned: Bs is a multiple of 256): the scalar body above
// keeps 2-7 four-byte loads in flight per thread and reaches about half of the copy bandwidth; 16-byte accesses with
// all rows of an element group requested together reach most of it.  Per-element arithmetic (and so every result bit)
// is that of d_combine.  Used by the launch-per-stage drivers; the single-CTA whole-solve kernels keep the scalar body.
#define CNF_V4(expr_x, expr_y, expr_z, expr_w) make_float4(expr_x, expr_y, expr_z, expr_w)
__global__ void __launch_bounds__(CNF_NT) k_combine4(const CnfCtrl* __restrict__ c, CnfBufs bf, int stage, int also_S) {
    if (c->done) return;
    const int nrows = (stage == 6) ? c->R : c->D;
    const long long n4 = (long long)nrows * bf.Bs / 4, nS4 = (long long)c->D * bf.Bs / 4;
    const float4* u = reinterpret_cast<const float4*>(bf.slots[cnf_slot(c, c->naccept)]);
    float4* out = reinterpret_cast<float4*>((stage == 6) ? bf.slots[cnf_slot(c, c->naccept + 1)] : bf.S);
    float4* S4 = reinterpret_cast<float4*>(bf.S);
    const float4* k1 = reinterpret_cast<const float4*>(bf.K[c->k1idx]);
    const float4* kj[5];
#pragma unroll
    for (int j = 1; j < 6; ++j) kj[j - 1] = reinterpret_cast<const float4*>(bf.K[j]);
    const float dt = c->dt, h0 = c->tdir * c->dt0init;
    float a[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) a[j] = (stage >= 1 && j < stage) ? c_A[stage][j] : 0.0f;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n4; e += (long long)gridDim.x * blockDim.x) {
        const float4 uv = u[e], k = k1[e];
        float4 kv[5];
#pragma unroll
        for (int j = 1; j < 6; ++j)
            if (j < stage) kv[j - 1] = kj[j - 1][e];
        float4 v;
        if (stage < 0) {
            v = CNF_V4(fmaf(h0, k.x, uv.x), fmaf(h0, k.y, uv.y), fmaf(h0, k.z, uv.z), fmaf(h0, k.w, uv.w));
        } else {
            float4 acc = CNF_V4(a[0] * k.x, a[0] * k.y, a[0] * k.z, a[0] * k.w);
#pragma unroll
            for (int j = 1; j < 6; ++j)
                if (j < stage)
                    acc = CNF_V4(fmaf(a[j], kv[j - 1].x, acc.x), fmaf(a[j], kv[j - 1].y, acc.y), fmaf(a[j], kv[j - 1].z, acc.z),
                                 fmaf(a[j], kv[j - 1].w, acc.w));
            v = CNF_V4(fmaf(dt, acc.x, uv.x), fmaf(dt, acc.y, uv.y), fmaf(dt, acc.z, uv.z), fmaf(dt, acc.w, uv.w));
        }
        out[e] = v;
        if (stage >= 0 && also_S && e < nS4) S4[e] = v;   // z rows of the new state
    }
}
void launch_combine(cnf_handle* h, const CnfBufs& bf, int stage, int also_S) {
    int nrows = (stage == 6) ? h->cur_R : h->cfg.D;
    k_combine4<<<grid_for((long long)nrows * bf.Bs / 4), CNF_NT, 0, h->stream>>>(h->d_ctrl, bf, stage, also_S);
    h->launches++;
}

// ---------------------------------------------------------------- embedded error norm (A.3)
__device__ __forceinline__ void d_errnorm(CnfCtrl* c, const CnfBufs& bf) {
    if (c->done) return;
    __shared__ double sh[32];
    const int R = c->R;
    const float* u = bf.slots[cnf_slot(c, c->naccept)];
    const float* un = bf.slots[cnf_slot(c, c->naccept + 1)];
    const float* k1 = bf.K[c->k1idx];
    const float* k7 = bf.K[6 - c->k1idx];
    const float dt = c->dt, reltol = c->reltol, abstol = c->abstol;
    double acc = 0.0, su = 0.0, sd = 0.0;
    const long long n = (long long)R * bf.Bs;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        long long b = e % bf.Bs;
        if (b >= bf.B) continue;
        float a = c_BT[0] * k1[e];
        for (int j = 1; j < 6; ++j) a = fmaf(c_BT[j], bf.K[j][e], a);
        a = fmaf(c_BT[6], k7[e], a);
        {   // stiffness estimate (A.1): k7 - k6 and u_new - g6 = dt (sum_j (a_7j - a_6j) k_j + a_76 k_6)
            float gd = (c_A[6][0] - c_A[5][0]) * k1[e];
            for (int j = 1; j < 5; ++j) gd = fmaf(c_A[6][j] - c_A[5][j], bf.K[j][e], gd);
            gd = dt * fmaf(c_A[6][5], bf.K[5][e], gd);
            const float kd = k7[e] - bf.K[5][e];
            su += (double)kd * (double)kd;
            sd += (double)gd * (double)gd;
        }
        float ut = dt * a;
        float sc = fmaf(fmaxf(fabsf(u[e]), fabsf(un[e])), reltol, abstol);
        float r = ut / sc;
        if (un[e] != un[e]) r = un[e];
        acc += (double)r * (double)r;
    }
    acc = cnf_block_sum(acc, sh);
    su = cnf_block_sum(su, sh);
    sd = cnf_block_sum(sd, sh);
    if (threadIdx.x == 0) { atomicAdd(&c->errsum, acc); atomicAdd(&c->stiffu, su); atomicAdd(&c->stiffd, sd); }
}
__global__ void k_errnorm(CnfCtrl* c, CnfBufs bf) { d_errnorm(c, bf); }
// four samples at a time, like k_combine4 (nine rows are read per element: the scalar body is latency-bound).  The
// per-element quantities are those of d_errnorm; a thread adds its elements in a different order, which moves the
// FP64 sums in their last bits (as the order of the block atomics already does).
__global__ void __launch_bounds__(CNF_NT) k_errnorm4(CnfCtrl* c, CnfBufs bf) {
    if (c->done) return;
    __shared__ double sh[32];
    const int R = c->R;
    const float4* u = reinterpret_cast<const float4*>(bf.slots[cnf_slot(c, c->naccept)]);
    const float4* un = reinterpret_cast<const float4*>(bf.slots[cnf_slot(c, c->naccept + 1)]);
    const float4* k1 = reinterpret_cast<const float4*>(bf.K[c->k1idx]);
    const float4* k7 = reinterpret_cast<const float4*>(bf.K[6 - c->k1idx]);
    const float4* kj[5];
#pragma unroll
    for (int j = 1; j < 6; ++j) kj[j - 1] = reinterpret_cast<const float4*>(bf.K[j]);
    const float dt = c->dt, reltol = c->reltol, abstol = c->abstol;
    double acc = 0.0, su = 0.0, sd = 0.0;
    const long long n4 = (long long)R * bf.Bs / 4, Bs4 = bf.Bs / 4;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n4; e += (long long)gridDim.x * blockDim.x) {
        const long long b0 = (e % Bs4) * 4;
        if (b0 >= bf.B) continue;
        const float4 uv = u[e], nv = un[e], q1 = k1[e], q7 = k7[e];
        float4 kv[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) kv[j] = kj[j][e];
        const float U[4] = {uv.x, uv.y, uv.z, uv.w}, N[4] = {nv.x, nv.y, nv.z, nv.w}, K1[4] = {q1.x, q1.y, q1.z, q1.w},
                    K7[4] = {q7.x, q7.y, q7.z, q7.w};
        float KJ[5][4];
#pragma unroll
        for (int j = 0; j < 5; ++j) { KJ[j][0] = kv[j].x; KJ[j][1] = kv[j].y; KJ[j][2] = kv[j].z; KJ[j][3] = kv[j].w; }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (b0 + i >= bf.B) continue;
            float a = c_BT[0] * K1[i];
#pragma unroll
            for (int j = 1; j < 6; ++j) a = fmaf(c_BT[j], KJ[j - 1][i], a);
            a = fmaf(c_BT[6], K7[i], a);
            float gd = (c_A[6][0] - c_A[5][0]) * K1[i];
#pragma unroll
            for (int j = 1; j < 5; ++j) gd = fmaf(c_A[6][j] - c_A[5][j], KJ[j - 1][i], gd);
            gd = dt * fmaf(c_A[6][5], KJ[4][i], gd);
            const float kd = K7[i] - KJ[4][i];
            su += (double)kd * (double)kd;
            sd += (double)gd * (double)gd;
            const float ut = dt * a;
            const float sc = fmaf(fmaxf(fabsf(U[i]), fabsf(N[i])), reltol, abstol);
            float r = ut / sc;
            if (N[i] != N[i]) r = N[i];
            acc += (double)r * (double)r;
        }
    }
    acc = cnf_block_sum(acc, sh);
    su = cnf_block_sum(su, sh);
    sd = cnf_block_sum(sd, sh);
    if (threadIdx.x == 0) { atomicAdd(&c->errsum, acc); atomicAdd(&c->stiffu, su); atomicAdd(&c->stiffd, sd); }
}
void launch_errnorm(cnf_handle* h, const CnfBufs& bf) {
    k_errnorm4<<<grid_for((long long)h->cur_R * bf.Bs / 4), CNF_NT, 0, h->stream>>>(h->d_ctrl, bf);
    h->launches++;
}

// ---------------------------------------------------------------- controller (A.5)
// Clips the next dt to dtmax / the end of the interval and checks the failure exits; runs
// before every attempt (end of the previous controller, or end of the initial-dt routine).
__device__ void begin_attempt(CnfCtrl* c) {
    if (c->done) return;
    if (c->t == c->t1) { c->done = 1; return; }
    if (++c->iter > c->maxiters) { c->retcode = CNF_RET_MAXITERS; c->done = 1; return; }
    float adt = fminf(fabsf(c->dt), c->dtmax);
    adt = fminf(adt, fabsf(c->t1 - c->t));
    c->dt = c->tdir * adt;
    if (c->t + c->dt == c->t || (c->dtmin > 0.0f && adt < c->dtmin)) {
        c->retcode = CNF_RET_DTLESSTHANMIN;
        c->done = 1;
    }
}

__device__ __forceinline__ void d_controller(CnfCtrl* c, float* dts) {
    c->accepted = 0;
    if (c->done) return;
    const double es = c->errsum;
    c->errsum = 0.0;
    const double stu = c->stiffu, std_ = c->stiffd;
    c->stiffu = 0.0; c->stiffd = 0.0;
    c->nf += 6;
    const float EEst = (float)sqrt(es / c->nglobal);
    if (EEst != EEst) { c->retcode = CNF_RET_UNSTABLE; c->done = 1; return; }
    double q11 = 0.0, q;
    if (EEst == 0.0f) {
        q = 1.0 / CNF_QMAX;
    } else {
        q11 = pow((double)EEst, CNF_BETA1);
        q = q11 / pow(c->qold, CNF_BETA2);
        q = fmax(1.0 / CNF_QMAX, fmin(1.0 / CNF_QMIN, q / CNF_GAMMA));
    }
    if (EEst <= 1.0f) {
        if (std_ > 0.0) {   // kernel paths that accumulate the two norms (k_errnorm)
            const float stiff = (float)(fabs((double)c->dt) * sqrt(stu / std_) / CNF_TSIT5_STABILITY);
            c->stiff_max = fmaxf(c->stiff_max, stiff);
            c->stiff_cur = stiff > CNF_STIFFTOL ? c->stiff_cur + 1 : 0;
            c->stiff_run = max(c->stiff_run, c->stiff_cur);
        }
        if (c->naccept < CNF_MAX_SLOTS) dts[c->naccept] = c->dt;
        float tn = c->t + c->dt;
        if (fabsf(tn - c->t1) < 100.0f * 1.1920929e-7f * fmaxf(fabsf(c->t), fabsf(c->t1))) tn = c->t1;
        c->t = tn;
        c->k1idx = 6 - c->k1idx;   // FSAL: k7 becomes k1
        c->qold = fmax((double)EEst, CNF_QOLDINIT);
        c->dt_last = c->dt;
        c->dt = (float)((double)c->dt / q);
        c->naccept++;              // the candidate slot becomes the current state
        c->accepted = 1;
        if (c->use_tape && c->naccept >= CNF_MAX_SLOTS - 1 && c->t != c->t1) {
            c->retcode = CNF_RET_MAXITERS;
            c->done = 1;
            return;
        }
    } else {
        c->dt = (float)((double)c->dt / fmin(1.0 / CNF_QMIN, q11 / CNF_GAMMA));
        c->nreject++;
    }
    begin_attempt(c);
}
// Stage tape for the reverse sweep: after an accepted step the z rows of k_1..k_5 (all the stage
// inputs s_2..s_6 of that step are u_n + dt * sum a_ij k_j) are kept, so the sweep does not have to
// recompute five vector-field evaluations per step.  ktape[n] = 5 x D x Bs floats, or null (then the
// sweep recomputes that step).
__device__ __forceinline__ void d_tape_stages(const CnfCtrl* c, const CnfBufs& bf, float* const* ktape, long long nz, long long cap) {
    if (!c->accepted) return;
    const long long n = c->naccept - 1;
    if (n < 0 || n >= cap) return;
    float* dst = ktape[n];
    if (!dst) return;
    const float* k1 = bf.K[6 - c->k1idx];   // the controller has already swapped k1/k7 for the next step
    const long long n4 = nz >> 2;            // Bs is a multiple of 256
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n4; e += (long long)gridDim.x * blockDim.x) {
        reinterpret_cast<float4*>(dst)[e] = reinterpret_cast<const float4*>(k1)[e];
#pragma unroll
        for (int j = 1; j < 5; ++j)
            reinterpret_cast<float4*>(dst + j * nz)[e] = reinterpret_cast<const float4*>(bf.K[j])[e];
    }
}
__global__ void k_tape_stages(const CnfCtrl* c, CnfBufs bf, float* const* ktape, long long nz, long long cap) {
    d_tape_stages(c, bf, ktape, nz, cap);
}
void launch_tape_stages(cnf_handle* h, const CnfBufs& bf) {
    const long long nz = (long long)h->cfg.D * bf.Bs;
    k_tape_stages<<<grid_for(nz >> 2), CNF_NT, 0, h->stream>>>(h->d_ctrl, bf, h->d_ktape, nz, (long long)h->ktape.size());
    h->launches++;
}

__global__ void k_controller(CnfCtrl* c, float* dts) { d_controller(c, dts); }

// Controller fused with the multi-GPU exchange of the error-norm partial (one warp; lane r talks to rank r):
// store {partial, sequence number} into the slot [parity][my rank] of every rank's mailbox (peer memory over
// NVLink), wait until the nranks slots of the own mailbox carry this sequence number, add them in rank order
// (the same sum, bit for bit, on every rank: the replicated controller needs no broadcast) and run the
// controller.  Slots alternate by parity: a rank can be at most one exchange ahead of another, because
// finishing exchange s needs every rank's store of s.  Nothing is exchanged once `done` is set (every rank
// sets it at the same attempt), so the speculative attempts after the end of a solve stay no-ops.
__global__ void k_controller_p2p(CnfCtrl* c, float* dts, CnfXSlot* mine, CnfXSlot* const* peers,
                                 unsigned long long* xseq, int rank, int nranks) {
    __shared__ double sh_v[32];
    const int lane = threadIdx.x;
    if (c->done) {
        if (lane == 0) c->accepted = 0;
        return;
    }
    const unsigned long long seq = *xseq + 1;
    const int par = (int)(seq & 1);
    const double v = c->errsum;
    __syncwarp();
    if (lane < nranks) {
        volatile CnfXSlot* dst = peers[lane] + par * nranks + rank;
        dst->v = v;
        __threadfence_system();
        dst->seq = seq;
        volatile CnfXSlot* src = mine + par * nranks + lane;
        unsigned long long t0 = 0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while (src->seq != seq) {
            __nanosleep(100);
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 60000000000ull) { asm volatile("trap;"); }   // a peer that never arrives: fail, do not hang
        }
        __threadfence_system();
        sh_v[lane] = src->v;
    }
    __syncwarp();
    if (lane == 0) {
        double s = 0.0;
        for (int r = 0; r < nranks; ++r) s += sh_v[r];
        c->errsum = s;
        *xseq = seq;
        d_controller(c, dts);
    }
}
void launch_controller(cnf_handle* h) {
    if (h->nranks > 1 && h->p2p_ok)
        k_controller_p2p<<<1, 32, 0, h->stream>>>(h->d_ctrl, h->d_dts, h->d_xslots, h->d_xpeers, h->d_xseq, h->rank, h->nranks);
    else
        k_controller<<<1, 1, 0, h->stream>>>(h->d_ctrl, h->d_dts);
    h->launches++;
}

// ---------------------------------------------------------------- Hairer initial step (A.5)
// which 0: initsum[0] += sum (u/sk)^2, initsum[1] += sum (k1/sk)^2 ; which 1: initsum[2] += sum ((f1-k1)/sk)^2
__global__ void k_initnorm(CnfCtrl* __restrict__ c, CnfBufs bf, int which) {
    __shared__ double sh[32];
    const float* u = bf.slots[cnf_slot(c, 0)];
    const float* k1 = bf.K[c->k1idx];
    const float* f1 = bf.K[1];
    const float reltol = c->reltol, abstol = c->abstol;
    double a0 = 0.0, a1 = 0.0;
    const long long n = (long long)c->R * bf.Bs;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        if (e % bf.Bs >= bf.B) continue;
        float sk = fmaf(fabsf(u[e]), reltol, abstol);
        if (which == 0) {
            float r0 = u[e] / sk, r1 = k1[e] / sk;
            a0 += (double)r0 * (double)r0;
            a1 += (double)r1 * (double)r1;
        } else {
            float r = (f1[e] - k1[e]) / sk;
            a0 += (double)r * (double)r;
        }
    }
    a0 = cnf_block_sum(a0, sh);
    a1 = cnf_block_sum(a1, sh);
    if (threadIdx.x == 0) {
        if (which == 0) { atomicAdd(&c->initsum[0], a0); atomicAdd(&c->initsum[1], a1); }
        else atomicAdd(&c->initsum[2], a0);
    }
}
void launch_initnorm(cnf_handle* h, const CnfBufs& bf, int which) {
    k_initnorm<<<grid_for((long long)h->cur_R * bf.Bs), CNF_NT, 0, h->stream>>>(h->d_ctrl, bf, which);
    h->launches++;
}

__global__ void k_initdt(CnfCtrl* c, int phase) {
    if (phase == 0) {
        double d0 = sqrt(c->initsum[0] / c->nglobal), d1 = sqrt(c->initsum[1] / c->nglobal);
        double dta = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
        if (dta > (double)c->dtmax) dta = (double)c->dtmax;
        c->dt0init = (float)dta;
        return;
    }
    if (phase == 1) {
        double d1 = sqrt(c->initsum[1] / c->nglobal);
        double d2 = sqrt(c->initsum[2] / c->nglobal) / (double)c->dt0init;
        double mx = d1 > d2 ? d1 : d2;
        double dt1 = (mx <= 1e-15) ? fmax(1e-6, (double)c->dt0init * 1e-3)
                                   : pow(10.0, -(2.0 + log10(mx)) / 5.0);
        double dtb = fmin(fmin(100.0 * (double)c->dt0init, dt1), (double)c->dtmax);
        c->dt = c->tdir * (float)dtb;
        c->nf += 2;
    }
    begin_attempt(c);
}
void launch_initdt(cnf_handle* h, int phase) {
    k_initdt<<<1, 1, 0, h->stream>>>(h->d_ctrl, phase);
    h->launches++;
}

// ---------------------------------------------------------------- final-state epilogue
// inference_sol (reference src/base_icnf.jl:167-189): logpz = logpdf(MvNormal(0,I), z),
// logp = logpz - dlogp, A = ||z_aug|| ; loss terms of src/icnf.jl:489 summed into lossacc.
// Outputs are in the caller's sample-major layout.
__global__ void k_epilogue(const CnfCtrl* __restrict__ c, CnfBufs bf, int mode, int nvars,
                           unsigned flags, float l1, float l2, float l3,
                           float* __restrict__ logp, float* __restrict__ regs,
                           float* __restrict__ z_out, double* __restrict__ lossacc) {
    __shared__ double sh[32];
    const int D = c->D;
    const float* u = bf.slots[cnf_slot(c, c->naccept)];
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double term = 0.0;
    if (b < bf.B) {
        float zz = 0.0f, za = 0.0f;
        for (int i = 0; i < D; ++i) {
            float v = u[(long long)i * bf.Bs + b];
            zz = fmaf(v, v, zz);
            if (i >= nvars) za = fmaf(v, v, za);
            if (z_out) z_out[b * D + i] = v;
        }
        const float log2pi = 1.8378770664093453f;
        float logpz = -0.5f * zz - 0.5f * (float)D * log2pi;
        float lp = logpz - u[(long long)D * bf.Bs + b];
        if (logp) logp[b] = lp;
        const bool useA = (flags & CNF_FLAG_NORM_Z_AUG) && (flags & CNF_FLAG_AUGMENTED);
        float A = useA ? sqrtf(za) : 0.0f;
        if (mode == CNF_MODE_TRAIN) {
            float E = u[(long long)(D + 1) * bf.Bs + b], n = u[(long long)(D + 2) * bf.Bs + b];
            if (regs) { regs[b * 3] = E; regs[b * 3 + 1] = n; regs[b * 3 + 2] = A; }
            term = (double)(-lp + l1 * E + l2 * n + l3 * A);
        } else {
            if (regs) regs[b] = A;
            term = (double)(-lp);
        }
    }
    term = cnf_block_sum(term, sh);
    if (threadIdx.x == 0 && lossacc) atomicAdd(lossacc, term);
}
void launch_epilogue(cnf_handle* h, const CnfBufs& bf, int mode, float* logp, float* regs,
                     float* z_out, long long Bglobal) {
    (void)Bglobal;
    k_epilogue<<<(unsigned)((bf.B + CNF_NT - 1) / CNF_NT), CNF_NT, 0, h->stream>>>(
        h->d_ctrl, bf, mode, h->cfg.nvars, h->cfg.flags, h->cfg.lambda1, h->cfg.lambda2,
        h->cfg.lambda3, logp, regs, z_out, h->d_lossacc);
    h->launches++;
}

// adjoint seed at t1: d loss / d z = (z + lambda3 z_aug/||z_aug||) / B_global   (D x Bs)
__global__ void k_init_ubar(const CnfCtrl* __restrict__ c, CnfBufs bf, int nvars, unsigned flags,
                            float l3, float invB, float* __restrict__ ubar) {
    const int D = c->D;
    const float* u = bf.slots[cnf_slot(c, c->naccept)];
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= bf.Bs) return;
    const bool live = b < bf.B;
    float za = 0.0f;
    for (int i = nvars; i < D; ++i) { float v = u[(long long)i * bf.Bs + b]; za = fmaf(v, v, za); }
    const bool useA = (flags & CNF_FLAG_NORM_Z_AUG) && (flags & CNF_FLAG_AUGMENTED);
    float cA = (useA && za > 0.0f) ? l3 / sqrtf(za) : 0.0f;
    for (int i = 0; i < D; ++i) {
        float v = u[(long long)i * bf.Bs + b];
        ubar[(long long)i * bf.Bs + b] = live ? (v + (i >= nvars ? cA * v : 0.0f)) * invB : 0.0f;
    }
}
void launch_init_ubar(cnf_handle* h, const CnfBufs& bf, float* ubar, long long Bglobal, int mode) {
    // the TestMode loss has no regulariser terms (src/base_icnf.jl:489-508)
    k_init_ubar<<<(unsigned)((bf.Bs + CNF_NT - 1) / CNF_NT), CNF_NT, 0, h->stream>>>(
        h->d_ctrl, bf, h->cfg.nvars, mode == CNF_MODE_TRAIN ? h->cfg.flags : 0u, h->cfg.lambda3,
        1.0f / (float)Bglobal, ubar);
    h->launches++;
}
