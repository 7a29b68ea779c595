// cnf_vector_generic.cu — row N4, the rest of VectorMode: every sample its own adaptive solve for ANY network, mode and
// width the MatrixMode stage kernels serve (TrainMode Hutchinson in both compute modes, conditional flows, 1..8 layers,
// the tensor-core paths for wide nets, the fused multi-layer kernel).
//
// The reference's VectorMode (inference_prob / inference_sol for <:VectorMode, src/base_icnf.jl:145-165, 213-264; the
// Distributions adapter broadcasts it over columns, src/exts/dist_ext/core_icnf.jl:24-25) solves each column as its own
// ODE problem: own dt, own RMS error norm over its R entries, own accept/reject history.  augmented_f is autonomous
// and column-wise (src/icnf.jl:330-349), so one evaluation of the whole batch serves every column whatever its time
// and step size are.  Only the controller changes: t, dt, qold, the step counters and `done` become per-sample arrays;
// the RK combination reads dt[b]; the error norm is reduced per sample (one thread owns a sample's R rows, coalesced
// across the batch); accepted samples copy u_new -> u and k7 -> k1 (FSAL), rejected ones keep both.  Samples that have
// reached t1 keep riding along (their state is frozen) until the slowest sample is done — the batch-level analogue of
// the idle lanes of k_solve_vector (cnf_vector.cu, the one-launch kernel kept for the narrow exact-trace shapes).
// Arithmetic follows the MatrixMode controller (cnf_solver.cu) operation by operation, so a sample's result equals the
// MatrixMode solve of a batch holding only that sample.
#include "cnf_internal.h"

struct VgArrays {
    float* t;
    float* dt;
    float* d1;        // Hairer d1, then unused
    double* qold;
    int* done;
    int* naccept;
    int* nreject;
    int* iter;
};
struct VgParams {
    float t0, t1, tdir, dtmax, dtmin, reltol, abstol, dt0;
    long long maxiters, Bs;
    int B, R, D;
};
// summary in the control block: [0] max naccept, [1] max nreject, [2] worst retcode, [3] active samples
__device__ __forceinline__ void vg_begin_attempt(const VgParams& p, const VgArrays& a, long long b, int* summary) {
    if (a.done[b]) return;
    const float t = a.t[b];
    if (t == p.t1) { a.done[b] = 1; return; }
    if (++a.iter[b] > p.maxiters) { atomicMax(&summary[2], CNF_RET_MAXITERS); a.done[b] = 1; return; }
    float adt = fminf(fabsf(a.dt[b]), p.dtmax);
    adt = fminf(adt, fabsf(p.t1 - t));
    const float dt = p.tdir * adt;
    a.dt[b] = dt;
    if (t + dt == t || (p.dtmin > 0.0f && adt < p.dtmin)) { atomicMax(&summary[2], CNF_RET_DTLESSTHANMIN); a.done[b] = 1; }
}

// Hairer-Wanner initial step, first half, per sample: d0, d1 -> dt0; S = z rows of u0 + tdir dt0 k1
__global__ void k_vg_init0(VgParams p, VgArrays a, const float* __restrict__ u, const float* __restrict__ k1, float* __restrict__ S) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.Bs) return;
    a.t[b] = p.t0; a.qold[b] = CNF_QOLDINIT; a.naccept[b] = 0; a.nreject[b] = 0; a.iter[b] = 0;
    a.done[b] = b < p.B ? 0 : 1;
    double s0 = 0.0, s1 = 0.0;
    for (int r = 0; r < p.R; ++r) {
        const float uv = u[(long long)r * p.Bs + b], kv = k1[(long long)r * p.Bs + b];
        const float sk = fmaf(fabsf(uv), p.reltol, p.abstol);
        const float x0 = uv / sk, x1 = kv / sk;
        s0 += (double)x0 * (double)x0;
        s1 += (double)x1 * (double)x1;
    }
    const double d0 = sqrt(s0 / (double)p.R), d1 = sqrt(s1 / (double)p.R);
    double dta = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
    if (dta > (double)p.dtmax) dta = (double)p.dtmax;
    const float dt0 = p.dt0 > 0.0f ? p.dt0 : (float)dta;
    a.dt[b] = dt0;
    a.d1[b] = (float)d1;
    for (int r = 0; r < p.D; ++r)
        S[(long long)r * p.Bs + b] = fmaf(p.tdir * dt0, k1[(long long)r * p.Bs + b], u[(long long)r * p.Bs + b]);
}
// second half: d2 from f1 = f(u0 + tdir dt0 k1) -> dt; then the first attempt's clipping
__global__ void k_vg_init1(VgParams p, VgArrays a, const float* __restrict__ u, const float* __restrict__ k1, const float* __restrict__ f1,
                           int* summary) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= p.Bs) return;
    if (p.dt0 > 0.0f) {
        a.dt[b] = p.tdir * p.dt0;
    } else {
        const float dt0 = a.dt[b];
        double s2 = 0.0;
        for (int r = 0; r < p.R; ++r) {
            const float uv = u[(long long)r * p.Bs + b];
            const float sk = fmaf(fabsf(uv), p.reltol, p.abstol);
            const float x = (f1[(long long)r * p.Bs + b] - k1[(long long)r * p.Bs + b]) / sk;
            s2 += (double)x * (double)x;
        }
        const double d1 = (double)a.d1[b];
        const double d2 = sqrt(s2 / (double)p.R) / (double)dt0;
        const double mx = d1 > d2 ? d1 : d2;
        const double dt1 = (mx <= 1e-15) ? fmax(1e-6, (double)dt0 * 1e-3) : pow(10.0, -(2.0 + log10(mx)) / 5.0);
        a.dt[b] = p.tdir * (float)fmin(fmin(100.0 * (double)dt0, dt1), (double)p.dtmax);
    }
    vg_begin_attempt(p, a, b, summary);
}

// stage input i (1..5): S = z rows of u + dt[b] sum_j a_ij k_j; stage 6: all R rows of the candidate state u_new
__global__ void k_vg_combine(VgParams p, VgArrays a, CnfBufs bf, const float* __restrict__ u, int stage, float* __restrict__ out, int rows) {
    const long long n = (long long)rows * p.Bs;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const long long b = e % p.Bs;
        float acc = c_A[stage][0] * bf.K[0][e];
        for (int j = 1; j < stage; ++j) acc = fmaf(c_A[stage][j], bf.K[j][e], acc);
        out[e] = fmaf(a.dt[b], acc, u[e]);
    }
}

// embedded error (A.3) over the sample's own R entries, PI controller (A.5), accept / reject, FSAL, next attempt
__global__ void k_vg_control(VgParams p, VgArrays a, CnfBufs bf, float* __restrict__ u, const float* __restrict__ un, int* summary) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    int active = 0;
    if (b < p.B && !a.done[b]) {
        const float dt = a.dt[b];
        double acc = 0.0;
        bool bad = false;
        for (int r = 0; r < p.R; ++r) {
            const long long e = (long long)r * p.Bs + b;
            float s = c_BT[0] * bf.K[0][e];
            for (int j = 1; j < 7; ++j) s = fmaf(c_BT[j], bf.K[j][e], s);
            const float ut = dt * s;
            const float sc = fmaf(fmaxf(fabsf(u[e]), fabsf(un[e])), p.reltol, p.abstol);
            const float q = ut / sc;
            if (un[e] != un[e]) bad = true;
            acc += (double)q * (double)q;
        }
        const float EEst = (float)sqrt(acc / (double)p.R);
        if (EEst != EEst || bad) {
            atomicMax(&summary[2], CNF_RET_UNSTABLE);
            a.done[b] = 1;
        } else {
            double q11 = 0.0, q;
            const double qold = a.qold[b];
            if (EEst == 0.0f) {
                q = 1.0 / CNF_QMAX;
            } else {
                q11 = pow((double)EEst, CNF_BETA1);
                q = q11 / pow(qold, CNF_BETA2);
                q = fmax(1.0 / CNF_QMAX, fmin(1.0 / CNF_QMIN, q / CNF_GAMMA));
            }
            if (EEst <= 1.0f) {
                const float t = a.t[b];
                float tn = t + dt;
                if (fabsf(tn - p.t1) < 100.0f * 1.1920929e-7f * fmaxf(fabsf(t), fabsf(p.t1))) tn = p.t1;
                a.t[b] = tn;
                for (int r = 0; r < p.R; ++r) {          // u <- u_new, k1 <- k7 (FSAL) for this sample only
                    const long long e = (long long)r * p.Bs + b;
                    u[e] = un[e];
                    bf.K[0][e] = bf.K[6][e];
                }
                a.qold[b] = fmax((double)EEst, CNF_QOLDINIT);
                a.dt[b] = (float)((double)dt / q);
                a.naccept[b]++;
            } else {
                a.dt[b] = (float)((double)dt / fmin(1.0 / CNF_QMIN, q11 / CNF_GAMMA));
                a.nreject[b]++;
            }
            vg_begin_attempt(p, a, b, summary);
        }
        if (a.done[b]) {
            atomicMax(&summary[0], a.naccept[b]);
            atomicMax(&summary[1], a.nreject[b]);
        } else {
            active = 1;
        }
    }
    active = __syncthreads_or(active);
    if (threadIdx.x == 0 && active) atomicAdd(&summary[3], 1);
}
__global__ void k_vg_clear_active(int* summary) { summary[3] = 0; }

static int vg_eval(cnf_handle* h, const CnfBufs& bf, int mode, const float* src, float* kout) {
    stage_timer_begin(h);
    int rc;
    if (tc_supported(h->cfg, mode) && h->tc_state) {
        h->tc_mode = mode;
        h->path_used = 3;
        rc = tc_eval(h, bf, src, -1, kout, 0);       // (the fused multi-layer kernel where it applies: path 7)
    } else {
        rc = launch_eval_generic(h, bf, mode, src, kout);
    }
    stage_timer_end(h);
    return rc;
}

// slot 0 holds [x; 0] (R x Bs); on return slot 0 holds the final state of every sample and `out` a control block the
// epilogue kernels can read (naccept = 0 -> slot 0); out->iter = the largest accepted-step count over the samples
int vector_solve_generic(cnf_handle* h, const CnfBufs& bf, int mode, int R, float t0, float t1, CnfCtrl* out) {
    const cnf_config& c = h->cfg;
    int rc;
    // per-sample controller arrays (grown lazily)
    const size_t per = 4 * sizeof(float) + sizeof(double) + 4 * sizeof(int);
    if (h->vg_cap < bf.Bs) {
        CNF_CUDA(cudaStreamSynchronize(h->stream));
        cudaFree(h->d_vg);
        h->d_vg = nullptr;
        CNF_CUDA(cudaMalloc(&h->d_vg, per * (size_t)bf.Bs));
        h->vg_cap = bf.Bs;
    }
    VgArrays a;
    {
        char* base = (char*)h->d_vg;
        a.qold = (double*)base; base += sizeof(double) * bf.Bs;
        a.t = (float*)base; base += sizeof(float) * bf.Bs;
        a.dt = (float*)base; base += sizeof(float) * bf.Bs;
        a.d1 = (float*)base; base += sizeof(float) * bf.Bs;
        a.done = (int*)base; base += sizeof(int) * bf.Bs;
        a.naccept = (int*)base; base += sizeof(int) * bf.Bs;
        a.nreject = (int*)base; base += sizeof(int) * bf.Bs;
        a.iter = (int*)base;
    }
    VgParams p;
    p.t0 = t0; p.t1 = t1;
    p.tdir = (t1 >= t0) ? 1.0f : -1.0f;
    p.dtmax = c.dtmax > 0 ? c.dtmax : fabsf(t1 - t0);
    p.dtmin = c.dtmin;
    p.reltol = c.reltol > 0 ? c.reltol : CNF_DEFAULT_RELTOL;
    p.abstol = c.abstol > 0 ? c.abstol : CNF_DEFAULT_ABSTOL;
    p.dt0 = c.dt0 > 0 ? c.dt0 : 0.0f;
    p.maxiters = c.maxiters > 0 ? c.maxiters : CNF_DEFAULT_MAXITERS;
    if (p.maxiters > 0x7fffffffLL) p.maxiters = 0x7fffffffLL;
    p.Bs = bf.Bs; p.B = bf.B; p.R = R; p.D = c.D;
    CnfCtrl ic;
    memset(&ic, 0, sizeof(ic));
    ic.R = R; ic.D = c.D;
    ic.t = t1; ic.t0 = t0; ic.t1 = t1;
    ic.done = 1;
    ic.stiff_max = -1.0f;
    CNF_CUDA(cudaMemcpyAsync(h->d_ctrl, &ic, sizeof(CnfCtrl), cudaMemcpyHostToDevice, h->stream));
    CNF_CUDA(cudaEventRecord(h->ev1, h->stream));
    int* summary = reinterpret_cast<int*>(&h->d_ctrl->initsum[0]);   // 4 ints of scratch inside the control block
    float* u = h->slots[0];
    float* un = h->slots[1];
    const unsigned gb = (unsigned)((bf.Bs + CNF_NT - 1) / CNF_NT);
    auto grid_rows = [&](int rows) {
        long long g = ((long long)rows * bf.Bs + CNF_NT - 1) / CNF_NT;
        if (g > 148LL * 16) g = 148LL * 16;
        return (unsigned)g;
    };
    h->path_used = 8;
    long long nf = 0;
    if ((rc = vg_eval(h, bf, mode, u, h->d_K[0]))) return rc;   // k1 = f(u0)
    nf++;
    k_vg_init0<<<gb, CNF_NT, 0, h->stream>>>(p, a, u, h->d_K[0], h->d_S);
    h->launches++;
    if (!(p.dt0 > 0.0f)) {
        if ((rc = vg_eval(h, bf, mode, h->d_S, h->d_K[1]))) return rc;
        nf += 2;   // f(u1) and the integrator's own re-evaluation of f(u0) (reused), as the MatrixMode driver counts
    }
    k_vg_init1<<<gb, CNF_NT, 0, h->stream>>>(p, a, u, h->d_K[0], h->d_K[1], summary);
    h->launches++;
    long long launched = 0;
    int pend = -1;
    for (;;) {
        for (int i = 1; i <= 5; ++i) {
            k_vg_combine<<<grid_rows(c.D), CNF_NT, 0, h->stream>>>(p, a, bf, u, i, h->d_S, c.D);
            h->launches++;
            if ((rc = vg_eval(h, bf, mode, h->d_S, h->d_K[i]))) return rc;
        }
        k_vg_combine<<<grid_rows(R), CNF_NT, 0, h->stream>>>(p, a, bf, u, 6, un, R);
        h->launches++;
        if ((rc = vg_eval(h, bf, mode, un, h->d_K[6]))) return rc;
        k_vg_clear_active<<<1, 1, 0, h->stream>>>(summary);
        k_vg_control<<<gb, CNF_NT, 0, h->stream>>>(p, a, bf, u, un, summary);
        h->launches += 2;
        nf += 6;
        launched++;
        const int cur = (int)(launched & 1);
        CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[cur], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
        CNF_CUDA(cudaEventRecord(h->evchk[cur], h->stream));
        if (pend >= 0) {   // one attempt of look-ahead, like the MatrixMode driver: the GPU never idles on the host
            CNF_CUDA(cudaEventSynchronize(h->evchk[pend]));
            if (reinterpret_cast<const int*>(&h->h_ctrl[pend].initsum[0])[3] == 0) break;
        }
        pend = cur;
        if (launched > p.maxiters + 2) break;
    }
    CNF_CUDA(cudaEventRecord(h->ev2, h->stream));
    CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[0], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    *out = h->h_ctrl[0];
    const int* sm = reinterpret_cast<const int*>(&out->initsum[0]);
    out->naccept = 0;            // the epilogue reads slot(naccept) = slot 0
    out->nreject = sm[1];
    out->nf = nf;
    out->retcode = sm[2];
    out->iter = sm[0];           // reported as naccept: the maximum over the samples
    h->path_used = 8;            // reported: VectorMode on the MatrixMode stage kernels (whatever stage kernel served it)
    CNF_CUDA(cudaGetLastError());
    return CNF_OK;
}
