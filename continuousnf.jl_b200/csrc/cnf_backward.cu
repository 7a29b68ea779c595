// cnf_backward.cu — discretise-then-optimise backward pass (generic FP32 SIMT tile kernels).
// Exact reverse mode of the accepted Tsit5 steps with dt_n frozen; stage checkpointing: only
// u_n per accepted step is kept (the state slots double as the tape), the six stage inputs of
// a step are recomputed in the reverse sweep.  Replaces the Enzyme + SciMLSensitivity gradient
// of the reference (call sites src/exts/mlj_ext/core_icnf.jl:59-63, test/call_tests.jl:239-251).
// The per-stage kernel is the VJP of TrainMode augmented_f (src/icnf.jl:318-350 VecJac,
// :384-420 JacVec), i.e. second order through the tanh chain; maths in DESIGN.md.
#include "cnf_internal.h"

struct BwLayout {
    int hoff[CNF_MAX_LAYERS + 1];
    int goff[CNF_MAX_LAYERS + 1];  // l = 1..L, dims[l] rows: g_l (VecJac) / tau_l (JacVec)
    int qoff[CNF_MAX_LAYERS + 1];  // l = 1..L, dims[l] rows: q_{l+1} (VecJac) / r_l (JacVec), later db_l
    int w[6];                      // work buffers, maxd rows each
    int e, ybar, vbar, red;
    int z0;                        // zeros: h_L inside 1 - h^2 when the last layer is linear
    int total;
};
static BwLayout bw_layout(const cnf_config& cfg, int TB) {
    BwLayout L;
    int TBP = TB + 1, o = 0, maxd = 0;
    for (int l = 0; l <= cfg.n_layers; ++l) {
        L.hoff[l] = o; o += cfg.dims[l] * TBP;
        if (cfg.dims[l] > maxd) maxd = cfg.dims[l];
    }
    L.goff[0] = L.qoff[0] = 0;
    for (int l = 1; l <= cfg.n_layers; ++l) { L.goff[l] = o; o += cfg.dims[l] * TBP; }
    for (int l = 1; l <= cfg.n_layers; ++l) { L.qoff[l] = o; o += cfg.dims[l] * TBP; }
    for (int j = 0; j < 6; ++j) { L.w[j] = o; o += maxd * TBP; }
    L.e = o; o += cfg.D * TBP;
    L.ybar = o; o += cfg.D * TBP;
    L.vbar = o; o += cfg.D * TBP;
    L.red = o; o += 6 * TBP;
    L.z0 = o; o += cfg.D * TBP;
    L.total = o;
    return L;
}

// gW[i*dout + o] += sum_s a[o][s] * b[i][s]   (i < ncols); coalesced float atomics
template <int TB>
__device__ __forceinline__ void outer_acc_tile(float* __restrict__ gW, int dout, int ncols,
                                               const float* a, const float* b) {
    constexpr int TBP = TB + 1;
    const int n = dout * ncols;
    for (int p = threadIdx.x; p < n; p += CNF_NT) {
        int o = p % dout, i = p / dout;
        const float* ao = a + o * TBP;
        const float* bi = b + i * TBP;
        float acc = 0.0f;
#pragma unroll
        for (int s = 0; s < TB; ++s) acc = fmaf(ao[s], bi[s], acc);
        atomicAdd(gW + p, acc);
    }
}

template <int TB>
__device__ __forceinline__ void bias_acc_tile(float* __restrict__ gb, int dout, const float* a) {
    constexpr int TBP = TB + 1;
    for (int o = threadIdx.x; o < dout; o += CNF_NT) {
        float acc = 0.0f;
#pragma unroll
        for (int s = 0; s < TB; ++s) acc += a[o * TBP + s];
        atomicAdd(gb + o, acc);
    }
}

#define EW_LOOP(rows) for (int idx = threadIdx.x; idx < (rows) * TB; idx += CNF_NT)
#define EW_AT (((idx / TB) * TBP) + (idx % TB))

template <int TB>
__device__ __forceinline__ void d_vjp_generic(const CnfBufs& bf, const CnfNet& net, const BwLayout& lay, const float* src,
                                              const float* abar, int stage, const float* dtp, float c_l,
                                              float c_E, float c_n, float* ubar, float* kb0, float* kb1, float* kb2,
                                              float* kb3, float* kb4, float* grad, int unit) {
    extern __shared__ float sm[];
    constexpr int TBP = TB + 1;
    const int D = net.D, L = net.L;
    const long long Bs = bf.Bs;
    const long long b0 = (long long)blockIdx.x * TB;
    const int tid = threadIdx.x;
    const float dt = *dtp;
    const float wgt = dt * c_A[6][stage];
    const float alpha = wgt * c_l, beta = wgt * c_E, gamma = wgt * c_n;
    const bool jvp = net.trace_mode == CNF_TRACE_JVP;
    const bool lin = (net.flags & CNF_FLAG_LAST_LINEAR) != 0;
    for (int idx = tid; idx < D * TBP; idx += CNF_NT) sm[lay.z0 + idx] = 0.0f;
    // h_l as the argument of the activation derivative (and of its derivative -2 h (1 - h^2))
    auto hder = [&](int l) -> const float* { return (lin && l == L) ? sm + lay.z0 : sm + lay.hoff[l]; };
    // ---- forward activations
    {
        GenLayout gl;
        for (int l = 0; l <= L; ++l) gl.hoff[l] = lay.hoff[l];
        gen_forward<TB>(net, gl, sm, src, bf.ys, Bs, b0, bf.B);
    }
    float* e = sm + lay.e;
    float* ybar = sm + lay.ybar;
    float* vbar = sm + lay.vbar;
    float* red = sm + lay.red;
    EW_LOOP(D) {
        long long b = b0 + idx % TB;
        int i = idx / TB;
        bool live = b < bf.B;
        // unit >= 0 (TestMode): eps = e_unit, the exact trace is the sum over the D unit vectors of the
        // Hutchinson term; the zdot cotangent enters once (with unit 0)
        e[EW_AT] = live ? (unit >= 0 ? (i == unit ? 1.0f : 0.0f) : bf.eps[(long long)i * Bs + b]) : 0.0f;
        ybar[EW_AT] = (live && unit <= 0) ? abar[(long long)i * Bs + b] : 0.0f;
    }
    __syncthreads();
    const float* hL = sm + lay.hoff[L];
    const float* v;
    // ---- the eps chain (kept per layer: needed by the second-order terms)
    if (!jvp) {
        float* gL = sm + lay.goff[L];
        const float* hLd = hder(L);
        EW_LOOP(D) { float hh = hLd[EW_AT]; gL[EW_AT] = e[EW_AT] * fmaf(-hh, hh, 1.0f); }
        __syncthreads();
        for (int l = L; l >= 1; --l) {
            int din = net.dims[l - 1], dout = net.dims[l];
            if (l > 1) {
                float* q = sm + lay.qoff[l - 1];
                dense_tile<TB>(net.W[l - 1], dout, nullptr, dout, din, sm + lay.goff[l], q, EPI_NONE, nullptr);
                __syncthreads();
                const float* hm = sm + lay.hoff[l - 1];
                float* g = sm + lay.goff[l - 1];
                EW_LOOP(din) { float hh = hm[EW_AT]; g[EW_AT] = q[EW_AT] * fmaf(-hh, hh, 1.0f); }
                __syncthreads();
            } else {
                dense_tile<TB>(net.W[0], dout, nullptr, dout, D, sm + lay.goff[1], sm + lay.w[0], EPI_NONE, nullptr);
                __syncthreads();
            }
        }
        v = sm + lay.w[0];
    } else {
        for (int l = 1; l <= L; ++l) {
            int din = net.dims[l - 1], dout = net.dims[l];
            float* r = sm + lay.qoff[l];
            dense_tile<TB>(net.Wt[l - 1], din, nullptr, l == 1 ? D : din, dout,
                           l == 1 ? e : sm + lay.goff[l - 1], r, EPI_NONE, nullptr);
            __syncthreads();
            const float* hm = hder(l);
            float* t = sm + lay.goff[l];
            EW_LOOP(dout) { float hh = hm[EW_AT]; t[EW_AT] = r[EW_AT] * fmaf(-hh, hh, 1.0f); }
            __syncthreads();
        }
        v = sm + lay.goff[L];
    }
    // ---- cotangents on y = zdot and on v
    if (tid < TB) {
        float ny = 0.0f, nv = 0.0f;
        for (int i = 0; i < D; ++i) {
            float y = hL[i * TBP + tid], vv = v[i * TBP + tid];
            ny = fmaf(y, y, ny);
            nv = fmaf(vv, vv, nv);
        }
        bool live = (b0 + tid) < bf.B;
        float nys = sqrtf(ny), nvs = sqrtf(nv);
        red[tid] = (live && (net.flags & CNF_FLAG_NORM_Z) && nys > 0.0f) ? beta / nys : 0.0f;
        red[TBP + tid] = (live && (net.flags & CNF_FLAG_NORM_J) && nvs > 0.0f) ? gamma / nvs : 0.0f;
        red[2 * TBP + tid] = live ? alpha : 0.0f;
    }
    __syncthreads();
    EW_LOOP(D) {
        int s = idx % TB;
        ybar[EW_AT] = fmaf(red[s], hL[EW_AT], ybar[EW_AT]);
        vbar[EW_AT] = fmaf(red[TBP + s], v[EW_AT], -red[2 * TBP + s] * e[EW_AT]);
    }
    __syncthreads();
    float* hb = ybar;  // adjoint of h_l, walks down the layers
    if (!jvp) {
        // adjoint of the reverse chain runs forward: gb_l = W_l qb_l, qb_1 = [vbar; 0]
        const float* qb = vbar;
        for (int l = 1; l <= L; ++l) {
            int din = net.dims[l - 1], dout = net.dims[l];
            int used = (l == 1) ? D : din;
            float* gb = sm + lay.w[1];
            dense_tile<TB>(net.Wt[l - 1], din, nullptr, used, dout, qb, gb, EPI_NONE, nullptr);
            outer_acc_tile<TB>(grad + net.w_off[l - 1], dout, used, sm + lay.goff[l], qb);
            __syncthreads();
            float* q = sm + lay.qoff[l];  // q_{l+1} -> db_l
            if (l < L) {
                float* qn = sm + lay.w[2 + (l & 1)];
                const float* hm = sm + lay.hoff[l];
                EW_LOOP(dout) {
                    float g = gb[EW_AT], hh = hm[EW_AT];
                    qn[EW_AT] = g * fmaf(-hh, hh, 1.0f);
                    q[EW_AT] = g * q[EW_AT];
                }
                qb = qn;
            } else {
                EW_LOOP(dout) { q[EW_AT] = gb[EW_AT] * e[EW_AT]; }
            }
            __syncthreads();
        }
    }
    const float* taub = vbar;
    for (int l = L; l >= 1; --l) {
        int din = net.dims[l - 1], dout = net.dims[l];
        int used = (l == 1) ? D : din;
        const float* hm = hder(l);
        float* db = sm + lay.qoff[l];
        float* pb = sm + lay.w[1];
        if (jvp) {
            float* rb = sm + lay.w[0];
            EW_LOOP(dout) {
                float tb = taub[EW_AT], hh = hm[EW_AT];
                rb[EW_AT] = tb * fmaf(-hh, hh, 1.0f);
                db[EW_AT] = tb * db[EW_AT];   // r_l -> db_l
            }
            __syncthreads();
            outer_acc_tile<TB>(grad + net.w_off[l - 1], dout, used, rb, l == 1 ? e : sm + lay.goff[l - 1]);
            if (l > 1) {
                float* tn = sm + lay.w[2 + (l & 1)];
                dense_tile<TB>(net.W[l - 1], dout, nullptr, dout, din, rb, tn, EPI_NONE, nullptr);
                taub = tn;
            }
        }
        EW_LOOP(dout) {
            float hh = hm[EW_AT];
            float hbk = fmaf(-2.0f * hh, db[EW_AT], hb[EW_AT]);
            pb[EW_AT] = hbk * fmaf(-hh, hh, 1.0f);
        }
        __syncthreads();
        outer_acc_tile<TB>(grad + net.w_off[l - 1], dout, din, pb, sm + lay.hoff[l - 1]);
        if (!(lin && l == L)) bias_acc_tile<TB>(grad + net.b_off[l - 1], dout, pb);
        float* hn = sm + lay.w[4 + (l & 1)];
        dense_tile<TB>(net.W[l - 1], dout, nullptr, dout, used, pb, hn, EPI_NONE, nullptr);
        hb = hn;
        __syncthreads();
    }
    // ---- zbar = hb[0:D]: propagate into ubar and the earlier stages' cotangents
    float* kb[5] = {kb0, kb1, kb2, kb3, kb4};
    EW_LOOP(D) {
        long long b = b0 + idx % TB;
        int i = idx / TB;
        if (b < Bs) {
            float zb = hb[EW_AT];
            long long at = (long long)i * Bs + b;
            ubar[at] += zb;
            for (int j = 0; j < stage; ++j) kb[j][at] = fmaf(dt * c_A[stage][j], zb, kb[j][at]);
        }
    }
}

template <int TB>
__global__ void __launch_bounds__(CNF_NT)
k_vjp_generic(CnfBufs bf, CnfNet net, BwLayout lay, const float* src, const float* abar, int stage, const float* dtp, float c_l,
              float c_E, float c_n, float* ubar, float* kb0, float* kb1, float* kb2, float* kb3, float* kb4, float* grad,
              int unit) {
    d_vjp_generic<TB>(bf, net, lay, src, abar, stage, dtp, c_l, c_E, c_n, ubar, kb0, kb1, kb2, kb3, kb4, grad, unit);
}

// recompute helper: out = u + dt * sum_{j<stage} a[stage][j] K[j]   (D rows)
__device__ __forceinline__ void d_combine_bw(const float* u, const CnfBufs& bf, int stage, const float* dtp, float* out, long long n) {
    const float dt = *dtp;
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        float acc = c_A[stage][0] * bf.K[0][e];
        for (int j = 1; j < stage; ++j) acc = fmaf(c_A[stage][j], bf.K[j][e], acc);
        out[e] = fmaf(dt, acc, u[e]);
    }
}

__global__ void k_combine_bw(const float* u, CnfBufs bf, int stage, const float* dtp, float* out, long long n) {
    d_combine_bw(u, bf, stage, dtp, out, n);
}

// kb_j = dt * a7j * ubar, j = 0..5
__device__ __forceinline__ void d_kbar_init(const float* ubar, const float* dtp, float* kb0, float* kb1, float* kb2, float* kb3,
                                            float* kb4, float* kb5, long long n) {
    const float dt = *dtp;
    float* kb[6] = {kb0, kb1, kb2, kb3, kb4, kb5};
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x) {
        float ub = ubar[e];
#pragma unroll
        for (int j = 0; j < 6; ++j) kb[j][e] = dt * c_A[6][j] * ub;
    }
}

__global__ void k_kbar_init(const float* ubar, const float* dtp, float* kb0, float* kb1, float* kb2, float* kb3, float* kb4,
                            float* kb5, long long n) {
    d_kbar_init(ubar, dtp, kb0, kb1, kb2, kb3, kb4, kb5, n);
}

template <int TB>
static void launch_vjp_tb(cnf_handle* h, const CnfBufs& bf, const BwLayout& lay, int smem,
                          const float* src, const float* abar, int stage, const float* dtp,
                          float c_l, float c_E, float c_n, float* ubar, float** kb, float* grad, int unit) {
    static bool attr_dev[64] = {};   // per device
    bool& attr_set = attr_dev[h->device & 63];
    if (!attr_set) {
        cudaFuncSetAttribute(k_vjp_generic<TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        attr_set = true;
    }
    unsigned grid = (unsigned)((bf.Bs + TB - 1) / TB);
    k_vjp_generic<TB><<<grid, CNF_NT, smem, h->stream>>>(bf, h->net, lay, src, abar, stage, dtp, c_l,
                                                        c_E, c_n, ubar, kb[0], kb[1], kb[2], kb[3],
                                                        kb[4], grad, unit);
    h->launches++;
}

static int launch_vjp_generic_stage(cnf_handle* h, const CnfBufs& bf, const float* src,
                                    const float* abar, int stage, const float* dtp, float c_l,
                                    float c_E, float c_n, float* ubar, float** kb, float* grad, int unit = -1) {
    BwLayout l1 = bw_layout(h->cfg, 0);
    int smem = 0;
    int TB = pick_tb(l1.total, &smem);
    if (!TB) { h->err = "generic backward kernel: network too wide for shared memory"; return CNF_ERR_UNSUPPORTED; }
    BwLayout lay = bw_layout(h->cfg, TB);
    switch (TB) {
        case 32: launch_vjp_tb<32>(h, bf, lay, smem, src, abar, stage, dtp, c_l, c_E, c_n, ubar, kb, grad, unit); break;
        case 16: launch_vjp_tb<16>(h, bf, lay, smem, src, abar, stage, dtp, c_l, c_E, c_n, ubar, kb, grad, unit); break;
        case 8: launch_vjp_tb<8>(h, bf, lay, smem, src, abar, stage, dtp, c_l, c_E, c_n, ubar, kb, grad, unit); break;
        default: launch_vjp_tb<4>(h, bf, lay, smem, src, abar, stage, dtp, c_l, c_E, c_n, ubar, kb, grad, unit); break;
    }
    return CNF_OK;
}

// forward re-evaluation used by the stage checkpointing: tensor path when available
static int bw_eval(cnf_handle* h, const CnfBufs& bf, int mode, const float* src, float* kout) {
    if (mode == CNF_MODE_TEST) return launch_eval_generic(h, bf, CNF_MODE_TEST, src, kout);
    if (tc_bw_supported(h)) return tc_eval_z(h, bf, src, kout);   // only the zdot rows feed the stage inputs
    if (tc_supported(h->cfg, CNF_MODE_TRAIN) && h->tc_state) { h->tc_mode = CNF_MODE_TRAIN; return tc_eval(h, bf, src, -1, kout, 0); }
    return launch_eval_generic(h, bf, CNF_MODE_TRAIN, src, kout);
}

static unsigned bw_grid(long long n) {
    long long g = (n + CNF_NT - 1) / CNF_NT;
    if (g > 148LL * 16) g = 148LL * 16;
    return (unsigned)(g < 1 ? 1 : g);
}

// eps = e_unit for every sample (exact-trace reverse sweep on the tensor path)
__global__ void k_onehot_rows(float* __restrict__ e, long long Bs, int D, int unit) {
    const long long n = (long long)D * Bs;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (long long)gridDim.x * blockDim.x)
        e[p] = (p / Bs == unit) ? 1.0f : 0.0f;
}

// reverse sweep over the accepted steps; d_bw: [0..4] stage inputs s_2..s_6, [5..10] kb_1..kb_6,
// [11] ubar.  On return ubar holds d loss / d z(t0) and h->d_grad the parameter gradient.
static int backward_sweep(cnf_handle* h, const CnfBufs& bf, int mode, long long naccept, long long Bglobal) {
    const cnf_config& c = h->cfg;
    const long long nz = (long long)c.D * bf.Bs;
    float** w = h->d_bw;
    float* ubar = w[11];
    float* kb[6] = {w[5], w[6], w[7], w[8], w[9], w[10]};
    const float invB = 1.0f / (float)Bglobal;
    const bool train = mode == CNF_MODE_TRAIN;
    const float c_l = invB, c_E = train ? c.lambda1 * invB : 0.0f, c_n = train ? c.lambda2 * invB : 0.0f;
    int rc;
    if (tile_sweep(h, bf, mode, naccept, Bglobal)) return CNF_OK;   // one-tile batch: the whole sweep in one launch
    const bool tcb = train && tc_bw_supported(h);
    // TestMode (exact trace) on wide nets: the D unit-vector VJPs on the tensor-core sweep as well (eps = e_unit in
    // d_bw[13], zero zdot cotangent in d_bw[14] for unit > 0)
    const bool tct = !train && tc_supported(c, CNF_MODE_TEST) && tc_bw_supported(h);
    if (tct) {
        if (h->cap_bw13 < nz) {
            for (int j = 13; j < 15; ++j) {
                cudaFree(w[j]);
                w[j] = nullptr;
                CNF_CUDA(cudaMalloc(&w[j], (size_t)nz * sizeof(float)));
            }
            h->cap_bw13 = nz;
        }
        CNF_CUDA(cudaMemsetAsync(w[14], 0, (size_t)nz * sizeof(float), h->stream));
    }
    for (long long n = naccept - 1; n >= 0; --n) {
        const float* un = h->slots[n];
        const float* dtp = h->d_dts + n;
        // the stage inputs s_2..s_6 of step n: from the taped k_1..k_5 of the forward solve when the step
        // has a tape entry, else stage checkpointing (recompute k_1..k_5 from u_n)
        const bool taped = !h->ktape_off && h->ktape_elems == nz && n < (long long)h->ktape.size();
        CnfBufs bk = bf;
        if (taped)
            for (int j = 0; j < 5; ++j) bk.K[j] = h->ktape[(size_t)n] + (long long)j * nz;
        else if ((rc = bw_eval(h, bf, mode, un, h->d_K[0]))) return rc;
        for (int i = 1; i <= 5; ++i) {
            k_combine_bw<<<bw_grid(nz), CNF_NT, 0, h->stream>>>(un, bk, i, dtp, w[i - 1], nz);
            h->launches++;
            if (!taped && i < 5 && (rc = bw_eval(h, bf, mode, w[i - 1], h->d_K[i]))) return rc;
        }
        k_kbar_init<<<bw_grid(nz), CNF_NT, 0, h->stream>>>(ubar, dtp, kb[0], kb[1], kb[2], kb[3], kb[4], kb[5], nz);
        h->launches++;
        for (int i = 5; i >= 0; --i) {
            const float* src = (i == 0) ? un : w[i - 1];
            if (tcb) rc = tc_vjp_stage(h, bf, src, kb[i], i, dtp, c_l, c_E, c_n, ubar, kb, h->d_grad);
            else if (train) rc = launch_vjp_generic_stage(h, bf, src, kb[i], i, dtp, c_l, c_E, c_n, ubar, kb, h->d_grad);
            else if (tct)
                for (int unit = 0; unit < c.D && !rc; ++unit) {
                    k_onehot_rows<<<bw_grid(nz), CNF_NT, 0, h->stream>>>(w[13], bf.Bs, c.D, unit);
                    h->launches++;
                    CnfBufs bu = bf;
                    bu.eps = w[13];
                    rc = tc_vjp_stage(h, bu, src, unit == 0 ? kb[i] : w[14], i, dtp, c_l, 0.0f, 0.0f, ubar, kb, h->d_grad);
                }
            else   // exact trace: one Hutchinson-shaped VJP per unit vector (reference src/utils.jl:1-36 seeds)
                for (int unit = 0; unit < c.D && !rc; ++unit)
                    rc = launch_vjp_generic_stage(h, bf, src, kb[i], i, dtp, c_l, 0.0f, 0.0f, ubar, kb, h->d_grad, unit);
            if (rc) return rc;
        }
    }
    return CNF_OK;
}
