// cnf_stage_generic.cu — generic FP32 SIMT stage kernels for any Dense/tanh chain.
// One CTA owns a tile of TB samples; all activations of the tile live in shared memory and the
// whole forward + VJP/JVP chain + trace/regulariser rows run inside the kernel (no global
// activation workspace).  Weights stream from L2 through the read-only path.
// Restates augmented_f: reference src/icnf.jl:148-164 (TestMode, exact trace via
// src/utils.jl:1-36), :318-350 (TrainMode VecJac), :384-420 (TrainMode JacVec).
#include "cnf_internal.h"

enum { EPI_NONE = 0, EPI_TANH = 1, EPI_MULD = 2 };
enum { SRC_S = 0, SRC_UNEW = 1, SRC_SLOT0 = 2, SRC_PTR = 3 };

// out[o][s] = epi( (bias? bias[o]:0) + sum_{k<K} Wrow[o*ldw + k] * in[k][s] ), o < N
// EPI_MULD multiplies by (1 - hm[o][s]^2).  Accumulation order: ascending k from the bias
// (the oracle's order).  4 outputs per thread share each activation load.
template <int TB>
__device__ __forceinline__ void dense_tile(const float* __restrict__ Wrow, int ldw,
                                           const float* __restrict__ bias, int K, int N,
                                           const float* in, float* out, int epi, const float* hm) {
    constexpr int TBP = TB + 1;
    constexpr int G = CNF_NT / TB;
    const int s = threadIdx.x % TB, og = threadIdx.x / TB;
    for (int o0 = og; o0 < N; o0 += 4 * G) {
        float acc[4];
        const float* w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int o = o0 + j * G;
            int oc = o < N ? o : o0;
            w[j] = Wrow + (long long)oc * ldw;
            acc[j] = bias ? __ldg(bias + oc) : 0.0f;
        }
        for (int k = 0; k < K; ++k) {
            float x = in[k * TBP + s];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j] = fmaf(__ldg(w[j] + k), x, acc[j]);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int o = o0 + j * G;
            if (o < N) {
                float v = acc[j];
                if (epi == EPI_TANH) v = cnf_tanh(v);
                else if (epi == EPI_MULD) { float hh = hm[o * TBP + s]; v *= fmaf(-hh, hh, 1.0f); }
                out[o * TBP + s] = v;
            }
        }
    }
}

struct GenLayout {
    int hoff[CNF_MAX_LAYERS + 1];  // float offsets of h_l
    int gA, gB, e, red;            // ping-pong chain buffers, eps tile, per-sample scalars
    int z0;                        // zeros: stands in for h_L in 1 - h^2 when the last layer is linear
    int total;                     // floats
};
static GenLayout gen_layout(const cnf_config& cfg, int TB) {
    GenLayout L;
    int TBP = TB + 1, o = 0, maxd = 0;
    for (int l = 0; l <= cfg.n_layers; ++l) {
        L.hoff[l] = o;
        o += cfg.dims[l] * TBP;
        if (cfg.dims[l] > maxd) maxd = cfg.dims[l];
    }
    L.gA = o; o += maxd * TBP;
    L.gB = o; o += maxd * TBP;
    L.e = o; o += cfg.D * TBP;
    L.red = o; o += 4 * TBP;
    L.z0 = o; o += cfg.D * TBP;
    L.total = o;
    return L;
}

// forward chain into smem h_0..h_L ; h_0 = [z; ys] (CondLayer, reference src/layers/cond_layer.jl:1-9)
template <int TB>
__device__ void gen_forward(const CnfNet& net, const GenLayout& lay, float* sm, const float* src,
                            const float* ys, long long Bs, long long b0, int B) {
    constexpr int TBP = TB + 1;
    const int d0 = net.dims[0];
    for (int idx = threadIdx.x; idx < d0 * TB; idx += CNF_NT) {
        int i = idx / TB, s = idx % TB;
        long long b = b0 + s;
        float v = 0.0f;
        if (b < B) v = (i < net.D) ? src[(long long)i * Bs + b] : ys[(long long)(i - net.D) * Bs + b];
        sm[lay.hoff[0] + i * TBP + s] = v;
    }
    __syncthreads();
    for (int l = 0; l < net.L; ++l) {
        // CNF_FLAG_LAST_LINEAR: the last layer is a plain product (PlanarLayer's u * act(.), planar_layer.jl:70-88)
        const bool lin = (net.flags & CNF_FLAG_LAST_LINEAR) && l == net.L - 1;
        dense_tile<TB>(net.Wt[l], net.dims[l], lin ? nullptr : net.b[l], net.dims[l], net.dims[l + 1],
                       sm + lay.hoff[l], sm + lay.hoff[l + 1], lin ? EPI_NONE : EPI_TANH, nullptr);
        __syncthreads();
    }
}

// h_l as the argument of the activation derivative 1 - h^2 (zeros for a linear last layer: derivative 1)
__device__ __forceinline__ const float* gen_hder(const CnfNet& net, const GenLayout& lay, const float* sm, int l) {
    return ((net.flags & CNF_FLAG_LAST_LINEAR) && l == net.L) ? sm + lay.z0 : sm + lay.hoff[l];
}

// reverse chain from cotangent g_L (already multiplied by 1-h_L^2) in buffer `cur`;
// returns the buffer holding q_1 = (W_1^T g_1)[0:D]
template <int TB>
__device__ float* gen_vjp_chain(const CnfNet& net, const GenLayout& lay, float* sm, float* cur,
                                float* oth) {
    for (int l = net.L; l >= 1; --l) {
        int din = net.dims[l - 1], dout = net.dims[l];
        int used = (l == 1) ? net.D : din;
        // q_l[i] = sum_o W[o,i] g[o] : col-major W == row-major [din][dout]
        dense_tile<TB>(net.W[l - 1], dout, nullptr, dout, used, cur, oth,
                       l > 1 ? EPI_MULD : EPI_NONE, sm + lay.hoff[l - 1]);
        __syncthreads();
        float* t = cur; cur = oth; oth = t;
    }
    return cur;
}

// forward tangent chain from tau_1 (already multiplied by 1-h_1^2) in `cur`; returns tau_L buffer
template <int TB>
__device__ float* gen_jvp_chain(const CnfNet& net, const GenLayout& lay, float* sm, float* cur,
                                float* oth) {
    for (int l = 2; l <= net.L; ++l) {
        dense_tile<TB>(net.Wt[l - 1], net.dims[l - 1], nullptr, net.dims[l - 1], net.dims[l], cur,
                       oth, EPI_MULD, gen_hder(net, lay, sm, l));
        __syncthreads();
        float* t = cur; cur = oth; oth = t;
    }
    return cur;
}

template <int TB>
__device__ __forceinline__ void d_eval_generic(const CnfCtrl* c, const CnfBufs& bf, const CnfNet& net, const GenLayout& lay, int mode,
                                               int src_sel, const float* src_ptr, int kidx, float* kout_ptr, int check_done) {
    if (check_done && c->done) return;
    extern __shared__ float sm[];
    constexpr int TBP = TB + 1;
    const int D = net.D, L = net.L;
    const long long Bs = bf.Bs;
    const float* src = src_ptr;
    if (src_sel == SRC_S) src = bf.S;
    else if (src_sel == SRC_UNEW) src = bf.slots[cnf_slot(c, c->naccept + 1)];
    else if (src_sel == SRC_SLOT0) src = bf.slots[cnf_slot(c, 0)];
    float* kout = kout_ptr;
    if (!kout) kout = (kidx == 6) ? bf.K[6 - c->k1idx] : (kidx == 0 ? bf.K[c->k1idx] : bf.K[kidx]);
    const long long b0 = (long long)blockIdx.x * TB;
    for (int idx = threadIdx.x; idx < D * TBP; idx += CNF_NT) sm[lay.z0 + idx] = 0.0f;
    gen_forward<TB>(net, lay, sm, src, bf.ys, Bs, b0, bf.B);
    float* hL = sm + lay.hoff[L];
    const float* hLd = gen_hder(net, lay, sm, L);   // h_L where 1 - h_L^2 is meant
    float* gA = sm + lay.gA;
    float* gB = sm + lay.gB;
    float* e = sm + lay.e;
    float* red = sm + lay.red;  // [0]=ldot [1]=||v||^2
    const int tid = threadIdx.x;
    if (tid < TB) { red[tid] = 0.0f; red[TBP + tid] = 0.0f; }
    if (mode == CNF_MODE_TEST) {
        for (int i = 0; i < D; ++i) {   // jacobian_batched: one one-hot sweep per dimension
            float* res;
            if (net.trace_mode == CNF_TRACE_VJP) {
                for (int idx = tid; idx < D * TB; idx += CNF_NT) {
                    int o = idx / TB, s = idx % TB;
                    float hh = hLd[o * TBP + s];
                    gA[o * TBP + s] = (o == i) ? fmaf(-hh, hh, 1.0f) : 0.0f;
                }
                __syncthreads();
                res = gen_vjp_chain<TB>(net, lay, sm, gA, gB);
            } else {
                const int d1 = net.dims[1];
                const float* h1 = gen_hder(net, lay, sm, 1);
                for (int idx = tid; idx < d1 * TB; idx += CNF_NT) {
                    int o = idx / TB, s = idx % TB;
                    float hh = h1[o * TBP + s];
                    gA[o * TBP + s] = __ldg(net.Wt[0] + (long long)o * net.dims[0] + i) * fmaf(-hh, hh, 1.0f);
                }
                __syncthreads();
                res = gen_jvp_chain<TB>(net, lay, sm, gA, gB);
            }
            if (tid < TB) red[tid] -= res[i * TBP + tid];
            __syncthreads();
        }
    } else {
        for (int idx = tid; idx < D * TB; idx += CNF_NT) {
            int i = idx / TB, s = idx % TB;
            long long b = b0 + s;
            e[i * TBP + s] = (b < bf.B) ? bf.eps[(long long)i * Bs + b] : 0.0f;
        }
        __syncthreads();
        float* res;
        if (net.trace_mode == CNF_TRACE_VJP) {
            for (int idx = tid; idx < D * TB; idx += CNF_NT) {
                int o = idx / TB, s = idx % TB;
                float hh = hLd[o * TBP + s];
                gA[o * TBP + s] = e[o * TBP + s] * fmaf(-hh, hh, 1.0f);
            }
            __syncthreads();
            res = gen_vjp_chain<TB>(net, lay, sm, gA, gB);
        } else {
            dense_tile<TB>(net.Wt[0], net.dims[0], nullptr, D, net.dims[1], e, gA, EPI_MULD,
                           gen_hder(net, lay, sm, 1));
            __syncthreads();
            res = gen_jvp_chain<TB>(net, lay, sm, gA, gB);
        }
        if (tid < TB) {
            float ld = 0.0f, nv = 0.0f;
            for (int i = 0; i < D; ++i) {
                float v = res[i * TBP + tid];
                ld = fmaf(-v, e[i * TBP + tid], ld);
                nv = fmaf(v, v, nv);
            }
            red[tid] = ld;
            red[TBP + tid] = nv;
        }
        __syncthreads();
    }
    // write du rows (coalesced along samples)
    for (int idx = tid; idx < D * TB; idx += CNF_NT) {
        int i = idx / TB, s = idx % TB;
        long long b = b0 + s;
        if (b < Bs) kout[(long long)i * Bs + b] = hL[i * TBP + s];
    }
    if (tid < TB) {
        long long b = b0 + tid;
        if (b < Bs) {
            kout[(long long)D * Bs + b] = red[tid];
            if (mode == CNF_MODE_TRAIN) {
                float ny = 0.0f;
                for (int i = 0; i < D; ++i) { float y = hL[i * TBP + tid]; ny = fmaf(y, y, ny); }
                kout[(long long)(D + 1) * Bs + b] = (net.flags & CNF_FLAG_NORM_Z) ? sqrtf(ny) : 0.0f;
                kout[(long long)(D + 2) * Bs + b] = (net.flags & CNF_FLAG_NORM_J) ? sqrtf(red[TBP + tid]) : 0.0f;
            }
        }
    }
}

template <int TB>
__global__ void __launch_bounds__(CNF_NT)
k_eval_generic(const CnfCtrl* c, CnfBufs bf, CnfNet net, GenLayout lay, int mode,
               int src_sel, const float* src_ptr, int kidx, float* kout_ptr, int check_done) {
    d_eval_generic<TB>(c, bf, net, lay, mode, src_sel, src_ptr, kidx, kout_ptr, check_done);
}

static int pick_tb(int total_floats_per_tbp /*rows*/, int* smem_bytes) {
    const int cand[4] = {32, 16, 8, 4};
    for (int i = 0; i < 4; ++i) {
        long long bytes = (long long)total_floats_per_tbp * (cand[i] + 1) * 4;
        if (bytes <= 227 * 1024) { *smem_bytes = (int)bytes; return cand[i]; }
    }
    return 0;
}

template <int TB>
static int launch_eval_tb(cnf_handle* h, const CnfBufs& bf, const GenLayout& lay, int smem, int mode,
                          int src_sel, const float* src_ptr, int kidx, float* kout_ptr, int check_done) {
    static bool attr_dev[64] = {};   // the opt-in is per device: handles may live on different GPUs of one process
    bool& attr_set = attr_dev[h->device & 63];
    if (!attr_set) {
        cudaFuncSetAttribute(k_eval_generic<TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        attr_set = true;
    }
    unsigned grid = (unsigned)((bf.Bs + TB - 1) / TB);
    k_eval_generic<TB><<<grid, CNF_NT, smem, h->stream>>>(h->d_ctrl, bf, h->net, lay, mode, src_sel,
                                                         src_ptr, kidx, kout_ptr, check_done);
    h->launches++;
    return 0;
}

// src_sel/kidx form used by the solver loop; explicit pointers when src_ptr/kout_ptr are given
int launch_eval_generic_sel(cnf_handle* h, const CnfBufs& bf, int mode, int src_sel,
                            const float* src_ptr, int kidx, float* kout_ptr, int check_done) {
    GenLayout l1 = gen_layout(h->cfg, 0);  // TB=0 -> TBP=1: rows count
    int smem = 0;
    int TB = pick_tb(l1.total, &smem);
    if (!TB) { h->err = "generic stage kernel: network too wide for shared memory"; return CNF_ERR_UNSUPPORTED; }
    GenLayout lay = gen_layout(h->cfg, TB);
    switch (TB) {
        case 32: return launch_eval_tb<32>(h, bf, lay, smem, mode, src_sel, src_ptr, kidx, kout_ptr, check_done);
        case 16: return launch_eval_tb<16>(h, bf, lay, smem, mode, src_sel, src_ptr, kidx, kout_ptr, check_done);
        case 8: return launch_eval_tb<8>(h, bf, lay, smem, mode, src_sel, src_ptr, kidx, kout_ptr, check_done);
        default: return launch_eval_tb<4>(h, bf, lay, smem, mode, src_sel, src_ptr, kidx, kout_ptr, check_done);
    }
}
int launch_eval_generic(cnf_handle* h, const CnfBufs& bf, int mode, const float* src, float* kout) {
    return launch_eval_generic_sel(h, bf, mode, SRC_PTR, src, -1, kout, 0);
}
