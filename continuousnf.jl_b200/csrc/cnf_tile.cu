// cnf_tile.cu — whole-solve kernels for batches that fit ONE tile of the generic stage kernel.
//
// The reference's default training configuration is small: MLJ `fit` with `batch_size = 32`
// (src/exts/mlj_ext/core_icnf.jl:17-24, README example).  At that size one loss+gradient through the
// step driver is ~860 dependent launches of 3-14 us single-CTA kernels (scripts/small_batch.py): the
// GPU idles between them and the host round trip per attempted step dominates.  Here the same device
// bodies (d_combine, d_eval_generic, d_errnorm, d_controller, d_tape_stages; d_combine_bw, d_kbar_init,
// d_vjp_generic) run inside two single-CTA kernels with __syncthreads() where the launch boundaries
// were: k_solve_tile loops over attempted steps until the controller says done, k_sweep_tile loops
// over the accepted steps in reverse.  Same code, same order of operations, same buffers: results are
// identical to the multi-launch path.  Conditions: one rank, generic path, B <= tile width, stage tape.
#include "cnf_internal.h"

#include <cooperative_groups.h>

// launch boundary of the per-stage path: a block barrier for one tile, a grid barrier (cooperative launch,
// one CTA per tile, all co-resident) for several
template <bool COOP>
__device__ __forceinline__ void tile_sync() {
    if (COOP) cooperative_groups::this_grid().sync();
    else __syncthreads();
}

template <int TB, bool COOP>
__global__ void __launch_bounds__(CNF_NT)
k_solve_tile(CnfCtrl* c, CnfBufs bf, CnfNet net, GenLayout lay, int mode, float* dts, float* const* ktape, long long nz,
             long long ktape_cap, long long slot_cap, int max_attempts) {
    for (int a = 0; a < max_attempts; ++a) {
        if (c->done) break;
        if (c->use_tape && c->naccept + 2 >= slot_cap) break;   // the host grows the tape and relaunches
        for (int i = 1; i <= 5; ++i) {
            d_combine(c, bf, i, 0);
            tile_sync<COOP>();
            d_eval_generic<TB>(c, bf, net, lay, mode, SRC_S, nullptr, i, nullptr, 1);
            tile_sync<COOP>();
        }
        d_combine(c, bf, 6, 0);
        tile_sync<COOP>();
        d_eval_generic<TB>(c, bf, net, lay, mode, SRC_UNEW, nullptr, 6, nullptr, 1);
        tile_sync<COOP>();
        d_errnorm(c, bf);
        tile_sync<COOP>();
        if (blockIdx.x == 0 && threadIdx.x == 0) d_controller(c, dts);
        tile_sync<COOP>();
        if (ktape) {
            d_tape_stages(c, bf, ktape, nz, ktape_cap);
            tile_sync<COOP>();
        }
    }
}

// the solve kernel also holds the error-norm reduction's static shared memory
#define TILE_SMEM_MAX (227 * 1024 - 1024)

struct SweepArgs {
    float* w[5];      // stage inputs s_2..s_6
    float* kb[6];     // cotangents of k_1..k_6
    float* ubar;
    float* grad;
    const float* dts;
    float* const* ktape;
    long long nz;
    long long naccept;
    float c_l, c_E, c_n;
    int train, D;
};

template <int TB, bool COOP>
__global__ void __launch_bounds__(CNF_NT)
k_sweep_tile(CnfBufs bf, CnfNet net, BwLayout lay, SweepArgs sa) {
    for (long long n = sa.naccept - 1; n >= 0; --n) {
        const float* un = bf.slots[n];
        const float* dtp = sa.dts + n;
        CnfBufs bk = bf;
        for (int j = 0; j < 5; ++j) bk.K[j] = sa.ktape[n] + (long long)j * sa.nz;
        for (int i = 1; i <= 5; ++i) d_combine_bw(un, bk, i, dtp, sa.w[i - 1], sa.nz);
        d_kbar_init(sa.ubar, dtp, sa.kb[0], sa.kb[1], sa.kb[2], sa.kb[3], sa.kb[4], sa.kb[5], sa.nz);
        tile_sync<COOP>();
        for (int i = 5; i >= 0; --i) {
            const float* src = (i == 0) ? un : sa.w[i - 1];
            if (sa.train) {
                d_vjp_generic<TB>(bf, net, lay, src, sa.kb[i], i, dtp, sa.c_l, sa.c_E, sa.c_n, sa.ubar, sa.kb[0], sa.kb[1],
                                  sa.kb[2], sa.kb[3], sa.kb[4], sa.grad, -1);
                tile_sync<COOP>();
            } else {
                for (int unit = 0; unit < sa.D; ++unit) {
                    d_vjp_generic<TB>(bf, net, lay, src, sa.kb[i], i, dtp, sa.c_l, 0.0f, 0.0f, sa.ubar, sa.kb[0], sa.kb[1],
                                      sa.kb[2], sa.kb[3], sa.kb[4], sa.grad, unit);
                    tile_sync<COOP>();
                }
            }
        }
    }
}

// tile width the generic forward / reverse kernels would pick for this network (0: does not fit)
static int tile_tb_eval(const cnf_config& c, int* smem) { return pick_tb(gen_layout(c, 0).total, smem); }
static int tile_tb_bw(const cnf_config& c, int* smem) { return pick_tb(bw_layout(c, 0).total, smem); }

// the whole-solve kernels pay for their generality in registers: they only win while a stage kernel is
// launch-latency-bound, i.e. for small networks (measured: hidden 256 is 20 % slower in one kernel)
static bool tile_small_net(const cnf_config& c) {
    long long P = 0;
    for (int l = 0; l < c.n_layers; ++l) P += (long long)c.dims[l] * c.dims[l + 1];
    return P <= 8192;
}

static int tile_sms() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    return sms;
}

// launch `kern` on `tiles` CTAs: plain launch for one tile, cooperative launch (grid barriers) for more;
// false if the CTAs cannot all be resident
template <typename K>
static bool tile_launch(cnf_handle* h, K kern, int tiles, int smem, void** args) {
    if (tiles == 1) return cudaLaunchKernel((const void*)kern, dim3(1), dim3(CNF_NT), args, (size_t)smem, h->stream) == cudaSuccess;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, CNF_NT, (size_t)smem) != cudaSuccess) return false;
    if ((long long)per_sm * tile_sms() < tiles) return false;
    return cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)tiles), dim3(CNF_NT), args, (size_t)smem, h->stream) ==
           cudaSuccess;
}

template <typename K>
static long long tile_capacity(K kern, int smem) {
    int per_sm = 0;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX);
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, CNF_NT, (size_t)smem) != cudaSuccess) return 0;
    return (long long)per_sm * tile_sms();
}

#define TILE_DISPATCH(TBV, EXPR1, EXPRN)                         \
    switch (TBV) {                                               \
        case 32: { constexpr int TB = 32; if (one) { EXPR1; } else { EXPRN; } } break; \
        case 16: { constexpr int TB = 16; if (one) { EXPR1; } else { EXPRN; } } break; \
        case 8: { constexpr int TB = 8; if (one) { EXPR1; } else { EXPRN; } } break;   \
        default: { constexpr int TB = 4; if (one) { EXPR1; } else { EXPRN; } } break;  \
    }

// forward: B samples in ceil(B / TB) tiles, all resident at once
bool tile_solve_ok(cnf_handle* h, int mode, long long B) {
    if (h->tile_off || h->cfg.kernel_path == 1 || h->nranks > 1 || h->time_stages || !tile_small_net(h->cfg)) return false;
    if (use_narrow(h, mode) || use_tc(h, mode)) return false;
    int smem = 0;
    const int tb = tile_tb_eval(h->cfg, &smem);
    if (tb <= 0 || smem > TILE_SMEM_MAX) return false;
    const long long tiles = (B + tb - 1) / tb;
    if (tiles == 1) return true;
    const bool one = false;
    long long cap = 0;
    TILE_DISPATCH(tb, (void)0, cap = tile_capacity(k_solve_tile<TB, true>, smem));
    return tiles <= cap;
}

// the attempt loop of solve_device in one launch (relaunched only if the tape has to grow); the initial
// k1 / dt launches have been enqueued
int tile_solve(cnf_handle* h, const CnfBufs& bf, int mode, bool use_tape, long long maxiters, CnfCtrl* out) {
    int smem = 0;
    const int tbv = tile_tb_eval(h->cfg, &smem);
    GenLayout lay = gen_layout(h->cfg, tbv);
    const int tiles = (int)((bf.B + tbv - 1) / tbv);
    const bool one = tiles == 1;
    int rc;
    size_t want = 66;   // tape entries for the first launch; grown only when a launch stops at the cap
    for (long long launched = 0;;) {
        if (use_tape) {
            if ((rc = ensure_slots(h, want))) return rc;
            if ((rc = ensure_ktape(h, want, (long long)h->cfg.D * bf.Bs))) return rc;
        }
        int chunk = 4096, md = mode;
        CnfCtrl* dc = h->d_ctrl;
        CnfBufs b = bf;
        CnfNet net = h->net;
        float* dts = h->d_dts;
        float* const* kt = (use_tape && !h->ktape.empty()) ? h->d_ktape : nullptr;
        long long nz = (long long)h->cfg.D * bf.Bs, kcap = (long long)h->ktape.size(), scap = (long long)h->slots.size();
        void* args[] = {&dc, &b, &net, &lay, &md, &dts, &kt, &nz, &kcap, &scap, &chunk};
        bool ok = false;
        bool attr[2][33] = {};   // (set on every call: the attribute is per device and the call is cheap)
        TILE_DISPATCH(tbv,
            { if (!attr[0][TB]) { cudaFuncSetAttribute(k_solve_tile<TB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX); attr[0][TB] = true; }
              ok = tile_launch(h, k_solve_tile<TB, false>, 1, smem, args); },
            { if (!attr[1][TB]) { cudaFuncSetAttribute(k_solve_tile<TB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX); attr[1][TB] = true; }
              ok = tile_launch(h, k_solve_tile<TB, true>, tiles, smem, args); });
        if (!ok) { h->err = std::string("whole-solve kernel launch failed: ") + cudaGetErrorString(cudaGetLastError()); return CNF_ERR_CUDA; }
        h->launches++;
        launched += chunk;
        CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[0], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
        CNF_CUDA(cudaStreamSynchronize(h->stream));
        if (h->h_ctrl[0].done) break;
        if (launched > maxiters + 2) break;
        if (use_tape && h->slots.size() >= CNF_MAX_SLOTS) break;   // the controller flags MaxIters at the slot limit
        want = h->slots.size() + 256;
    }
    *out = h->h_ctrl[0];
    CNF_CUDA(cudaGetLastError());
    return CNF_OK;
}

// reverse sweep in one launch; false = conditions not met (the caller uses the launch-per-stage sweep)
bool tile_sweep(cnf_handle* h, const CnfBufs& bf, int mode, long long naccept, long long Bglobal) {
    if (h->tile_off || h->cfg.kernel_path == 1 || h->nranks > 1 || h->time_stages || !tile_small_net(h->cfg)) return false;
    if (use_narrow(h, mode) || use_tc(h, mode)) return false;
    const cnf_config& c = h->cfg;
    const long long nz = (long long)c.D * bf.Bs;
    if (h->ktape_off || h->ktape_elems != nz || naccept > (long long)h->ktape.size()) return false;
    int smem = 0;
    const int tbv = tile_tb_bw(c, &smem);
    if (tbv <= 0 || smem > TILE_SMEM_MAX) return false;
    const int tiles = (int)((bf.B + tbv - 1) / tbv);
    const bool one = tiles == 1;
    BwLayout lay = bw_layout(c, tbv);
    const bool train = mode == CNF_MODE_TRAIN;
    const float invB = 1.0f / (float)Bglobal;
    SweepArgs sa;
    for (int j = 0; j < 5; ++j) sa.w[j] = h->d_bw[j];
    for (int j = 0; j < 6; ++j) sa.kb[j] = h->d_bw[5 + j];
    sa.ubar = h->d_bw[11];
    sa.grad = h->d_grad;
    sa.dts = h->d_dts;
    sa.ktape = h->d_ktape;
    sa.nz = nz;
    sa.naccept = naccept;
    sa.c_l = invB;
    sa.c_E = train ? c.lambda1 * invB : 0.0f;
    sa.c_n = train ? c.lambda2 * invB : 0.0f;
    sa.train = train ? 1 : 0;
    sa.D = c.D;
    CnfBufs b = bf;
    CnfNet net = h->net;
    void* args[] = {&b, &net, &lay, &sa};
    bool ok = false;
    bool attr[2][33] = {};   // (set on every call: the attribute is per device and the call is cheap)
    TILE_DISPATCH(tbv,
        { if (!attr[0][TB]) { cudaFuncSetAttribute(k_sweep_tile<TB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX); attr[0][TB] = true; }
          ok = tile_launch(h, k_sweep_tile<TB, false>, 1, smem, args); },
        { if (!attr[1][TB]) { cudaFuncSetAttribute(k_sweep_tile<TB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX); attr[1][TB] = true; }
          ok = tile_launch(h, k_sweep_tile<TB, true>, tiles, smem, args); });
    if (!ok) { cudaGetLastError(); return false; }   // e.g. more tiles than can be resident: launch-per-stage sweep
    h->launches++;
    return true;
}
