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

template <int TB>
__global__ void __launch_bounds__(CNF_NT)
k_solve_tile(CnfCtrl* c, CnfBufs bf, CnfNet net, GenLayout lay, int mode, float* dts, float* const* ktape, long long nz,
             long long ktape_cap, long long slot_cap, int max_attempts) {
    for (int a = 0; a < max_attempts; ++a) {
        if (c->done) break;
        if (c->use_tape && c->naccept + 2 >= slot_cap) break;   // the host grows the tape and relaunches
        for (int i = 1; i <= 5; ++i) {
            d_combine(c, bf, i, 0);
            __syncthreads();
            d_eval_generic<TB>(c, bf, net, lay, mode, SRC_S, nullptr, i, nullptr, 1);
            __syncthreads();
        }
        d_combine(c, bf, 6, 0);
        __syncthreads();
        d_eval_generic<TB>(c, bf, net, lay, mode, SRC_UNEW, nullptr, 6, nullptr, 1);
        __syncthreads();
        d_errnorm(c, bf);
        __syncthreads();
        if (threadIdx.x == 0) d_controller(c, dts);
        __syncthreads();
        if (ktape) {
            d_tape_stages(c, bf, ktape, nz, ktape_cap);
            __syncthreads();
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

template <int TB>
__global__ void __launch_bounds__(CNF_NT)
k_sweep_tile(CnfBufs bf, CnfNet net, BwLayout lay, SweepArgs sa) {
    for (long long n = sa.naccept - 1; n >= 0; --n) {
        const float* un = bf.slots[n];
        const float* dtp = sa.dts + n;
        CnfBufs bk = bf;
        for (int j = 0; j < 5; ++j) bk.K[j] = sa.ktape[n] + (long long)j * sa.nz;
        for (int i = 1; i <= 5; ++i) d_combine_bw(un, bk, i, dtp, sa.w[i - 1], sa.nz);
        d_kbar_init(sa.ubar, dtp, sa.kb[0], sa.kb[1], sa.kb[2], sa.kb[3], sa.kb[4], sa.kb[5], sa.nz);
        __syncthreads();
        for (int i = 5; i >= 0; --i) {
            const float* src = (i == 0) ? un : sa.w[i - 1];
            if (sa.train) {
                d_vjp_generic<TB>(bf, net, lay, src, sa.kb[i], i, dtp, sa.c_l, sa.c_E, sa.c_n, sa.ubar, sa.kb[0], sa.kb[1],
                                  sa.kb[2], sa.kb[3], sa.kb[4], sa.grad, -1);
                __syncthreads();
            } else {
                for (int unit = 0; unit < sa.D; ++unit) {
                    d_vjp_generic<TB>(bf, net, lay, src, sa.kb[i], i, dtp, sa.c_l, 0.0f, 0.0f, sa.ubar, sa.kb[0], sa.kb[1],
                                      sa.kb[2], sa.kb[3], sa.kb[4], sa.grad, unit);
                    __syncthreads();
                }
            }
        }
    }
}

// tile width the generic forward / reverse kernels would pick for this network (0: does not fit)
static int tile_tb_eval(const cnf_config& c, int* smem) { return pick_tb(gen_layout(c, 0).total, smem); }
static int tile_tb_bw(const cnf_config& c, int* smem) { return pick_tb(bw_layout(c, 0).total, smem); }

bool tile_solve_ok(cnf_handle* h, int mode, long long B) {
    if (h->tile_off || h->nranks > 1 || h->time_stages) return false;
    if (use_narrow(h, mode) || use_tc(h, mode)) return false;
    int smem = 0;
    const int tb = tile_tb_eval(h->cfg, &smem);
    return tb > 0 && B <= tb && smem <= TILE_SMEM_MAX;
}

template <int TB>
static void tile_launch_solve(cnf_handle* h, const CnfBufs& bf, const GenLayout& lay, int smem, int mode, int use_tape,
                              int max_attempts) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k_solve_tile<TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX);
        attr_set = true;
    }
    const bool tape = use_tape && !h->ktape.empty();
    k_solve_tile<TB><<<1, CNF_NT, smem, h->stream>>>(h->d_ctrl, bf, h->net, lay, mode, h->d_dts, tape ? h->d_ktape : nullptr,
                                                    (long long)h->cfg.D * bf.Bs, (long long)h->ktape.size(),
                                                    (long long)h->slots.size(), max_attempts);
    h->launches++;
}

// the attempt loop of solve_device for a one-tile batch; the initial k1 / dt launches have been enqueued
int tile_solve(cnf_handle* h, const CnfBufs& bf, int mode, bool use_tape, long long maxiters, CnfCtrl* out) {
    int smem = 0;
    const int TB = tile_tb_eval(h->cfg, &smem);
    const GenLayout lay = gen_layout(h->cfg, TB);
    int rc;
    size_t want = 66;   // tape entries for the first launch; grown only when a launch stops at the cap
    for (long long launched = 0;;) {
        if (use_tape) {
            if ((rc = ensure_slots(h, want))) return rc;
            if ((rc = ensure_ktape(h, want, (long long)h->cfg.D * bf.Bs))) return rc;
        }
        const int chunk = 4096;
        switch (TB) {
            case 32: tile_launch_solve<32>(h, bf, lay, smem, mode, use_tape, chunk); break;
            case 16: tile_launch_solve<16>(h, bf, lay, smem, mode, use_tape, chunk); break;
            case 8: tile_launch_solve<8>(h, bf, lay, smem, mode, use_tape, chunk); break;
            default: tile_launch_solve<4>(h, bf, lay, smem, mode, use_tape, chunk); break;
        }
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

template <int TB>
static void tile_launch_sweep(cnf_handle* h, const CnfBufs& bf, const BwLayout& lay, int smem, const SweepArgs& sa) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(k_sweep_tile<TB>, cudaFuncAttributeMaxDynamicSharedMemorySize, TILE_SMEM_MAX);
        attr_set = true;
    }
    k_sweep_tile<TB><<<1, CNF_NT, smem, h->stream>>>(bf, h->net, lay, sa);
    h->launches++;
}

// reverse sweep of a one-tile batch in one launch; false = conditions not met (caller uses the launch-per-stage sweep)
bool tile_sweep(cnf_handle* h, const CnfBufs& bf, int mode, long long naccept, long long Bglobal) {
    if (!tile_solve_ok(h, mode, bf.B)) return false;
    const cnf_config& c = h->cfg;
    const long long nz = (long long)c.D * bf.Bs;
    if (h->ktape_off || h->ktape_elems != nz || naccept > (long long)h->ktape.size()) return false;
    int smem = 0;
    const int TB = tile_tb_bw(c, &smem);
    if (TB <= 0 || bf.B > TB || smem > TILE_SMEM_MAX) return false;
    const BwLayout lay = bw_layout(c, TB);
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
    switch (TB) {
        case 32: tile_launch_sweep<32>(h, bf, lay, smem, sa); break;
        case 16: tile_launch_sweep<16>(h, bf, lay, smem, sa); break;
        case 8: tile_launch_sweep<8>(h, bf, lay, smem, sa); break;
        default: tile_launch_sweep<4>(h, bf, lay, smem, sa); break;
    }
    return true;
}
