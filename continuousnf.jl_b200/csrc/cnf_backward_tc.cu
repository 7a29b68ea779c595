// cnf_backward_tc.cu — reverse sweep of the discrete adjoint on the tensor path (wide nets,
// TrainMode VecJac Hutchinson).  Replaces, per RK stage, the generic SIMT k_vjp_generic by
//   3L+L layer GEMMs (k_tc_gemm<1,true>: forward chain, eps chain, gbar chain, hbar chain; BF16x3)
//   + L weight-gradient GEMMs (k_tc_wgrad: dW_l += pbar_l^T h_{l-1} + g_l^T qbar_l, db_l += sum pbar_l).
// Same maths as cnf_backward.cu / DESIGN.md "Backward"; reference call sites of the gradient:
// src/exts/mlj_ext/core_icnf.jl:59-63, test/call_tests.jl:239-251 (Enzyme + SciMLSensitivity there).
//
// Weight-gradient GEMM: the reduction runs over the batch, so both operands are read straight
// from the row-major [sample x width] planes as MN-major tcgen05 operands (TMA box 64 samples x
// 64 columns, SWIZZLE_128B; LBO = one box = 8 KB, SBO = one 8-sample group = 1 KB) — no
// transposed copies.  One CTA = one 128 x BN tile of dW_l and a slice of the batch (split-K over
// the SMs); partial tiles are added to the gradient with coalesced FP32 reductions.  The bias
// gradient rides along as an N = 16 MMA of pbar_l against a tile of ones.
#include "cnf_internal.h"

#define WG_THREADS 192   // warp 0 TMA, warp 1 MMA, warps 2..5 epilogue
#define WG_BOX 8192      // one TMA box: 64 samples x 64 bf16

struct WgMaps { CUtensorMap a[2][2], b[2][2]; };   // [term][plane]
struct WgArgs {
    int nterm, BN, m_tiles, n_tiles, ksplit, kblocks;
    int dout, din;
    float* gW;            // grad + w_off: column-major dout x din
    float* gb;            // grad + b_off or nullptr
    uint32_t lbo, sbo;    // MN-major descriptor strides (bytes)
};
struct WgSmemCtl {
    uint64_t full[4], empty[4], tfull;
    uint32_t tmem_base;
};

namespace tc {
// MN-major operand, SWIZZLE_128B: 64 contiguous elements along M/N (128 B) x 8 along K per atom;
// LBO = byte stride between atoms along M/N, SBO = byte stride between atoms along K
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// kind::f16, BF16 x BF16 -> FP32; a_mn / b_mn: operand is MN-major
__device__ __forceinline__ uint32_t umma_idesc_bf16_major(int M, int N, int a_mn, int b_mn) {
    return umma_idesc_bf16(M, N) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16);
}
__device__ __forceinline__ void tmem_ld_32x1(uint32_t taddr, uint32_t& v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(v) : "r"(taddr) : "memory");
}
}  // namespace tc

__global__ void __launch_bounds__(WG_THREADS, 1)
k_tc_wgrad(const __grid_constant__ WgMaps mp, WgArgs ar) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int BN = ar.BN;
    const uint32_t a_bytes = 2 * WG_BOX;                 // one plane of A: 128 columns
    const uint32_t b_bytes = (uint32_t)(BN / 64) * WG_BOX;
    const uint32_t stage_bytes = 2 * (a_bytes + b_bytes);
    const int nstage = 2;
    uint8_t* ones = smem;                                // 2 KB of bf16 1.0
    uint8_t* stages = smem + 2048;
    WgSmemCtl* ctl = reinterpret_cast<WgSmemCtl*>(stages + (size_t)nstage * stage_bytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    const int tile = blockIdx.x / ar.ksplit, ks = blockIdx.x % ar.ksplit;
    const int mb = tile / ar.n_tiles, nb = tile % ar.n_tiles;
    const int kb0 = (int)((long long)ks * ar.kblocks / ar.ksplit);
    const int kb1 = (int)((long long)(ks + 1) * ar.kblocks / ar.ksplit);
    const bool has_bias = ar.gb != nullptr && nb == 0;
    if (kb0 >= kb1) return;

    for (int i = threadIdx.x; i < 512; i += WG_THREADS) reinterpret_cast<uint32_t*>(ones)[i] = 0x3F803F80u;
    tc::fence_proxy_async();
    if (warp == 0 && lane == 0) {
        for (int t = 0; t < ar.nterm; ++t)
            for (int q = 0; q < 2; ++q) { tc::tma_prefetch_desc(&mp.a[t][q]); tc::tma_prefetch_desc(&mp.b[t][q]); }
        for (int s = 0; s < nstage; ++s) { tc::mbar_init(&ctl->full[s], 1); tc::mbar_init(&ctl->empty[s], 1); }
        tc::mbar_init(&ctl->tfull, 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(&ctl->tmem_base, 512);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;

    if (warp == 0) {
        if (lane == 0) {   // ===== TMA producer
            int stage = 0; uint32_t phase = 0;
            for (int t = 0; t < ar.nterm; ++t)
                for (int kb = kb0; kb < kb1; ++kb) {
                    tc::mbar_wait(&ctl->empty[stage], phase ^ 1, 0x201);
                    uint8_t* s = stages + (size_t)stage * stage_bytes;
                    tc::mbar_arrive_expect_tx(&ctl->full[stage], stage_bytes);
                    for (int q = 0; q < 2; ++q) {
                        for (int j = 0; j < 2; ++j)
                            tc::tma_load_2d(&mp.a[t][q], &ctl->full[stage], s + q * a_bytes + j * WG_BOX,
                                            mb * 128 + j * 64, kb * 64);
                        for (int j = 0; j < BN / 64; ++j)
                            tc::tma_load_2d(&mp.b[t][q], &ctl->full[stage], s + 2 * a_bytes + q * b_bytes + j * WG_BOX,
                                            nb * BN + j * 64, kb * 64);
                    }
                    if (++stage == nstage) { stage = 0; phase ^= 1; }
                }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ===== MMA issuer
            const uint32_t idesc = tc::umma_idesc_bf16_major(128, BN, 1, 1);
            const uint32_t idesc_b = tc::umma_idesc_bf16_major(128, 16, 1, 0);
            const uint64_t ones_desc = tc::umma_desc_sw128(tc::smem_u32(ones));
            int stage = 0; uint32_t phase = 0;
            uint32_t first = 1, first_b = 1;
            for (int t = 0; t < ar.nterm; ++t)
                for (int kb = kb0; kb < kb1; ++kb) {
                    tc::mbar_wait(&ctl->full[stage], phase, 0x202);
                    tc::tc_fence_after();
                    const uint32_t s = tc::smem_u32(stages + (size_t)stage * stage_bytes);
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {   // 16 samples per MMA = two 8-sample atoms
                        uint64_t a[2], b[2];
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            a[q] = tc::umma_desc_mn_sw128(s + q * a_bytes + kk * 2 * ar.sbo, ar.lbo, ar.sbo);
                            b[q] = tc::umma_desc_mn_sw128(s + 2 * a_bytes + q * b_bytes + kk * 2 * ar.sbo, ar.lbo, ar.sbo);
                        }
                        tc::umma_bf16(tmem, a[0], b[0], idesc, first ^ 1);
                        first = 0;
                        tc::umma_bf16(tmem, a[1], b[0], idesc, 1);
                        tc::umma_bf16(tmem, a[0], b[1], idesc, 1);
                        if (has_bias && t == 0) {
                            tc::umma_bf16(tmem + 256, a[0], ones_desc, idesc_b, first_b ^ 1);
                            first_b = 0;
                            tc::umma_bf16(tmem + 256, a[1], ones_desc, idesc_b, 1);
                        }
                    }
                    tc::umma_commit(&ctl->empty[stage]);
                    if (++stage == nstage) { stage = 0; phase ^= 1; }
                }
            tc::umma_commit(&ctl->tfull);
        }
    } else {               // ===== epilogue: lane = output feature o, columns = input features i
        const int quad = warp & 3;
        tc::mbar_wait(&ctl->tfull, 0, 0x203);
        tc::tc_fence_after();
        const int o = mb * 128 + quad * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t)(quad * 32) << 16);
        for (int c0 = 0; c0 < BN; c0 += 32) {
            uint32_t v[32];
            tc::tmem_ld_32x32(tbase + c0, v);
            tc::tmem_ld_wait();
            if (o < ar.dout) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int i = nb * BN + c0 + j;
                    if (i < ar.din) atomicAdd(ar.gW + (long long)i * ar.dout + o, __uint_as_float(v[j]));
                }
            }
        }
        if (has_bias) {
            uint32_t bv;
            tc::tmem_ld_32x1(tbase + 256, bv);
            tc::tmem_ld_wait();
            if (o < ar.dout) atomicAdd(ar.gb + o, __uint_as_float(bv));
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

// zbar (D x Bs scratch written by the TCB_ZBAR epilogue) -> ubar += zbar, kbar_j += dt a_{stage,j} zbar
__global__ void k_zbar_apply(const float* __restrict__ zbar, const float* __restrict__ dtp, int stage,
                             float* __restrict__ ubar, float* kb0, float* kb1, float* kb2, float* kb3, float* kb4,
                             long long n) {
    const float dt = *dtp;
    float* kb[5] = {kb0, kb1, kb2, kb3, kb4};
    for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
        const float zb = zbar[e];
        ubar[e] += zb;
#pragma unroll
        for (int s = 0; s < 5; ++s)
            if (s < stage) kb[s][e] = fmaf(dt * c_A[stage][s], zb, kb[s][e]);
    }
}

// ------------------------------------------------------------------ host side
bool tc_bw_supported(cnf_handle* h) {
    // VecJac always; JacVec needs one more plane pair (QB[L + 1]), so L <= CNF_MAX_LAYERS - 1
    return tc_supported(h->cfg, CNF_MODE_TRAIN) && h->tc_state != nullptr && !h->tc_bw_off &&
           (h->cfg.trace_mode == CNF_TRACE_VJP || h->cfg.n_layers < CNF_MAX_LAYERS);
}

static int tc_ensure_bw(cnf_handle* h, TcState* s) {
    if (s->bw_ok) return CNF_OK;
    const cnf_config& c = h->cfg;
    const int L = c.n_layers;
    int rc;
    for (int l = 1; l <= L; ++l) {
        if ((rc = tc_alloc_plane(h, s->QB[l], s->Mc, pad_to(c.dims[l - 1], TC_BK), TC_BM))) return rc;
        if ((rc = tc_alloc_plane(h, s->PB[l], s->Mc, pad_to(c.dims[l], TC_BK), TC_BM))) return rc;
        if (l < L && (rc = tc_alloc_f32(h, s->Q[l], s->Mc, pad_to(c.dims[l], TC_BK)))) return rc;
    }
    if ((rc = tc_alloc_f32(h, s->Y, s->Mc, pad_to(c.D, TC_BK)))) return rc;
    if ((rc = tc_alloc_f32(h, s->YB, s->Mc, pad_to(c.D, TC_BK)))) return rc;
    // JacVec: sbar_l lives in QB[l + 1] (width of layer l), so sbar_L needs one more pair
    if (c.trace_mode == CNF_TRACE_JVP && (rc = tc_alloc_plane(h, s->QB[L + 1], s->Mc, pad_to(c.dims[L], TC_BK), TC_BM))) return rc;
    static bool attr_dev[64] = {};   // per device
    bool& attr = attr_dev[h->device & 63];
    if (!attr) {
        CNF_CUDA(cudaFuncSetAttribute(k_tc_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr = true;
    }
    s->bw_ok = true;
    return CNF_OK;
}

// dW += A1^T B1 (+ A2^T B2), db += colsum(A1) over the mc samples of this chunk
static void tc_launch_wgrad(cnf_handle* h, TcState* s, const TcPlane& A1, const TcPlane& B1, const TcPlane* A2,
                            const TcPlane* B2, int mc, int dout, int din, float* gW, float* gb) {
    WgMaps mp;
    WgArgs ar;
    memset(&ar, 0, sizeof(ar));
    for (int q = 0; q < 2; ++q) {
        mp.a[0][q] = A1.mapk[q]; mp.b[0][q] = B1.mapk[q];
        mp.a[1][q] = A2 ? A2->mapk[q] : A1.mapk[q];
        mp.b[1][q] = B2 ? B2->mapk[q] : B1.mapk[q];
    }
    const int n64 = pad_to(din, 64);
    ar.nterm = A2 ? 2 : 1;
    ar.BN = n64 < 256 ? n64 : 256;
    ar.n_tiles = (n64 + ar.BN - 1) / ar.BN;
    ar.m_tiles = (dout + 127) / 128;
    ar.kblocks = mc / 64;
    const int tiles = ar.m_tiles * ar.n_tiles;
    int ksplit = s->sms / tiles;
    if (ksplit < 1) ksplit = 1;
    if (ksplit > ar.kblocks) ksplit = ar.kblocks;
    ar.ksplit = ksplit;
    ar.dout = dout; ar.din = din; ar.gW = gW; ar.gb = gb;
    ar.lbo = h->wg_lbo ? (uint32_t)h->wg_lbo : WG_BOX;
    ar.sbo = h->wg_sbo ? (uint32_t)h->wg_sbo : 1024;
    const int stage_bytes = 2 * (2 * WG_BOX + (ar.BN / 64) * WG_BOX);
    const int smem = 2048 + 2 * stage_bytes + 1024 + (int)sizeof(WgSmemCtl) + 64;
    k_tc_wgrad<<<tiles * ksplit, WG_THREADS, smem, h->stream>>>(mp, ar);
    h->launches++;
}

// JacVec (reference src/icnf.jl:384-420 under the outer gradient): the Hutchinson term comes from the forward-mode chain
//   tau_0 = [eps; 0], s_l = W_l tau_{l-1}, tau_l = s_l .* d_l (d_l = 1 - h_l^2);  ldot = -eps . tau_L, ndot = |tau_L|.
// Reverse of one evaluation (cotangents abar of zdot, c_l / c_E / c_n of the quadrature rows, all scaled by dt a_7i):
//   ybar = abar + beta y/|y|,  taubar_L = -alpha eps + gamma tau_L/|tau_L|
//   l = L..1:  sbar_l = taubar_l .* d_l,  dbar_l = taubar_l .* s_l,  dW_l += sbar_l tau_{l-1}^T,  taubar_{l-1} = W_l^T sbar_l
//   l = L..1:  abar_l = (hbar_l - 2 h_l dbar_l) .* d_l,  dW_l += abar_l h_{l-1}^T,  db_l += abar_l,  hbar_{l-1} = W_l^T abar_l
// The same layer-GEMM kernels as the VecJac sweep with the roles of the two weight orientations exchanged: the tangent
// chain runs on the forward planes (epilogue TCB_BWD_HIDDEN keeps s_l in fp32 and writes tau_l), its adjoint on the
// transposed planes (TCB_GBAR_HIDDEN: dbar_l in place of s_l, sbar_l planes); only the last layer needs its own
// epilogue (TCB_JVP_LAST).  Planes: tau_l in G[l] (G[0] = seed), sbar_l in QB[l + 1], abar_l in PB[l], s_l / dbar_l in Q[l].
static int tc_vjp_stage_jvp(cnf_handle* h, TcState* s, const CnfBufs& bf, const float* src, TcEpi base, long long row0, int mc,
                            float* grad) {
    const cnf_config& c = h->cfg;
    const int L = c.n_layers;
    k_tc_prep<1><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, src, c.D, bf.ys, c.C, row0, mc, s->H[0].p[0],
                                                      s->H[0].p[1], s->H[0].ld, -1, nullptr, nullptr, 0);
    k_tc_prep<1><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, bf.eps, c.D, nullptr, 0, row0, mc, s->G[0].p[0],
                                                      s->G[0].p[1], s->G[0].ld, 0, nullptr, nullptr, 0);
    h->launches += 2;
    tc_forward_hidden<1>(h, s, base, mc);
    {   // y, ybar (the g_L planes this kind also writes are not used here)
        TcEpi ep = base;
        ep.kind = TCB_FWD_LAST;
        ep.bias = h->d_theta + h->b_off[L - 1];
        ep.N = c.D;
        for (int q = 0; q < 2; ++q) ep.out[q] = s->G[L].p[q];
        ep.ld_out = s->G[L].ld;
        ep.f32a = s->Y.p; ep.f32b = s->YB.p; ep.ld_f = s->Y.ld;
        tc_launch_gemm<1, true>(h, s, s->H[L - 1], s->wf[L - 1], mc, ep);
    }
    for (int l = 1; l <= L; ++l) {   // tangent chain: s_l = tau_{l-1} W_l^T
        TcEpi ep = base;
        if (l < L) {
            ep.kind = TCB_BWD_HIDDEN;    // s_l kept in fp32 (Q[l]), tau_l = s_l (1 - h_l^2) planes (G[l])
            ep.N = c.dims[l];
            ep.aux_p = &s->H[l];
            for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l].p[q]; ep.out[q] = s->G[l].p[q]; }
            ep.ld_h = s->H[l].ld; ep.ld_out = s->G[l].ld;
            ep.f32a = s->Q[l].p; ep.ld_f = s->Q[l].ld;
        } else {
            ep.kind = TCB_JVP_LAST;
            ep.N = c.D;
            for (int q = 0; q < 2; ++q) { ep.out[q] = s->QB[L + 1].p[q]; ep.out2[q] = s->PB[L].p[q]; }
            ep.ld_out = s->QB[L + 1].ld; ep.ld_out2 = s->PB[L].ld;
            ep.f32a = s->Y.p; ep.f32b = s->YB.p; ep.ld_f = s->Y.ld;
        }
        tc_launch_gemm<1, true>(h, s, s->G[l - 1], s->wf[l - 1], mc, ep);
    }
    for (int l = L; l >= 2; --l) {   // adjoint of the tangent chain: taubar_{l-1} = sbar_l W_l
        TcEpi ep = base;
        ep.kind = TCB_GBAR_HIDDEN;       // dbar_{l-1} = taubar_{l-1} s_{l-1} (in place of s_{l-1}), sbar_{l-1} = taubar_{l-1} d_{l-1}
        ep.N = c.dims[l - 1];
        ep.aux_p = &s->H[l - 1]; ep.aux_f = &s->Q[l - 1];
        for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l - 1].p[q]; ep.out[q] = s->QB[l].p[q]; }
        ep.ld_h = s->H[l - 1].ld; ep.ld_out = s->QB[l].ld;
        ep.f32a = s->Q[l - 1].p; ep.ld_f = s->Q[l - 1].ld;
        tc_launch_gemm<1, true>(h, s, s->QB[l + 1], s->wb[l - 1], mc, ep);
    }
    for (int l = L; l >= 1; --l) {   // adjoint of the forward chain: hbar_{l-1} = abar_l W_l; both weight-gradient terms of layer l
        tc_launch_wgrad(h, s, s->PB[l], s->H[l - 1], &s->QB[l + 1], &s->G[l - 1], mc, c.dims[l], c.dims[l - 1],
                        grad + h->w_off[l - 1], grad + h->b_off[l - 1]);
        TcEpi ep = base;
        if (l > 1) {
            ep.kind = TCB_HBAR_HIDDEN;
            ep.N = c.dims[l - 1];
            ep.aux_p = &s->H[l - 1]; ep.aux_f = &s->Q[l - 1];
            for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l - 1].p[q]; ep.out[q] = s->PB[l - 1].p[q]; }
            ep.ld_h = s->H[l - 1].ld; ep.ld_out = s->PB[l - 1].ld;
            ep.f32a = s->Q[l - 1].p; ep.ld_f = s->Q[l - 1].ld;
        } else {
            ep.kind = TCB_ZBAR;
            ep.N = c.D;
        }
        tc_launch_gemm<1, true>(h, s, s->PB[l], s->wb[l - 1], mc, ep);
    }
    return CNF_OK;
}

// VJP of TrainMode augmented_f at the stage input src against (abar; c_l, c_E, c_n):
// ubar += zbar, kb[j<stage] += dt a_{stage,j} zbar, grad += parameter gradient.
int tc_vjp_stage(cnf_handle* h, const CnfBufs& bf, const float* src, const float* abar, int stage,
                 const float* dtp, float c_l, float c_E, float c_n, float* ubar, float** kb, float* grad) {
    TcState* s = (TcState*)h->tc_state;
    const cnf_config& c = h->cfg;
    const int L = c.n_layers;
    int rc;
    if ((rc = tc_ensure_acts(h, s, bf.Bs))) return rc;
    if ((rc = tc_ensure_bw(h, s))) return rc;
    for (long long row0 = 0; row0 < bf.Bs; row0 += s->Mc) {
        const int mc = (int)((bf.Bs - row0) < s->Mc ? (bf.Bs - row0) : s->Mc);
        TcEpi base = tc_base_epi(h, s, bf, row0, 0);
        base.kout_ind = nullptr;
        base.abar = abar; base.ubar = h->d_bw[12]; base.dtp = dtp; base.stage = stage;   // ubar field = zbar scratch
        base.c_l = c_l; base.c_E = c_E; base.c_n = c_n;
        for (int j = 0; j < 5; ++j) base.kb[j] = kb[j];
        if (c.trace_mode == CNF_TRACE_JVP) {
            if ((rc = tc_vjp_stage_jvp(h, s, bf, src, base, row0, mc, grad))) return rc;
            continue;
        }
        k_tc_prep<1><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, src, c.D, bf.ys, c.C, row0, mc, s->H[0].p[0],
                                                          s->H[0].p[1], s->H[0].ld, -1, nullptr, nullptr, 0);
        h->launches++;
        tc_forward_hidden<1>(h, s, base, mc);
        {   // y, g_L, ybar
            TcEpi ep = base;
            ep.kind = TCB_FWD_LAST;
            ep.bias = h->d_theta + h->b_off[L - 1];
            ep.N = c.D;
            for (int q = 0; q < 2; ++q) ep.out[q] = s->G[L].p[q];
            ep.ld_out = s->G[L].ld;
            ep.f32a = s->Y.p; ep.f32b = s->YB.p; ep.ld_f = s->Y.ld;
            tc_launch_gemm<1, true>(h, s, s->H[L - 1], s->wf[L - 1], mc, ep);
        }
        for (int l = L; l >= 1; --l) {   // eps chain: q_l = g_l W_l, g_{l-1} = q_l (1-h_{l-1}^2); v = q_1[0:D]
            TcEpi ep = base;
            if (l > 1) {
                ep.kind = TCB_BWD_HIDDEN;
                ep.N = c.dims[l - 1];
                ep.aux_p = &s->H[l - 1];
                for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l - 1].p[q]; ep.out[q] = s->G[l - 1].p[q]; }
                ep.ld_h = s->H[l - 1].ld; ep.ld_out = s->G[l - 1].ld;
                ep.f32a = s->Q[l - 1].p; ep.ld_f = s->Q[l - 1].ld;
            } else {
                ep.kind = TCB_BWD_LAST;
                ep.N = c.D;
                for (int q = 0; q < 2; ++q) ep.out[q] = s->QB[1].p[q];
                ep.ld_out = s->QB[1].ld;
            }
            tc_launch_gemm<1, true>(h, s, s->G[l], s->wb[l - 1], mc, ep);
        }
        for (int l = 1; l <= L; ++l) {   // adjoint of the eps chain runs forward: gbar_l = qbar_l W_l^T
            TcEpi ep = base;
            if (l < L) {
                ep.kind = TCB_GBAR_HIDDEN;
                ep.N = c.dims[l];
                ep.aux_p = &s->H[l]; ep.aux_f = &s->Q[l];
                for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l].p[q]; ep.out[q] = s->QB[l + 1].p[q]; }
                ep.ld_h = s->H[l].ld; ep.ld_out = s->QB[l + 1].ld;
                ep.f32a = s->Q[l].p; ep.ld_f = s->Q[l].ld;
            } else {
                ep.kind = TCB_GBAR_LAST;
                ep.N = c.D;
                for (int q = 0; q < 2; ++q) ep.out[q] = s->PB[L].p[q];
                ep.ld_out = s->PB[L].ld;
                ep.f32a = s->Y.p; ep.f32b = s->YB.p; ep.ld_f = s->Y.ld;
            }
            tc_launch_gemm<1, true>(h, s, s->QB[l], s->wf[l - 1], mc, ep);
        }
        for (int l = L; l >= 1; --l) {   // adjoint of the forward chain: hbar_{l-1} = pbar_l W_l
            tc_launch_wgrad(h, s, s->PB[l], s->H[l - 1], &s->G[l], &s->QB[l], mc, c.dims[l], c.dims[l - 1],
                            grad + h->w_off[l - 1], grad + h->b_off[l - 1]);
            TcEpi ep = base;
            if (l > 1) {
                ep.kind = TCB_HBAR_HIDDEN;
                ep.N = c.dims[l - 1];
                ep.aux_p = &s->H[l - 1]; ep.aux_f = &s->Q[l - 1];
                for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l - 1].p[q]; ep.out[q] = s->PB[l - 1].p[q]; }
                ep.ld_h = s->H[l - 1].ld; ep.ld_out = s->PB[l - 1].ld;
                ep.f32a = s->Q[l - 1].p; ep.ld_f = s->Q[l - 1].ld;
            } else {
                ep.kind = TCB_ZBAR;
                ep.N = c.D;
            }
            tc_launch_gemm<1, true>(h, s, s->PB[l], s->wb[l - 1], mc, ep);
        }
    }
    {
        const long long nz = (long long)c.D * bf.Bs;
        long long g = (nz + 255) / 256;
        if (g > 148LL * 16) g = 148LL * 16;
        k_zbar_apply<<<(unsigned)g, 256, 0, h->stream>>>(h->d_bw[12], dtp, stage, ubar, kb[0], kb[1], kb[2], kb[3], kb[4], nz);
        h->launches++;
    }
    return CNF_OK;
}
