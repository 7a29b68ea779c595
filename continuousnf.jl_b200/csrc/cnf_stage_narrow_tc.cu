// cnf_stage_narrow_tc.cu — RK stage kernel for narrow exact-trace nets (D -> 64 -> 64 -> D, TestMode,
// JacVec; reference src/icnf.jl:148-164 through src/utils.jl:19-36) with the one product that holds
// 94 % of the flops — layer 2 applied to the activation h1 AND to the D tangents tau1_v — on the
// tensor cores.  The register-resident kernel (cnf_stage_narrow.cu) is bound by the FP32-FMA pipe
// (13 k FFMA per sample); here a group of 256 threads owns a tile of 128 samples, two threads per
// sample (each takes one half of the hidden units / accumulator columns):
//   1. SIMT: stage input u + dt*sum a_ij k_j, layer 1 and the D one-hot tangents in registers,
//      written as IEEE-half hi/lo planes straight into shared memory in the canonical K-major
//      SWIZZLE_128B operand layout (no TMA: the operand never exists in global memory);
//   2. the issuer thread launches 2 x 4 x 3 tcgen05.mma (M = 128, K = 16; hi*hi + lo*hi + hi*lo = FP32-level
//      products, as in cnf_stage_tc.cu): h1 against the resident W2 planes (N = 64) and dn = h1^2 - 1 against
//      the resident TANGENT planes (N = 64 D); accumulators in TMEM.  The tangents tau1_v = W1[:, v] .* (1 - h1^2)
//      are never formed: the per-unit factors on both sides of layer 2 are folded into the weights on the host,
//      T_v[o][k] = W3[v][o] W2[o][k] C W1[k][v], so ONE operand vector (dn) serves all D tangents, the D products
//      are one MMA stacked along N, and the epilogue needs one FMA per tangent column;
//   3. SIMT epilogue from TMEM: bias + tanh, the trace dot products (only the diagonal of the D x D Jacobian
//      product is needed), layer 3; the two halves of a sample meet through shared memory; output rows; at the
//      7th stage also the new state and the embedded-error partial sum.
// Two such groups per CTA (one persistent CTA per SM, 16 worker warps) alternate, so the tensor pipe works
// on one tile while the other group's SIMT phases run; a 17th warp issues the MMAs of both groups as
// soon as a tile's 256 writers have arrived on its mbarrier.  The SIMT phases use the packed f32x2
// instructions (FFMA2/FMUL2/FADD2): ~2.4 k instructions per sample instead of 17.7 k.
#include "cnf_internal.h"
#include "cnf_tc_ptx.cuh"
#include <cmath>

#define NTC_H 64            // hidden width (one 128-byte swizzle row of halves)
#define NTC_GROUPS 2
#define NTC_GT 256          // threads per group: 128 sample rows x 2 halves
#define NTC_WORKERS (NTC_GT * NTC_GROUPS)
#ifndef NTC_INWARP
#define NTC_INWARP 0                    // 0: a 17th warp issues the MMAs of both groups (96 registers per thread); 1: warp 0 of a group issues its own (16 warps, 128 registers, no spill, double-buffered loads everywhere) — measured 0.122 ms per stage against 0.116: the issuing warp's detour is on its group's critical path
#endif
#define NTC_THREADS (NTC_WORKERS + (NTC_INWARP ? 0 : 32))
#define NTC_PLANE_A (128 * NTC_H * 2)   // 16 KB: one plane of one operand vector
#define NTC_PLANE_B (NTC_H * NTC_H * 2) // 8 KB: one plane of W2 (or of one tangent matrix T_v)
#define NTC_NVA 2                       // operand vectors written per tile: h1 and dn = h1^2 - 1
#define NTC_C 2.8853900817779268f       // 2 log2(e): tanh(x) = 1 - 2 / (exp2(C x) + 1)

// small weights per PAIR of hidden units (o, o+1) — the SIMT phases run on packed f32x2 instructions
// (FFMA2 / FMUL2 / FADD2: one issue slot for two hidden units) — read with broadcast LDS.128 / LDS.64:
// l1w[p] = C * {W1[o][0], W1[o+1][0], W1[o][1], W1[o+1][1]}, l1b[p] = C * {b1[o], b1[o+1]};
// l3w[p] = {C * b2[o], C * b2[o+1], W3[0][o], W3[0][o+1]}, l3v[p] = {W3[1][o], W3[1][o+1]}
struct alignas(16) NtcW {
    float4 l1w[NTC_H / 2];
    float4 l3w[NTC_H / 2];
    float2 l1b[NTC_H / 2];
    float2 l3v[NTC_H / 2];
    float b3[4];   // b3[0..D); b3[3] = 1 / (C S): undoes the factor C and the power-of-two scale S of the tangent planes
};

struct NtcState {
    NtcW* w_host = nullptr;
    NtcW* d_w = nullptr;
    __half* d_w2 = nullptr;            // W2 hi, W2 lo, T_1..T_D hi, T_1..T_D lo: 8 KB planes, pre-swizzled K-major SWIZZLE_128B
    int D = 0;
    int sms = 148;
};

template <int D>
constexpr int ntc_smem_bytes() {
    return 1024 + (1 + D) * 2 * NTC_PLANE_B + NTC_GROUPS * NTC_NVA * 2 * NTC_PLANE_A + (int)sizeof(NtcW) + NTC_GROUPS * 128 * 16 + 64;
}

// byte offset of the 16-byte chunk `ch` (8 halves along K) of row `r` inside a SWIZZLE_128B K-major tile
__device__ __forceinline__ uint32_t ntc_swz(int r, int ch) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((ch ^ (r & 7)) << 4));
}
__device__ __forceinline__ void ntc_sts128(uint32_t addr, const uint32_t (&v)[4]) {
    asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]) : "memory");
}
__device__ __forceinline__ void ntc_bar_sync(int id) { asm volatile("bar.sync %0, 256;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void ntc_bar_arrive(int id) { asm volatile("bar.arrive %0, 256;" ::"r"(id) : "memory"); }
// wait with back-off: a spinning try_wait loop takes issue slots from the other group's warps on the same scheduler
__device__ __forceinline__ void ntc_wait_sleep(uint64_t* bar, uint32_t parity, int nosleep = 0) {
    for (uint32_t it = 0; !tc::mbar_try_wait(bar, parity); ++it) {
        if (!nosleep) __nanosleep(64);
        if (it > (1u << 24)) tc::trap_with_record(0x601, parity);
    }
}
// packed FP32 pairs (sm_100 f32x2 instructions); ptxas keeps the pairs in aligned registers, no moves
__device__ __forceinline__ float2 ntc_fma2(float2 a, float2 b, float2 c) {
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tmov.b64 rc, {%6,%7};\n\t"
        "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
    return d;
}
__device__ __forceinline__ float2 ntc_mul2(float2 a, float2 b) {
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tmul.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
__device__ __forceinline__ float2 ntc_add2(float2 a, float2 b) {
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n\tmov.b64 ra, {%2,%3};\n\tmov.b64 rb, {%4,%5};\n\tadd.rn.f32x2 rd, ra, rb;\n\tmov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}
// tanh of a pair given C * x: h = 1 - 2 / (exp2(C x) + 1); also dn = h^2 - 1 = -(1 - h^2)
__device__ __forceinline__ void ntc_tanh2(float2 cx, float2& h, float2& dn) {
    float2 t, rr;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(cx.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(cx.y));
    t = ntc_add2(t, make_float2(1.0f, 1.0f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rr.x) : "f"(t.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rr.y) : "f"(t.y));
    h = ntc_fma2(rr, make_float2(-2.0f, -2.0f), make_float2(1.0f, 1.0f));
    dn = ntc_fma2(h, h, make_float2(-1.0f, -1.0f));
}
// a pair -> packed IEEE-half hi and lo words
__device__ __forceinline__ void ntc_split2(float2 x, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(x.x, x.y);
    const float2 l = ntc_fma2(__half22float2(h), make_float2(-1.0f, -1.0f), x);
    const __half2 l2 = __floats2half2_rn(l.x, l.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l2);
}
// tanh of x given cx = C * x
__device__ __forceinline__ float ntc_tanh_scaled(float cx) {
    float t, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(cx));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}

// stage numbering as in k_stage_narrow_test_jvp: 0 k1 = f(u0); -1 initial-dt probe -> K[1];
// 1..5 -> K[stage]; 6 -> new state, k7 and the error partial sum
// the layer-2 products of one tile: h1 x W2 -> columns [0, 64) (first commit), dn x [T_1; ..; T_D] -> columns
// [64, 64 + 64 D) (second commit); 3 MMAs per k-step (hi*hi + lo*hi + hi*lo); issued by ONE thread
template <int D>
__device__ __forceinline__ void ntc_issue_tile(uint32_t sAq, uint32_t sB, uint32_t tq, uint64_t* bar0, uint64_t* bar1) {
    const uint32_t idesc = tc::umma_idesc_f16(128, NTC_H, 0);
    const uint32_t idesc_t = tc::umma_idesc_f16(128, NTC_H * D, 0);   // the D tangent products stacked along N
    const uint64_t dhi = (uint64_t)(tc::umma_desc_sw128(0) >> 32) << 32;
    const uint32_t blo0 = (uint32_t)tc::umma_desc_sw128(sB);
    const uint32_t alo0 = (uint32_t)tc::umma_desc_sw128(sAq);
#pragma unroll
    for (int v = 0; v < NTC_NVA; ++v) {
        const uint32_t bhi0 = v == 0 ? 0u : 2u * NTC_PLANE_B;
        const uint32_t blo0b = v == 0 ? (uint32_t)NTC_PLANE_B : (2u + D) * NTC_PLANE_B;
        const uint32_t id = v == 0 ? idesc : idesc_t;
#pragma unroll
        for (int kk = 0; kk < NTC_H / 16; ++kk) {
            const uint64_t ahi = dhi | (alo0 + (uint32_t)(((v * 2) * NTC_PLANE_A + kk * 32) >> 4));
            const uint64_t alo = dhi | (alo0 + (uint32_t)(((v * 2 + 1) * NTC_PLANE_A + kk * 32) >> 4));
            const uint64_t bhi = dhi | (blo0 + (uint32_t)((bhi0 + kk * 32) >> 4));
            const uint64_t blo = dhi | (blo0 + (uint32_t)((blo0b + kk * 32) >> 4));
            tc::umma_bf16(tq + v * NTC_H, ahi, bhi, id, kk != 0);
            tc::umma_bf16(tq + v * NTC_H, alo, bhi, id, 1);
            tc::umma_bf16(tq + v * NTC_H, ahi, blo, id, 1);
        }
        // two commits: the epilogue's tanh half starts on the layer-2 pre-activations while the (twice as long)
        // tangent product is still running
        tc::umma_commit(v == 0 ? bar0 : bar1);
    }
}

template <int D>
__global__ void __maxnreg__(NTC_INWARP ? 128 : 96)
k_stage_narrow_tc(const NtcW* __restrict__ wsmall, const __half* __restrict__ w2planes, CnfCtrl* __restrict__ c,
                  CnfBufs bf, int stage, int check_done, int ntiles, int dbg) {
    constexpr int R = D + 1;
    constexpr int NV = 1 + D;                       // accumulator vectors: layer-2 pre-activation, D folded tangents
    constexpr int NB_BYTES = (1 + D) * 2 * NTC_PLANE_B;
    extern __shared__ uint8_t ntc_raw[];
    const uint32_t sbase = (tc::smem_u32(ntc_raw) + 1023u) & ~1023u;
    uint8_t* sgen = ntc_raw + (sbase - tc::smem_u32(ntc_raw));
    const uint32_t sB = sbase;                                       // W2 hi, lo; T hi (D planes), T lo (D planes)
    const uint32_t sA = sbase + NB_BYTES;                            // [group][vector h1 / dn][plane]
    uint8_t* after = sgen + NB_BYTES + NTC_GROUPS * NTC_NVA * 2 * NTC_PLANE_A;
    const NtcW* sw = reinterpret_cast<const NtcW*>(after);
    float4* xch = reinterpret_cast<float4*>(after + sizeof(NtcW));   // [group][row]: partial sums of the upper half
    uint64_t* bars = reinterpret_cast<uint64_t*>(after + sizeof(NtcW) + NTC_GROUPS * 128 * 16);
    uint64_t* full = bars + NTC_GROUPS;   // operand planes of a tile written (256 arrivals)
    uint64_t* bars0 = bars + 2 * NTC_GROUPS;   // the h1 x W2 product of a tile is complete (bars: the tangent product too)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3 * NTC_GROUPS);
    __shared__ double sh_red[32];

    const int tid = threadIdx.x, warp = tid >> 5;
    const int g = tid >> 8;              // group
    const int gt = tid & (NTC_GT - 1);
    const int hw = gt >> 7;              // which half of the hidden units / accumulator columns
    const int r = gt & 127;              // row of the tile = TMEM lane
    // ---- one-time setup: W2 planes and the small weights into shared memory, barriers, TMEM
    {
        const uint4* src = reinterpret_cast<const uint4*>(w2planes);
        uint4* dst = reinterpret_cast<uint4*>(sgen);
        for (int i = tid; i < NB_BYTES / 16; i += NTC_THREADS) dst[i] = src[i];
        const uint4* src2 = reinterpret_cast<const uint4*>(wsmall);
        uint4* dst2 = reinterpret_cast<uint4*>(after);
        for (int i = tid; i < (int)sizeof(NtcW) / 16; i += NTC_THREADS) dst2[i] = src2[i];
    }
    if (tid == 0) {
        for (int i = 0; i < NTC_GROUPS; ++i) { tc::mbar_init(&bars[i], 1); tc::mbar_init(&bars0[i], 1); tc::mbar_init(&full[i], NTC_GT); }
        tc::fence_barrier_init();
    }
    if (warp == 0) tc::tmem_alloc(tmem_slot, 512);
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = *tmem_slot + (uint32_t)g * 256u;          // this group's accumulators: NV x 64 columns
    const uint32_t tlane = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)hw * 32u;
    const uint32_t myA = sA + (uint32_t)g * (NTC_NVA * 2 * NTC_PLANE_A);
    const int barB = 1 + g;

    // Programmatic dependent launch: everything above touches only this launch's own resources (weights, barriers,
    // tensor memory), so it may run while the previous kernel of the stream is still finishing on other SMs; the
    // controller block and the stage rows are read only after the wait.  (No-ops when launched without the attribute.)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const bool skip = check_done && c->done;

    double acc = 0.0;
    const int tile0 = blockIdx.x * NTC_GROUPS + g, tstride = gridDim.x * NTC_GROUPS;
    if (skip) {
        // the solve has finished: a speculative launch
    } else if (!NTC_INWARP && warp == NTC_WORKERS / 32) {
        // ---- MMA issuer: layer 2 for all operand vectors of a tile as soon as its planes are written;
        // 3 MMAs per k-step (hi*hi + lo*hi + hi*lo), one commit per tile
        if ((tid & 31) == 0) {
            int left[NTC_GROUPS];
            uint32_t ph[NTC_GROUPS];
#pragma unroll
            for (int q = 0; q < NTC_GROUPS; ++q) {
                const int t0 = blockIdx.x * NTC_GROUPS + q;
                left[q] = t0 < ntiles ? (ntiles - 1 - t0) / tstride + 1 : 0;
                ph[q] = 0;
            }
            uint32_t spins = 0;
            while (left[0] > 0 || left[1] > 0) {
#pragma unroll
                for (int q = 0; q < NTC_GROUPS; ++q) {
                    // test_wait, not try_wait: the issuer watches both groups and must not be put to sleep on one
                    // group's barrier while the other group's tile becomes ready (option tc_dbg & 256: the old try_wait)
                    if (left[q] > 0 && ((dbg & 256) ? tc::mbar_try_wait(&full[q], ph[q]) : tc::mbar_test_wait(&full[q], ph[q]))) {
                        tc::tc_fence_after();
                        ntc_issue_tile<D>(sA + (uint32_t)q * (NTC_NVA * 2 * NTC_PLANE_A), sB, *tmem_slot + (uint32_t)q * 256u, &bars0[q], &bars[q]);
                        ph[q] ^= 1;
                        left[q]--;
                        spins = 0;
                    }
                }
                __nanosleep(32);   // a busy poll would take issue slots from the four worker warps of this scheduler
                if (++spins > (1u << 24)) tc::trap_with_record(0x602, 0);
            }
        }
    } else {
    const long long Bs = bf.Bs;
    const float* u = bf.slots[cnf_slot(c, stage <= 0 ? 0 : c->naccept)];
    const float* k1 = bf.K[c->k1idx];
    float* kout = (stage == 0) ? bf.K[c->k1idx] : (stage == -1 ? bf.K[1] : (stage == 6 ? bf.K[6 - c->k1idx] : bf.K[stage]));
    float* unew = bf.slots[cnf_slot(c, c->naccept + 1)];
    const float dt = c->dt;
    const float hh = c->tdir * c->dt0init;
    const float reltol = c->reltol, abstol = c->abstol;
    // rows D.. only at the 7th stage (pure quadrature rows), and only in the thread that finishes the sample
    const int nrow = (stage == 6 && hw == 0) ? R : D;

    // raw stage-input operands of one tile: issued as independent loads one tile ahead (an in-order
    // warp that loads at the point of use pays one DRAM latency per k_j)
    float ru[R], rk[R][6];
    auto load_raw = [&](int tile) {
        const long long b = (long long)tile * 128 + r;
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            if (rr < nrow) {
                const long long at = (long long)rr * Bs + b;
                ru[rr] = u[at];
                if (stage != 0) rk[rr][0] = k1[at];
#pragma unroll
                for (int j = 1; j < 6; ++j)
                    if (j < stage) rk[rr][j] = bf.K[j][at];
            }
        }
    };

    if (tile0 < ntiles) load_raw(tile0);
    uint32_t it = 0;
    for (int tile = tile0; tile < ntiles; tile += tstride, ++it) {
        const long long b = (long long)tile * 128 + r;
        // ---- stage input (and, at the 7th stage, the new state and the k1..k6 part of the error estimate)
        float z[D], ep[R] = {}, uo[R] = {}, un[R] = {};
#pragma unroll
        for (int rr = 0; rr < R; ++rr) {
            if (rr < nrow) {
                float v;
                if (stage == 0) v = ru[rr];
                else if (stage == -1) v = fmaf(hh, rk[rr][0], ru[rr]);
                else {
                    float a = c_A[stage][0] * rk[rr][0];
#pragma unroll
                    for (int j = 1; j < 6; ++j)
                        if (j < stage) a = fmaf(c_A[stage][j], rk[rr][j], a);
                    v = fmaf(dt, a, ru[rr]);
                    if (stage == 6 && hw == 0) {
                        unew[(long long)rr * Bs + b] = v;
                        float e = c_BT[0] * rk[rr][0];
#pragma unroll
                        for (int j = 1; j < 6; ++j) e = fmaf(c_BT[j], rk[rr][j], e);
                        ep[rr] = e; uo[rr] = ru[rr]; un[rr] = v;
                    }
                }
                if (rr < D) z[rr < D ? rr : 0] = v;
            }
        }
        // ---- layer 1 of this thread's 32 hidden units, two units per instruction: h1 and dn = h1^2 - 1 (the operand of
        // every tangent product, see the header) -> hi/lo operand planes in shared memory
        const float2 z0 = make_float2(z[0], z[0]), z1 = make_float2(z[D > 1 ? 1 : 0], z[D > 1 ? 1 : 0]);
#pragma unroll
        for (int cq = 0; cq < 4; ++cq) {
            const int cb = hw * 4 + cq;
            float2 x[NTC_NVA][4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 w = sw->l1w[cb * 4 + q];
                float2 a = ntc_fma2(make_float2(w.x, w.y), z0, sw->l1b[cb * 4 + q]);
                if (D > 1) a = ntc_fma2(make_float2(w.z, w.w), z1, a);
                ntc_tanh2(a, x[0][q], x[1][q]);
            }
            const uint32_t off = ntc_swz(r, cb);
#pragma unroll
            for (int v = 0; v < NTC_NVA; ++v) {
                uint32_t hi[4], lo[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) ntc_split2(x[v][q], hi[q], lo[q]);
                ntc_sts128(myA + (uint32_t)(v * 2) * NTC_PLANE_A + off, hi);
                ntc_sts128(myA + (uint32_t)(v * 2 + 1) * NTC_PLANE_A + off, lo);
            }
        }
        tc::fence_proxy_async();     // generic-proxy writes -> visible to the tensor core's async proxy
        tc::tc_fence_before();       // the previous tile's tcgen05.ld are ordered before the arrival
        tc::mbar_arrive(&full[g]);   // the issuer warp starts this tile's MMAs after all 256 arrivals
        // ---- next tile's operands: in flight during the MMA wait and the epilogue
        if (tile + tstride < ntiles) load_raw(tile + tstride);
        if (NTC_INWARP && (gt >> 5) == 0) {   // warp 0 of the group issues its MMAs once all 256 writers have arrived
            ntc_wait_sleep(&full[g], it & 1, 1);
            tc::tc_fence_after();
            if ((tid & 31) == 0) ntc_issue_tile<D>(myA, sB, tmem, &bars0[g], &bars[g]);
            __syncwarp();
        }
        // ---- epilogue on this thread's 32 columns: tanh, layer 3, and the trace dot products: accumulator column o
        // of tangent v already holds W3[v][o] (W2 diag(C W1[:, v]) dn1)[o], so the diagonal of the Jacobian product is
        // its dot product with dn2 (both dn carry a minus sign: their product does not; the factor C leaves at the end)
        float2 p3a = make_float2(0.0f, 0.0f), p3b = p3a, r3a = p3a, r3b = p3a;   // even / odd hidden units
        float2 dn2[16];              // dn of this thread's 32 layer-2 units, kept for the tangent half
        ntc_wait_sleep(&bars0[g], it & 1, dbg & 32);
        __syncwarp();
        tc::tc_fence_after();
        {
            uint32_t va[2][8];       // chunk ch + 1 is in flight while chunk ch is consumed
            tc::tmem_ld_32x8(tlane, va[0]);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                const int cur = ch & 1;
                tc::tmem_ld_wait();
                if (ch + 1 < 4) tc::tmem_ld_32x8(tlane + (ch + 1) * 8, va[cur ^ 1]);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 w = sw->l3w[hw * 16 + ch * 4 + q];
                    float2 h;
                    ntc_tanh2(ntc_fma2(make_float2(__uint_as_float(va[cur][2 * q]), __uint_as_float(va[cur][2 * q + 1])),
                                       make_float2(NTC_C, NTC_C), make_float2(w.x, w.y)), h, dn2[ch * 4 + q]);
                    p3a = ntc_fma2(make_float2(w.z, w.w), h, p3a);
                    if (D > 1) p3b = ntc_fma2(sw->l3v[hw * 16 + ch * 4 + q], h, p3b);
                }
            }
        }
        ntc_wait_sleep(&bars[g], it & 1, dbg & 32);
        __syncwarp();
        tc::tc_fence_after();
        {
#if NTC_INWARP
            uint32_t va[2][D][8];    // double-buffered like the tanh half (the registers are there with 16 warps)
#pragma unroll
            for (int v = 0; v < D; ++v) tc::tmem_ld_32x8(tlane + (1 + v) * NTC_H, va[0][v]);
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
                const int cur = ch & 1;
                tc::tmem_ld_wait();
                if (ch + 1 < 4) {
#pragma unroll
                    for (int v = 0; v < D; ++v) tc::tmem_ld_32x8(tlane + (1 + v) * NTC_H + (ch + 1) * 8, va[cur ^ 1][v]);
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    r3a = ntc_fma2(make_float2(__uint_as_float(va[cur][0][2 * q]), __uint_as_float(va[cur][0][2 * q + 1])), dn2[ch * 4 + q], r3a);
                    if (D > 1)
                        r3b = ntc_fma2(make_float2(__uint_as_float(va[cur][D > 1 ? 1 : 0][2 * q]),
                                                   __uint_as_float(va[cur][D > 1 ? 1 : 0][2 * q + 1])), dn2[ch * 4 + q], r3b);
                }
            }
#else
            // (single-buffered: by now the tangent product has long finished, and 16 more registers would spill at 96)
            uint32_t va[D][8];
#pragma unroll
            for (int ch = 0; ch < 4; ++ch) {
#pragma unroll
                for (int v = 0; v < D; ++v) tc::tmem_ld_32x8(tlane + (1 + v) * NTC_H + ch * 8, va[v]);
                tc::tmem_ld_wait();
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    r3a = ntc_fma2(make_float2(__uint_as_float(va[0][2 * q]), __uint_as_float(va[0][2 * q + 1])), dn2[ch * 4 + q], r3a);
                    if (D > 1)
                        r3b = ntc_fma2(make_float2(__uint_as_float(va[D > 1 ? 1 : 0][2 * q]),
                                                   __uint_as_float(va[D > 1 ? 1 : 0][2 * q + 1])), dn2[ch * 4 + q], r3b);
                }
            }
#endif
        }
        float p3[D], r3[D];
        p3[0] = p3a.x + p3a.y; r3[0] = r3a.x + r3a.y;
        if (D > 1) { p3[D > 1 ? 1 : 0] = p3b.x + p3b.y; r3[D > 1 ? 1 : 0] = r3b.x + r3b.y; }
        if (hw == 1) {
            xch[g * 128 + r] = make_float4(p3[0], r3[0], D > 1 ? p3[D > 1 ? 1 : 0] : 0.0f, D > 1 ? r3[D > 1 ? 1 : 0] : 0.0f);
            ntc_bar_arrive(barB);
            continue;
        }
        ntc_bar_sync(barB);
        {
            const float4 o = xch[g * 128 + r];
            p3[0] += o.x; r3[0] += o.y;
            if (D > 1) { p3[D > 1 ? 1 : 0] += o.z; r3[D > 1 ? 1 : 0] += o.w; }
        }
        float y[D], tr = 0.0f;
#pragma unroll
        for (int j = 0; j < D; ++j) {
            y[j] = cnf_tanh_fastpath(p3[j] + sw->b3[j]);
            tr -= fmaf(-y[j], y[j], 1.0f) * (r3[j] * sw->b3[3]);
        }
#pragma unroll
        for (int j = 0; j < D; ++j) kout[(long long)j * Bs + b] = y[j];
        kout[(long long)D * Bs + b] = tr;
        // ---- 7th stage: embedded error partial sum over this thread's R elements
        if (stage == 6 && b < bf.B) {
#pragma unroll
            for (int rr = 0; rr < R; ++rr) {
                const float a = fmaf(c_BT[6], rr < D ? y[rr < D ? rr : 0] : tr, ep[rr]);
                const float ut = dt * a;
                const float sc = fmaf(fmaxf(fabsf(uo[rr]), fabsf(un[rr])), reltol, abstol);
                float q = ut / sc;
                if (un[rr] != un[rr]) q = un[rr];
                acc += (double)q * (double)q;
            }
        }
    }
    }   // workers
    tc::tc_fence_before();
    if (stage == 6 && !skip) {
        acc = cnf_block_sum(acc, sh_red);
        if (tid == 0) atomicAdd(&c->errsum, acc);
    }
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(*tmem_slot, 512);
}

// ------------------------------------------------------------------ host side
static void ntc_pack(const cnf_handle* h, const float* theta, NtcW* out, __half* planes) {
    const int D = h->cfg.D;
    memset(out, 0, sizeof(*out));
    const float* W1 = theta + h->w_off[0];   // col-major H x D
    const float* W2 = theta + h->w_off[1];   // col-major H x H: W2[i * H + o]
    const float* W3 = theta + h->w_off[2];   // col-major D x H
    for (int p = 0; p < NTC_H / 2; ++p) {
        const int o = 2 * p;
        const float* b1 = theta + h->b_off[0];
        const float* b2 = theta + h->b_off[1];
        out->l1w[p] = make_float4(NTC_C * W1[o], NTC_C * W1[o + 1], D > 1 ? NTC_C * W1[NTC_H + o] : 0.0f,
                                  D > 1 ? NTC_C * W1[NTC_H + o + 1] : 0.0f);
        out->l1b[p] = make_float2(NTC_C * b1[o], NTC_C * b1[o + 1]);
        out->l3w[p] = make_float4(NTC_C * b2[o], NTC_C * b2[o + 1], W3[o * D], W3[(o + 1) * D]);
        out->l3v[p] = D > 1 ? make_float2(W3[o * D + 1], W3[(o + 1) * D + 1]) : make_float2(0.0f, 0.0f);
    }
    for (int j = 0; j < D; ++j) out->b3[j] = theta[h->b_off[2] + j];
    // B operands: row n = output unit, K = input unit; K-major SWIZZLE_128B (same layout as ntc_swz).  Plane order:
    // W2 hi, W2 lo, T_1..T_D hi, T_1..T_D lo with T_v[n][k] = W3[v][n] W2[n][k] C W1[k][v] (the D hi planes are
    // contiguous, so they are ONE operand of 64 D rows: SBO = 1024 bytes per 8 rows runs across the plane boundary)
    // The products are small (three weights): scaled by a power of two S so that the largest is in [4096, 8192) and the
    // lo words stay clear of the half subnormals (absolute resolution 6e-8 would cost 3-4 bits of T otherwise)
    const size_t PH = NTC_PLANE_B / 2;   // halves per plane
    double tmax = 0.0;
    for (int n = 0; n < NTC_H; ++n)
        for (int k = 0; k < NTC_H; ++k)
            for (int v = 0; v < D; ++v)
                tmax = fmax(tmax, fabs((double)W3[n * D + v] * (double)W2[k * NTC_H + n] * ((double)NTC_C * (double)W1[v * NTC_H + k])));
    const double S = (tmax > 0.0 && std::isfinite(tmax)) ? exp2(floor(log2(8192.0 / tmax))) : 1.0;
    out->b3[3] = (float)(1.0 / ((double)NTC_C * S));
    for (int n = 0; n < NTC_H; ++n)
        for (int k = 0; k < NTC_H; ++k) {
            const size_t off = (size_t)(n >> 3) * 512 + (n & 7) * 64 + (((k >> 3) ^ (n & 7)) << 3) + (k & 7);   // in halves
            const float w = W2[k * NTC_H + n];
            const __half hi = __float2half_rn(w);
            planes[off] = hi;
            planes[PH + off] = __float2half_rn(w - __half2float(hi));
            for (int v = 0; v < D; ++v) {
                const float t = (float)(S * (double)W3[n * D + v] * (double)w * ((double)NTC_C * (double)W1[v * NTC_H + k]));
                const __half thi = __float2half_rn(t);
                planes[(2 + v) * PH + off] = thi;
                planes[(2 + D + v) * PH + off] = __float2half_rn(t - __half2float(thi));
            }
        }
}

static bool ntc_shape(const cnf_config& c) {
    return c.n_layers == 3 && c.C == 0 && !(c.flags & CNF_FLAG_LAST_LINEAR) && (c.D == 1 || c.D == 2) &&
           c.dims[1] == NTC_H && c.dims[2] == NTC_H;
}
bool ntc_supported(const cnf_handle* h, int mode) {
    return mode == CNF_MODE_TEST && ntc_shape(h->cfg) && h->ntc_state != nullptr;   // both trace modes: see narrow_supported
}

// called from set_params with a host copy of theta
int ntc_set_params(cnf_handle* h, const float* theta_host) {
    if (!ntc_shape(h->cfg)) return CNF_OK;
    NtcState* s = (NtcState*)h->ntc_state;
    if (!s) {
        s = new NtcState();
        s->D = h->cfg.D;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, h->device) == cudaSuccess) s->sms = prop.multiProcessorCount;
        s->w_host = new NtcW();
        h->ntc_state = s;
        CNF_CUDA(cudaMalloc(&s->d_w, sizeof(NtcW)));
        CNF_CUDA(cudaMalloc(&s->d_w2, (1 + 2) * 2 * NTC_PLANE_B));   // sized for D = 2
        CNF_CUDA(cudaFuncSetAttribute(k_stage_narrow_tc<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntc_smem_bytes<1>()));
        CNF_CUDA(cudaFuncSetAttribute(k_stage_narrow_tc<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, ntc_smem_bytes<2>()));
    }
    const size_t plane_bytes = (size_t)(1 + s->D) * 2 * NTC_PLANE_B;
    std::vector<__half> planes(plane_bytes / 2);
    ntc_pack(h, theta_host, s->w_host, planes.data());
    CNF_CUDA(cudaMemcpyAsync(s->d_w, s->w_host, sizeof(NtcW), cudaMemcpyHostToDevice, h->stream));
    CNF_CUDA(cudaMemcpyAsync(s->d_w2, planes.data(), plane_bytes, cudaMemcpyHostToDevice, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    return CNF_OK;
}
void ntc_free(cnf_handle* h) {
    NtcState* s = (NtcState*)h->ntc_state;
    if (!s) return;
    delete s->w_host;
    cudaFree(s->d_w);
    cudaFree(s->d_w2);
    delete s;
    h->ntc_state = nullptr;
}
int launch_stage_narrow_tc(cnf_handle* h, const CnfBufs& bf, int stage, int check_done) {
    NtcState* s = (NtcState*)h->ntc_state;
    if (!s) { h->err = "narrow tensor path: no parameters"; return CNF_ERR_UNSUPPORTED; }
    const int ntiles = (int)(bf.Bs / 128);
    int grid = (ntiles + NTC_GROUPS - 1) / NTC_GROUPS;
    if (grid > s->sms) grid = s->sms;
    // launched with programmatic stream serialization: the prologue of this launch (weights into shared memory,
    // barriers, tensor-memory allocation) overlaps the tail of the previous stage kernel (option tc_dbg & 64: off)
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof(lc));
    lc.gridDim = dim3((unsigned)grid);
    lc.blockDim = dim3(NTC_THREADS);
    lc.dynamicSmemBytes = (size_t)(s->D == 1 ? ntc_smem_bytes<1>() : ntc_smem_bytes<2>());
    lc.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at;
    lc.numAttrs = (h->tc_dbg & 64) ? 0 : 1;
    const NtcW* dw = s->d_w;
    const __half* dw2 = s->d_w2;
    if (s->D == 1)
        CNF_CUDA(cudaLaunchKernelEx(&lc, k_stage_narrow_tc<1>, dw, dw2, h->d_ctrl, bf, stage, check_done, ntiles, h->tc_dbg));
    else
        CNF_CUDA(cudaLaunchKernelEx(&lc, k_stage_narrow_tc<2>, dw, dw2, h->d_ctrl, bf, stage, check_done, ntiles, h->tc_dbg));
    h->launches++;
    return CNF_OK;
}
