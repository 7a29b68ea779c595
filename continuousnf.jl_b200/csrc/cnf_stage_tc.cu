// cnf_stage_tc.cu — tensor-core path for wide nets (hidden >= 128, TrainMode Hutchinson).
// Every layer product of the Dense/tanh chain, of its VJP / JVP chain and (cnf_backward_tc.cu) of
// the second-order reverse sweep is a TMA-fed tcgen05.mma GEMM with the batch as the M dimension
// (128 samples per tile, one TMEM lane per sample), FP32 accumulators in TMEM, and the layer's
// elementwise work fused into the TMEM->register epilogue (bias + tanh, the (1-h^2) Hadamard
// factor, the eps dot product and the un-squared norms of reference src/icnf.jl:335-348).
// Precision: every operand is split into two 16-bit planes x = hi + lo and a product costs three
// MMAs (hi*hi + lo*hi + hi*lo, FP32 accumulate).
//   PREC 0, forward solve: IEEE half planes (FP16x3).  hi + lo carries 22 significand bits for the
//     O(1) operands of this path (tanh outputs, Glorot-scale weights, the eps chain), so f has
//     FP32-level noise (rms 2e-7 vs 1.7e-7 for an FP32 SIMT evaluation; scripts/precision_fp16.py).
//     That level is required: BF16x3 (16 bits, 1e-5 noise in f) meets the 1e-4 value tolerance but
//     the noise is amplified by the abstol = 1e-6 rows of the error norm and changes the accepted-
//     step count by 5-7; plain TF32 misses the value tolerance as well.  (The first tensor path
//     used BF16x6 = three bf16 planes and six MMAs for the same accuracy: 1.5x the plane traffic,
//     2x the tensor work.)  Inputs are assumed to lie inside the half range (|h_0| < 6e4).
//   PREC 1, reverse sweep: bfloat16 planes (BF16x3, ~1e-5): cotangents scale with dt/B and need
//     the FP32 exponent range; the gradient tolerance is 1e-3.
// Warp roles: warp 0 = TMA producer, warp 1 = MMA issuer (one elected thread), warps 2..13 =
// epilogue (three warps per TMEM lane quadrant, every third 16-column chunk each); two TMEM
// accumulator buffers so the epilogue of tile i overlaps the MMAs of tile i+1.
#include "cnf_internal.h"
#include "cnf_tc_ptx.cuh"
#include <cuda.h>

#define TC_BM 128
#define TC_BK 64            // bf16 elements per 128-byte swizzled row
#define TC_THREADS 480      // 15 warps: TMA, MMA, 12 epilogue, epilogue-operand producer
#define TC_THREADS_WIDE 736 // TC_FWD_HIDDEN: 20 epilogue warps (its tanh + split epilogue is issue-latency-bound, 76 registers)
#define TC_EPI_WARP0 2

enum {
    TC_FWD_HIDDEN = 0, TC_FWD_LAST = 1, TC_BWD_HIDDEN = 2, TC_BWD_LAST = 3, TC_JVP_LAST = 4,
    // reverse-sweep kinds (VecJac; maths in DESIGN.md "Backward")
    TCB_FWD_LAST = 8,      // y, g_L = eps (1-y^2), ybar = abar + beta y/|y|
    TCB_BWD_HIDDEN = 9,    // q_{l+1} kept (fp32), g_l = q_{l+1} (1-h_l^2)
    TCB_BWD_LAST = 10,     // v -> vbar = -alpha eps + gamma v/|v|  (planes of qbar_1)
    TCB_GBAR_HIDDEN = 11,  // gbar_l: dbar_l = gbar_l q_{l+1} (fp32, in place), qbar_{l+1} = gbar_l (1-h_l^2)
    TCB_GBAR_LAST = 12,    // gbar_L: dbar_L = gbar_L eps, pbar_L = (ybar - 2 y dbar_L)(1-y^2)
    TCB_HBAR_HIDDEN = 13,  // hbar_l: pbar_l = (hbar_l - 2 h_l dbar_l)(1-h_l^2)
    TCB_ZBAR = 14,         // zbar: ubar += zbar, kbar_j += dt a_ij zbar
    // JacVec reverse sweep (forward-mode tangent chain tau_l = (tau_{l-1} W_l^T) .* d_l; maths in cnf_backward_tc.cu)
    TCB_JVP_LAST = 15      // s_L -> tau_L = s_L d_L, taubar_L = -alpha eps + gamma tau_L/|tau_L|; sbar_L = taubar_L d_L (planes),
                           // dbar_L = taubar_L s_L, abar_L = (ybar - 2 y dbar_L) d_L (second plane pair)
};

struct TcEpi {
    int kind;
    int N;                         // valid output columns
    const float* bias;             // [N]                       (FWD_*)
    __nv_bfloat16* out[2];         // [rows x ld_out] planes    (hidden kinds; FWD_LAST: g_L planes)
    int ld_out;
    __nv_bfloat16* out2[2];        // second plane pair (TCB_JVP_LAST: abar_L)
    int ld_out2;
    const __nv_bfloat16* hp[2];    // activations for d = 1-h^2 (BWD_HIDDEN)
    int ld_h;
    const float* eps;              // D x Bs (SoA)
    float* kout;                   // R x Bs (SoA); if kout_ind != nullptr the pointer is read from there
    float* const* kout_ind;        // device-resolved output buffer (k7 / FSAL index lives on the device)
    const CnfCtrl* ctrl;           // early exit when the solve is done (nullptr: never)
    long long Bs;
    long long row0;                // first sample of this chunk
    int D;
    unsigned flags;
    int vjp;                       // FWD_LAST: also emit g_L = eps .* (1-y^2) planes; 2: z rows only
    int test;                      // TestMode (exact trace): R = D + 1 rows, no regulariser rows; JVP_LAST adds pass i's -(J e_i)_i
    int pass;                      // index i of the one-hot tangent
    // ---- reverse sweep
    float* f32a;                   // fp32 side array [rows x ld_f] (q_{l+1} / dbar_l ; y)
    float* f32b;                   // second fp32 side array (ybar)
    int ld_f;
    const float* abar;             // D x Bs cotangent of this stage's zdot rows
    float* ubar;                   // D x Bs
    float* kb[5];                  // cotangents of the earlier stages
    const float* dtp;              // accepted dt of this step (device)
    int stage;                     // 0..5
    float c_l, c_E, c_n;           // cotangents of the quadrature rows (1/B, lambda1/B, lambda2/B)
    int B;                         // live samples (rows >= B are padding)
    int dbg;                       // debug switches of the ring kernels
    // ---- host only: whose tiles the TMA producer prefetches into L2 for the epilogue (TcPlane*, TcF32*)
    const void* aux_p;
    const void* aux_f;
};

// two floats -> packed (hi, hi) and (lo, lo) 16-bit pairs; element a in the low half-word
template <int PREC>
__device__ __forceinline__ void tc_split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    if (PREC == 0) {
        __half2 h = __floats2half2_rn(a, b);
        float2 f = __half22float2(h);
        __half2 l = __floats2half2_rn(a - f.x, b - f.y);
        hi = *reinterpret_cast<uint32_t*>(&h);
        lo = *reinterpret_cast<uint32_t*>(&l);
    } else {
        __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
        float2 f = __bfloat1622float2(h);
        __nv_bfloat162 l = __floats2bfloat162_rn(a - f.x, b - f.y);
        hi = *reinterpret_cast<uint32_t*>(&h);
        lo = *reinterpret_cast<uint32_t*>(&l);
    }
}
template <int PREC>
__device__ __forceinline__ float2 tc_unpack2(uint32_t w) {
    if (PREC == 0) return __half22float2(*reinterpret_cast<__half2*>(&w));
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}

struct TcMaps {
    CUtensorMap a[2], b[2];   // A planes, B planes
    CUtensorMap x[3];         // RING kernels: epilogue operands, hi / lo plane and fp32 array, boxes of 128 rows x 16 columns
};

struct TcSmemCtl {
    uint64_t full[8], empty[8], tfull[2], tempty[2];
    uint64_t xfull[12], xempty[12];   // rings of epilogue-operand chunks (RING kernels): [sub][depth]
    uint32_t tmem_base;
    float2 red[2][4][3][32];   // [tile parity][lane quadrant][warp of the quadrant][lane]: row reductions
};

// ---- epilogue helpers: one thread = one sample row, 16 accumulator columns per step.
// 12 epilogue warps: warp w may touch TMEM lane quadrant w & 3; the three warps of a quadrant take
// every third 16-column chunk, so each scheduler has three epilogue warps to hide latencies with.
// Whatever a chunk needs from global memory besides the accumulator (the h planes for 1-h^2, the
// fp32 side arrays, eps rows, bias) is loaded one chunk ahead into registers (TcAux) — for the
// first chunk of a tile before the wait on the MMA — with 256-bit loads; an in-order warp that
// loads at the point of use pays the DRAM latency once per chunk (ncu: 12 % of all stall samples
// on the first use of the h planes) or, behind possibly-aliasing stores, once per column.
#define TC_CW 16
#define TC_NSUB 3

struct TcAux {
    uint32_t p[2][8];   // 16 columns of the hi / lo plane
    float f[TC_CW];     // fp32 side array, eps rows or bias
};

__device__ __forceinline__ void tc_ld256_nc(const void* p, uint32_t (&r)[8]) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(p));
}
__device__ __forceinline__ void tc_ld256(const float* p, float* r) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "l"(p)
                 : "memory");
}
__device__ __forceinline__ void tc_ld256_ncf(const float* p, float* r) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]), "=f"(r[7])
                 : "l"(p));
}
__device__ __forceinline__ void tc_st256(void* p, const uint32_t (&r)[8]) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(r[0]), "r"(r[1]), "r"(r[2]),
                 "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}
__device__ __forceinline__ void tc_st256f(float* p, const float* r) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r[0]), "f"(r[1]), "f"(r[2]),
                 "f"(r[3]), "f"(r[4]), "f"(r[5]), "f"(r[6]), "f"(r[7])
                 : "memory");
}

// tanh(x) = 1 - 2 / (exp(2x) + 1) on the SFU (ex2.approx + rcp.approx): absolute error <= 4e-7,
// exact saturation; five instructions instead of the sixteen of the NNlib rational
__device__ __forceinline__ float tc_tanh(float x) {
    float t, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(x * 2.8853900817779268f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(t + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}

// 16 values -> one 32-byte store per plane
template <int PREC>
__device__ __forceinline__ void tc_store_planes16(const float (&x)[TC_CW], __nv_bfloat16* const (&dst)[2], long long off) {
    uint32_t pk[2][8];
#pragma unroll
    for (int j = 0; j < 8; ++j) tc_split2<PREC>(x[2 * j], x[2 * j + 1], pk[0][j], pk[1][j]);
    tc_st256(dst[0] + off, pk[0]);
    tc_st256(dst[1] + off, pk[1]);
}
template <int PREC>
__device__ __forceinline__ void tc_unpack_planes16(const uint32_t (&p)[2][8], float (&h)[TC_CW]) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float2 hi = tc_unpack2<PREC>(p[0][k]), lo = tc_unpack2<PREC>(p[1][k]);
        h[2 * k] = hi.x + lo.x;
        h[2 * k + 1] = hi.y + lo.y;
    }
}
__device__ __forceinline__ void tc_load_f32x16(const float* src, float (&x)[TC_CW]) {
    tc_ld256(src, x);
    tc_ld256(src + 8, x + 8);
}
__device__ __forceinline__ void tc_store_f32x16(float* dst, const float (&x)[TC_CW]) {
    tc_st256f(dst, x);
    tc_st256f(dst + 8, x + 8);
}

// 16 rows n0..n0+15 of a read-only SoA buffer (row stride Bs) for this thread's sample; rows >= nmax read as 0
__device__ __forceinline__ void tc_load_soa16(const float* __restrict__ base, long long Bs, long long grow, int n0,
                                              int nmax, float (&x)[TC_CW]) {
#pragma unroll
    for (int j = 0; j < TC_CW; ++j) x[j] = (n0 + j < nmax) ? __ldg(base + (long long)(n0 + j) * Bs + grow) : 0.0f;
}

// what chunk (nb, ch) of this thread's row needs besides the accumulator, by kind
template <int PREC, bool BWD, int KIND>
__device__ __forceinline__ void tc_aux_load(const TcEpi& ep, long long rloc, long long grow, int ng, TcAux& a) {
    constexpr int k = KIND;
    if (!BWD) {
        if (k == TC_FWD_HIDDEN) {
            const float* b = ep.bias + ng;
            if ((reinterpret_cast<uintptr_t>(b) & 31) == 0) { tc_ld256_ncf(b, a.f); tc_ld256_ncf(b + 8, a.f + 8); }
            else {
#pragma unroll
                for (int j = 0; j < TC_CW; ++j) a.f[j] = __ldg(b + j);
            }
        } else if (k == TC_BWD_HIDDEN) {
            tc_ld256_nc(ep.hp[0] + rloc * ep.ld_h + ng, a.p[0]);
            tc_ld256_nc(ep.hp[1] + rloc * ep.ld_h + ng, a.p[1]);
        } else if ((k != TC_FWD_LAST || ep.vjp == 1) && !ep.test) {
            tc_load_soa16(ep.eps, ep.Bs, grow, ng, ep.D, a.f);
        }
    } else {
        if (k == TCB_BWD_HIDDEN || k == TCB_GBAR_HIDDEN || k == TCB_HBAR_HIDDEN) {
            tc_ld256_nc(ep.hp[0] + rloc * ep.ld_h + ng, a.p[0]);
            tc_ld256_nc(ep.hp[1] + rloc * ep.ld_h + ng, a.p[1]);
            if (k != TCB_BWD_HIDDEN) tc_load_f32x16(ep.f32a + rloc * ep.ld_f + ng, a.f);
        } else if (k == TCB_FWD_LAST || k == TCB_GBAR_LAST) {
            tc_load_soa16(ep.eps, ep.Bs, grow, ng, ep.D, a.f);
        }
    }
}

// Epilogue-operand ring (RING kernels: the kinds whose epilogue reads h planes / an fp32 side array).
// A dedicated producer thread TMA-loads, chunk by chunk (128 rows x 16 columns), what the epilogue
// warps need into a ring of shared-memory slots, as far ahead as the ring is deep; the registers of
// 12 warps can only hold one chunk ahead (~48 KB in flight per SM), which left these kernels at
// ~3 TB/s waiting on long-scoreboard stalls.
// Each of the three warp groups (sub = chunk % 3) has its own ring of `d` slots, visited in order by
// its four warps: a warp that waits for round r of a slot has itself consumed round r-1 of it.  (With
// one ring shared by all groups a warp could reach a slot two phases early, and a parity wait on an
// mbarrier cannot tell "two phases behind" from "complete": stale operands in random rows.)
struct TcRing {
    const uint8_t* base;
    uint32_t j;      // chunks this warp has consumed so far
    int d;           // slots per warp group
    uint32_t zmask;  // always 0 (from a kernel argument): see tc_ring_read
};
template <bool BWD, int KIND> struct TcRingOf {
    static constexpr bool kind = BWD ? (KIND == TCB_BWD_HIDDEN || KIND == TCB_GBAR_HIDDEN || KIND == TCB_HBAR_HIDDEN)
                                     : (KIND == TC_BWD_HIDDEN);
    static constexpr bool f32 = BWD && (KIND == TCB_GBAR_HIDDEN || KIND == TCB_HBAR_HIDDEN);
    static constexpr uint32_t slot_bytes = f32 ? 16384u : 8192u;
};
__device__ __forceinline__ void tc_lds128(uint32_t addr, uint32_t* v) {
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(addr) : "memory");
}
template <bool F32>
__device__ __forceinline__ void tc_ring_read(TcSmemCtl* ctl, TcRing& ring, int sub, int row, int lane, TcAux& a) {
    const int slot = sub * ring.d + (int)(ring.j % (uint32_t)ring.d);
    tc::mbar_wait(&ctl->xfull[slot], (ring.j / (uint32_t)ring.d) & 1, 0x101);
    ring.j++;
    const uint32_t sp = tc::smem_u32(ring.base) + (uint32_t)slot * (F32 ? 16384u : 8192u);
    uint32_t dep = 0;
#pragma unroll
    for (int q = 0; q < 2; ++q) {
        tc_lds128(sp + q * 4096 + row * 32, a.p[q]);
        tc_lds128(sp + q * 4096 + row * 32 + 16, a.p[q] + 4);
        dep ^= a.p[q][0] ^ a.p[q][4];
    }
    if (F32) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            uint32_t w[4];
            tc_lds128(sp + 8192 + row * 64 + j * 16, w);
#pragma unroll
            for (int i = 0; i < 4; ++i) a.f[4 * j + i] = __uint_as_float(w[i]);
            dep ^= w[0];
        }
    }
    // The slot may be overwritten by the async proxy as soon as the fourth warp has arrived, so every
    // lane's reads must have RETURNED before the arrive is issued: the arrive's address is made to depend
    // on the loaded words (ptxas does not wait for outstanding loads ahead of a SYNCS.ARRIVE by itself).
    dep = __reduce_or_sync(0xffffffffu, dep & ring.zmask);   // zmask is 0 at run time, unknown to the compiler
    if (lane == 0) {
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(&ctl->xempty[slot]) + dep) : "memory");
    }
}

// sum of (a, b) over the warps of a TMEM lane quadrant (same lane = same sample row);
// double-buffered by tile parity, so one named barrier per reduction is enough
__device__ __forceinline__ float2 tc_quad_reduce(TcSmemCtl* ctl, int quad, int sub, int lane, int par, float a, float b) {
    ctl->red[par][quad][sub][lane] = make_float2(a, b);
    asm volatile("bar.sync %0, %1;" ::"r"(quad + 1), "n"(32 * TC_NSUB) : "memory");
    float2 r = make_float2(0.0f, 0.0f);
#pragma unroll
    for (int s = 0; s < TC_NSUB; ++s) {
        const float2 p = ctl->red[par][quad][s][lane];
        r.x += p.x; r.y += p.y;
    }
    return r;
}

// ---- forward-solve epilogue of one 128 x BN tile (kinds TC_*); cur = aux of this warp's first chunk
template <int PREC, int KIND, bool RING>
__device__ __forceinline__ void tc_epilogue_fwd(const TcEpi& ep, TcSmemCtl* ctl, uint32_t tbase, int mb, int nb, int BN,
                                                int quad, int sub, int nsub, int lane, int par, TcAux& cur, TcRing& ring) {
    const long long rloc = (long long)mb * TC_BM + quad * 32 + lane;   // row inside the chunk
    const long long grow = ep.row0 + rloc;                               // global sample
    constexpr bool last_kind = (KIND == TC_FWD_LAST || KIND == TC_BWD_LAST || KIND == TC_JVP_LAST);
    const int nchunk = (BN + TC_CW - 1) / TC_CW;
    const int nmax = last_kind ? ep.D : ep.N;
    float s0 = 0.0f, s1 = 0.0f;
    for (int ch = sub; ch < nchunk; ch += nsub) {
        const int c0 = ch * TC_CW, ng = nb * BN + c0;   // global column of v[0]
        if (ng >= nmax) break;
        TcAux nxt;
        if (RING) tc_ring_read<false>(ctl, ring, sub, quad * 32 + lane, lane, cur);
        else if (ch + nsub < nchunk && ng + nsub * TC_CW < nmax)
            tc_aux_load<PREC, false, KIND>(ep, rloc, grow, ng + nsub * TC_CW, nxt);
        uint32_t v[TC_CW];
        tc::tmem_ld_32x16(tbase + c0, v);
        tc::tmem_ld_wait();
        float x[TC_CW];
        if (KIND == TC_FWD_HIDDEN) {
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) x[j] = tc_tanh(__uint_as_float(v[j]) + cur.f[j]);
            tc_store_planes16<PREC>(x, ep.out, rloc * ep.ld_out + ng);
        } else if (KIND == TC_BWD_HIDDEN) {   // acc .* (1 - h^2)
            float h[TC_CW];
            tc_unpack_planes16<PREC>(cur.p, h);
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) x[j] = __uint_as_float(v[j]) * fmaf(-h[j], h[j], 1.0f);
            tc_store_planes16<PREC>(x, ep.out, rloc * ep.ld_out + ng);
        } else if (KIND == TC_FWD_LAST) {
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) {
                x[j] = 0.0f;
                if (c0 + j < ep.D) {
                    const float y = tc_tanh(__uint_as_float(v[j]) + __ldg(ep.bias + c0 + j));
                    ep.kout[(long long)(c0 + j) * ep.Bs + grow] = y;
                    s0 = fmaf(y, y, s0);
                    if (ep.vjp == 1) x[j] = cur.f[j] * fmaf(-y, y, 1.0f);
                }
            }
            if (ep.vjp == 1) tc_store_planes16<PREC>(x, ep.out, rloc * ep.ld_out + c0);
        } else {   // TC_BWD_LAST / TC_JVP_LAST: -eps . v and |v|^2
            float y[TC_CW];
            if (KIND == TC_JVP_LAST) tc_load_soa16(ep.kout, ep.Bs, grow, c0, ep.D, y);
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) {
                if (c0 + j < ep.D) {
                    float vv = __uint_as_float(v[j]);
                    if (KIND == TC_JVP_LAST) vv *= fmaf(-y[j], y[j], 1.0f);
                    s0 = fmaf(-vv, ep.test ? (c0 + j == ep.pass ? 1.0f : 0.0f) : cur.f[j], s0);
                    s1 = fmaf(vv, vv, s1);
                }
            }
        }
        if (!RING) cur = nxt;
    }
    if (!last_kind || (KIND == TC_FWD_LAST && ep.vjp == 2)) return;   // (vjp 2: zdot rows only)
    // per-sample reductions over all D columns: partial sums per warp, combined per quadrant
    const float2 r = tc_quad_reduce(ctl, quad, sub, lane, par, s0, s1);
    if (sub != 0) return;
    if (KIND == TC_FWD_LAST) {
        if (!ep.test) ep.kout[(long long)(ep.D + 1) * ep.Bs + grow] = (ep.flags & CNF_FLAG_NORM_Z) ? sqrtf(r.x) : 0.0f;
    } else if (ep.test) {
        // exact trace: with eps = e_i, -eps . (J eps) = -(J e_i)_i; the passes i = 0..D-1 add up to -tr J
        float* tr = ep.kout + (long long)ep.D * ep.Bs + grow;
        *tr = ep.pass ? *tr + r.x : r.x;
    } else {
        ep.kout[(long long)ep.D * ep.Bs + grow] = r.x;
        ep.kout[(long long)(ep.D + 2) * ep.Bs + grow] = (ep.flags & CNF_FLAG_NORM_J) ? sqrtf(r.y) : 0.0f;
    }
}

// ---- reverse-sweep epilogue of one 128 x BN tile (kinds TCB_*, bfloat16 planes)
template <int KIND, bool RING>
__device__ __forceinline__ void tc_epilogue_bwd(const TcEpi& ep, TcSmemCtl* ctl, uint32_t tbase, int mb, int nb, int BN,
                                                int quad, int sub, int lane, int par, TcAux& cur, TcRing& ring) {
    const long long rloc = (long long)mb * TC_BM + quad * 32 + lane;
    const long long grow = ep.row0 + rloc;
    const bool live = grow < (long long)ep.B;
    const int nchunk = (BN + TC_CW - 1) / TC_CW;
    const float dt = *ep.dtp;
    const float wgt = dt * c_A[6][ep.stage];
    if (KIND == TCB_FWD_LAST) {
        float s0 = 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            TcAux nxt;
            if (ch + TC_NSUB < nchunk && c0 + TC_NSUB * TC_CW < ep.D)
                tc_aux_load<1, true, KIND>(ep, rloc, grow, c0 + TC_NSUB * TC_CW, nxt);
            uint32_t v[TC_CW];
            tc::tmem_ld_32x16(tbase + c0, v);
            tc::tmem_ld_wait();
            float y[TC_CW], g[TC_CW];
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) {
                y[j] = 0.0f; g[j] = 0.0f;
                if (c0 + j < ep.D) {
                    y[j] = tc_tanh(__uint_as_float(v[j]) + __ldg(ep.bias + c0 + j));
                    s0 = fmaf(y[j], y[j], s0);
                    g[j] = cur.f[j] * fmaf(-y[j], y[j], 1.0f);
                }
            }
            tc_store_f32x16(ep.f32a + rloc * ep.ld_f + c0, y);
            tc_store_planes16<1>(g, ep.out, rloc * ep.ld_out + c0);
            cur = nxt;
        }
        const float2 r = tc_quad_reduce(ctl, quad, sub, lane, par, s0, 0.0f);
        const float inv = (live && (ep.flags & CNF_FLAG_NORM_Z) && r.x > 0.0f) ? wgt * ep.c_E / sqrtf(r.x) : 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            float y[TC_CW], yb[TC_CW], ab[TC_CW];
            tc_load_soa16(ep.abar, ep.Bs, grow, c0, ep.D, ab);
            tc_load_f32x16(ep.f32a + rloc * ep.ld_f + c0, y);
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) yb[j] = (live && c0 + j < ep.D) ? fmaf(inv, y[j], ab[j]) : 0.0f;
            tc_store_f32x16(ep.f32b + rloc * ep.ld_f + c0, yb);
        }
        return;
    }
    if (KIND == TCB_BWD_LAST) {
        float s1 = 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            uint32_t v[TC_CW];
            tc::tmem_ld_32x16(tbase + c0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < TC_CW; ++j)
                if (c0 + j < ep.D) s1 = fmaf(__uint_as_float(v[j]), __uint_as_float(v[j]), s1);
        }
        const float2 r = tc_quad_reduce(ctl, quad, sub, lane, par, s1, 0.0f);
        const float inv = (live && (ep.flags & CNF_FLAG_NORM_J) && r.x > 0.0f) ? wgt * ep.c_n / sqrtf(r.x) : 0.0f;
        const float al = live ? wgt * ep.c_l : 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            uint32_t v[TC_CW];
            float x[TC_CW], e[TC_CW];
            tc_load_soa16(ep.eps, ep.Bs, grow, c0, ep.D, e);
            tc::tmem_ld_32x16(tbase + c0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) x[j] = (c0 + j < ep.D) ? fmaf(inv, __uint_as_float(v[j]), -al * e[j]) : 0.0f;
            tc_store_planes16<1>(x, ep.out, rloc * ep.ld_out + c0);
        }
        return;
    }
    if (KIND == TCB_JVP_LAST) {
        // acc = s_L = tau_{L-1} W_L^T.  Pass 1: |tau_L|^2 with tau_L = s_L (1 - y^2); pass 2: the three cotangents.
        float s1 = 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            uint32_t v[TC_CW];
            float y[TC_CW];
            tc_load_f32x16(ep.f32a + rloc * ep.ld_f + c0, y);
            tc::tmem_ld_32x16(tbase + c0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < TC_CW; ++j)
                if (c0 + j < ep.D) {
                    const float t = __uint_as_float(v[j]) * fmaf(-y[j], y[j], 1.0f);
                    s1 = fmaf(t, t, s1);
                }
        }
        const float2 r = tc_quad_reduce(ctl, quad, sub, lane, par, s1, 0.0f);
        const float inv = (live && (ep.flags & CNF_FLAG_NORM_J) && r.x > 0.0f) ? wgt * ep.c_n / sqrtf(r.x) : 0.0f;
        const float al = live ? wgt * ep.c_l : 0.0f;
        for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
            const int c0 = ch * TC_CW;
            if (c0 >= ep.D) break;
            uint32_t v[TC_CW];
            float y[TC_CW], yb[TC_CW], e[TC_CW], sb[TC_CW], pb[TC_CW];
            tc_load_soa16(ep.eps, ep.Bs, grow, c0, ep.D, e);
            tc_load_f32x16(ep.f32a + rloc * ep.ld_f + c0, y);
            tc_load_f32x16(ep.f32b + rloc * ep.ld_f + c0, yb);
            tc::tmem_ld_32x16(tbase + c0, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) {
                sb[j] = 0.0f; pb[j] = 0.0f;
                if (c0 + j < ep.D) {
                    const float sl = __uint_as_float(v[j]);
                    const float dl = fmaf(-y[j], y[j], 1.0f);
                    const float tb = fmaf(inv, sl * dl, -al * e[j]);     // taubar_L
                    sb[j] = tb * dl;                                      // sbar_L
                    pb[j] = fmaf(-2.0f * y[j], tb * sl, yb[j]) * dl;      // abar_L = (ybar - 2 y dbar_L) d_L
                }
            }
            tc_store_planes16<1>(sb, ep.out, rloc * ep.ld_out + c0);
            tc_store_planes16<1>(pb, ep.out2, rloc * ep.ld_out2 + c0);
        }
        return;
    }
    // column-wise kinds
    for (int ch = sub; ch < nchunk; ch += TC_NSUB) {
        const int c0 = ch * TC_CW, ng = nb * BN + c0;
        if (ng >= ep.N) break;
        TcAux nxt;
        if (RING) tc_ring_read<TcRingOf<true, KIND>::f32>(ctl, ring, sub, quad * 32 + lane, lane, cur);
        else if (ch + TC_NSUB < nchunk && ng + TC_NSUB * TC_CW < ep.N)
            tc_aux_load<1, true, KIND>(ep, rloc, grow, ng + TC_NSUB * TC_CW, nxt);
        float y[TC_CW], yb[TC_CW];
        if (KIND == TCB_GBAR_LAST) {
            tc_load_f32x16(ep.f32a + rloc * ep.ld_f + ng, y);
            tc_load_f32x16(ep.f32b + rloc * ep.ld_f + ng, yb);
        }
        uint32_t v[TC_CW];
        tc::tmem_ld_32x16(tbase + c0, v);
        tc::tmem_ld_wait();
        float x[TC_CW];
        if (KIND == TCB_ZBAR) {   // zbar goes to an SoA scratch; k_zbar_apply streams it into ubar / kbar_j
#pragma unroll
            for (int j = 0; j < TC_CW; ++j)
                if (ng + j < ep.D) ep.ubar[(long long)(ng + j) * ep.Bs + grow] = __uint_as_float(v[j]);
            continue;
        }
        if (KIND == TCB_GBAR_LAST) {
#pragma unroll
            for (int j = 0; j < TC_CW; ++j) {
                x[j] = 0.0f;
                if (ng + j < ep.D) {
                    const float db = __uint_as_float(v[j]) * cur.f[j];
                    x[j] = fmaf(-2.0f * y[j], db, yb[j]) * fmaf(-y[j], y[j], 1.0f);
                }
            }
        } else {
            float h[TC_CW];
            tc_unpack_planes16<1>(cur.p, h);
            if (KIND == TCB_BWD_HIDDEN) {
                float q[TC_CW];
#pragma unroll
                for (int j = 0; j < TC_CW; ++j) {
                    q[j] = __uint_as_float(v[j]);
                    x[j] = q[j] * fmaf(-h[j], h[j], 1.0f);
                }
                tc_store_f32x16(ep.f32a + rloc * ep.ld_f + ng, q);
            } else if (KIND == TCB_GBAR_HIDDEN) {
                float q[TC_CW];
#pragma unroll
                for (int j = 0; j < TC_CW; ++j) {
                    const float gb = __uint_as_float(v[j]);
                    q[j] = cur.f[j] * gb;                      // dbar_l
                    x[j] = gb * fmaf(-h[j], h[j], 1.0f);      // qbar_{l+1}
                }
                tc_store_f32x16(ep.f32a + rloc * ep.ld_f + ng, q);
            } else {   // TCB_HBAR_HIDDEN
#pragma unroll
                for (int j = 0; j < TC_CW; ++j)
                    x[j] = fmaf(-2.0f * h[j], cur.f[j], __uint_as_float(v[j])) * fmaf(-h[j], h[j], 1.0f);
            }
        }
        tc_store_planes16<1>(x, ep.out, rloc * ep.ld_out + ng);
        if (!RING) cur = nxt;
    }
}

// BM x BN output tile per (m_blk, n_blk); K in blocks of 64 bf16; persistent over tiles.
// PREC = operand format of the two hi/lo planes (0: IEEE half, forward solve; 1: bfloat16, reverse sweep).
template <int PREC, bool BWD, int KIND, bool RING>
__global__ void __launch_bounds__((!BWD && KIND == TC_FWD_HIDDEN) ? TC_THREADS_WIDE : TC_THREADS, 1)
k_tc_gemm(const __grid_constant__ TcMaps mp, int m_tiles, int n_tiles, int BN, int nk, int nstage, int naux, TcEpi ep) {
    if (ep.ctrl && ep.ctrl->done) return;
    if (ep.kout_ind) ep.kout = *ep.kout_ind;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const uint32_t a_bytes = TC_BM * 128;          // one plane of A per stage
    const uint32_t b_bytes = (uint32_t)BN * 128;   // one plane of B per stage
    const uint32_t stage_bytes = 2 * (a_bytes + b_bytes);
    constexpr uint32_t slot_bytes = TcRingOf<BWD, KIND>::slot_bytes;
    uint8_t* ring_base = smem + (size_t)nstage * stage_bytes;
    TcSmemCtl* ctl = reinterpret_cast<TcSmemCtl*>(ring_base + (RING ? (size_t)TC_NSUB * naux * slot_bytes : 0));   // naux = slots per warp group
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int total = m_tiles * n_tiles;
    const int nsub = (((int)blockDim.x >> 5) - 3) >> 2;   // epilogue warps per TMEM lane quadrant (3, or 5 for TC_FWD_HIDDEN)
    const int aux_warp = TC_EPI_WARP0 + 4 * nsub;         // the last warp

    if (warp == 0 && lane == 0) {
        for (int q = 0; q < 2; ++q) { tc::tma_prefetch_desc(&mp.a[q]); tc::tma_prefetch_desc(&mp.b[q]); }
        for (int s = 0; s < nstage; ++s) { tc::mbar_init(&ctl->full[s], 1); tc::mbar_init(&ctl->empty[s], 1); }
        for (int a = 0; a < 2; ++a) { tc::mbar_init(&ctl->tfull[a], 1); tc::mbar_init(&ctl->tempty[a], 4 * nsub); }
        if (RING)
            for (int x = 0; x < TC_NSUB * naux; ++x) { tc::mbar_init(&ctl->xfull[x], 1); tc::mbar_init(&ctl->xempty[x], 4); }
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
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int mb = t / n_tiles, nb = t % n_tiles;
                for (int kb = 0; kb < nk; ++kb) {
                    tc::mbar_wait(&ctl->empty[stage], phase ^ 1, 0x102);
                    uint8_t* s = smem + (size_t)stage * stage_bytes;
                    tc::mbar_arrive_expect_tx(&ctl->full[stage], stage_bytes);
                    for (int q = 0; q < 2; ++q) {
                        tc::tma_load_2d(&mp.a[q], &ctl->full[stage], s + q * a_bytes, kb * TC_BK, mb * TC_BM);
                        tc::tma_load_2d(&mp.b[q], &ctl->full[stage], s + 2 * a_bytes + q * b_bytes, kb * TC_BK, nb * BN);
                    }
                    if (++stage == nstage) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {   // ===== MMA issuer
            const uint32_t idesc = tc::umma_idesc_f16(TC_BM, BN, PREC);
            int stage = 0; uint32_t phase = 0; int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
                const int acc = it & 1;
                tc::mbar_wait(&ctl->tempty[acc], ((it >> 1) & 1) ^ 1, 0x103);
                tc::tc_fence_after();
                const uint32_t dcol = tmem + (uint32_t)acc * 256;
                for (int kb = 0; kb < nk; ++kb) {
                    tc::mbar_wait(&ctl->full[stage], phase, 0x104);
                    tc::tc_fence_after();
                    const uint32_t s = tc::smem_u32(smem + (size_t)stage * stage_bytes);
#pragma unroll
                    for (int kk = 0; kk < TC_BK / 16; ++kk) {
                        uint64_t a[2], b[2];
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            a[q] = tc::umma_desc_sw128(s + q * a_bytes + kk * 32);
                            b[q] = tc::umma_desc_sw128(s + 2 * a_bytes + q * b_bytes + kk * 32);
                        }
                        tc::umma_bf16(dcol, a[0], b[0], idesc, (kb | kk) != 0);
                        tc::umma_bf16(dcol, a[1], b[0], idesc, 1);
                        tc::umma_bf16(dcol, a[0], b[1], idesc, 1);
                    }
                    tc::umma_commit(&ctl->empty[stage]);
                    if (++stage == nstage) { stage = 0; phase ^= 1; }
                }
                tc::umma_commit(&ctl->tfull[acc]);
            }
        }
    } else if (warp == aux_warp) {
        if (RING && lane == 0) {   // ===== epilogue-operand producer: chunk c of a tile goes to the warps with sub = c % 3
            uint32_t j[TC_NSUB] = {0, 0, 0};
            for (int t = blockIdx.x; t < total; t += gridDim.x) {
                const int mb = t / n_tiles, nb = t % n_tiles;
                const int ncol = min(BN, ep.N - nb * BN);
                for (int c = 0; c * TC_CW < ncol; ++c) {
                    const int k = c % TC_NSUB;
                    const uint32_t jj = (k == 0) ? j[0]++ : (k == 1 ? j[1]++ : j[2]++);
                    const int slot = k * naux + (int)(jj % (uint32_t)naux);
                    tc::mbar_wait(&ctl->xempty[slot], ((jj / (uint32_t)naux) & 1) ^ 1, 0x105);
                    uint8_t* sp = ring_base + (size_t)slot * slot_bytes;
                    tc::mbar_arrive_expect_tx(&ctl->xfull[slot], slot_bytes);
                    tc::tma_load_2d(&mp.x[0], &ctl->xfull[slot], sp, nb * BN + c * TC_CW, mb * TC_BM);
                    tc::tma_load_2d(&mp.x[1], &ctl->xfull[slot], sp + 4096, nb * BN + c * TC_CW, mb * TC_BM);
                    if (TcRingOf<BWD, KIND>::f32)
                        tc::tma_load_2d(&mp.x[2], &ctl->xfull[slot], sp + 8192, nb * BN + c * TC_CW, mb * TC_BM);
                }
            }
        }
    } else {               // ===== epilogue warps
        TcRing ring;
        ring.base = ring_base; ring.j = 0; ring.d = naux; ring.zmask = (uint32_t)ep.dbg & 0x80000000u;
        const int quad = warp & 3;                    // TMEM lane quadrant this warp may access
        const int sub = (warp - TC_EPI_WARP0) >> 2;   // 0..2: which chunks of the quadrant's columns
        int it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const int mb = t / n_tiles, nb = t % n_tiles;
            const int acc = it & 1;
            TcAux aux;   // global operands of this warp's first chunk: in flight while the MMAs finish
            {
                const long long rloc = (long long)mb * TC_BM + quad * 32 + lane;
                const int ng0 = nb * BN + sub * TC_CW;
                if (!RING && sub * TC_CW < BN && ng0 < ep.N) tc_aux_load<PREC, BWD, KIND>(ep, rloc, ep.row0 + rloc, ng0, aux);
            }
            tc::mbar_wait(&ctl->tfull[acc], (it >> 1) & 1, 0x106);
            tc::tc_fence_after();
            const uint32_t tbase = tmem + ((uint32_t)(quad * 32) << 16) + (uint32_t)acc * 256;
            if (BWD) tc_epilogue_bwd<KIND, RING>(ep, ctl, tbase, mb, nb, BN, quad, sub, lane, it & 1, aux, ring);
            else tc_epilogue_fwd<PREC, KIND, RING>(ep, ctl, tbase, mb, nb, BN, quad, sub, nsub, lane, it & 1, aux, ring);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&ctl->tempty[acc]);
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------ small helper kernels
// fp32 weights (Julia column-major dout x din) -> hi/lo planes [rows x ld], zero padded, in both
// formats (half for the forward solve, bfloat16 for the reverse sweep).
// transpose = 0: plane[o][i] = W[o,i] (forward B operand); 1: plane[i][o] = W[o,i] (VJP B operand)
__global__ void k_tc_weights(const float* __restrict__ W, int dout, int din, __nv_bfloat16* __restrict__ h0,
                             __nv_bfloat16* __restrict__ h1, __nv_bfloat16* __restrict__ b0,
                             __nv_bfloat16* __restrict__ b1, int rows, int ld, int transpose) {
    const long long n = (long long)rows * ld;
    unsigned short* o16[4] = {reinterpret_cast<unsigned short*>(h0), reinterpret_cast<unsigned short*>(h1),
                              reinterpret_cast<unsigned short*>(b0), reinterpret_cast<unsigned short*>(b1)};
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += (long long)gridDim.x * blockDim.x) {
        int r = (int)(p / ld), c = (int)(p % ld);
        int o = transpose ? c : r, i = transpose ? r : c;
        float w = (o < dout && i < din) ? W[(long long)i * dout + o] : 0.0f;
        uint32_t hi, lo;
        tc_split2<0>(w, 0.0f, hi, lo);
        o16[0][p] = (unsigned short)(hi & 0xffff);
        o16[1][p] = (unsigned short)(lo & 0xffff);
        tc_split2<1>(w, 0.0f, hi, lo);
        o16[2][p] = (unsigned short)(hi & 0xffff);
        o16[3][p] = (unsigned short)(lo & 0xffff);
    }
}

struct TcDyn { float* kout; };

// stage input rows (SoA, D x Bs) [+ condition rows] -> hi/lo planes [mc x ld] of this chunk;
// also resolves the device-side output buffer of this evaluation (kidx: 0 -> K[k1idx], 6 -> K[6-k1idx])
template <int PREC>
__global__ void k_tc_prep(const CnfCtrl* __restrict__ c, CnfBufs bf, const float* __restrict__ src, int nsrc,
                          const float* __restrict__ src2, int nsrc2, long long row0, int mc,
                          __nv_bfloat16* __restrict__ p0, __nv_bfloat16* __restrict__ p1, int ld, int kidx,
                          float* kout_fixed, TcDyn* dyn, int check_done) {
    if (check_done && c->done) return;
    if (blockIdx.x == 0 && threadIdx.x == 0 && dyn) {
        float* k = kout_fixed;
        if (!k) k = (kidx == 6) ? bf.K[6 - c->k1idx] : (kidx == 0 ? bf.K[c->k1idx] : bf.K[kidx]);
        dyn->kout = k;
    }
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= mc) return;
    const long long b = row0 + r;
    const int ntot = nsrc + nsrc2;
    for (int c0 = 0; c0 < ntot; c0 += 8) {
        uint32_t pk[2][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v[2];
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int i = c0 + 2 * j + t;
                v[t] = 0.0f;
                if (i < nsrc) v[t] = src[(long long)i * bf.Bs + b];
                else if (i < ntot) v[t] = src2[(long long)(i - nsrc) * bf.Bs + b];
            }
            tc_split2<PREC>(v[0], v[1], pk[0][j], pk[1][j]);
        }
        __nv_bfloat16* dst[2] = {p0, p1};
#pragma unroll
        for (int q = 0; q < 2; ++q)
            *reinterpret_cast<uint4*>(dst[q] + (long long)r * ld + c0) = make_uint4(pk[q][0], pk[q][1], pk[q][2], pk[q][3]);
    }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled tc_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess) fn = (PFN_encodeTiled)p;
    }
    return fn;
}
// 2D bf16 tensor [rows x cols] row-major (ld elements), box = [box_rows x 64 cols], SWIZZLE_128B
static bool tc_make_map(CUtensorMap* m, const void* base, long long rows, long long cols, long long ld, int box_rows,
                        int box_cols = TC_BK) {
    PFN_encodeTiled fn = tc_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, box_cols == TC_BK ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static bool tc_make_map_f32(CUtensorMap* m, const void* base, long long rows, long long ld) {
    PFN_encodeTiled fn = tc_encode_fn();
    if (!fn) return false;
    cuuint64_t dims[2] = {(cuuint64_t)ld, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    cuuint32_t box[2] = {TC_CW, TC_BM};
    cuuint32_t estr[2] = {1, 1};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

static inline int pad_to(int x, int m) { return (x + m - 1) / m * m; }

struct TcPlane {
    __nv_bfloat16* p[2] = {nullptr, nullptr};   // hi / lo plane (16-bit storage; half or bfloat16 by use)
    int rows = 0, ld = 0;
    CUtensorMap map[2];    // box [box_rows x 64 cols]: K-major operand of the layer GEMMs
    CUtensorMap mapk[2];   // box [ 64 rows x 64 cols]: MN-major operand of the weight-gradient GEMMs
    CUtensorMap mapx[2];   // box [128 rows x 16 cols], no swizzle: epilogue-operand ring
};
struct TcWeight {
    TcPlane ph, pb;   // half planes (forward solve), bfloat16 planes (reverse sweep); box rows = BN
    int N = 0, K = 0, BN = 0, n_tiles = 0;
    int BN128 = 0, n_tiles128 = 0;   // RING kernels: tiles of at most 128 columns
    CUtensorMap map128[2][2];        // [format][plane], box rows = BN128
};
struct TcF32 {
    float* p = nullptr;
    int rows = 0, ld = 0;
    CUtensorMap map;   // box [128 rows x 16 cols], no swizzle: epilogue-operand ring
};
struct TcState {
    int L = 0, Mc = 0;
    TcWeight wf[CNF_MAX_LAYERS], wb[CNF_MAX_LAYERS];
    TcPlane H[CNF_MAX_LAYERS + 1], G[CNF_MAX_LAYERS + 1];
    // reverse sweep (cnf_backward_tc.cu): qbar_l, pbar_l planes (l = 1..L), q_{l+1}/dbar_l (l = 1..L-1), y, ybar
    TcPlane QB[CNF_MAX_LAYERS + 1], PB[CNF_MAX_LAYERS + 1];
    TcF32 Q[CNF_MAX_LAYERS + 1], Y, YB;
    __nv_bfloat16* ones = nullptr;   // 2 KB of bf16 1.0: B operand of the bias-gradient MMAs
    bool bw_ok = false;
    TcDyn* dyn = nullptr;
    int sms = 148;
    bool weights_ok = false, acts_ok = false;
};

static void tc_free_plane(TcPlane& p) {
    for (int q = 0; q < 2; ++q) { cudaFree(p.p[q]); p.p[q] = nullptr; }
    p.rows = p.ld = 0;
}
static void tc_free_f32(TcF32& f) { cudaFree(f.p); f.p = nullptr; f.rows = f.ld = 0; }
static int tc_alloc_plane(cnf_handle* h, TcPlane& p, int rows, int ld, int box_rows, bool kmap = true) {
    tc_free_plane(p);
    p.rows = rows; p.ld = ld;
    size_t bytes = (size_t)rows * ld * sizeof(__nv_bfloat16);
    for (int q = 0; q < 2; ++q) {
        CNF_CUDA(cudaMalloc(&p.p[q], bytes));
        CNF_CUDA(cudaMemsetAsync(p.p[q], 0, bytes, h->stream));
        if (!tc_make_map(&p.map[q], p.p[q], rows, ld, ld, box_rows) ||
            (kmap && (!tc_make_map(&p.mapk[q], p.p[q], rows, ld, ld, 64) ||
                      !tc_make_map(&p.mapx[q], p.p[q], rows, ld, ld, TC_BM, TC_CW)))) {
            h->err = "cuTensorMapEncodeTiled failed";
            return CNF_ERR_CUDA;
        }
    }
    return CNF_OK;
}
static int tc_alloc_f32(cnf_handle* h, TcF32& f, int rows, int ld) {
    tc_free_f32(f);
    f.rows = rows; f.ld = ld;
    CNF_CUDA(cudaMalloc(&f.p, (size_t)rows * ld * sizeof(float)));
    CNF_CUDA(cudaMemsetAsync(f.p, 0, (size_t)rows * ld * sizeof(float), h->stream));
    if (!tc_make_map_f32(&f.map, f.p, rows, ld)) { h->err = "cuTensorMapEncodeTiled failed"; return CNF_ERR_CUDA; }
    return CNF_OK;
}

// the tensor path covers TrainMode (Hutchinson) on chains of >= 2 layers whose hidden widths are
// multiples of 32 and >= 128 (auto) — or any multiple of 32 when forced with kernel_path = 3
// TrainMode (Hutchinson rows, both trace modes) and TestMode (exact trace = D one-hot tangent chains on the
// same layer-GEMM kernels; D passes, so only for moderate D)
bool tc_supported(const cnf_config& c, int mode) {
    if ((mode != CNF_MODE_TRAIN && mode != CNF_MODE_TEST) || c.n_layers < 2 || c.D > 256) return false;
    if (mode == CNF_MODE_TEST && c.D > 64) return false;
    if (c.flags & CNF_FLAG_LAST_LINEAR) return false;   // generic kernels only
    bool wide = true;
    for (int l = 1; l < c.n_layers; ++l) {
        if (c.dims[l] % 32 != 0) return false;
        if (c.dims[l] < 128) wide = false;
    }
    return c.kernel_path == 3 || (c.kernel_path == 0 && wide);
}

void tc_free(cnf_handle* h) {
    TcState* s = (TcState*)h->tc_state;
    if (!s) return;
    for (int l = 0; l < CNF_MAX_LAYERS; ++l) {
        tc_free_plane(s->wf[l].ph); tc_free_plane(s->wf[l].pb); tc_free_plane(s->wb[l].ph); tc_free_plane(s->wb[l].pb);
    }
    for (int l = 0; l <= CNF_MAX_LAYERS; ++l) {
        tc_free_plane(s->H[l]); tc_free_plane(s->G[l]); tc_free_plane(s->QB[l]); tc_free_plane(s->PB[l]);
        tc_free_f32(s->Q[l]);
    }
    tc_free_f32(s->Y); tc_free_f32(s->YB);
    cudaFree(s->ones);
    cudaFree(s->dyn);
    delete s;
    h->tc_state = nullptr;
}

static int tc_setup_weight(cnf_handle* h, TcWeight& w, const float* W, int dout, int din, int transpose, int Nvalid) {
    // operand B: [N rows x K cols]; forward: N = dout, K = din; VJP: N = din (or D), K = dout
    w.N = Nvalid;
    w.K = pad_to(transpose ? dout : din, TC_BK);
    const int n16 = pad_to(Nvalid, 16);
    w.BN = n16 < 256 ? n16 : 256;   // two planes of A and B per stage: 96 KB at BN = 256
    w.n_tiles = (n16 + w.BN - 1) / w.BN;
    const int rows = w.n_tiles * w.BN;
    if (w.ph.rows != rows || w.ph.ld != w.K) {
        int rc;
        if ((rc = tc_alloc_plane(h, w.ph, rows, w.K, w.BN, false))) return rc;
        if ((rc = tc_alloc_plane(h, w.pb, rows, w.K, w.BN, false))) return rc;
    }
    w.BN128 = n16 < 128 ? n16 : 128;
    w.n_tiles128 = (n16 + w.BN128 - 1) / w.BN128;
    for (int q = 0; q < 2; ++q)
        if (!tc_make_map(&w.map128[0][q], w.ph.p[q], rows, w.K, w.K, w.BN128) ||
            !tc_make_map(&w.map128[1][q], w.pb.p[q], rows, w.K, w.K, w.BN128)) {
            h->err = "cuTensorMapEncodeTiled failed";
            return CNF_ERR_CUDA;
        }
    k_tc_weights<<<296, 256, 0, h->stream>>>(W, dout, din, w.ph.p[0], w.ph.p[1], w.pb.p[0], w.pb.p[1], rows, w.K, transpose);
    h->launches++;
    return CNF_OK;
}

// called from set_params (device theta is in place)
int tc_set_params(cnf_handle* h) {
    const cnf_config& c = h->cfg;
    if (!tc_supported(c, CNF_MODE_TRAIN)) return CNF_OK;
    if (!h->tc_state) {
        TcState* s = new TcState();
        h->tc_state = s;
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, h->device) == cudaSuccess) s->sms = prop.multiProcessorCount;
        CNF_CUDA(cudaMalloc(&s->dyn, sizeof(TcDyn)));
    }
    TcState* s = (TcState*)h->tc_state;
    s->L = c.n_layers;
    int rc;
    for (int l = 0; l < c.n_layers; ++l) {
        const float* W = h->d_theta + h->w_off[l];
        int dout = c.dims[l + 1], din = c.dims[l];
        if ((rc = tc_setup_weight(h, s->wf[l], W, dout, din, 0, dout))) return rc;
        if ((rc = tc_setup_weight(h, s->wb[l], W, dout, din, 1, l == 0 ? c.D : din))) return rc;
    }
    s->weights_ok = true;
    return CNF_OK;
}

// bytes of activation workspace per sample row: H/G planes + reverse-sweep planes and fp32 side arrays
static long long tc_row_bytes(const cnf_config& c) {
    long long b = 0;
    for (int l = 0; l <= c.n_layers; ++l) b += 2LL * pad_to(c.dims[l], TC_BK) * 2 * 2;
    for (int l = 1; l <= c.n_layers; ++l) b += (long long)(pad_to(c.dims[l - 1], TC_BK) + pad_to(c.dims[l], TC_BK)) * 2 * 2;
    for (int l = 1; l < c.n_layers; ++l) b += 4LL * pad_to(c.dims[l], TC_BK);
    b += 8LL * pad_to(c.D, TC_BK);
    return b;
}

static int tc_ensure_acts(cnf_handle* h, TcState* s, long long Bs) {
    const cnf_config& c = h->cfg;
    // chunk of samples per GEMM launch: as large as a 16 GB activation budget allows (whole waves of
    // 128-sample tiles), so the per-launch pipeline fill / TMEM allocation is amortised
    long long rows_budget = (16LL << 30) / tc_row_bytes(c);
    long long wave = (long long)s->sms * TC_BM;
    long long want = rows_budget / wave * wave;
    if (want < 2 * wave) want = 2 * wave;
    if (want > Bs) want = Bs;
    if (s->acts_ok && s->Mc == (int)want) return CNF_OK;
    int rc;
    s->Mc = (int)want;
    s->bw_ok = false;
    for (int l = 0; l < c.n_layers; ++l)
        if ((rc = tc_alloc_plane(h, s->H[l], s->Mc, pad_to(c.dims[l], TC_BK), TC_BM))) return rc;
    for (int l = 1; l <= c.n_layers; ++l)
        if ((rc = tc_alloc_plane(h, s->G[l], s->Mc, pad_to(c.dims[l], TC_BK), TC_BM))) return rc;
    if ((rc = tc_alloc_plane(h, s->G[0], s->Mc, pad_to(c.dims[0], TC_BK), TC_BM))) return rc;   // JVP seed planes [eps; 0]
    s->acts_ok = true;
    return CNF_OK;
}

// one kernel instantiation per (format, epilogue kind[, ring]): each carries only its own epilogue path
template <int PREC, bool BWD, int KIND, bool RING>
static void tc_launch_kind(cnf_handle* h, int grid, int smem, const TcMaps& mp, int m_tiles, int n_tiles, int BN, int nk,
                           int nstage, int naux, const TcEpi& ep) {
    constexpr bool valid = (BWD ? (KIND >= TCB_FWD_LAST && PREC == 1) : (KIND <= TC_JVP_LAST && (PREC == 0 || KIND == TC_FWD_HIDDEN))) &&
                           (!RING || TcRingOf<BWD, KIND>::kind);
    if constexpr (valid) {
        static bool attr_dev[64] = {};   // per device
        bool& attr = attr_dev[h->device & 63];
        if (!attr) {
            cudaFuncSetAttribute(k_tc_gemm<PREC, BWD, KIND, RING>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            attr = true;
        }
        constexpr int threads = (!BWD && KIND == TC_FWD_HIDDEN) ? TC_THREADS_WIDE : TC_THREADS;
        k_tc_gemm<PREC, BWD, KIND, RING><<<grid, threads, smem, h->stream>>>(mp, m_tiles, n_tiles, BN, nk, nstage, naux, ep);
    }
}

static bool tc_kind_has_ring(int kind) {
    return kind == TC_BWD_HIDDEN || kind == TCB_BWD_HIDDEN || kind == TCB_GBAR_HIDDEN || kind == TCB_HBAR_HIDDEN;
}

template <int PREC, bool BWD>
static void tc_launch_gemm(cnf_handle* h, TcState* s, const TcPlane& A, const TcWeight& B, int mc, const TcEpi& ep) {
    // RING kernels trade half of the tile width for a shared-memory ring of epilogue operands
    bool ring = tc_kind_has_ring(ep.kind) && ep.aux_p && !h->tc_noring;
    const bool f32kind = ep.kind == TCB_GBAR_HIDDEN || ep.kind == TCB_HBAR_HIDDEN;
    if ((h->tc_dbg & 8) && f32kind) ring = false;
    if ((h->tc_dbg & 16) && !f32kind) ring = false;
    const int BN = ring ? B.BN128 : B.BN, n_tiles = ring ? B.n_tiles128 : B.n_tiles;
    const int m_tiles = mc / TC_BM;
    const int nk = B.K / TC_BK;
    const int stage_bytes = 2 * (TC_BM * 128 + BN * 128);
    const int budget = 225 * 1024 - 2048 - (int)sizeof(TcSmemCtl);
    int nstage, naux = 0, slot = 0;
    if (ring) {
        slot = (ep.kind == TCB_GBAR_HIDDEN || ep.kind == TCB_HBAR_HIDDEN) ? 16384 : 8192;
        nstage = 2;
        // slots per warp group; the row-reduction scratch at the end of TcSmemCtl is not used by these kinds
        naux = (227 * 1024 - 1024 - 64 - (int)offsetof(TcSmemCtl, red) - nstage * stage_bytes) / (TC_NSUB * slot);
        if (naux > 4) naux = 4;
        if ((h->tc_dbg & 4) && naux > 1) naux = 1;
    } else {
        nstage = budget / stage_bytes;
        if (nstage > 6) nstage = 6;
        if (nstage < 2) nstage = 2;
    }
    const int smem = nstage * stage_bytes + TC_NSUB * naux * slot + 1024 + (ring ? (int)offsetof(TcSmemCtl, red) : (int)sizeof(TcSmemCtl)) + 64;
    int grid = m_tiles * n_tiles;
    if (grid > s->sms) grid = s->sms;
    const TcPlane& Bp = PREC == 0 ? B.ph : B.pb;
    TcMaps mp;
    for (int q = 0; q < 2; ++q) { mp.a[q] = A.map[q]; mp.b[q] = ring ? B.map128[PREC][q] : Bp.map[q]; }
    if (ring) {
        const TcPlane* xp = (const TcPlane*)ep.aux_p;
        mp.x[0] = xp->mapx[0]; mp.x[1] = xp->mapx[1];
        if (ep.aux_f) mp.x[2] = ((const TcF32*)ep.aux_f)->map;
    }
    switch (ep.kind) {
#define TC_CASE(K)                                                                                                  \
    case K:                                                                                                         \
        if (ring) tc_launch_kind<PREC, BWD, K, true>(h, grid, smem, mp, m_tiles, n_tiles, BN, nk, nstage, naux, ep);   \
        else tc_launch_kind<PREC, BWD, K, false>(h, grid, smem, mp, m_tiles, n_tiles, BN, nk, nstage, naux, ep);       \
        break;
        TC_CASE(TC_FWD_HIDDEN) TC_CASE(TC_FWD_LAST) TC_CASE(TC_BWD_HIDDEN) TC_CASE(TC_BWD_LAST) TC_CASE(TC_JVP_LAST)
        TC_CASE(TCB_FWD_LAST) TC_CASE(TCB_BWD_HIDDEN) TC_CASE(TCB_BWD_LAST) TC_CASE(TCB_GBAR_HIDDEN)
        TC_CASE(TCB_GBAR_LAST) TC_CASE(TCB_HBAR_HIDDEN) TC_CASE(TCB_ZBAR) TC_CASE(TCB_JVP_LAST)
#undef TC_CASE
        default: break;
    }
    h->launches++;
}

static TcEpi tc_base_epi(cnf_handle* h, TcState* s, const CnfBufs& bf, long long row0, int check_done) {
    TcEpi base;
    memset(&base, 0, sizeof(base));
    base.eps = bf.eps; base.Bs = bf.Bs; base.row0 = row0; base.D = h->cfg.D; base.flags = h->cfg.flags;
    base.kout_ind = &s->dyn->kout;
    base.ctrl = check_done ? h->d_ctrl : nullptr;
    base.B = bf.B;
    base.dbg = h->tc_dbg;
    return base;
}

// forward chain h_0 -> h_1 .. h_{L-1} (planes H[l]) of one chunk; the last layer is left to the caller
template <int PREC>
static void tc_forward_hidden(cnf_handle* h, TcState* s, const TcEpi& base, int mc) {
    const cnf_config& c = h->cfg;
    for (int l = 0; l < c.n_layers - 1; ++l) {
        TcEpi ep = base;
        ep.bias = h->d_theta + h->b_off[l];
        ep.N = c.dims[l + 1];
        ep.kind = TC_FWD_HIDDEN;
        for (int q = 0; q < 2; ++q) ep.out[q] = s->H[l + 1].p[q];
        ep.ld_out = s->H[l + 1].ld;
        tc_launch_gemm<PREC, false>(h, s, s->H[l], s->wf[l], mc, ep);
    }
}

// first layer of the i-th one-hot tangent chain of the exact trace: tau_1 = (e_i W_1^T) .* (1 - h_1^2) =
// W_1[:, i] .* (1 - h_1^2) — no GEMM and no seed planes; one thread = 16 columns of one sample row
template <int PREC>
__global__ void k_tc_onehot_l1(const CnfCtrl* __restrict__ ctrl, const float* __restrict__ wcol, int N,
                               const __nv_bfloat16* __restrict__ h0, const __nv_bfloat16* __restrict__ h1, int ld_h,
                               __nv_bfloat16* o0, __nv_bfloat16* o1, int ld_o, int mc) {
    if (ctrl && ctrl->done) return;
    const int nch = N / TC_CW;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long row = idx / nch;
    const int c0 = (int)(idx % nch) * TC_CW;
    if (row >= mc) return;
    uint32_t p[2][8];
    tc_ld256_nc(h0 + row * ld_h + c0, p[0]);
    tc_ld256_nc(h1 + row * ld_h + c0, p[1]);
    float h[TC_CW], x[TC_CW];
    tc_unpack_planes16<PREC>(p, h);
#pragma unroll
    for (int j = 0; j < TC_CW; ++j) x[j] = __ldg(wcol + c0 + j) * fmaf(-h[j], h[j], 1.0f);
    __nv_bfloat16* const out[2] = {o0, o1};
    tc_store_planes16<PREC>(x, out, row * ld_o + c0);
}

// du = augmented_f(src) on the tensor path: TrainMode (Hutchinson rows) or, with h->tc_mode = TestMode, the
// exact trace (reference src/icnf.jl:148-164: zdot rows and -tr J, J from jacobian_batched src/utils.jl:1-36).  src: D x Bs z rows (host-known pointer);
// output buffer either fixed (kout_fixed) or resolved on the device from kidx (FSAL index).
int tc_eval(cnf_handle* h, const CnfBufs& bf, const float* src, int kidx, float* kout_fixed, int check_done) {
    TcState* s = (TcState*)h->tc_state;
    if (!s || !s->weights_ok) { h->err = "tensor path: parameters not set"; return CNF_ERR_UNSUPPORTED; }
    const cnf_config& c = h->cfg;
    int rc;
    if (fused_supported(h, h->tc_mode)) return fused_eval(h, bf, src, kidx, kout_fixed, check_done);
    if ((rc = tc_ensure_acts(h, s, bf.Bs))) return rc;
    const int L = c.n_layers;
    const bool test = h->tc_mode == CNF_MODE_TEST;
    const bool jvp = c.trace_mode == CNF_TRACE_JVP || test;   // the exact trace pushes one-hot tangents forward
    const int npass = test ? c.D : 1;
    for (long long row0 = 0; row0 < bf.Bs; row0 += s->Mc) {
        const int mc = (int)((bf.Bs - row0) < s->Mc ? (bf.Bs - row0) : s->Mc);
        TcEpi base = tc_base_epi(h, s, bf, row0, check_done);
        base.test = test ? 1 : 0;
        // h_0 = [z; ys] planes (+ resolve the output buffer)
        k_tc_prep<0><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, src, c.D, bf.ys, c.C, row0, mc, s->H[0].p[0],
                                                          s->H[0].p[1], s->H[0].ld, kidx, kout_fixed, s->dyn, check_done);
        h->launches++;
        tc_forward_hidden<0>(h, s, base, mc);
        {
            TcEpi ep = base;
            ep.bias = h->d_theta + h->b_off[L - 1];
            ep.N = c.dims[L];
            ep.kind = TC_FWD_LAST;
            ep.vjp = jvp ? 0 : 1;
            for (int q = 0; q < 2; ++q) ep.out[q] = s->G[L].p[q];
            ep.ld_out = s->G[L].ld;
            tc_launch_gemm<0, false>(h, s, s->H[L - 1], s->wf[L - 1], mc, ep);
        }
        if (!jvp) {
            // reverse chain: g_{l} = (g_{l+1} W_{l+1}) .* (1 - h_l^2), v = (g_1 W_1)[0:D]
            for (int l = L - 1; l >= 0; --l) {
                TcEpi ep = base;
                if (l > 0) {
                    ep.kind = TC_BWD_HIDDEN;
                    ep.N = c.dims[l];
                    for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l].p[q]; ep.out[q] = s->G[l].p[q]; }
                    ep.aux_p = &s->H[l];
                    ep.ld_h = s->H[l].ld; ep.ld_out = s->G[l].ld;
                } else {
                    ep.kind = TC_BWD_LAST;
                    ep.N = c.D;
                }
                tc_launch_gemm<0, false>(h, s, s->G[l + 1], s->wb[l], mc, ep);
            }
        } else for (int pass = 0; pass < npass; ++pass) {
            // tangent chain: tau_0 = eps, tau_l = (tau_{l-1} W_l^T) .* (1 - h_l^2), J eps = tau_L
            // (TestMode: eps = e_pass, D passes over the same forward activations)
            if (test) {
                base.pass = pass;
                const int N1 = c.dims[1];
                const long long nthr = (long long)mc * (N1 / TC_CW);
                k_tc_onehot_l1<0><<<(unsigned)((nthr + 255) / 256), 256, 0, h->stream>>>(
                    check_done ? h->d_ctrl : nullptr, h->d_theta + h->w_off[0] + (long long)pass * N1, N1, s->H[1].p[0],
                    s->H[1].p[1], s->H[1].ld, s->G[1].p[0], s->G[1].p[1], s->G[1].ld, mc);
                h->launches++;
            } else {
            k_tc_prep<0><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, bf.eps, c.D, nullptr, 0, row0, mc,
                                                              s->G[0].p[0], s->G[0].p[1], s->G[0].ld, 0,
                                                              nullptr, nullptr,
                                                              check_done);
            h->launches++;
            }
            for (int l = test ? 1 : 0; l < L; ++l) {   // (TestMode: layer 1 of the chain was the elementwise kernel)
                TcEpi ep = base;
                ep.N = c.dims[l + 1];
                if (l < L - 1) {
                    ep.kind = TC_BWD_HIDDEN;
                    for (int q = 0; q < 2; ++q) { ep.hp[q] = s->H[l + 1].p[q]; ep.out[q] = s->G[l + 1].p[q]; }
                    ep.aux_p = &s->H[l + 1];
                    ep.ld_h = s->H[l + 1].ld; ep.ld_out = s->G[l + 1].ld;
                } else {
                    ep.kind = TC_JVP_LAST;
                }
                // the seed planes G[0] = [eps; 0] have the first layer's padded K width, so the
                // condition columns of W_1 meet zeros (Jacobian w.r.t. z only)
                tc_launch_gemm<0, false>(h, s, l == 0 ? s->G[0] : s->G[l], s->wf[l], mc, ep);
            }
        }
    }
    return CNF_OK;
}

// zdot rows only (forward chain, BF16x3): the stage-checkpoint recomputation of the reverse sweep
// needs k_i[0:D] to rebuild the stage inputs, not the trace / regulariser rows.
int tc_eval_z(cnf_handle* h, const CnfBufs& bf, const float* src, float* kout) {
    TcState* s = (TcState*)h->tc_state;
    if (!s || !s->weights_ok) { h->err = "tensor path: parameters not set"; return CNF_ERR_UNSUPPORTED; }
    const cnf_config& c = h->cfg;
    int rc;
    if ((rc = tc_ensure_acts(h, s, bf.Bs))) return rc;
    const int L = c.n_layers;
    for (long long row0 = 0; row0 < bf.Bs; row0 += s->Mc) {
        const int mc = (int)((bf.Bs - row0) < s->Mc ? (bf.Bs - row0) : s->Mc);
        TcEpi base = tc_base_epi(h, s, bf, row0, 0);
        k_tc_prep<0><<<(mc + 255) / 256, 256, 0, h->stream>>>(h->d_ctrl, bf, src, c.D, bf.ys, c.C, row0, mc, s->H[0].p[0],
                                                          s->H[0].p[1], s->H[0].ld, -1, kout, s->dyn, 0);
        h->launches++;
        tc_forward_hidden<0>(h, s, base, mc);
        TcEpi ep = base;
        ep.bias = h->d_theta + h->b_off[L - 1];
        ep.N = c.dims[L];
        ep.kind = TC_FWD_LAST;
        ep.vjp = 2;
        tc_launch_gemm<0, false>(h, s, s->H[L - 1], s->wf[L - 1], mc, ep);
    }
    return CNF_OK;
}
