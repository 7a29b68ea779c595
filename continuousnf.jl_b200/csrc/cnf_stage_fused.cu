// cnf_stage_fused.cu — fused multi-layer tensor-core evaluation for hidden width 128 (C4-class nets).
//
// One augmented_f evaluation in TrainMode / VecJac (reference src/icnf.jl:318-350: zdot = nn(z), eps^T J by
// value_and_pullback, ldot = -(eps^T J) . eps, Edot = |zdot|, ndot = |eps^T J|) as ONE kernel: a persistent CTA owns
// 128-sample tiles and chains all seven layer products of a tile on chip.  Only the z rows, the condition rows and
// the eps rows are read from HBM and only the R output rows are written ((2R + D) * 4 bytes per sample, SURVEY.md
// 8.4); the per-layer path (cnf_stage_tc.cu) moves every activation through HBM as hi/lo planes, 50 x as much.
//
//  * weights: the fp16 hi/lo planes of W1, W2, W3 stay resident in shared memory for the whole launch (TMA,
//    SWIZZLE_128B, the K-major image the forward products read).  The reverse products q = g W read THE SAME image
//    as an MN-major operand (an 8-row swizzle atom of the K-major image is an 8 x 64 atom of the MN-major canonical
//    layout), so no transposed copy exists.
//  * activations: the accumulator of layer l (TMEM, one lane per sample) is turned IN PLACE into the A operand of the
//    next product: the owner thread reads 16 fp32 columns (tcgen05.ld), applies bias + tanh (or the (1 - h^2)
//    Hadamard factor), splits into fp16 hi/lo pairs and stores 8 + 8 packed columns back over the same 16 columns
//    (tcgen05.st); the next tcgen05.mma takes A from tensor memory.  Two regions of 128 columns per tile ping-pong;
//    two tiles are in flight per SM (4 x 128 = 512 columns), so one tile's elementwise phases run under the other
//    tile's MMAs.  The two narrow operands (h_0 = [z; ys] and g_3 = eps (1 - y^2)) go through shared memory in the
//    no-swizzle core-matrix layout.
//  * VJP needs (1 - h_1^2) last, long after h_1's columns were recycled: layer 1 (K = d0 <= 64, 3-12 MMAs) is
//    recomputed instead of kept.
//  * precision: x = hi + lo in IEEE half, three MMAs per product (hi*hi + lo*hi + hi*lo), FP32 accumulate — the same
//    scheme as the per-layer path (DESIGN.md "Precision").
//
// Phases of a tile (W = its 8 worker warps, I = the single MMA-issuer thread; X, Y = its two TMEM regions):
//   W P0  h_0 -> smem                         I M1  Y = h_0 W1^T
//   W E1  Y: tanh(. + b1) -> planes           I M2  X = h_1 W2^T
//   W E2  X: tanh(. + b2) -> planes           I M3  Y[0:N3] = h_2 W3^T
//   W E3  y = tanh(. + b3) -> zdot rows, |y|; g_3 -> smem          I M4  Y = g_3 W3
//   W E4  Y: . (1 - h_2^2) -> planes (h_2 read back from X)        I M5  X = g_2 W2 ; M6  Y = h_0 W1^T (again)
//   W E7  X: . (1 - tanh(Y + b1)^2) -> planes                      I M8  Y[0:N3] = g_1 W1[:, 0:D]
//   W E9  v -> -eps . v, |v| rows
#include "cnf_internal.h"
#include "cnf_tc_ptx.cuh"

#define FU_WORKERS 8                      // warps per tile slot: two per TMEM lane quadrant (column halves)
#define FU_SLOTS 2
#define FU_THREADS (FU_SLOTS * FU_WORKERS * 32)   // 16 warps: with a 17th the register file rounds to 20 warps and caps a thread at 96 registers
#define FU_NPHASE 6
#define FU_NARROW 32                      // d0 and D are at most this (register-staged stage-input and eps rows)
#define FU_ISSUER 4                       // warp of the slot whose lane 0 issues the slot's MMAs (quadrant 0, column half 1)
#define FU_RCP_MUFU_EVERY 4               // of every 4 column pairs one takes its reciprocal on the MUFU pipe, three on the FMA pipe (balances the two)
#define FU_C 2.8853900817779268f          // 2 log2(e): tanh(x) = 1 - 2 / (exp2(C x) + 1)

namespace tc {
// registers -> TMEM: 32 lanes x 16 columns
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        :
        : "r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n"
        :
        : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// K-major operand without swizzle: core matrices of 8 rows x 16 bytes (128 contiguous bytes); SBO = byte stride
// between 8-row groups, LBO = byte stride between the two 16-byte K chunks of one K = 16 step
__device__ __forceinline__ uint64_t umma_desc_nosw(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
// kind::f16 with IEEE half A/B, FP32 accumulate; b_mn = 1: B is MN-major
__device__ __forceinline__ uint32_t umma_idesc_h(int M, int N, int b_mn) {
    return (1u << 4) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
}  // namespace tc

struct FusedMaps { CUtensorMap w[3][2]; };   // [layer][hi/lo] forward weight planes (IEEE half)

struct FusedArgs {
    const CnfCtrl* ctrl;
    CnfBufs bf;
    const float* src;        // D x Bs stage input (z rows)
    float* kout_fixed;
    const float *b1, *b2, *b3;
    int kidx, check_done;
    int D, C, K0, N3;        // K0 = pad16(D + C), N3 = pad16(D)
    int ntiles;
    unsigned flags;
    long long* trace;        // debug: clock64 stamps of CTA 0 (option tc_dbg & 16), [slot][tile < 8][16]
    int dbg;                 // timing experiments (option tc_dbg): 4 = no MMAs are issued (results are wrong)
};

struct FusedCtl {
    uint64_t wbar, ready[FU_SLOTS], done[FU_SLOTS], turn[FU_SLOTS][2];   // turn: two barriers per direction, used alternately
    uint32_t tmem_base;
};

template <int H>
struct FusedLayout {
    static constexpr int KB = H / 64;   // K blocks of the hidden width
    __host__ __device__ static int w1() { return 0; }
    __host__ __device__ static int w2() { return 2 * H * 128; }
    __host__ __device__ static int w3() { return w2() + 2 * KB * H * 128; }
    __host__ __device__ static int slots(int N3) { return w3() + 2 * KB * N3 * 128; }
    __host__ __device__ static int slot_bytes(int K0, int N3) { return 2 * K0 * 256 + 2 * N3 * 256; }
    __host__ __device__ static int bias(int K0, int N3) { return slots(N3) + FU_SLOTS * slot_bytes(K0, N3); }
    __host__ __device__ static int ctl(int K0, int N3) { return bias(K0, N3) + (2 * H + 64) * 4; }
    __host__ __device__ static int total(int K0, int N3) { return ctl(K0, N3) + (int)sizeof(FusedCtl) + 1024; }
};

// ---- packed-pair arithmetic of the elementwise phases (FFMA2 / FMUL2 / FADD2: one issue slot for two columns)
__device__ __forceinline__ float2 fu_f2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 fu_ex2(float2 x) {
    float2 t;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(x.x));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(x.y));
    return t;
}
__device__ __forceinline__ float2 fu_rcp(float2 x) {
    float2 t;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(x.x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(x.y));
    return t;
}
// reciprocal of d >= 1 on the FMA pipe: the MUFU pipe (16 lanes per clock per SM) is what bounds the tanh phases when
// both ex2 and rcp run there.  Magic-constant seed (relative error < 12 %), three Newton steps r <- r (2 - d r):
// 12 % -> 1.5 % -> 2.2e-4 -> 5e-8.  For d = inf (x > 44) the seed is ~0 and stays 0: tanh saturates at 1 exactly.
__device__ __forceinline__ float2 fu_rcp_fma(float2 d) {
    float2 r = make_float2(__uint_as_float(0x7EF311C7u - __float_as_uint(d.x)), __uint_as_float(0x7EF311C7u - __float_as_uint(d.y)));
    const float2 nd = ntc_mul2(d, fu_f2(-1.0f));
#pragma unroll
    for (int it = 0; it < 3; ++it) r = ntc_mul2(r, ntc_fma2(nd, r, fu_f2(2.0f)));
    return r;
}
// r = 1 / (exp2(C (a + b)) + 1) with cb = C b: tanh(a + b) = 1 - 2 r and 1 - tanh^2 = 4 r (1 - r).
// FMA = true: the reciprocal on the FMA pipe (fu_rcp_fma), else on the MUFU pipe
template <bool FMA>
__device__ __forceinline__ float2 fu_sig(float2 a, float2 cb) {
    float2 d = ntc_add2(fu_ex2(ntc_fma2(a, fu_f2(FU_C), cb)), fu_f2(1.0f));
    if (FMA) {
        d.x = fminf(d.x, 1.0e30f);      // inf - inf in the Newton step otherwise; 1e30 still gives r = 1e-30 -> tanh = 1
        d.y = fminf(d.y, 1.0e30f);
        return fu_rcp_fma(d);
    }
    return fu_rcp(d);
}
// (x0, x1) -> packed half hi pair and lo pair
__device__ __forceinline__ void fu_split2(float2 x, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(x.x, x.y);
    const float2 f = __half22float2(h);
    const float2 d = ntc_fma2(f, fu_f2(-1.0f), x);
    const __half2 l = __floats2half2_rn(d.x, d.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
}
__device__ __forceinline__ float2 fu_unpack2(uint32_t hi, uint32_t lo) {
    return ntc_add2(__half22float2(*reinterpret_cast<const __half2*>(&hi)), __half22float2(*reinterpret_cast<const __half2*>(&lo)));
}

// One wide elementwise phase of a worker thread: its columns [cbeg, cbeg + H/2) of region A, 16 at a time, in place:
//   KIND 0: tanh(A + bias)                          (E1, E2)      cbias = C * bias
//   KIND 1: A .* (1 - h^2), h = hi/lo planes in B   (E4)
//   KIND 2: A .* (1 - tanh(B + bias)^2)             (E7)
// The tcgen05.ld of block k + 1 is issued before block k is processed (tcgen05.wait::ld waits for every outstanding
// load of the thread, so the prefetch goes out right after the wait).
template <int KIND, int H>
__device__ __forceinline__ void fu_phase(uint32_t A, uint32_t B, int cbeg, const float* cbias) {
    constexpr int NB = H / 2 / 16;
    uint32_t a[2][16], b[2][16];
    tc::tmem_ld_32x16(A + cbeg, a[0]);
    if (KIND != 0) tc::tmem_ld_32x16(B + cbeg, b[0]);
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const int c0 = cbeg + 16 * k, cur = k & 1;
        tc::tmem_ld_wait();
        if (k + 1 < NB) {
            tc::tmem_ld_32x16(A + c0 + 16, a[cur ^ 1]);
            if (KIND != 0) tc::tmem_ld_32x16(B + c0 + 16, b[cur ^ 1]);
        }
        uint32_t o[16];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float2 av = make_float2(__uint_as_float(a[cur][2 * i]), __uint_as_float(a[cur][2 * i + 1]));
            float2 x;
            if (KIND == 0) {
                const float2 cb = *reinterpret_cast<const float2*>(cbias + c0 + 2 * i);
                const float2 r = (i % FU_RCP_MUFU_EVERY) ? fu_sig<true>(av, cb) : fu_sig<false>(av, cb);
                x = ntc_fma2(r, fu_f2(-2.0f), fu_f2(1.0f));
            } else if (KIND == 1) {
                const float2 h = fu_unpack2(b[cur][i], b[cur][8 + i]);
                x = ntc_mul2(av, ntc_fma2(ntc_mul2(h, h), fu_f2(-1.0f), fu_f2(1.0f)));
            } else {
                const float2 bv = make_float2(__uint_as_float(b[cur][2 * i]), __uint_as_float(b[cur][2 * i + 1]));
                const float2 cb = *reinterpret_cast<const float2*>(cbias + c0 + 2 * i);
                const float2 r = (i % FU_RCP_MUFU_EVERY) ? fu_sig<true>(bv, cb) : fu_sig<false>(bv, cb);
                x = ntc_mul2(ntc_mul2(ntc_mul2(av, fu_f2(4.0f)), r), ntc_fma2(r, fu_f2(-1.0f), fu_f2(1.0f)));
            }
            fu_split2(x, o[i], o[8 + i]);
        }
        tc::tmem_st_32x16(A + c0, o);
    }
}

// 3 MMAs of one K = 16 step with A in tensor memory (hi at acol, lo at acol + 8)
__device__ __forceinline__ void fu_mma3_ts(uint32_t d, uint32_t acol, uint64_t bhi, uint64_t blo, uint32_t idesc, uint32_t acc) {
    tc::umma_f16_ts(d, acol, bhi, idesc, acc);
    tc::umma_f16_ts(d, acol + 8, bhi, idesc, 1);
    tc::umma_f16_ts(d, acol, blo, idesc, 1);
}
__device__ __forceinline__ void fu_mma3_ss(uint32_t d, uint64_t ahi, uint64_t alo, uint64_t bhi, uint64_t blo, uint32_t idesc,
                                           uint32_t acc) {
    tc::umma_bf16(d, ahi, bhi, idesc, acc);
    tc::umma_bf16(d, alo, bhi, idesc, 1);
    tc::umma_bf16(d, ahi, blo, idesc, 1);
}

// the MMAs of phase p of one tile slot (X, Y: its TMEM regions; a0, g3: its narrow shared-memory operands)
template <int H>
__device__ __forceinline__ void fu_issue(int p, uint32_t X, uint32_t Y, const uint32_t (&a0)[2], const uint32_t (&g3)[2],
                                         const uint32_t (&w1p)[2], const uint32_t (&w2p)[2], const uint32_t (&w3p)[2], int K0, int N3) {
    const uint32_t idH = tc::umma_idesc_h(128, H, 0), idHm = tc::umma_idesc_h(128, H, 1);
    const uint32_t id3 = tc::umma_idesc_h(128, N3, 0), id3m = tc::umma_idesc_h(128, N3, 1);
    if (p == 0 || p == 4) {
        if (p == 4) {   // M5: X = g_2 W2 (MN-major view of the resident W2 image)
#pragma unroll
            for (int kk = 0; kk < H / 16; ++kk)
                fu_mma3_ts(X, Y + 16 * kk, tc::umma_desc_mn_sw128(w2p[0] + kk * 2048, H * 128, 1024),
                           tc::umma_desc_mn_sw128(w2p[1] + kk * 2048, H * 128, 1024), idHm, kk != 0);
        }
        // M1 / M6: Y = h_0 W1^T
        for (int kk = 0; kk < K0 / 16; ++kk)
            fu_mma3_ss(Y, tc::umma_desc_nosw(a0[0] + kk * 4096, 2048, 128), tc::umma_desc_nosw(a0[1] + kk * 4096, 2048, 128),
                       tc::umma_desc_sw128(w1p[0] + kk * 32), tc::umma_desc_sw128(w1p[1] + kk * 32), idH, kk != 0);
    } else if (p == 1) {   // M2: X = h_1 W2^T
#pragma unroll
        for (int kk = 0; kk < H / 16; ++kk) {
            const uint32_t off = (uint32_t)(kk >> 2) * (H * 128) + (uint32_t)(kk & 3) * 32;
            fu_mma3_ts(X, Y + 16 * kk, tc::umma_desc_sw128(w2p[0] + off), tc::umma_desc_sw128(w2p[1] + off), idH, kk != 0);
        }
    } else if (p == 2) {   // M3: Y[0:N3] = h_2 W3^T
#pragma unroll
        for (int kk = 0; kk < H / 16; ++kk) {
            const uint32_t off = (uint32_t)(kk >> 2) * (N3 * 128) + (uint32_t)(kk & 3) * 32;
            fu_mma3_ts(Y, X + 16 * kk, tc::umma_desc_sw128(w3p[0] + off), tc::umma_desc_sw128(w3p[1] + off), id3, kk != 0);
        }
    } else if (p == 3) {   // M4: Y = g_3 W3 (MN-major view of W3: N' = hidden, K' = output rows)
        for (int kk = 0; kk < N3 / 16; ++kk)
            fu_mma3_ss(Y, tc::umma_desc_nosw(g3[0] + kk * 4096, 2048, 128), tc::umma_desc_nosw(g3[1] + kk * 4096, 2048, 128),
                       tc::umma_desc_mn_sw128(w3p[0] + kk * 2048, N3 * 128, 1024),
                       tc::umma_desc_mn_sw128(w3p[1] + kk * 2048, N3 * 128, 1024), idHm, kk != 0);
    } else {               // M8: Y[0:N3] = g_1 W1[:, 0:D]
#pragma unroll
        for (int kk = 0; kk < H / 16; ++kk)
            fu_mma3_ts(Y, X + 16 * kk, tc::umma_desc_mn_sw128(w1p[0] + kk * 2048, H * 128, 1024),
                       tc::umma_desc_mn_sw128(w1p[1] + kk * 2048, H * 128, 1024), id3m, kk != 0);
    }
}

template <int H>
__global__ void __launch_bounds__(FU_THREADS, 1) k_tc_fused(const __grid_constant__ FusedMaps mp, FusedArgs ar) {
    if (ar.check_done && ar.ctrl->done) return;
    using LY = FusedLayout<H>;
    constexpr int KB = LY::KB;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int K0 = ar.K0, N3 = ar.N3, D = ar.D;
    const uint32_t sb = tc::smem_u32(smem);
    const uint32_t w1p[2] = {sb + LY::w1(), sb + LY::w1() + H * 128};
    const uint32_t w2p[2] = {sb + LY::w2(), sb + LY::w2() + KB * H * 128};
    const uint32_t w3p[2] = {sb + LY::w3(), sb + LY::w3() + KB * N3 * 128};
    float* sbias = reinterpret_cast<float*>(smem + LY::bias(K0, N3));   // C b1[H], C b2[H], b3[64]
    FusedCtl* ctl = reinterpret_cast<FusedCtl*>(smem + LY::ctl(K0, N3));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long Bs = ar.bf.Bs;

    float* kout = ar.kout_fixed;
    if (!kout) {
        const int k1 = ar.ctrl->k1idx;
        kout = (ar.kidx == 6) ? ar.bf.K[6 - k1] : (ar.kidx == 0 ? ar.bf.K[k1] : ar.bf.K[ar.kidx]);
    }

    if (threadIdx.x == 0) {
        for (int l = 0; l < 3; ++l)
            for (int q = 0; q < 2; ++q) tc::tma_prefetch_desc(&mp.w[l][q]);
        tc::mbar_init(&ctl->wbar, 1);
        for (int s = 0; s < FU_SLOTS; ++s) { tc::mbar_init(&ctl->ready[s], FU_WORKERS); tc::mbar_init(&ctl->done[s], 1); tc::mbar_init(&ctl->turn[s][0], 1); tc::mbar_init(&ctl->turn[s][1], 1); }
        tc::fence_barrier_init();
    }
    if (warp == 0) tc::tmem_alloc(&ctl->tmem_base, 512);
    for (int i = threadIdx.x; i < 2 * H + 64; i += blockDim.x) {
        float v = 0.0f;
        if (i < H) v = FU_C * ar.b1[i];
        else if (i < 2 * H) v = FU_C * ar.b2[i - H];
        else if (i - 2 * H < D) v = ar.b3[i - 2 * H];
        sbias[i] = v;
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem = ctl->tmem_base;
    const int my_tiles = (ar.ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;   // tiles of this CTA
    if (threadIdx.x == 0) {   // resident weights: one TMA box per plane and K block
        const uint32_t wbytes = 2u * (H * 128 + KB * H * 128 + KB * N3 * 128);
        tc::mbar_arrive_expect_tx(&ctl->wbar, wbytes);
        for (int q = 0; q < 2; ++q) {
            tc::tma_load_2d(&mp.w[0][q], &ctl->wbar, smem + LY::w1() + q * H * 128, 0, 0);
            for (int kb = 0; kb < KB; ++kb) {
                tc::tma_load_2d(&mp.w[1][q], &ctl->wbar, smem + LY::w2() + q * KB * H * 128 + kb * H * 128, kb * 64, 0);
                tc::tma_load_2d(&mp.w[2][q], &ctl->wbar, smem + LY::w3() + q * KB * N3 * 128 + kb * N3 * 128, kb * 64, 0);
            }
        }
    }

    // ============================================================== worker warps (all 16)
    // Column halves: warps with hf = 0 own columns [0, H/2) of the wide phases and the stage-input work (P0; they
    // prefetch the NEXT tile's z / ys rows into registers while the hf = 1 warps run E3); warps with hf = 1 own columns
    // [H/2, H) and the narrow output phases E3 / E9 (their eps rows are loaded into registers while the hf = 0 warps
    // run P0).  No global-memory latency sits on the tile's dependent chain.  Lane 0 of warp FU_ISSUER of the slot
    // issues the slot's MMAs once the slot's eight warps have arrived.
    const int s = warp / FU_WORKERS;
    const int wl = warp % FU_WORKERS;
    const int quad = warp & 3;                  // TMEM lane quadrant of this warp
    const int hf = (warp >> 2) & 1;             // column half
    const int row = quad * 32 + lane;
    const uint32_t Xi = tmem + (uint32_t)s * (2 * H), Yi = Xi + H;                 // issuer view (lane field 0)
    const uint32_t X = Xi + ((uint32_t)(quad * 32) << 16), Y = X + H;             // this warp's lanes
    const uint32_t a0[2] = {sb + LY::slots(N3) + s * LY::slot_bytes(K0, N3), sb + LY::slots(N3) + s * LY::slot_bytes(K0, N3) + K0 * 256};
    const uint32_t g3[2] = {a0[0] + 2 * K0 * 256, a0[0] + 2 * K0 * 256 + N3 * 256};
    const float* sb1 = sbias;
    const float* sb2 = sbias + H;
    const float* sb3 = sbias + 2 * H;
    uint32_t dpar = 0, rpar = 0;
    const int cbeg = hf * (H / 2);
    if (wl == FU_ISSUER) tc::mbar_wait(&ctl->wbar, 0, 0x740);
    // hand the slot's operands to the tensor core: arrive; the issuer warp waits for all eight warps, issues phase p
    // and commits; everybody then waits for the MMAs
#define FU_STAMP(k_)                                                                              \
    if (ar.trace && blockIdx.x == 0 && wl == 0 && lane == 0 && (j / FU_SLOTS) < 8)                \
        ar.trace[(s * 8 + j / FU_SLOTS) * 16 + (k_)] = clock64();
#define FU_HANDOFF(p_)                                                                            \
    {                                                                                             \
        FU_STAMP(2 * (p_))                                                                        \
        FU_BATON_PASS()                                                                           \
        __syncwarp();                                                                             \
        if (lane == 0) tc::mbar_arrive(&ctl->ready[s]);                                           \
        if (wl == FU_ISSUER) {                                                                    \
            tc::mbar_wait(&ctl->ready[s], rpar, 0x720 + (p_));                                    \
            rpar ^= 1;                                                                            \
            tc::tc_fence_after();                                                                 \
            if (lane == 0) {                                                                      \
                if (!(ar.dbg & 4)) fu_issue<H>(p_, Xi, Yi, a0, g3, w1p, w2p, w3p, K0, N3);        \
                tc::umma_commit(&ctl->done[s]);                                                   \
            }                                                                                     \
            __syncwarp();                                                                         \
        }                                                                                         \
        tc::mbar_wait(&ctl->done[s], dpar, 0x710 + (p_));                                                      \
        dpar ^= 1;                                                                                \
        tc::tc_fence_after();                                                                     \
        if (!((p_) == 5 && j + FU_SLOTS >= my_tiles)) FU_BATON_WAIT()                             \
        FU_STAMP(2 * (p_) + 1)                                                                    \
    }
#define FU_SAMPLE_OF(j_) (((long long)blockIdx.x + (long long)(j_) * gridDim.x) * 128 + row)
// stage-input rows [z; ys] of sample g_ into nreg[0 .. K0): 8 at a time, only the chunks the net has
#define FU_LOAD_Z(g_)                                                                                         \
    _Pragma("unroll") for (int kc_ = 0; kc_ < FU_NARROW / 8; ++kc_) {                                         \
        if (kc_ * 8 < K0) {                                                                                   \
            _Pragma("unroll") for (int i_ = 0; i_ < 8; ++i_) {                                                \
                const int c_ = kc_ * 8 + i_;                                                                  \
                float v_ = 0.0f;                                                                              \
                if (c_ < D) v_ = ar.src[(long long)c_ * Bs + (g_)];                                           \
                else if (c_ < D + ar.C) v_ = __ldg(ar.bf.ys + (long long)(c_ - D) * Bs + (g_));               \
                nreg[c_] = v_;                                                                                \
            }                                                                                                 \
        }                                                                                                     \
    }
    float nreg[FU_NARROW];     // one array, two roles: hf = 0: stage input rows of the next tile; hf = 1: eps rows of this tile
#pragma unroll
    for (int c = 0; c < FU_NARROW; ++c) nreg[c] = 0.0f;
    if (hf == 0 && s < my_tiles) {
        const long long g0 = FU_SAMPLE_OF(s);
        FU_LOAD_Z(g0)
    }
    // The two slots take turns on the SIMT side ("baton"): a slot runs an elementwise section only while the other slot
    // waits for its MMAs, so the MMAs + hand-off latency of one tile hide under the other tile's elementwise work and the
    // two tiles do not fight for issue slots and the MUFU pipe at the same time (left alone the slots run in lockstep:
    // measured, both in their elementwise phase together, then both waiting for the tensor pipe).  Section k of slot 0
    // follows section k - 1 of slot 1; section k of slot 1 follows section k of slot 0.  A section = the work between
    // two hand-offs (E9 of a tile and P0 of the next one are one section); 6 per tile.
    // Pass k of a slot goes to barrier k & 1 of the other slot's pair and is awaited with parity (k >> 1) & 1.  With ONE
    // barrier per direction the scheme deadlocked about once in 10^4 launches: warp 0 passes the baton on its own, at
    // the start of its hand-off, and in the sections where the hf = 0 warps have no work the two slots' warp 0 can
    // complete two phases of that barrier within a few hundred cycles; a sibling warp still on its way to its wait
    // (cold instruction cache: the first launches after another model's kernels) then tests a parity that has already
    // come round again — a parity wait cannot tell "two phases ahead" from "not yet" — and waits for ever (found through
    // the trap record, tc::trap_with_record, site 0x730).  With two barriers used alternately a barrier completes a
    // second time only after passes k + 1 and k + 2 of BOTH slots, and pass k + 2 of this slot needs hand-off k + 1 of
    // this slot, at which all eight of its warps have arrived, i.e. have left the wait for pass k.
    const bool baton = (ar.dbg & 8) == 0;
    const int n_other6 = 6 * ((my_tiles - (1 - s) + 1) / 2);   // sections of the other slot
    uint32_t ksec = 0;
#define FU_BATON_WAIT()                                                                              \
    if (baton) {                                                                                     \
        if (s == 0) { if (ksec >= 1 && (int)ksec - 1 < n_other6) tc::mbar_wait(&ctl->turn[0][(ksec - 1) & 1], ((ksec - 1) >> 1) & 1, 0x730); } \
        else tc::mbar_wait(&ctl->turn[1][ksec & 1], (ksec >> 1) & 1, 0x731);                          \
    }
#define FU_BATON_PASS()                                                                              \
    {                                                                                                \
        if (baton && wl == 0 && lane == 0) tc::mbar_arrive(&ctl->turn[1 - s][ksec & 1]);             \
        ++ksec;                                                                                      \
    }
    if (s < my_tiles) FU_BATON_WAIT()
    for (int j = s; j < my_tiles; j += FU_SLOTS) {
        const long long g = FU_SAMPLE_OF(j);
        FU_STAMP(12)
        // ---- P0: h_0 = [z; ys] -> shared memory (core-matrix layout: chunk kc of row r at kc * 2048 + r * 16)
        if (hf == 0) {
#pragma unroll
            for (int kc = 0; kc < FU_NARROW / 8; ++kc) {
                if (kc * 8 < K0) {
                    uint32_t hi[4], lo[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) tc_split2<0>(nreg[kc * 8 + 2 * i], nreg[kc * 8 + 2 * i + 1], hi[i], lo[i]);
                    tc::st_shared_v4(a0[0] + kc * 2048 + row * 16, hi[0], hi[1], hi[2], hi[3]);
                    tc::st_shared_v4(a0[1] + kc * 2048 + row * 16, lo[0], lo[1], lo[2], lo[3]);
                }
            }
            tc::fence_proxy_async();
        } else {
#pragma unroll
            for (int kc = 0; kc < FU_NARROW / 8; ++kc) {
                if (kc * 8 < D) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) nreg[kc * 8 + i] = (kc * 8 + i < D) ? __ldg(ar.bf.eps + (long long)(kc * 8 + i) * Bs + g) : 0.0f;
                }
            }
        }
        tc::tc_fence_before();   // this thread's tcgen05.ld of the previous tile precede the MMAs that recycle its columns
        FU_HANDOFF(0)
        // ---- E1: Y: h_1 = tanh(acc + b1) -> A planes in place
        fu_phase<0, H>(Y, 0, cbeg, sb1);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        FU_HANDOFF(1)
        // ---- E2: X: h_2 = tanh(acc + b2) -> A planes in place
        fu_phase<0, H>(X, 0, cbeg, sb2);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        FU_HANDOFF(2)
        // ---- E3: y = tanh(acc3 + b3): zdot rows, |y|; g_3 = eps (1 - y^2) -> shared memory
        if (hf == 1) {
            float s0 = 0.0f;
#pragma unroll
            for (int cc = 0; cc < FU_NARROW / 16; ++cc) {
                const int c0 = cc * 16;
                if (c0 < N3) {
                    uint32_t v[16];
                    tc::tmem_ld_32x16(Y + c0, v);
                    tc::tmem_ld_wait();
                    float gq[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        gq[i] = 0.0f;
                        if (c0 + i < D) {
                            const float y = tc_tanh(__uint_as_float(v[i]) + sb3[c0 + i]);
                            kout[(long long)(c0 + i) * Bs + g] = y;
                            s0 = fmaf(y, y, s0);
                            gq[i] = nreg[c0 + i] * fmaf(-y, y, 1.0f);
                        }
                    }
#pragma unroll
                    for (int kc = 0; kc < 2; ++kc) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) tc_split2<0>(gq[kc * 8 + 2 * i], gq[kc * 8 + 2 * i + 1], hi[i], lo[i]);
                        tc::st_shared_v4(g3[0] + (c0 / 8 + kc) * 2048 + row * 16, hi[0], hi[1], hi[2], hi[3]);
                        tc::st_shared_v4(g3[1] + (c0 / 8 + kc) * 2048 + row * 16, lo[0], lo[1], lo[2], lo[3]);
                    }
                }
            }
            kout[(long long)(D + 1) * Bs + g] = (ar.flags & CNF_FLAG_NORM_Z) ? sqrtf(s0) : 0.0f;
            tc::fence_proxy_async();
        } else if (j + FU_SLOTS < my_tiles) {
            const long long gn = FU_SAMPLE_OF(j + FU_SLOTS);     // in flight under E3 .. E9 of this tile
            FU_LOAD_Z(gn)
        }
        tc::tc_fence_before();
        FU_HANDOFF(3)
        // ---- E4: Y: g_2 = q_3 (1 - h_2^2) -> A planes in place (h_2 planes are still in X)
        fu_phase<1, H>(Y, X, cbeg, nullptr);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        FU_HANDOFF(4)
        // ---- E7: X: g_1 = q_2 (1 - h_1^2), h_1 = tanh(Y + b1) recomputed
        fu_phase<2, H>(X, Y, cbeg, sb1);
        tc::tmem_st_wait();
        tc::tc_fence_before();
        FU_HANDOFF(5)
        // ---- E9: v = eps^T J: ldot = -v . eps, ndot = |v|
        if (hf == 1) {
            float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
            for (int cc = 0; cc < FU_NARROW / 16; ++cc) {
                const int c0 = cc * 16;
                if (c0 < N3) {
                    uint32_t v[16];
                    tc::tmem_ld_32x16(Y + c0, v);
                    tc::tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (c0 + i < D) {
                            const float vv = __uint_as_float(v[i]);
                            s0 = fmaf(-vv, nreg[c0 + i], s0);
                            s1 = fmaf(vv, vv, s1);
                        }
                }
            }
            kout[(long long)D * Bs + g] = s0;
            kout[(long long)(D + 2) * Bs + g] = (ar.flags & CNF_FLAG_NORM_J) ? sqrtf(s1) : 0.0f;
        }
        FU_STAMP(13)
    }
#undef FU_HANDOFF
#undef FU_STAMP
#undef FU_BATON_WAIT
#undef FU_BATON_PASS
#undef FU_SAMPLE_OF
#undef FU_LOAD_Z
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------ host side
static inline int fu_pad16(int x) { return (x + 15) / 16 * 16; }

// TrainMode / VecJac on three Dense(tanh) layers with both hidden widths 128, d0 <= 64, D <= 64
bool fused_supported(const cnf_handle* h, int mode) {
    const cnf_config& c = h->cfg;
    if (h->fused_off || c.kernel_path == 3 || c.kernel_path == 1) return false;
    if (mode != CNF_MODE_TRAIN || c.trace_mode != CNF_TRACE_VJP) return false;
    if (c.n_layers != 3 || c.dims[1] != 128 || c.dims[2] != 128) return false;
    if (c.dims[0] > FU_NARROW || c.D > FU_NARROW) return false;
    if (c.flags & CNF_FLAG_LAST_LINEAR) return false;
    return FusedLayout<128>::total(fu_pad16(c.dims[0]), fu_pad16(c.D)) <= 227 * 1024;
}

int fused_eval(cnf_handle* h, const CnfBufs& bf, const float* src, int kidx, float* kout_fixed, int check_done) {
    TcState* s = (TcState*)h->tc_state;
    if (!s || !s->weights_ok) { h->err = "fused tensor path: parameters not set"; return CNF_ERR_UNSUPPORTED; }
    const cnf_config& c = h->cfg;
    FusedMaps mp;
    for (int l = 0; l < 3; ++l)
        for (int q = 0; q < 2; ++q) mp.w[l][q] = s->wf[l].ph.map[q];
    FusedArgs ar;
    memset(&ar, 0, sizeof(ar));
    ar.ctrl = h->d_ctrl;
    ar.bf = bf;
    ar.src = src;
    ar.kout_fixed = kout_fixed;
    ar.b1 = h->d_theta + h->b_off[0];
    ar.b2 = h->d_theta + h->b_off[1];
    ar.b3 = h->d_theta + h->b_off[2];
    ar.kidx = kidx;
    ar.check_done = check_done;
    ar.D = c.D; ar.C = c.C;
    ar.K0 = fu_pad16(c.dims[0]);
    ar.N3 = fu_pad16(c.D);
    ar.ntiles = (int)(bf.Bs / 128);
    ar.flags = c.flags;
    ar.dbg = h->tc_dbg;
    ar.trace = nullptr;
    if (h->tc_dbg & 16) {
        static long long* d_trace = nullptr;
        if (!d_trace) cudaMalloc(&d_trace, 2 * 8 * 16 * sizeof(long long));
        ar.trace = d_trace;
    }
    const int smem = FusedLayout<128>::total(ar.K0, ar.N3);
    static bool attr_dev[64] = {};
    bool& attr = attr_dev[h->device & 63];
    if (!attr) {
        CNF_CUDA(cudaFuncSetAttribute(k_tc_fused<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr = true;
    }
    int grid = ar.ntiles < s->sms ? ar.ntiles : s->sms;
    k_tc_fused<128><<<grid, FU_THREADS, smem, h->stream>>>(mp, ar);
    h->launches++;
    h->path_used = 7;
    if (ar.trace) {   // debug: print the phase stamps of CTA 0 (cycles relative to the first stamp of each tile)
        long long t[2 * 8 * 16];
        cudaStreamSynchronize(h->stream);
        cudaMemcpy(t, ar.trace, sizeof(t), cudaMemcpyDeviceToHost);
        for (int sl = 0; sl < 2; ++sl)
            for (int k = 0; k < 4; ++k) {
                const long long* r = t + (sl * 8 + k) * 16;
                fprintf(stderr, "slot %d tile %d start %lld:", sl, k, r[12] - t[12]);
                const int order[14] = {12, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 13};
                for (int q = 1; q < 14; ++q) fprintf(stderr, " %lld", r[order[q]] - r[order[q - 1]]);
                fprintf(stderr, "\n");
            }
    }
    return CNF_OK;
}
