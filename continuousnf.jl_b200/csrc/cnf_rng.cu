// cnf_rng.cu — "next" row N1: the Hutchinson probe eps and the base sample z0 drawn ON the device.
//
// The reference draws both on the host once per call: `randn!(rng, ϵ)` in inference_prob /
// generate_prob (src/base_icnf.jl:277-278, 367-372) and `rand(rng, basedist, n)` (:365-366), then
// ships D x B floats to the device.  With a seed registered through cnfb200_seed_rng a NULL `eps`
// (TrainMode) or NULL `z0` (generate) means "draw it here": Philox4x32-10 (Salmon et al., SC'11)
// keyed by the seed, counter = (global sample index, row group, draw index), Box-Muller on 23-bit
// uniforms, written straight into the SoA buffers the stage kernels read.  One counter value per
// (sample, 4 rows), so a draw does not depend on how the batch is sharded over ranks
// (sample_offset = first global column of the shard).  Opt-in only: results are reproducible
// against oracle/cnf_oracle.c's restatement of the same generator, not against Julia's Xoshiro.
#include "cnf_internal.h"

#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ void philox4x32_10(uint32_t& c0, uint32_t& c1, uint32_t& c2, uint32_t& c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

// uniform in (0, 1), exactly representable: ((x >> 9) + 1/2) 2^-23
__device__ __forceinline__ float philox_u01(uint32_t x) { return ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-07f; }

// out[r * Bs + b], r < rows, b < B: standard normals.  One thread = one sample x one group of 4 rows.
__global__ void k_philox_normal(float* __restrict__ out, long long Bs, long long B, int rows, unsigned long long seed,
                                unsigned long long sample_offset, uint32_t draw) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int groups = (rows + 3) >> 2;
    if (i >= B * groups) return;
    const long long b = i % B;
    const int g = (int)(i / B);
    const unsigned long long gs = (unsigned long long)b + sample_offset;
    uint32_t c0 = (uint32_t)gs, c1 = (uint32_t)(gs >> 32), c2 = (uint32_t)g, c3 = draw;
    philox4x32_10(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
    float n[4];
    {
        const float r0 = sqrtf(-2.0f * logf(philox_u01(c0))), r1 = sqrtf(-2.0f * logf(philox_u01(c2)));
        float s0, q0, s1, q1;
        sincospif(2.0f * philox_u01(c1), &s0, &q0);
        sincospif(2.0f * philox_u01(c3), &s1, &q1);
        n[0] = r0 * q0; n[1] = r0 * s0; n[2] = r1 * q1; n[3] = r1 * s1;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
        if (4 * g + j < rows) out[(long long)(4 * g + j) * Bs + b] = n[j];
}

void launch_philox_normal(cnf_handle* h, float* out, long long Bs, long long B, int rows, uint32_t draw) {
    const long long n = B * ((rows + 3) / 4);
    const int threads = 256;
    k_philox_normal<<<(unsigned)((n + threads - 1) / threads), threads, 0, h->stream>>>(out, Bs, B, rows, h->rng_seed,
                                                                                       h->rng_offset, draw);
    h->launches++;
}
