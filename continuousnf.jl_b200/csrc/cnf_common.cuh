// cnf_common.cuh — shared device/host definitions of libcnfb200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>
#include "../../include/cnfb200.h"
#include "../../include/cnf_tsit5.h"

#define CNF_MAX_SLOTS 4096   // accepted-step tape capacity (training)
#define CNF_NT 256           // threads per CTA of the streaming / generic kernels

// Device-resident controller state: the host never reads the ODE state, only this block
// (copied asynchronously to pinned memory to learn `done`).  SURVEY.md A.3-A.5.
struct CnfCtrl {
    double errsum;       // sum (utilde/sc)^2 of the current attempt (all ranks after allreduce)
    double nanflag;      // >0 if a NaN was produced (double so one allreduce covers both)
    double initsum[4];   // initial-dt partial sums: d0^2, d1^2, d2^2 numerators
    double qold;
    double nglobal;      // R * B_global
    long long iter, maxiters, naccept, nreject, nf;
    float t, dt, t0, t1, tdir, dtmax, dtmin, reltol, abstol, dt_last, dt0init;
    int done, retcode, k1idx, use_tape, R, D;
    int accepted;        // the attempt the controller has just judged was accepted (k_tape_stages)
    // stiffness estimate of the default algorithm's auto-switch (reported only)
    double stiffu, stiffd;   // sum (k7 - k6)^2, sum (u_new - g6)^2 of the current attempt
    float stiff_max;
    int stiff_run, stiff_cur;
};

// Network description passed by value to kernels.
struct CnfNet {
    int L, D, C;
    int dims[CNF_MAX_LAYERS + 1];
    const float* W[CNF_MAX_LAYERS];   // col-major d_{l+1} x d_l  (== row-major d_l x d_{l+1})
    const float* Wt[CNF_MAX_LAYERS];  // row-major d_{l+1} x d_l  (transposed copy)
    const float* b[CNF_MAX_LAYERS];
    long long w_off[CNF_MAX_LAYERS], b_off[CNF_MAX_LAYERS];  // offsets into theta / grad
    unsigned flags;
    int trace_mode;
};

// Everything a stage kernel needs to find its buffers.  State and stage derivatives are SoA:
// row r of buffer X is X + r*Bs, sample b at [r*Bs + b].
struct CnfBufs {
    float* const* slots;  // device table of state slots (R x Bs each); slot(n) = tape ? n : n&1
    float* K[7];          // k1..k7 (R x Bs); k1/k7 swap through ctrl->k1idx
    float* S;             // stage input, z rows only (D x Bs)   [generic path]
    const float* eps;     // D x Bs
    const float* ys;      // C x Bs
    long long Bs;         // row stride
    int B;                // local samples
};

__device__ __forceinline__ int cnf_slot(const CnfCtrl* c, long long n) {
    return c->use_tape ? (int)n : (int)(n & 1);
}

// NNlib.tanh_fast rational with IEEE division (generic path: oracle-exact).
__device__ __forceinline__ float cnf_tanh(float x) {
    float x2 = x * x;
    float n = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, CNF_TANH_N4, CNF_TANH_N3), CNF_TANH_N2), CNF_TANH_N1), 1.0f);
    float d = fmaf(x2, fmaf(x2, fmaf(x2, fmaf(x2, CNF_TANH_D4, CNF_TANH_D3), CNF_TANH_D2), CNF_TANH_D1), 1.0f);
    float y = x * __fdiv_rn(n, d);    // IEEE division: bit-identical to the CPU oracle
    return x2 < CNF_TANH_CLAMP ? y : copysignf(1.0f, x);
}

// single translation unit (cnf_lib.cu includes every .cu), so plain definitions
__constant__ float c_A[7][6];
__constant__ float c_BT[7];

__device__ __forceinline__ double cnf_block_sum(double v, double* sh) {
    // sh: at least 32 doubles of shared memory
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sh[w] = v;
    __syncthreads();
    int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.0;
    if (w == 0)
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __syncthreads();
    return v;  // valid in thread 0
}
