// cnf_vector.cu — "next" row N4: VectorMode semantics at batch scale.
//
// The reference's other compute-mode family solves every sample as its own ODE problem
// (inference_prob / inference_sol for <:VectorMode, src/base_icnf.jl:145-165, 213-264; the
// Distributions adapter broadcasts it over columns, src/exts/dist_ext/core_icnf.jl:24-25): state
// vector [z; dlogp] of length R = D + 1, its own dt, its own error norm (RMS over those R entries),
// its own accept/reject history.  Here that is ONE kernel launch for the whole batch: a thread owns
// a sample and runs the complete adaptive Tsit5 solve in registers — Hairer initial step, the seven
// stages through narrow_f (cnf_stage_narrow.cu), embedded error, PI controller — with no global
// coupling, no host round trip and no state traffic between stages.  Lanes of a warp that finish
// early idle until the slowest lane of the warp is done.
//
// Arithmetic follows the MatrixMode kernels operation by operation (same fma order in the stage
// combination and the error estimate, the controller in double), so a sample's result equals the
// MatrixMode solve of a batch holding only that sample.  TestMode, exact trace, narrow 3-layer
// shapes (narrow_find); everything else raises CNF_ERR_UNSUPPORTED.
#include "cnf_internal.h"

struct VecParams {
    float t0, t1, tdir, dtmax, dtmin, reltol, abstol, dt0;
    long long maxiters;
    long long Bs;
    int B;
};

// per-batch summary: [0] max naccept, [1] max nreject, [2] max nf, [3] worst retcode
template <int D, int H1, int H2>
__global__ void __launch_bounds__(128)
k_solve_vector(const __grid_constant__ NarrowW p, VecParams vp, const float* __restrict__ u0, float* __restrict__ uout,
               int* __restrict__ summary) {
    constexpr int R = D + 1;
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = b < vp.B;
    const long long Bs = vp.Bs;
    float u[R], k[7][R];   // k is indexed by the (uniform) stage counter: it lives in local memory, L1-resident
#pragma unroll
    for (int r = 0; r < R; ++r) u[r] = live ? u0[(long long)r * Bs + b] : 0.0f;
    const double nglobal = (double)R;
    long long nf = 1, iter = 0, naccept = 0, nreject = 0;
    int retcode = 0;
    bool done = !live;
    float t = vp.t0, dt = 0.0f, hh = 0.0f, dt0init = 0.0f;
    double qold = CNF_QOLDINIT, d1 = 0.0;
    float sk[R], un[R];
#pragma unroll
    for (int r = 0; r < R; ++r) { sk[r] = 1.0f; un[r] = 0.0f; }
    // begin_attempt of cnf_solver.cu, per sample
    auto begin_attempt = [&]() {
        if (done) return;
        if (t == vp.t1) { done = true; return; }
        if (++iter > vp.maxiters) { retcode = CNF_RET_MAXITERS; done = true; return; }
        float adt = fminf(fabsf(dt), vp.dtmax);
        adt = fminf(adt, fabsf(vp.t1 - t));
        dt = vp.tdir * adt;
        if (t + dt == t || (vp.dtmin > 0.0f && adt < vp.dtmin)) { retcode = CNF_RET_DTLESSTHANMIN; done = true; }
    };
    // One call site of narrow_f inside one loop (the function is ~40 KB of straight-line FFMAs: eight
    // inlined copies would not fit the instruction cache).  phase 0: k1 = f(u0); 1: Hairer probe
    // f(u0 + tdir*dt0init*k1); 2..7: stage s = phase - 1 of the current attempt (uniform over the grid).
    int phase = 0;
#pragma unroll 1
    for (;;) {
        float z[D], y[D], tr;
        const int s = phase - 1;
        if (phase == 0) {
#pragma unroll
            for (int i = 0; i < D; ++i) z[i] = u[i];
        } else if (phase == 1) {
#pragma unroll
            for (int i = 0; i < D; ++i) z[i] = fmaf(hh, k[0][i], u[i]);
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if (s < 6 && r >= D) continue;
                float acc = c_A[s][0] * k[0][r];
                for (int j = 1; j < s; ++j) acc = fmaf(c_A[s][j], k[j][r], acc);
                const float v = fmaf(dt, acc, u[r]);
                if (s == 6) un[r] = v;
                if (r < D) z[r < D ? r : 0] = v;
            }
        }
        narrow_f<D, H1, H2>(p, z, y, tr);
        if (phase == 0) {
#pragma unroll
            for (int i = 0; i < D; ++i) k[0][i] = y[i];
            k[0][D] = tr;
            if (vp.dt0 > 0.0f) {
                dt = vp.tdir * vp.dt0;
                begin_attempt();
                phase = 2;
            } else {   // Hairer's initial step (k_initnorm / k_initdt with nglobal = R)
                double a0 = 0.0, a1 = 0.0;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    sk[r] = fmaf(fabsf(u[r]), vp.reltol, vp.abstol);
                    const float r0 = u[r] / sk[r], r1 = k[0][r] / sk[r];
                    a0 += (double)r0 * (double)r0;
                    a1 += (double)r1 * (double)r1;
                }
                const double d0 = sqrt(a0 / nglobal);
                d1 = sqrt(a1 / nglobal);
                double dta = (d0 < 1e-5 || d1 < 1e-5) ? 1e-6 : 0.01 * d0 / d1;
                if (dta > (double)vp.dtmax) dta = (double)vp.dtmax;
                dt0init = (float)dta;
                hh = vp.tdir * dt0init;
                phase = 1;
            }
        } else if (phase == 1) {
            double a2 = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float f1 = r < D ? y[r < D ? r : 0] : tr;
                const float q = (f1 - k[0][r]) / sk[r];
                a2 += (double)q * (double)q;
            }
            const double d2 = sqrt(a2 / nglobal) / (double)dt0init;
            const double mx = d1 > d2 ? d1 : d2;
            const double dt1 = (mx <= 1e-15) ? fmax(1e-6, (double)dt0init * 1e-3) : pow(10.0, -(2.0 + log10(mx)) / 5.0);
            const double dtb = fmin(fmin(100.0 * (double)dt0init, dt1), (double)vp.dtmax);
            dt = vp.tdir * (float)dtb;
            nf += 2;
            begin_attempt();
            phase = 2;
        } else {
#pragma unroll
            for (int i = 0; i < D; ++i) k[s][i] = y[i];
            k[s][D] = tr;
            if (s < 6) {
                phase++;
                continue;
            }
            // embedded error (k_errnorm / the 7th narrow stage)
            double es = 0.0;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                float a = c_BT[0] * k[0][r];
#pragma unroll
                for (int j = 1; j < 7; ++j) a = fmaf(c_BT[j], k[j][r], a);
                const float ut = dt * a;
                const float sc = fmaf(fmaxf(fabsf(u[r]), fabsf(un[r])), vp.reltol, vp.abstol);
                float q = ut / sc;
                if (un[r] != un[r]) q = un[r];
                es += (double)q * (double)q;
            }
            if (!done) {   // k_controller, per sample
                nf += 6;
                const float EEst = (float)sqrt(es / nglobal);
                if (EEst != EEst) {
                    retcode = CNF_RET_UNSTABLE;
                    done = true;
                } else {
                    double q11 = 0.0, q;
                    if (EEst == 0.0f) {
                        q = 1.0 / CNF_QMAX;
                    } else {
                        q11 = pow((double)EEst, CNF_BETA1);
                        q = q11 / pow(qold, CNF_BETA2);
                        q = fmax(1.0 / CNF_QMAX, fmin(1.0 / CNF_QMIN, q / CNF_GAMMA));
                    }
                    if (EEst <= 1.0f) {
                        float tn = t + dt;
                        if (fabsf(tn - vp.t1) < 100.0f * 1.1920929e-7f * fmaxf(fabsf(t), fabsf(vp.t1))) tn = vp.t1;
                        t = tn;
#pragma unroll
                        for (int r = 0; r < R; ++r) { u[r] = un[r]; k[0][r] = k[6][r]; }   // FSAL
                        qold = fmax((double)EEst, CNF_QOLDINIT);
                        dt = (float)((double)dt / q);
                        naccept++;
                    } else {
                        dt = (float)((double)dt / fmin(1.0 / CNF_QMIN, q11 / CNF_GAMMA));
                        nreject++;
                    }
                    begin_attempt();
                }
            }
            phase = 2;
        }
        if (phase == 2 && !__any_sync(0xffffffffu, !done)) break;
    }
    if (live) {
#pragma unroll
        for (int r = 0; r < R; ++r) uout[(long long)r * Bs + b] = u[r];
    }
    // batch summary: warp maxima, then one atomic per warp
    int m0 = live ? (int)naccept : 0, m1 = live ? (int)nreject : 0, m2 = live ? (int)nf : 0, m3 = live ? retcode : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m0 = max(m0, __shfl_xor_sync(0xffffffffu, m0, o));
        m1 = max(m1, __shfl_xor_sync(0xffffffffu, m1, o));
        m2 = max(m2, __shfl_xor_sync(0xffffffffu, m2, o));
        m3 = max(m3, __shfl_xor_sync(0xffffffffu, m3, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMax(&summary[0], m0); atomicMax(&summary[1], m1); atomicMax(&summary[2], m2); atomicMax(&summary[3], m3);
    }
}

template <int D, int H1, int H2>
static void vector_launch(cnf_handle* h, const VecParams& vp, const float* u0, float* uout, int* summary) {
    const unsigned grid = (unsigned)((vp.B + 127) / 128);
    k_solve_vector<D, H1, H2><<<grid, 128, 0, h->stream>>>(*(const NarrowW*)h->narrow_w, vp, u0, uout, summary);
    h->launches++;
}

bool vector_narrow_ok(const cnf_handle* h, int mode) {
    return mode == CNF_MODE_TEST && h->cfg.kernel_path != 1 && narrow_find(h->cfg) != nullptr && h->narrow_w != nullptr;
}

// slot 0 holds [x; 0] (R x Bs); on return slot 0 holds the final state of every sample and `out`
// a control block the epilogue kernels can read (naccept = 0 -> slot 0)
int vector_solve(cnf_handle* h, const CnfBufs& bf, float t0, float t1, CnfCtrl* out) {
    const cnf_config& c = h->cfg;
    const NarrowEntry* e = narrow_find(c);
    if (!e || !h->narrow_w) {
        h->err = "vector mode: only TestMode on the narrow 3-layer shapes (hidden <= 64) is built";
        return CNF_ERR_UNSUPPORTED;
    }
    VecParams vp;
    vp.t0 = t0; vp.t1 = t1;
    vp.tdir = (t1 >= t0) ? 1.0f : -1.0f;
    vp.dtmax = c.dtmax > 0 ? c.dtmax : fabsf(t1 - t0);
    vp.dtmin = c.dtmin;
    vp.reltol = c.reltol > 0 ? c.reltol : CNF_DEFAULT_RELTOL;
    vp.abstol = c.abstol > 0 ? c.abstol : CNF_DEFAULT_ABSTOL;
    vp.dt0 = c.dt0 > 0 ? c.dt0 : 0.0f;
    vp.maxiters = c.maxiters > 0 ? c.maxiters : CNF_DEFAULT_MAXITERS;
    vp.Bs = bf.Bs;
    vp.B = bf.B;
    int* summary = reinterpret_cast<int*>(&h->d_ctrl->initsum[0]);   // 4 ints of scratch inside the control block
    CnfCtrl ic;
    memset(&ic, 0, sizeof(ic));
    ic.R = c.D + 1;
    ic.D = c.D;
    ic.t = t1; ic.t0 = t0; ic.t1 = t1;
    ic.done = 1;
    CNF_CUDA(cudaMemcpyAsync(h->d_ctrl, &ic, sizeof(CnfCtrl), cudaMemcpyHostToDevice, h->stream));
    CNF_CUDA(cudaEventRecord(h->ev1, h->stream));
    float* s0 = h->slots[0];
#define VEC_CASE(DD, A, B_) if (e->D == DD && e->H1 == A && e->H2 == B_) vector_launch<DD, A, B_>(h, vp, s0, s0, summary);
    VEC_CASE(2, 64, 64) VEC_CASE(2, 32, 32) VEC_CASE(1, 64, 64) VEC_CASE(3, 64, 64) VEC_CASE(4, 32, 32)
#undef VEC_CASE
    CNF_CUDA(cudaGetLastError());
    CNF_CUDA(cudaEventRecord(h->ev2, h->stream));
    CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[0], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    *out = h->h_ctrl[0];
    const int* sm = reinterpret_cast<const int*>(&out->initsum[0]);
    const int na = sm[0], nr = sm[1], nfm = sm[2], rcw = sm[3];
    out->naccept = 0;            // the epilogue reads slot(naccept) = slot 0
    out->nreject = nr;
    out->nf = nfm;
    out->retcode = rcw;
    out->iter = na;              // fill_stats_vector reports it as naccept
    h->path_used = 4;
    return CNF_OK;
}
