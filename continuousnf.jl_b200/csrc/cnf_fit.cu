// cnf_fit.cu — "next" row N2: the caller's training loop, device-resident.
//
// The reference's fit (src/exts/mlj_ext/core_icnf.jl:30-90) wraps the data in
// MLUtils.DataLoader((x,); batchsize = model.batch_size (default 32), shuffle = true, partial = true)
// and runs `epochs = n_epochs` of Optimization.solve with Optimisers.Lion() (Optimisers.jl 0.4, a
// dependency that is not vendored in the reference tree): per mini-batch one loss + gradient and one
// optimiser step.  Through the C ABI that is one call, one H2D of the batch, one D2H of the
// gradient and one H2D of theta per mini-batch, which dominates once the solve itself is fast.
//
// cnfb200_fit_epochs keeps the data, theta, the Lion momentum and the loss history on the device for
// the whole fit: per mini-batch a gather of the shuffled columns into the solver's SoA state, eps
// drawn by the library's Philox stream (N1), the forward solve + reverse sweep of cnfb200_loss_grad,
// a fused Lion step, and the refresh of the derived weight forms.  The host only enqueues (and polls
// the solver's `done` flag as every solve does).
//
// Lion (Chen et al. 2023, "Symbolic Discovery of Optimization Algorithms", Alg. 2, no weight decay;
// Optimisers.Lion(eta = 0.001, beta = (0.9, 0.999))):
//     theta <- theta - eta * sign(beta1 * m + (1 - beta1) * g);   m <- beta2 * m + (1 - beta2) * g
#include "cnf_internal.h"

#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

// dst[r * Bs + j] = src[r * Ns + idx[j]], r < rows, j < B
__global__ void k_gather_cols(const float* __restrict__ src, long long Ns, const long long* __restrict__ idx, int rows,
                              long long B, float* __restrict__ dst, long long Bs) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= B * rows) return;
    const long long j = i % B;
    const int r = (int)(i / B);
    dst[(long long)r * Bs + j] = __ldg(src + (long long)r * Ns + __ldg(idx + j));
}

__global__ void k_lion_step(float* __restrict__ theta, float* __restrict__ m, const float* __restrict__ g, long long n,
                            float eta, float b1, float b2) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float gi = g[i], mi = m[i];
    const float c = fmaf(b1, mi, (1.0f - b1) * gi);
    theta[i] -= eta * (c > 0.0f ? 1.0f : (c < 0.0f ? -1.0f : 0.0f));
    m[i] = fmaf(b2, mi, (1.0f - b2) * gi);
}

// hist[e] += loss sum of this mini-batch (FP64, already summed over ranks)
__global__ void k_fit_acc(const double* __restrict__ lossacc, double* __restrict__ hist, int e) { hist[e] += *lossacc; }

static uint64_t splitmix64(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

extern "C" int32_t cnfb200_set_steer(cnf_handle* h, float steer_rate) {
    if (!h || !(steer_rate >= 0.0f)) return CNF_ERR_BADARG;
    h->steer_rate = steer_rate;
    return CNF_OK;
}

extern "C" int32_t cnfb200_fit_epochs(cnf_handle* h, const float* X, const float* Y, int64_t N, int64_t batch_size,
                                      int32_t n_epochs, const int64_t* perm, float t0, float t1, float eta, float beta1,
                                      float beta2, float* theta_out, float* loss_history, cnf_stats* stats) {
    if (!h) return CNF_ERR_BADARG;
    int rc = check_call(h, CNF_MODE_TRAIN, N);
    if (rc) return rc;
    if (!X || n_epochs < 1 || batch_size < 1) { h->err = "fit_epochs: bad X / n_epochs / batch_size"; return CNF_ERR_BADARG; }
    if (!h->rng_on) { h->err = "fit_epochs draws eps on the device: call cnfb200_seed_rng first"; return CNF_ERR_BADARG; }
    const cnf_config& c = h->cfg;
    if (c.C > 0 && !Y) { h->err = "conditional flow: Y is required"; return CNF_ERR_BADARG; }
    CNF_CUDA(cudaSetDevice(h->device));
    const int mode = CNF_MODE_TRAIN, R = c.D + 3;
    if (batch_size > N) batch_size = N;
    const long long Ns = round_up(N, 256);
    h->cur_R = R;
    h->launches = 0;
    long long total_launches = 0;
    if ((rc = ensure_workspace(h, batch_size, R))) return rc;
    if ((rc = ensure_bw(h, round_up(batch_size, 256)))) return rc;

    // resident copies: data in SoA (rows x Ns), indices of one epoch, Lion momentum, loss history
    float *dX = nullptr, *dY = nullptr, *dM = nullptr, *stage = nullptr;
    long long* dIdx = nullptr;
    double* dHist = nullptr;
    auto cleanup = [&]() { cudaFree(dX); cudaFree(dY); cudaFree(dM); cudaFree(stage); cudaFree(dIdx); cudaFree(dHist); };
#define FIT_CUDA(x)                                                            \
    do {                                                                       \
        cudaError_t e_ = (x);                                                  \
        if (e_ != cudaSuccess) { h->err = std::string("fit_epochs: ") + cudaGetErrorString(e_); cleanup(); return CNF_ERR_CUDA; } \
    } while (0)
    FIT_CUDA(cudaMalloc(&dX, (size_t)c.nvars * Ns * sizeof(float)));
    if (c.C > 0) FIT_CUDA(cudaMalloc(&dY, (size_t)c.C * Ns * sizeof(float)));
    FIT_CUDA(cudaMalloc(&dM, (size_t)h->nparams * sizeof(float)));
    FIT_CUDA(cudaMalloc(&dIdx, (size_t)N * sizeof(long long)));
    FIT_CUDA(cudaMalloc(&dHist, (size_t)n_epochs * sizeof(double)));
    FIT_CUDA(cudaMemsetAsync(dM, 0, (size_t)h->nparams * sizeof(float), h->stream));
    FIT_CUDA(cudaMemsetAsync(dHist, 0, (size_t)n_epochs * sizeof(double), h->stream));
    {
        const int rmax = c.nvars > c.C ? c.nvars : c.C;
        const bool need_stage = !is_device_ptr(X) || (Y && !is_device_ptr(Y));
        if (need_stage) FIT_CUDA(cudaMalloc(&stage, (size_t)rmax * N * sizeof(float)));
        const float* src = X;
        if (!is_device_ptr(X)) {
            FIT_CUDA(cudaMemcpyAsync(stage, X, (size_t)c.nvars * N * sizeof(float), cudaMemcpyHostToDevice, h->stream));
            src = stage;
        }
        launch_pack_rows(h, src, c.nvars, N, dX, Ns);
        if (c.C > 0) {
            src = Y;
            if (!is_device_ptr(Y)) {
                FIT_CUDA(cudaMemcpyAsync(stage, Y, (size_t)c.C * N * sizeof(float), cudaMemcpyHostToDevice, h->stream));
                src = stage;
            }
            launch_pack_rows(h, src, c.C, N, dY, Ns);
        }
    }
    FIT_CUDA(cudaEventRecord(h->ev0, h->stream));

    std::vector<long long> idx((size_t)N);
    uint64_t sm = h->rng_seed ^ 0x5DEECE66Dull;
    uint64_t steer_state = h->rng_seed ^ 0xA5A5A5A55A5A5A5Aull;   // the steer stream (cnfb200_set_steer)
    CnfCtrl fin;
    memset(&fin, 0, sizeof(fin));
    int worst = CNF_OK;
    for (int e = 0; e < n_epochs; ++e) {
        if (perm) {
            for (long long i = 0; i < N; ++i) {
                const long long v = perm[(long long)e * N + i];
                if (v < 0 || v >= N) { h->err = "fit_epochs: perm entry outside [0, N)"; cleanup(); return CNF_ERR_BADARG; }
                idx[(size_t)i] = v;
            }
        } else {   // Fisher-Yates, host side: index data only
            if (e == 0) for (long long i = 0; i < N; ++i) idx[(size_t)i] = i;
            for (long long i = N - 1; i > 0; --i) {
                const long long j = (long long)(splitmix64(sm) % (uint64_t)(i + 1));
                const long long t = idx[(size_t)i]; idx[(size_t)i] = idx[(size_t)j]; idx[(size_t)j] = t;
            }
        }
        FIT_CUDA(cudaStreamSynchronize(h->stream));   // the previous epoch's gathers have read dIdx
        FIT_CUDA(cudaMemcpyAsync(dIdx, idx.data(), (size_t)N * sizeof(long long), cudaMemcpyHostToDevice, h->stream));
        for (long long lo = 0; lo < N; lo += batch_size) {   // partial = true: the last batch may be short
            const long long B = (N - lo < batch_size) ? (N - lo) : batch_size;
            CnfBufs bf = make_bufs(h, B);
            long long Bglobal = global_batch(h, B, &rc);
            if (rc) { cleanup(); return rc; }
            float* slot0 = h->slots[0];
            FIT_CUDA(cudaMemsetAsync(slot0, 0, (size_t)R * bf.Bs * sizeof(float), h->stream));
            const int threads = 256;
            k_gather_cols<<<(unsigned)((B * c.nvars + threads - 1) / threads), threads, 0, h->stream>>>(dX, Ns, dIdx + lo, c.nvars,
                                                                                                        B, slot0, bf.Bs);
            if (c.C > 0)
                k_gather_cols<<<(unsigned)((B * c.C + threads - 1) / threads), threads, 0, h->stream>>>(dY, Ns, dIdx + lo, c.C, B,
                                                                                                      h->d_ys, bf.Bs);
            launch_philox_normal(h, h->d_eps, bf.Bs, B, c.D, h->rng_draws++);
            h->eps_B = B;
            // steer_tspan (src/base_icnf.jl:108-121) runs inside every TrainMode loss call of the reference's fit, so t1 is
            // re-drawn per mini-batch: t1' = muladd(|t1 - t0|, r, t1), r ~ U(-steer_rate, steer_rate)
            float t1b = t1;
            if (h->steer_rate != 0.0f) {
                const double u = (double)(splitmix64(steer_state) >> 40) * (1.0 / 16777216.0);
                const float r = (float)((2.0 * u - 1.0) * (double)h->steer_rate);
                t1b = fmaf(fabsf(t1 - t0), r, t1);
            }
            if ((rc = loss_grad_core(h, bf, mode, R, Bglobal, t0, t1b, true, &fin))) { cleanup(); return rc; }
            if (fin.retcode != CNF_OK) worst = fin.retcode;
            k_fit_acc<<<1, 1, 0, h->stream>>>(h->d_lossacc, dHist, e);
            k_lion_step<<<(unsigned)((h->nparams + threads - 1) / threads), threads, 0, h->stream>>>(h->d_theta, dM, h->d_grad,
                                                                                                   h->nparams, eta, beta1, beta2);
            h->launches += 4;
            if ((rc = refresh_params(h, false))) { cleanup(); return rc; }
            total_launches += h->launches;
            h->launches = 0;
        }
    }
    if ((rc = refresh_params(h, true))) { cleanup(); return rc; }
    FIT_CUDA(cudaEventRecord(h->ev3, h->stream));
    if (theta_out)
        FIT_CUDA(cudaMemcpyAsync(theta_out, h->d_theta, (size_t)h->nparams * sizeof(float), cudaMemcpyDefault, h->stream));
    std::vector<double> hist((size_t)n_epochs);
    FIT_CUDA(cudaMemcpyAsync(hist.data(), dHist, (size_t)n_epochs * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    FIT_CUDA(cudaStreamSynchronize(h->stream));
    FIT_CUDA(cudaGetLastError());
    if (loss_history) {
        // per-epoch mean over the samples seen (all ranks): sum of per-batch loss sums / global N
        int rcg = CNF_OK;
        const long long Nglobal = global_batch(h, N, &rcg);
        for (int e = 0; e < n_epochs; ++e) loss_history[e] = (float)(hist[(size_t)e] / (double)Nglobal);
    }
    h->launches = total_launches;
    fill_stats(h, fin, stats);
    cleanup();
#undef FIT_CUDA
    return worst;
}
