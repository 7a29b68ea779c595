/* cnfb200.h — C ABI of libcnfb200.so, the B200-native batched augmented-dynamics solve
 * of ContinuousNormalizingFlows.jl ("MatrixMode" hot path).
 *
 * Each entry point cites the reference interface it replaces (paths relative to the
 * reference checkout).  Plain C: pointers + sizes only; the Julia side binds these with
 * `ccall`, the in-repo Python host mirror binds them with ctypes.
 *
 * Layouts: every matrix is Julia column-major `rows x B`, i.e. sample-major contiguous
 * (`xs[b*nvars + i]`).  Pointers may be host or device pointers (detected with
 * cudaPointerGetAttributes); host buffers are copied in/out inside the call.
 * No pointer is retained after a call returns.
 *
 * Return codes: 0 ok; >0 solver retcodes (outputs are still written, like the reference,
 * which reads the last state without checking: src/base_icnf.jl:141-142);
 * <0 hard failure, text via cnfb200_last_error().
 */
#ifndef CNFB200_H
#define CNFB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define CNF_API __attribute__((visibility("default")))
#else
#define CNF_API
#endif

#define CNF_MAX_LAYERS 8

/* flags: the type-level booleans of `ICNF{...}` (src/base_icnf.jl:42-60, src/icnf.jl:69-104) */
#define CNF_FLAG_COND        (1u << 0) /* COND: CondLayer wraps nn (src/layers/cond_layer.jl:1-9) */
#define CNF_FLAG_AUGMENTED   (1u << 1) /* naugmented != 0 */
#define CNF_FLAG_NORM_Z      (1u << 2) /* lambda1 != 0: integrate ||zdot||_2 (src/icnf.jl:335-341) */
#define CNF_FLAG_NORM_J      (1u << 3) /* lambda2 != 0: integrate ||eps^T J||_2 (src/icnf.jl:342-348) */
#define CNF_FLAG_NORM_Z_AUG  (1u << 4) /* lambda3 != 0: ||z_aug(t1)||_2 (src/base_icnf.jl:179-187) */
#define CNF_FLAG_LAST_LINEAR (1u << 5) /* the last layer has no activation and no bias: with dims = [d0, 1, D] this is
                                          * PlanarLayer, u * act(w'z + b) (src/layers/planar_layer.jl:70-88; row N3);
                                          * b_off of that layer is ignored */

/* mode: src/types.jl Mode {TestMode, TrainMode} */
#define CNF_MODE_TEST  0 /* exact trace, R = D+1 rows (src/icnf.jl:148-184) */
#define CNF_MODE_TRAIN 1 /* Hutchinson,  R = D+3 rows (src/icnf.jl:318-456) */

/* trace_mode: src/types.jl DIVecJacMatrixMode / DIJacVecMatrixMode */
#define CNF_TRACE_VJP 0
#define CNF_TRACE_JVP 1

/* status / retcodes */
#define CNF_OK               0
#define CNF_RET_MAXITERS     1
#define CNF_RET_DTLESSTHANMIN 2
#define CNF_RET_UNSTABLE     3
#define CNF_ERR_BADCONFIG   (-1)
#define CNF_ERR_CUDA        (-2)
#define CNF_ERR_NCCL        (-3)
#define CNF_ERR_OOM         (-4)
#define CNF_ERR_BADARG      (-5)
#define CNF_ERR_UNSUPPORTED (-6)

/* mirrors the fields of the `ICNF` struct the path reads (src/icnf.jl:69-104) plus the
 * `sol_kwargs` the solver consumes (src/base_icnf.jl:26,141). */
typedef struct cnf_config {
    int32_t D;          /* nvars + naugmented */
    int32_t C;          /* condition rows (0 if unconditional) */
    int32_t nvars;
    int32_t n_layers;   /* L Dense(tanh) layers */
    int32_t dims[CNF_MAX_LAYERS + 1]; /* d0 = D+C, ..., dL = D */
    int32_t trace_mode; /* CNF_TRACE_* */
    uint32_t flags;     /* CNF_FLAG_* */
    float lambda1, lambda2, lambda3;
    float reltol, abstol;     /* 0 => solver default (1e-3, 1e-6) */
    int64_t maxiters;         /* 0 => 1e5 */
    float dt0, dtmax, dtmin;  /* 0 => solver default */
    int32_t device;           /* CUDA device ordinal of this process */
    int32_t kernel_path;      /* 0 auto; 1 force generic SIMT; 2 force register-resident narrow (instead of 6); 3 force tcgen05 layer GEMMs */
} cnf_config;

typedef struct cnf_stats {
    int64_t naccept, nreject, nf; /* nf = augmented_f evaluations */
    int32_t retcode;
    int32_t kernel_path;          /* path actually used: 1 generic SIMT, 2 register-resident narrow, 3 tcgen05 layer GEMMs, 4 VectorMode (one-launch register kernel), 5 whole-solve single-launch kernels, 6 narrow nets with layer 2 on tcgen05, 7 fused multi-layer tcgen05 kernel (hidden 128), 8 VectorMode on the MatrixMode stage kernels */
    float t_final, dt_last;
    double ms_total;              /* device time of the whole call (CUDA events) */
    double ms_solve;              /* device time of the RK loop only */
    int64_t gpu_launches;         /* kernels launched by this call */
    double ms_stage;              /* summed device time of the RK stage kernels (option time_stages) */
    int64_t n_stage;              /* number of stage-kernel launches timed */
    /* Reported, never acted on: the reference calls solve() without `alg` (src/base_icnf.jl:141), so its
     * default composite algorithm starts on Tsit5 and would leave it after 10 consecutive steps whose stiffness estimate
     * |dt| * ||k7 - k6|| / ||u_new - g6|| / 3.5068 (Tsit5's stability size) exceeds 9/10.  stiff_max = the largest
     * estimate over the accepted steps (-1: not computed on this kernel path), stiff_run = the longest run of
     * consecutive accepted steps above 9/10.  stiff_run >= 10 means the real reference would have switched solver and
     * this Tsit5-only path is not comparable on that problem.  Local shard only when the batch is sharded. */
    float stiff_max;
    int32_t stiff_run;
} cnf_stats;

typedef struct cnf_handle cnf_handle;

CNF_API const char* cnfb200_version(void);

/* replaces construct(...) bookkeeping on the device side (src/base_icnf.jl:1-77) */
CNF_API int32_t cnfb200_create(const cnf_config* cfg, cnf_handle** out);
CNF_API int32_t cnfb200_destroy(cnf_handle* h);
/* runtime options: "time_stages" (0 off; N >= 1: CUDA-event timing of every N-th stage-kernel launch ->
 * cnf_stats.ms_stage / n_stage; N = 1 times all of them, at the cost of two event records per launch),
 * "kernel_path" (0 auto, 1 generic, 2 narrow, 3 tcgen05) */
CNF_API int32_t cnfb200_set_option(cnf_handle* h, const char* key, int64_t value);
CNF_API const char* cnfb200_last_error(const cnf_handle* h);

/* flat ComponentArray parameter vector `ps` + per-layer weight/bias offsets (in floats);
 * weight l is d_{l+1} x d_l column-major (Lux Dense; src/exts/mlj_ext/core_icnf.jl:37-38) */
CNF_API int32_t cnfb200_set_params(cnf_handle* h, const float* theta, int64_t n,
                           const int64_t* w_off, const int64_t* b_off);

/* batch sharding over one-process-per-GPU ranks: rank 0 makes an id, every rank inits.
 * Afterwards every solve allreduces the scalar error-norm partials each attempted step
 * (whole-batch RMS semantics of the matrix-valued ODE state, src/base_icnf.jl:280-285)
 * and cnfb200_loss_grad allreduces loss and gradient.  B_global = sum of local B. */
CNF_API int32_t cnfb200_comm_unique_id(char id_out[128]);
CNF_API int32_t cnfb200_comm_init(cnf_handle* h, int32_t nranks, int32_t rank, const char id[128]);
/* 1 if the per-attempt error-norm exchange runs over NVLink peer memory inside the controller kernel
 * (mailboxes shared through CUDA IPC at cnfb200_comm_init), 0 if it is an ncclAllReduce launch
 * (single rank, set-up refused on some rank, or CNFB200_P2P=0 in the environment) */
CNF_API int32_t cnfb200_comm_p2p(const cnf_handle* h);

/* one evaluation du = augmented_f(u, p, t, icnf, mode, nn, st, eps)
 * (src/icnf.jl:148-164 Test; :318-350 Train/VecJac; :384-420 Train/JacVec).  u, du: R x B. */
CNF_API int32_t cnfb200_augmented_f(cnf_handle* h, int32_t mode, const float* u, const float* ys,
                            const float* eps, int64_t B, float* du);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * cnfb200_augmented_f on the generic FP32 kernels, but the call allocates nothing: its SoA work planes live in one
 * file-scope device array of CNFB200_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit
 * (csrc/cnf_noalloc.cu).  u, ys, eps and du must be DEVICE pointers (eps is required in TrainMode);
 * (2 D + C + R) * round_up(B, 256) floats must fit the array, larger batches return CNF_ERR_BADARG with the limit in
 * cnfb200_last_error.  All handles on a device share the array and it keeps the last call's contents (every plane a
 * call reads is rewritten first); the call returns after its work is done, but calls that overlap on the GPU (two
 * handles from two host threads) race. */
#define CNFB200_STATIC_SCRATCH_BYTES (16u << 20)
CNF_API int32_t cnfb200_augmented_f_static(cnf_handle* h, int32_t mode, const float* u, const float* ys,
                                   const float* eps, int64_t B, float* du);

/* replaces inference(icnf, mode, xs[, ys], ps, st) = inference_prob + inference_sol
 * (src/base_icnf.jl:266-309, 167-189, 407-426).  xs: nvars x B; ys: C x B or NULL;
 * eps: D x B (may be NULL in TestMode, where it is unused);
 * logp: B; regs: 3 x B rows (E, n, A) in TrainMode, 1 x B (A) in TestMode, or NULL;
 * z_out: D x B final latent state or NULL. */
CNF_API int32_t cnfb200_logdensity(cnf_handle* h, int32_t mode, const float* xs, const float* ys,
                           const float* eps, int64_t B, float t0, float t1,
                           float* logp, float* regs, float* z_out, cnf_stats* stats);

/* replaces loss(icnf, TrainMode(), xs[, ys], ps, st) (src/icnf.jl:481-502) and its gradient
 * w.r.t. ps / xs (call sites src/exts/mlj_ext/core_icnf.jl:59-63, test/call_tests.jl:239-251)
 * by a discretise-then-optimise reverse sweep.  grad_theta: n floats (layout of theta) or NULL;
 * grad_x: nvars x B or NULL.  With a communicator, loss and grad_theta are global means/sums. */
CNF_API int32_t cnfb200_loss_grad(cnf_handle* h, const float* xs, const float* ys, const float* eps,
                          int64_t B, float t0, float t1, float* loss, float* grad_theta,
                          float* grad_x, cnf_stats* stats);

/* the same for either mode: mode = CNF_MODE_TEST is loss(icnf, TestMode(), ...) = -mean(logp)
 * (src/base_icnf.jl:489-508; differentiated w.r.t. ps and xs in test/call_tests.jl:239-251), whose
 * reverse sweep differentiates the exact trace (eps is ignored and may be NULL). */
CNF_API int32_t cnfb200_loss_grad_mode(cnf_handle* h, int32_t mode, const float* xs, const float* ys,
                               const float* eps, int64_t B, float t0, float t1, float* loss,
                               float* grad_theta, float* grad_x, cnf_stats* stats);

/* replaces generate(...) = generate_prob + generate_sol (src/base_icnf.jl:358-405, 202-211,
 * 447-466): integrates z0 (D x n, drawn by the caller from basedist; NULL = drawn on the device, see
 * cnfb200_seed_rng) from t_from to t_to and returns rows 1:nvars.  xs_out: nvars x n. */
CNF_API int32_t cnfb200_generate(cnf_handle* h, int32_t mode, const float* z0, const float* ys,
                         const float* eps, int64_t n, float t_from, float t_to,
                         float* xs_out, cnf_stats* stats);

/* "next" row N1: draw the Hutchinson probe and the base sample on the device.  The reference draws
 * both on the host once per call (randn!(rng, eps): src/base_icnf.jl:277-278, 367-372;
 * rand(rng, basedist, n): :365-366) and the caller ships D x B floats.  After cnfb200_seed_rng a NULL
 * `eps` (TrainMode calls) or NULL `z0` (cnfb200_generate) is drawn here: Philox4x32-10 keyed by
 * `seed`, counter = (sample_offset + column, row group, draw index), Box-Muller; `sample_offset` is
 * the first global column of this rank's shard, so a draw does not depend on the sharding.  Opt-in:
 * the stream is reproducible against oracle/cnf_oracle.c (cnf_oracle_philox_normal), not against
 * Julia's RNG.  Each drawn matrix consumes one draw index, starting from 0 after seeding. */
CNF_API int32_t cnfb200_seed_rng(cnf_handle* h, uint64_t seed, uint64_t sample_offset);

/* the eps (D x B, caller's layout) the last TrainMode call of batch B used, drawn or passed */
CNF_API int32_t cnfb200_last_eps(cnf_handle* h, float* eps_out, int64_t B);

/* "next" row N2: the caller's training loop (MLJModelInterface.fit, src/exts/mlj_ext/core_icnf.jl:30-90:
 * DataLoader((x,); batchsize, shuffle = true, partial = true), `epochs = n_epochs` of loss + gradient +
 * Optimisers.Lion step per mini-batch) as ONE call with data, theta, the Lion momentum and the loss
 * history resident on the device.  X: nvars x N, Y: C x N (conditional flows, else NULL), column =
 * sample, host or device.  perm: n_epochs x N 0-based column indices (epoch e visits
 * perm[e*N .. e*N+N) in mini-batches of batch_size, the last one short), or NULL = shuffled by the
 * library.  eps is drawn per mini-batch by the device stream (cnfb200_seed_rng is required), draw
 * index advancing by one per mini-batch.  Starts from the theta of cnfb200_set_params; Lion:
 * theta -= eta*sign(beta1*m + (1-beta1)*g), m = beta2*m + (1-beta2)*g, m0 = 0.  theta_out: nparams
 * (may be NULL; the handle keeps the fitted theta either way); loss_history: n_epochs per-epoch means
 * of the mini-batch losses weighted by their sizes.  With several ranks every rank passes its own
 * shard (same number of mini-batches per epoch on every rank). */
/* steer_tspan inside the fit loop (src/base_icnf.jl:108-121: every TrainMode loss call of the reference's fit re-draws
 * t1' = muladd(|t1 - t0|, r, t1), r ~ Uniform(-steer_rate, steer_rate)).  The single-call entry points take the steered
 * (t0, t1) from the caller, who draws r from icnf.rng like the reference; cnfb200_fit_epochs runs many loss calls, so it
 * draws r itself, once per mini-batch, from a host stream: state = seed (of cnfb200_seed_rng) ^ 0xA5A5A5A55A5A5A5A,
 * u_k = (splitmix64(state) >> 40) * 2^-24, r_k = (2 u_k - 1) * steer_rate.  steer_rate = 0 (default) keeps t1 fixed. */
CNF_API int32_t cnfb200_set_steer(cnf_handle* h, float steer_rate);

CNF_API int32_t cnfb200_fit_epochs(cnf_handle* h, const float* X, const float* Y, int64_t N, int64_t batch_size,
                           int32_t n_epochs, const int64_t* perm, float t0, float t1, float eta, float beta1,
                           float beta2, float* theta_out, float* loss_history, cnf_stats* stats);

/* measurement helper (bench only): sustained FP32 FFMA rate of this GPU in TFLOP/s,
 * the roofline denominator MEASURED_PEAKS.json does not provide. */
CNF_API int32_t cnfb200_measure_fp32_peak(int32_t device, double* tflops_out);

/* page-locked host memory for the caller's input / output matrices: the copies inside the entry points
 * above run at full PCIe rate from / into such buffers (a pageable destination costs ~1 ms per 8 MB).
 * The Julia shim wraps them with unsafe_wrap(Array, ...); the library never keeps the pointers. */
CNF_API int32_t cnfb200_host_alloc(size_t bytes, void** out);
CNF_API int32_t cnfb200_host_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* CNFB200_H */
