// cnf_internal.h — host-side declarations shared by the .cu files of libcnfb200.
#pragma once
#include "cnf_common.cuh"
#include <string>
#include <vector>

struct CnfXSlot { double v; unsigned long long seq; };   // one mailbox slot of the peer-memory error-norm exchange

struct cnf_handle {
    cnf_config cfg;
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr, ev3 = nullptr, evchk[2] = {nullptr, nullptr};
    std::string err;
    // parameters
    long long nparams = 0;
    float* d_theta = nullptr;   // flat theta (layout of the caller)
    float* d_wt = nullptr;      // transposed weight copies, layer after layer
    float* d_grad = nullptr;    // gradient accumulator (layout of theta)
    long long w_off[CNF_MAX_LAYERS], b_off[CNF_MAX_LAYERS], wt_off[CNF_MAX_LAYERS];
    bool have_params = false;
    CnfNet net;
    // controller
    CnfCtrl* d_ctrl = nullptr;
    CnfCtrl* h_ctrl = nullptr;  // pinned, 2 entries
    float* d_dts = nullptr;     // accepted dt tape (CNF_MAX_SLOTS)
    double* d_lossacc = nullptr;
    // workspaces (grown lazily, never shrunk)
    long long cap_B = 0;        // samples the K/S/eps/ys buffers are sized for
    int cap_R = 0;
    int cur_R = 0;              // rows of the state in the current call
    float* d_K[7] = {};
    float* d_S = nullptr;
    float* d_eps = nullptr;
    // N1: on-device eps / z0 (cnf_rng.cu)
    bool tile_off = false;       // option tile_solve = 0: never use the single-CTA whole-solve kernels (cnf_tile.cu)
    bool vector_mode = false;    // N4: per-sample step control (cnf_vector.cu, cnf_vector_generic.cu)
    void* d_vg = nullptr;        // per-sample controller arrays of the generic VectorMode driver
    long long vg_cap = 0;
    bool rng_on = false;
    unsigned long long rng_seed = 0, rng_offset = 0;
    float steer_rate = 0.0f;     // cnfb200_set_steer: fit_epochs re-draws t1 per mini-batch (src/base_icnf.jl:108-121)
    uint32_t rng_draws = 0;      // draws made so far (the next draw's index)
    long long eps_B = 0;         // batch of the eps now in d_eps (cnfb200_last_eps)
    float* d_ys = nullptr;
    float* d_io = nullptr;      // staging for host<->device copies of sample-major I/O
    long long cap_io = 0;
    std::vector<float*> slots;  // state slots allocated so far
    long long slot_elems = 0;
    float** d_slots = nullptr;  // device table [CNF_MAX_SLOTS]
    std::vector<float*> ktape;  // per accepted step: z rows of k_1..k_5 (5 x D x Bs), reverse sweep only
    float** d_ktape = nullptr;  // device table [CNF_MAX_SLOTS]
    long long ktape_elems = 0;  // D * Bs the entries were sized for
    bool ktape_off = false;     // option stage_tape = 0: always recompute
    // backward workspaces
    float* d_bw[16] = {};
    long long cap_bw = 0;
    long long cap_bw13 = 0;     // d_bw[13], d_bw[14]: one-hot eps / zero cotangent of the exact-trace tensor sweep (allocated on first use)
    // communicator (NCCL via dlopen)
    void* nccl_comm = nullptr;
    int nranks = 1, rank = 0;
    long long bglobal_B = -1, bglobal = 0;   // cached global batch (sum of the ranks' B) for local batch bglobal_B
    // error-norm exchange over NVLink peer memory (cnf_nccl.cu): every rank owns a mailbox of
    // [2 parities][nranks] slots that its peers write into; the controller kernel does the exchange itself
    CnfXSlot* d_xslots = nullptr;
    CnfXSlot** d_xpeers = nullptr;       // device table: base of rank r's mailbox (peer-mapped)
    unsigned long long* d_xseq = nullptr;        // exchanges done so far
    std::vector<void*> xpeer_maps;               // cudaIpcOpenMemHandle mappings to close
    bool p2p_ok = false;
    long long launches = 0;
    // optional per-stage-kernel timing (cnfb200_set_option "time_stages")
    int time_stages = 0;         // 0 off; N: every N-th stage launch is timed
    unsigned long long stage_seq = 0;
    bool stage_timed_now = false;
    std::vector<cudaEvent_t> stage_ev;
    size_t stage_ev_used = 0;
    double ms_stage = 0.0;
    long long n_stage = 0;
    int path_used = 0;
    void* tc_state = nullptr;   // tensor-core path state (cnf_stage_tc.cu)
    int tc_dbg = 0;
    int tc_mode = CNF_MODE_TRAIN;   // mode of the solve tc_eval serves: TrainMode Hutchinson rows, or TestMode exact trace
    bool tc_noring = false;     // option "tc_ring" = 0: epilogue operands through registers instead of the shared-memory ring
    bool fused_off = false;     // option "fused" = 0: never use the fused multi-layer kernel (cnf_stage_fused.cu)
    bool tc_bw_off = false;     // option "tc_backward" = 0: reverse sweep on the generic SIMT kernels
    long long wg_lbo = 0, wg_sbo = 0;   // debug override of the MN-major descriptor strides (0 = default)
    void* narrow_w = nullptr;   // host copy of the packed kernel-parameter weights (narrow path)
    void* ntc_state = nullptr;  // narrow tensor-core path state (cnf_stage_narrow_tc.cu)
};

// " [bounded wait gave up: site 0x..., block ..., thread ..., parity ...]" when a kernel trapped out of a wedged wait
// (the record lives in page-locked host memory and outlives the context), else "" (cnf_api.cu)
std::string cnf_trap_note();

#define CNF_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e__ = (call);                                                        \
        if (e__ != cudaSuccess) {                                                        \
            h->err = std::string(#call) + ": " + cudaGetErrorString(e__) + cnf_trap_note();   \
            return (e__ == cudaErrorMemoryAllocation) ? CNF_ERR_OOM : CNF_ERR_CUDA;      \
        }                                                                                \
    } while (0)

// ---- stage timing helpers (cnf_api.cu)
void stage_timer_begin(cnf_handle* h);
void stage_timer_end(cnf_handle* h);

// ---- cnf_solver.cu
void cnf_upload_constants();
bool narrow_shape(const cnf_config& c);
bool tile_solve_ok(cnf_handle* h, int mode, long long B);
int tile_solve(cnf_handle* h, const CnfBufs& bf, int mode, bool use_tape, long long maxiters, CnfCtrl* out);
bool tile_sweep(cnf_handle* h, const CnfBufs& bf, int mode, long long naccept, long long Bglobal);
int vector_solve(cnf_handle* h, const CnfBufs& bf, float t0, float t1, CnfCtrl* out);
bool vector_narrow_ok(const cnf_handle* h, int mode);   // the one-launch register kernel serves this shape / mode
int vector_solve_generic(cnf_handle* h, const CnfBufs& bf, int mode, int R, float t0, float t1, CnfCtrl* out);
void launch_tape_stages(cnf_handle* h, const CnfBufs& bf);
void launch_philox_normal(cnf_handle* h, float* out, long long Bs, long long B, int rows, uint32_t draw);
void launch_pack_rows(cnf_handle* h, const float* src, int rows, long long B, float* dst, long long Bs);
void launch_unpack_rows(cnf_handle* h, const float* src, int rows, long long B, long long Bs, float* dst);
void launch_combine(cnf_handle* h, const CnfBufs& bf, int stage, int also_S = 0);   // stage 1..5 -> S, 6 -> unew, -1 -> init probe
void launch_errnorm(cnf_handle* h, const CnfBufs& bf);
void launch_controller(cnf_handle* h);
void launch_initnorm(cnf_handle* h, const CnfBufs& bf, int which);
void launch_initdt(cnf_handle* h, int phase);  // 0: dt0 from d0,d1; 1: final dt; 2: fixed dt0 begin
void launch_epilogue(cnf_handle* h, const CnfBufs& bf, int mode, float* logp, float* regs,
                     float* z_out, long long Bglobal);
void launch_init_ubar(cnf_handle* h, const CnfBufs& bf, float* ubar, long long Bglobal, int mode);

// ---- cnf_stage_generic.cu
// du = augmented_f(stage input) ; src: D x Bs z rows ; kout: R x Bs
int launch_eval_generic(cnf_handle* h, const CnfBufs& bf, int mode, const float* src, float* kout);
// VJP of TrainMode augmented_f: cotangent (abar D x Bs, scal[3] * w) -> zbar, grads
int launch_vjp_generic(cnf_handle* h, const CnfBufs& bf, const float* src, const float* abar,
                       int stage, float c_l, float c_E, float c_n, const float* dtp,
                       float* zbar, float* grad);

// ---- cnf_stage_narrow.cu : register-resident one-sample-per-thread kernels
bool narrow_supported(const cnf_config& cfg, int mode);
void narrow_set_params(cnf_handle* h, const float* theta_host);
void narrow_free(cnf_handle* h);
// fused: stage input from (u, k1..k_i) -> k_{i+1}; stage 6 also writes unew + error partials
int launch_stage_narrow(cnf_handle* h, const CnfBufs& bf, int stage, int check_done);

// ---- cnf_stage_narrow_tc.cu : narrow exact-trace nets with layer 2 on tcgen05 (D <= 2, hidden 64)
bool ntc_supported(const cnf_handle* h, int mode);
int ntc_set_params(cnf_handle* h, const float* theta_host);
void ntc_free(cnf_handle* h);
int launch_stage_narrow_tc(cnf_handle* h, const CnfBufs& bf, int stage, int check_done);

// ---- cnf_stage_tc.cu : tcgen05 tensor path (wide nets, TrainMode)
bool tc_supported(const cnf_config& cfg, int mode);
int tc_set_params(cnf_handle* h);
void tc_free(cnf_handle* h);
int tc_eval(cnf_handle* h, const CnfBufs& bf, const float* src, int kidx, float* kout_fixed, int check_done);
int tc_eval_z(cnf_handle* h, const CnfBufs& bf, const float* src, float* kout);   // zdot rows only (BF16x3)

// ---- cnf_stage_fused.cu : all layer products of a 128-sample tile chained on chip (hidden width 128)
bool fused_supported(const cnf_handle* h, int mode);
int fused_eval(cnf_handle* h, const CnfBufs& bf, const float* src, int kidx, float* kout_fixed, int check_done);

// ---- cnf_backward_tc.cu : reverse sweep on the tensor path (VecJac)
bool tc_bw_supported(cnf_handle* h);
int tc_vjp_stage(cnf_handle* h, const CnfBufs& bf, const float* src, const float* abar, int stage,
                 const float* dtp, float c_l, float c_E, float c_n, float* ubar, float** kb, float* grad);

// ---- cnf_nccl.cu
void cnf_p2p_free(cnf_handle* h);
void cnf_nccl_destroy(cnf_handle* h);
int cnf_nccl_allreduce_f64(cnf_handle* h, double* dptr, int n);
int cnf_nccl_allreduce_f32(cnf_handle* h, float* dptr, long long n);
