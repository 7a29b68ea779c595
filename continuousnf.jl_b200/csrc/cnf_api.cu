// cnf_api.cu — C-ABI entry points of libcnfb200.so (include/cnfb200.h) and the host-side
// orchestration: the host only enqueues kernels and polls the device controller block.
#include "cnf_internal.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

int launch_eval_generic_sel(cnf_handle* h, const CnfBufs& bf, int mode, int src_sel,
                            const float* src_ptr, int kidx, float* kout_ptr, int check_done);

static bool is_device_ptr(const void* p) {
    if (!p) return false;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}
static inline long long round_up(long long x, long long m) { return (x + m - 1) / m * m; }

// ---- trap record (cnf_tc_ptx.cuh: trap_with_record): one page-locked word per process, its address published to
// the device global of every device a handle is created on
static unsigned long long* g_trap_host = nullptr;
static bool g_trap_dev[64] = {};
static void trap_record_setup(int device) {
    if (device < 0 || device >= 64 || g_trap_dev[device]) return;
    if (!g_trap_host) {
        if (cudaHostAlloc((void**)&g_trap_host, sizeof(unsigned long long), cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) {
            cudaGetLastError();
            g_trap_host = nullptr;
            return;
        }
        *g_trap_host = 0ull;
    }
    unsigned long long* dptr = nullptr;
    if (cudaHostGetDevicePointer((void**)&dptr, g_trap_host, 0) == cudaSuccess &&
        cudaMemcpyToSymbol(tc::g_cnf_trap_record, &dptr, sizeof(dptr)) == cudaSuccess)
        g_trap_dev[device] = true;
    cudaGetLastError();
}
std::string cnf_trap_note() {
    if (!g_trap_host) return "";
    const unsigned long long v = *reinterpret_cast<volatile unsigned long long*>(g_trap_host);
    if (!(v & 1ull)) return "";
    char buf[160];
    snprintf(buf, sizeof(buf), " [bounded wait gave up: site 0x%llx, block %llu, thread %llu, parity %llu]", v >> 48,
             (v >> 32) & 0xffffull, (v >> 16) & 0xffffull, (v >> 8) & 0xffull);
    return buf;
}

extern "C" const char* cnfb200_version(void) { return "cnfb200 0.1.0 (sm_100a)"; }

extern "C" const char* cnfb200_last_error(const cnf_handle* h) { return h ? h->err.c_str() : "null handle"; }

static int validate_config(const cnf_config& c, std::string& err) {
    if (c.n_layers < 1 || c.n_layers > CNF_MAX_LAYERS) { err = "n_layers out of range"; return CNF_ERR_BADCONFIG; }
    if (c.D < 1 || c.nvars < 1 || c.nvars > c.D || c.C < 0) { err = "bad D/nvars/C"; return CNF_ERR_BADCONFIG; }
    if (c.dims[0] != c.D + c.C) { err = "dims[0] must equal D + C"; return CNF_ERR_BADCONFIG; }
    if (c.dims[c.n_layers] != c.D) { err = "dims[L] must equal D"; return CNF_ERR_BADCONFIG; }
    for (int l = 0; l <= c.n_layers; ++l)
        if (c.dims[l] < 1) { err = "layer width < 1"; return CNF_ERR_BADCONFIG; }
    if (c.trace_mode != CNF_TRACE_VJP && c.trace_mode != CNF_TRACE_JVP) { err = "bad trace_mode"; return CNF_ERR_BADCONFIG; }
    if (((c.flags & CNF_FLAG_COND) != 0) != (c.C > 0)) { err = "COND flag and C disagree"; return CNF_ERR_BADCONFIG; }
    return CNF_OK;
}

extern "C" int32_t cnfb200_create(const cnf_config* cfg, cnf_handle** out) {
    if (!cfg || !out) return CNF_ERR_BADARG;
    *out = nullptr;
    cnf_handle* h = new (std::nothrow) cnf_handle();
    if (!h) return CNF_ERR_OOM;
    *out = h;  // returned even on failure so the caller can read last_error, then destroy
    h->cfg = *cfg;
    int rc = validate_config(h->cfg, h->err);
    if (rc) return rc;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        h->err = "no CUDA device: libcnfb200 has no CPU fallback";
        return CNF_ERR_CUDA;
    }
    h->device = cfg->device;
    CNF_CUDA(cudaSetDevice(h->device));
    cudaDeviceProp prop;
    CNF_CUDA(cudaGetDeviceProperties(&prop, h->device));
    if (prop.major != 10) {
        h->err = "libcnfb200 is built for sm_100a only";
        return CNF_ERR_UNSUPPORTED;
    }
    CNF_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    trap_record_setup(h->device);
    CNF_CUDA(cudaEventCreate(&h->ev0));
    CNF_CUDA(cudaEventCreate(&h->ev1));
    CNF_CUDA(cudaEventCreate(&h->ev2));
    CNF_CUDA(cudaEventCreate(&h->ev3));
    CNF_CUDA(cudaEventCreateWithFlags(&h->evchk[0], cudaEventDisableTiming));
    CNF_CUDA(cudaEventCreateWithFlags(&h->evchk[1], cudaEventDisableTiming));
    CNF_CUDA(cudaMalloc(&h->d_ctrl, sizeof(CnfCtrl)));
    CNF_CUDA(cudaMallocHost(&h->h_ctrl, 2 * sizeof(CnfCtrl)));
    CNF_CUDA(cudaMalloc(&h->d_dts, CNF_MAX_SLOTS * sizeof(float)));
    CNF_CUDA(cudaMalloc(&h->d_lossacc, sizeof(double)));
    CNF_CUDA(cudaMalloc(&h->d_slots, CNF_MAX_SLOTS * sizeof(float*)));
    CNF_CUDA(cudaMemset(h->d_slots, 0, CNF_MAX_SLOTS * sizeof(float*)));
    CNF_CUDA(cudaMalloc(&h->d_ktape, CNF_MAX_SLOTS * sizeof(float*)));
    CNF_CUDA(cudaMemset(h->d_ktape, 0, CNF_MAX_SLOTS * sizeof(float*)));
    cnf_upload_constants();
    CNF_CUDA(cudaGetLastError());
    return CNF_OK;
}

static void free_ktape(cnf_handle* h) {
    for (float* p : h->ktape) cudaFree(p);
    if (!h->ktape.empty() && h->d_ktape) cudaMemset(h->d_ktape, 0, h->ktape.size() * sizeof(float*));
    h->ktape.clear();
    h->ktape_elems = 0;
}

// stage tape entries for steps [0, want): sized D x Bs x 5; stops growing (the reverse sweep then
// recomputes the later steps) when less than a quarter of the device memory would stay free
static int ensure_ktape(cnf_handle* h, size_t want, long long nz) {
    if (h->ktape_off) return CNF_OK;
    if (h->ktape_elems != nz) {
        CNF_CUDA(cudaStreamSynchronize(h->stream));
        free_ktape(h);
        h->ktape_elems = nz;
    }
    if (want > CNF_MAX_SLOTS) want = CNF_MAX_SLOTS;
    if (h->ktape.size() >= want) return CNF_OK;
    const size_t old = h->ktape.size();
    const size_t bytes = (size_t)5 * (size_t)nz * sizeof(float);
    while (h->ktape.size() < want) {
        size_t fr = 0, tot = 0;
        if (cudaMemGetInfo(&fr, &tot) != cudaSuccess || fr < bytes + tot / 4) break;
        float* p = nullptr;
        if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); break; }
        h->ktape.push_back(p);
    }
    if (h->ktape.size() > old)
        CNF_CUDA(cudaMemcpyAsync(h->d_ktape + old, h->ktape.data() + old, (h->ktape.size() - old) * sizeof(float*),
                                 cudaMemcpyHostToDevice, h->stream));
    return CNF_OK;
}

static void free_workspaces(cnf_handle* h) {
    for (int j = 0; j < 7; ++j) { cudaFree(h->d_K[j]); h->d_K[j] = nullptr; }
    cudaFree(h->d_S); h->d_S = nullptr;
    cudaFree(h->d_eps); h->d_eps = nullptr;
    cudaFree(h->d_ys); h->d_ys = nullptr;
    for (float* p : h->slots) cudaFree(p);
    h->slots.clear();
    free_ktape(h);
    for (int j = 0; j < 16; ++j) { cudaFree(h->d_bw[j]); h->d_bw[j] = nullptr; }
    h->cap_bw = 0;
    h->cap_bw13 = 0;
    h->cap_B = 0;
    h->cap_R = 0;
}

extern "C" int32_t cnfb200_destroy(cnf_handle* h) {
    if (!h) return CNF_OK;
    if (h->stream) {
        cudaSetDevice(h->device);
        cudaStreamSynchronize(h->stream);
    }
    free_workspaces(h);
    narrow_free(h);
    ntc_free(h);
    cnf_nccl_destroy(h);
    tc_free(h);
    cudaFree(h->d_io);
    cudaFree(h->d_vg);
    cudaFree(h->d_theta); cudaFree(h->d_wt); cudaFree(h->d_grad);
    cudaFree(h->d_ctrl); cudaFree(h->d_dts); cudaFree(h->d_lossacc); cudaFree(h->d_slots); cudaFree(h->d_ktape);
    if (h->h_ctrl) cudaFreeHost(h->h_ctrl);
    for (cudaEvent_t e : h->stage_ev) cudaEventDestroy(e);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev2) cudaEventDestroy(h->ev2);
    if (h->ev3) cudaEventDestroy(h->ev3);
    if (h->evchk[0]) cudaEventDestroy(h->evchk[0]);
    if (h->evchk[1]) cudaEventDestroy(h->evchk[1]);
    if (h->stream) cudaStreamDestroy(h->stream);
    cudaGetLastError();
    delete h;
    return CNF_OK;
}

// W (col-major dout x din) -> Wt (row-major dout x din): Wt[o*din + i] = W[i*dout + o]
__global__ void k_transpose_w(const float* __restrict__ W, int dout, int din, float* __restrict__ Wt) {
    long long n = (long long)dout * din;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < n;
         p += (long long)gridDim.x * blockDim.x) {
        int o = (int)(p / din), i = (int)(p % din);
        Wt[p] = W[(long long)i * dout + o];
    }
}

// every derived form of theta (transposed weights, the narrow kernels' packed copy, the tensor path's
// hi/lo planes) from h->d_theta; `narrow` = also the narrow-path copy (needs a host round trip)
static int refresh_params(cnf_handle* h, bool narrow) {
    const cnf_config& c = h->cfg;
    const long long n = h->nparams;
    CnfNet& net = h->net;
    net.L = c.n_layers; net.D = c.D; net.C = c.C;
    net.flags = c.flags; net.trace_mode = c.trace_mode;
    for (int l = 0; l <= c.n_layers; ++l) net.dims[l] = c.dims[l];
    for (int l = 0; l < c.n_layers; ++l) {
        net.W[l] = h->d_theta + h->w_off[l];
        net.b[l] = h->d_theta + h->b_off[l];
        net.Wt[l] = h->d_wt + h->wt_off[l];
        net.w_off[l] = h->w_off[l];
        net.b_off[l] = h->b_off[l];
        long long wsz = (long long)c.dims[l] * c.dims[l + 1];
        unsigned g = (unsigned)((wsz + 255) / 256);
        if (g > 1184) g = 1184;
        k_transpose_w<<<g, 256, 0, h->stream>>>(net.W[l], c.dims[l + 1], c.dims[l], h->d_wt + h->wt_off[l]);
    }
    CNF_CUDA(cudaGetLastError());
    if (narrow && narrow_shape(c)) {
        std::vector<float> th((size_t)n);
        CNF_CUDA(cudaMemcpyAsync(th.data(), h->d_theta, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        CNF_CUDA(cudaStreamSynchronize(h->stream));
        narrow_set_params(h, th.data());
        int rcn = ntc_set_params(h, th.data());
        if (rcn) return rcn;
    }
    int rc = tc_set_params(h);
    if (rc) return rc;
    if (narrow) CNF_CUDA(cudaStreamSynchronize(h->stream));
    return CNF_OK;
}

extern "C" int32_t cnfb200_set_params(cnf_handle* h, const float* theta, int64_t n,
                                      const int64_t* w_off, const int64_t* b_off) {
    if (!h || !theta || !w_off || !b_off || n <= 0) return CNF_ERR_BADARG;
    CNF_CUDA(cudaSetDevice(h->device));
    const cnf_config& c = h->cfg;
    long long wt_total = 0;
    for (int l = 0; l < c.n_layers; ++l) {
        long long wsz = (long long)c.dims[l] * c.dims[l + 1];
        const bool nobias = (c.flags & CNF_FLAG_LAST_LINEAR) && l == c.n_layers - 1;
        if (w_off[l] < 0 || w_off[l] + wsz > n || (!nobias && (b_off[l] < 0 || b_off[l] + c.dims[l + 1] > n))) {
            h->err = "set_params: offsets outside theta";
            return CNF_ERR_BADARG;
        }
        h->w_off[l] = w_off[l];
        h->b_off[l] = nobias ? 0 : b_off[l];
        h->wt_off[l] = wt_total;
        wt_total += wsz;
    }
    if (n != h->nparams) {
        cudaFree(h->d_theta); cudaFree(h->d_wt); cudaFree(h->d_grad);
        h->d_theta = h->d_wt = h->d_grad = nullptr;
        CNF_CUDA(cudaMalloc(&h->d_theta, n * sizeof(float)));
        CNF_CUDA(cudaMalloc(&h->d_wt, wt_total * sizeof(float)));
        CNF_CUDA(cudaMalloc(&h->d_grad, n * sizeof(float)));
        h->nparams = n;
    }
    CNF_CUDA(cudaMemcpyAsync(h->d_theta, theta, n * sizeof(float), cudaMemcpyDefault, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));  // theta may be a pageable host buffer
    {
        int rc = refresh_params(h, true);
        if (rc) return rc;
    }
    h->have_params = true;
    return CNF_OK;
}

// time_stages = N: every N-th stage launch of the handle is bracketed by a pair of events (N = 1: all of them).  A
// pair of records costs a few microseconds of stream time, so a benchmark that times every launch measures its own
// instrumentation; N = 7 walks through all stage indices of the 6-stage attempts and costs one pair per attempt.
void stage_timer_begin(cnf_handle* h) {
    h->stage_timed_now = false;
    if (!h->time_stages) return;
    if (++h->stage_seq % (unsigned long long)h->time_stages != 0) return;
    h->stage_timed_now = true;
    if (h->stage_ev_used + 2 > h->stage_ev.size()) {
        for (int i = 0; i < 64; ++i) { cudaEvent_t e; cudaEventCreate(&e); h->stage_ev.push_back(e); }
    }
    cudaEventRecord(h->stage_ev[h->stage_ev_used++], h->stream);
}
void stage_timer_end(cnf_handle* h) {
    if (!h->stage_timed_now) return;
    h->stage_timed_now = false;
    cudaEventRecord(h->stage_ev[h->stage_ev_used++], h->stream);
}
static void stage_timer_collect(cnf_handle* h) {
    h->ms_stage = 0.0;
    h->n_stage = 0;
    for (size_t i = 0; i + 1 < h->stage_ev_used; i += 2) {
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, h->stage_ev[i], h->stage_ev[i + 1]) == cudaSuccess) {
            h->ms_stage += ms;
            h->n_stage++;
        }
    }
    h->stage_ev_used = 0;
    cudaGetLastError();
}

extern "C" int32_t cnfb200_set_option(cnf_handle* h, const char* key, int64_t value) {
    if (!h || !key) return CNF_ERR_BADARG;
    if (!strcmp(key, "time_stages")) { h->time_stages = value > 0 ? (int)(value > 1000 ? 1000 : value) : 0; return CNF_OK; }
    if (!strcmp(key, "kernel_path")) { h->cfg.kernel_path = (int32_t)value; return CNF_OK; }
    if (!strcmp(key, "tile_solve")) { h->tile_off = value == 0; return CNF_OK; }
    if (!strcmp(key, "stage_tape")) { h->ktape_off = value == 0; return CNF_OK; }
    if (!strcmp(key, "vector_mode")) { h->vector_mode = value != 0; return CNF_OK; }
    if (!strcmp(key, "tc_dbg")) { h->tc_dbg = (int)value; return CNF_OK; }
    if (!strcmp(key, "tc_ring")) { h->tc_noring = value == 0; return CNF_OK; }
    if (!strcmp(key, "fused")) { h->fused_off = value == 0; return CNF_OK; }
    if (!strcmp(key, "tc_backward")) { h->tc_bw_off = value == 0; return CNF_OK; }
    if (!strcmp(key, "wg_lbo")) { h->wg_lbo = value; return CNF_OK; }
    if (!strcmp(key, "wg_sbo")) { h->wg_sbo = value; return CNF_OK; }
    h->err = std::string("unknown option ") + key;
    return CNF_ERR_BADARG;
}

// ------------------------------------------------------------------ workspaces
// grows the slot table towards `want` entries; `need` (<= want) is what the caller cannot do without: running
// short of memory between the two is not an error (the look-ahead just stops)
static int ensure_slots(cnf_handle* h, size_t want, size_t need = 0) {
    if (want > CNF_MAX_SLOTS) want = CNF_MAX_SLOTS;
    if (need == 0 || need > want) need = want;
    if (h->slots.size() >= want) return CNF_OK;
    size_t old = h->slots.size();
    const size_t bytes = (size_t)h->slot_elems * sizeof(float);
    while (h->slots.size() < want) {
        // keep 1/32 of the device free (the tape entries stop earlier, at 1/4): an out-of-memory here is reported
        // as CNF_ERR_OOM before any kernel of the attempt is enqueued
        size_t fr = 0, tot = 0;
        const bool tight = h->slots.size() >= 2 && cudaMemGetInfo(&fr, &tot) == cudaSuccess && fr < bytes + tot / 32;
        float* p = nullptr;
        if (tight || cudaMalloc(&p, bytes) != cudaSuccess) {
            cudaGetLastError();
            if (h->slots.size() >= need) break;
            if (!h->ktape.empty()) {
                // the stage tape is optional (the reverse sweep recomputes untaped steps): its last entry goes back to
                // the device before a solve is given up.  Kernels in flight may still write it: wait for them first.
                CNF_CUDA(cudaStreamSynchronize(h->stream));
                float* last = h->ktape.back();
                h->ktape.pop_back();
                CNF_CUDA(cudaMemsetAsync(h->d_ktape + h->ktape.size(), 0, sizeof(float*), h->stream));
                cudaFree(last);
                continue;
            }
            h->err = "state slots of the step tape do not fit the device memory (shard the batch over more GPUs)";
            return CNF_ERR_OOM;
        }
        h->slots.push_back(p);
    }
    if (h->slots.size() > old)
        CNF_CUDA(cudaMemcpyAsync(h->d_slots + old, h->slots.data() + old,
                                 (h->slots.size() - old) * sizeof(float*), cudaMemcpyHostToDevice, h->stream));
    return CNF_OK;
}

static int ensure_workspace(cnf_handle* h, long long B, int R) {
    long long Bs = round_up(B, 256);
    if (Bs <= h->cap_B && R <= h->cap_R) return CNF_OK;
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    long long capB = Bs > h->cap_B ? Bs : h->cap_B;
    int capR = R > h->cap_R ? R : h->cap_R;
    free_workspaces(h);
    const cnf_config& c = h->cfg;
    for (int j = 0; j < 7; ++j) CNF_CUDA(cudaMalloc(&h->d_K[j], (size_t)capR * capB * sizeof(float)));
    CNF_CUDA(cudaMalloc(&h->d_S, (size_t)c.D * capB * sizeof(float)));
    CNF_CUDA(cudaMalloc(&h->d_eps, (size_t)c.D * capB * sizeof(float)));
    CNF_CUDA(cudaMalloc(&h->d_ys, (size_t)(c.C > 0 ? c.C : 1) * capB * sizeof(float)));
    h->cap_B = capB;
    h->cap_R = capR;
    h->slot_elems = (long long)capR * capB;
    return ensure_slots(h, 2);
}

static int ensure_io(cnf_handle* h, long long nfloats) {
    if (nfloats <= h->cap_io) return CNF_OK;
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    cudaFree(h->d_io);
    cudaFree(h->d_vg);
    h->d_io = nullptr;
    CNF_CUDA(cudaMalloc(&h->d_io, (size_t)nfloats * sizeof(float)));
    h->cap_io = nfloats;
    return CNF_OK;
}

// bring a caller matrix (rows x B, sample-major; host or device) into an SoA buffer
static int stage_in(cnf_handle* h, const float* src, int rows, long long B, float* dst, long long Bs,
                    long long io_off) {
    const float* dsrc = src;
    if (!is_device_ptr(src)) {
        CNF_CUDA(cudaMemcpyAsync(h->d_io + io_off, src, (size_t)rows * B * sizeof(float),
                                 cudaMemcpyHostToDevice, h->stream));
        dsrc = h->d_io + io_off;
    }
    launch_pack_rows(h, dsrc, rows, B, dst, Bs);
    return CNF_OK;
}

static CnfBufs make_bufs(cnf_handle* h, long long B) {
    CnfBufs bf;
    bf.slots = h->d_slots;
    for (int j = 0; j < 7; ++j) bf.K[j] = h->d_K[j];
    bf.S = h->d_S;
    bf.eps = h->d_eps;
    bf.ys = h->d_ys;
    bf.Bs = round_up(B, 256);
    bf.B = (int)B;
    return bf;
}

// ------------------------------------------------------------------ the solve (device-driven)
static bool use_narrow(cnf_handle* h, int mode) {
    if (h->cfg.kernel_path == 1) return false;
    return narrow_supported(h->cfg, mode) && h->narrow_w != nullptr;
}

static bool use_tc(cnf_handle* h, int mode) {
    return tc_supported(h->cfg, mode) && h->tc_state != nullptr;
}
static int tc_stage(cnf_handle* h, const CnfBufs& bf, const float* src, int kidx, float* kout_fixed, int check_done) {
    stage_timer_begin(h);
    int rc = tc_eval(h, bf, src, kidx, kout_fixed, check_done);
    stage_timer_end(h);
    return rc;
}

// kernel_path 2 forces the register-resident SIMT kernel; 0 prefers the tensor-core variant where built
static bool use_ntc(cnf_handle* h, int mode) {
    return h->cfg.kernel_path != 2 && ntc_supported(h, mode);
}

static int narrow_stage(cnf_handle* h, const CnfBufs& bf, int stage, int check_done) {
    stage_timer_begin(h);
    int rc = h->path_used == 6 ? launch_stage_narrow_tc(h, bf, stage, check_done)
                               : launch_stage_narrow(h, bf, stage, check_done);
    stage_timer_end(h);
    return rc;
}

static int eval_stage(cnf_handle* h, const CnfBufs& bf, int mode, int src_sel, int kidx, int check_done) {
    stage_timer_begin(h);
    int rc = launch_eval_generic_sel(h, bf, mode, src_sel, nullptr, kidx, nullptr, check_done);
    stage_timer_end(h);
    return rc;
}

static int enqueue_attempt(cnf_handle* h, const CnfBufs& bf, int mode) {
    int rc;
    if (use_narrow(h, mode)) {   // fused path: 6 stage launches (the 7th stage carries the error norm)
        for (int i = 1; i <= 6; ++i)
            if ((rc = narrow_stage(h, bf, i, 1))) return rc;
        if (h->nranks > 1 && !h->p2p_ok && (rc = cnf_nccl_allreduce_f64(h, &h->d_ctrl->errsum, 1))) return rc;
        launch_controller(h);
        return CNF_OK;
    }
    if (use_tc(h, mode)) {       // tensor path: layer GEMMs on tcgen05, RK combination streaming
        for (int i = 1; i <= 5; ++i) {
            launch_combine(h, bf, i);
            if ((rc = tc_stage(h, bf, bf.S, i, nullptr, 1))) return rc;
        }
        launch_combine(h, bf, 6, 1);
        if ((rc = tc_stage(h, bf, bf.S, 6, nullptr, 1))) return rc;
        launch_errnorm(h, bf);
        if (h->nranks > 1 && !h->p2p_ok && (rc = cnf_nccl_allreduce_f64(h, &h->d_ctrl->errsum, 1))) return rc;
        launch_controller(h);
        return CNF_OK;
    }
    for (int i = 1; i <= 5; ++i) {
        launch_combine(h, bf, i);
        if ((rc = eval_stage(h, bf, mode, 0 /*SRC_S*/, i, 1))) return rc;
    }
    launch_combine(h, bf, 6);
    if ((rc = eval_stage(h, bf, mode, 1 /*SRC_UNEW*/, 6, 1))) return rc;
    launch_errnorm(h, bf);
    if (h->nranks > 1 && !h->p2p_ok && (rc = cnf_nccl_allreduce_f64(h, &h->d_ctrl->errsum, 1))) return rc;
    launch_controller(h);
    return CNF_OK;
}

// A rank that cannot enqueue the attempt its peers are about to run (out of memory for the next state slot) must not
// simply return: with the peer-memory exchange its peers would spin in k_controller_p2p until the 60 s trap, with NCCL
// they would block in the allreduce.  It contributes a NaN partial to that attempt's error-norm exchange instead, so
// the replicated controllers of ALL ranks end the solve at the same attempt (retcode Unstable on the peers), and then
// reports its own error.  In the NCCL variant the peers run one more (speculative, no-op) attempt before they learn
// `done`; its allreduce is matched here.
__global__ void k_poison_errsum(CnfCtrl* c) { c->errsum = __longlong_as_double(0x7ff8000000000000LL); }
static void poison_attempt(cnf_handle* h) {
    if (h->nranks <= 1) return;
    // was the solve already over after the last attempt this rank did enqueue?  (The controller is replicated: every
    // rank holds the same `done`.)  Then the peers stop after the attempt they are enqueuing now and there is no
    // further allreduce to match.
    int was_done = 0;
    cudaStreamSynchronize(h->stream);
    cudaMemcpy(&was_done, &h->d_ctrl->done, sizeof(int), cudaMemcpyDeviceToHost);
    k_poison_errsum<<<1, 1, 0, h->stream>>>(h->d_ctrl);
    if (!h->p2p_ok) cnf_nccl_allreduce_f64(h, &h->d_ctrl->errsum, 1);
    launch_controller(h);
    if (!h->p2p_ok && !was_done) cnf_nccl_allreduce_f64(h, &h->d_ctrl->errsum, 1);
    cudaStreamSynchronize(h->stream);
    cudaGetLastError();
}
// test hook: CNFB200_TEST_SLOT_FAIL="<rank>:<attempt>" makes that rank fail like an out-of-memory state slot would
static bool test_slot_fail(const cnf_handle* h, long long attempt) {
    const char* e = getenv("CNFB200_TEST_SLOT_FAIL");
    if (!e) return false;
    int r = -1;
    long long a = -1;
    return sscanf(e, "%d:%lld", &r, &a) == 2 && r == h->rank && a == attempt;
}

// state already packed into slot 0; runs the adaptive solve; fills stats (host copy of ctrl)
static int solve_device(cnf_handle* h, const CnfBufs& bf, int mode, int R, long long Bglobal,
                        float t0, float t1, bool use_tape, CnfCtrl* out) {
    const cnf_config& c = h->cfg;
    CnfCtrl ic;
    memset(&ic, 0, sizeof(ic));
    ic.qold = CNF_QOLDINIT;
    ic.nglobal = (double)R * (double)Bglobal;
    ic.maxiters = c.maxiters > 0 ? c.maxiters : CNF_DEFAULT_MAXITERS;
    ic.t = ic.t0 = t0;
    ic.t1 = t1;
    ic.tdir = (t1 >= t0) ? 1.0f : -1.0f;
    ic.dtmax = c.dtmax > 0 ? c.dtmax : fabsf(t1 - t0);
    ic.dtmin = c.dtmin;
    ic.reltol = c.reltol > 0 ? c.reltol : CNF_DEFAULT_RELTOL;
    ic.abstol = c.abstol > 0 ? c.abstol : CNF_DEFAULT_ABSTOL;
    ic.use_tape = use_tape ? 1 : 0;
    ic.R = R;
    ic.D = c.D;
    ic.nf = 1;
    ic.stiff_max = -1.0f;   // stays -1 on kernel paths that do not accumulate the estimate (narrow register kernels)
    if (c.dt0 > 0) ic.dt = ic.tdir * c.dt0;
    h->h_ctrl[0] = ic;
    CNF_CUDA(cudaMemcpyAsync(h->d_ctrl, &h->h_ctrl[0], sizeof(CnfCtrl), cudaMemcpyHostToDevice, h->stream));
    CNF_CUDA(cudaEventRecord(h->ev1, h->stream));
    int rc;
    const bool nar = use_narrow(h, mode);
    const bool tcp = !nar && use_tc(h, mode);
    h->tc_mode = mode;
    h->path_used = nar ? (use_ntc(h, mode) ? 6 : 2) : (tcp ? 3 : 1);
    if (nar) rc = narrow_stage(h, bf, 0, 0);
    else if (tcp) rc = tc_stage(h, bf, h->slots[0], 0, nullptr, 0);
    else rc = eval_stage(h, bf, mode, 2 /*SRC_SLOT0*/, 0, 0);   // k1 = f(u0)
    if (rc) return rc;
    if (c.dt0 > 0) {
        launch_initdt(h, 2);
    } else {
        launch_initnorm(h, bf, 0);
        if (h->nranks > 1 && (rc = cnf_nccl_allreduce_f64(h, h->d_ctrl->initsum, 2))) return rc;
        launch_initdt(h, 0);
        if (nar) {
            if ((rc = narrow_stage(h, bf, -1, 0))) return rc;
        } else if (tcp) {
            launch_combine(h, bf, -1);
            if ((rc = tc_stage(h, bf, bf.S, -1, bf.K[1], 0))) return rc;
        } else {
            launch_combine(h, bf, -1);
            if ((rc = launch_eval_generic_sel(h, bf, mode, 0 /*SRC_S*/, nullptr, -1, bf.K[1], 0))) return rc;
        }
        launch_initnorm(h, bf, 1);
        if (h->nranks > 1 && (rc = cnf_nccl_allreduce_f64(h, h->d_ctrl->initsum + 2, 1))) return rc;
        launch_initdt(h, 1);
    }
    if (!nar && !tcp && tile_solve_ok(h, mode, bf.B)) {
        // one-tile batch: the whole attempt loop runs inside one single-CTA kernel (cnf_tile.cu)
        CnfCtrl tmp;
        if ((rc = tile_solve(h, bf, mode, use_tape, ic.maxiters, &tmp))) return rc;
        h->path_used = 5;
    } else {
    // attempts are enqueued speculatively; kernels of attempts after `done` return immediately
    long long launched = 0;
    int pend = -1;  // index of the pinned snapshot in flight
    for (;;) {
        if (use_tape && (long long)h->slots.size() < launched + 2 && (rc = ensure_slots(h, (size_t)(launched + 10), (size_t)(launched + 2)))) {
            poison_attempt(h);
            return rc;
        }
        if (test_slot_fail(h, launched)) {
            poison_attempt(h);
            h->err = "state slots of the step tape do not fit the device memory (simulated: CNFB200_TEST_SLOT_FAIL)";
            return CNF_ERR_OOM;
        }
        if (use_tape && (long long)h->ktape.size() < launched + 2 &&
            (rc = ensure_ktape(h, (size_t)(launched + 10), (long long)c.D * bf.Bs))) return rc;
        if ((rc = enqueue_attempt(h, bf, mode))) return rc;
        if (use_tape && !h->ktape.empty()) launch_tape_stages(h, bf);
        launched++;
        int cur = (int)(launched & 1);
        CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[cur], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
        CNF_CUDA(cudaEventRecord(h->evchk[cur], h->stream));
        if (pend >= 0) {
            CNF_CUDA(cudaEventSynchronize(h->evchk[pend]));
            if (h->h_ctrl[pend].done) break;
        }
        pend = cur;
        if (launched > ic.maxiters + 2) break;
    }
    }
    CNF_CUDA(cudaEventRecord(h->ev2, h->stream));
    CNF_CUDA(cudaMemcpyAsync(&h->h_ctrl[0], h->d_ctrl, sizeof(CnfCtrl), cudaMemcpyDeviceToHost, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    *out = h->h_ctrl[0];
    CNF_CUDA(cudaGetLastError());
    if (getenv("CNFB200_DEBUG_DTS")) {   // debug: the accepted step sizes of this solve
        std::vector<float> dts((size_t)(out->naccept < CNF_MAX_SLOTS ? out->naccept : CNF_MAX_SLOTS));
        cudaMemcpy(dts.data(), h->d_dts, dts.size() * sizeof(float), cudaMemcpyDeviceToHost);
        fprintf(stderr, "cnfb200 dts (naccept %lld nreject %lld):", (long long)out->naccept, (long long)out->nreject);
        for (float d : dts) fprintf(stderr, " %.7g", d);
        fprintf(stderr, "\n");
    }
    return CNF_OK;
}

static void fill_stats(cnf_handle* h, const CnfCtrl& c, cnf_stats* st) {
    if (!st) return;
    memset(st, 0, sizeof(*st));
    st->naccept = c.naccept;
    st->nreject = c.nreject;
    st->nf = c.nf;
    st->retcode = c.retcode;
    st->kernel_path = h->path_used;
    st->t_final = c.t;
    st->dt_last = c.dt_last;
    st->stiff_max = c.stiff_max;
    st->stiff_run = c.stiff_run;
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev3) == cudaSuccess) st->ms_total = ms;
    if (cudaEventElapsedTime(&ms, h->ev1, h->ev2) == cudaSuccess) st->ms_solve = ms;
    cudaGetLastError();
    st->gpu_launches = h->launches;
    stage_timer_collect(h);
    st->ms_stage = h->ms_stage;
    st->n_stage = h->n_stage;
}

// ------------------------------------------------------------------ shared front/back ends
struct IoPlan {
    long long off_xs, off_ys, off_eps, off_logp, off_regs, off_z, total;
};
static IoPlan plan_io(const cnf_config& c, long long B, int rows_x) {
    IoPlan p;
    long long o = 0;
    p.off_xs = o; o += (long long)rows_x * B;
    p.off_ys = o; o += (long long)c.C * B;
    p.off_eps = o; o += (long long)c.D * B;
    p.off_logp = o; o += B;
    p.off_regs = o; o += 3 * B;
    p.off_z = o; o += (long long)c.D * B;
    p.total = o;
    return p;
}

// pack [x; 0] into slot 0, eps and ys into SoA buffers
static int load_problem(cnf_handle* h, const CnfBufs& bf, const IoPlan& io, int mode, int R,
                        const float* x, int rows_x, const float* ys, const float* eps, long long B) {
    const cnf_config& c = h->cfg;
    int rc;
    float* slot0 = h->slots[0];
    CNF_CUDA(cudaMemsetAsync(slot0, 0, (size_t)R * bf.Bs * sizeof(float), h->stream));
    if (x) {
        if ((rc = stage_in(h, x, rows_x, B, slot0, bf.Bs, io.off_xs))) return rc;
    } else {   // generate with z0 = NULL: the base sample is drawn here (N1)
        if (!h->rng_on) { h->err = "null z0 and no device RNG seeded (cnfb200_seed_rng)"; return CNF_ERR_BADARG; }
        launch_philox_normal(h, slot0, bf.Bs, B, rows_x, h->rng_draws++);
    }
    if (c.C > 0) {
        if (!ys) { h->err = "conditional flow: ys is required"; return CNF_ERR_BADARG; }
        if ((rc = stage_in(h, ys, c.C, B, h->d_ys, bf.Bs, io.off_ys))) return rc;
    }
    if (mode == CNF_MODE_TRAIN) {
        if (eps) {
            if ((rc = stage_in(h, eps, c.D, B, h->d_eps, bf.Bs, io.off_eps))) return rc;
        } else {
            if (!h->rng_on) { h->err = "TrainMode: eps is required (or seed the device RNG: cnfb200_seed_rng)"; return CNF_ERR_BADARG; }
            launch_philox_normal(h, h->d_eps, bf.Bs, B, c.D, h->rng_draws++);
        }
        h->eps_B = B;
    }
    return CNF_OK;
}

static int check_call(cnf_handle* h, int mode, long long B) {
    if (!h) return CNF_ERR_BADARG;
    if (!h->have_params) { h->err = "set_params has not been called"; return CNF_ERR_BADARG; }
    if (mode != CNF_MODE_TEST && mode != CNF_MODE_TRAIN) { h->err = "bad mode"; return CNF_ERR_BADARG; }
    if (B < 1 || B > 0x7fffff00LL) { h->err = "bad batch size"; return CNF_ERR_BADARG; }
    return CNF_OK;
}

static long long global_batch(cnf_handle* h, long long B, int* rc) {
    *rc = CNF_OK;
    if (h->nranks <= 1) return B;
    if (h->bglobal_B == B && h->bglobal > 0) return h->bglobal;   // same shard size as the last call on this communicator
    // sum of B and of B^2 over the ranks: n * sum(B^2) == sum(B)^2 iff every rank holds the same B, and only then
    // may every rank skip the exchange next time (all ranks must take the cached branch together)
    double* d = &h->d_ctrl->initsum[2];
    double v[2] = {(double)B, (double)B * (double)B};
    cudaMemcpyAsync(d, v, 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream);
    cudaStreamSynchronize(h->stream);
    *rc = cnf_nccl_allreduce_f64(h, d, 2);
    cudaMemcpyAsync(v, d, 2 * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    cudaStreamSynchronize(h->stream);
    const bool uniform = (double)h->nranks * v[1] == v[0] * v[0];
    h->bglobal_B = uniform ? B : -1;
    h->bglobal = (long long)(v[0] + 0.5);
    return h->bglobal;
}

extern "C" int32_t cnfb200_augmented_f(cnf_handle* h, int32_t mode, const float* u, const float* ys,
                                       const float* eps, int64_t B, float* du) {
    int rc = check_call(h, mode, B);
    if (rc) return rc;
    if (!u || !du) { h->err = "null u/du"; return CNF_ERR_BADARG; }
    CNF_CUDA(cudaSetDevice(h->device));
    const cnf_config& c = h->cfg;
    const int R = c.D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    h->cur_R = R;
    if ((rc = ensure_workspace(h, B, R))) return rc;
    IoPlan io = plan_io(c, B, R);
    if ((rc = ensure_io(h, io.total + (long long)R * B))) return rc;
    CnfBufs bf = make_bufs(h, B);
    if ((rc = load_problem(h, bf, io, mode, R, u, R, ys, eps, B))) return rc;
    h->path_used = 1;
    if (use_tc(h, mode)) {
        h->path_used = 3;
        h->tc_mode = mode;
        if ((rc = tc_eval(h, bf, h->slots[0], -1, h->d_K[0], 0))) return rc;
    } else if ((rc = launch_eval_generic(h, bf, mode, h->slots[0], h->d_K[0]))) return rc;
    float* dout = is_device_ptr(du) ? du : h->d_io + io.total;
    launch_unpack_rows(h, h->d_K[0], R, B, bf.Bs, dout);
    if (dout != du)
        CNF_CUDA(cudaMemcpyAsync(du, dout, (size_t)R * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    CNF_CUDA(cudaGetLastError());
    return CNF_OK;
}

// common body of logdensity / generate: x has rows_x rows (nvars for inference, D for generate)
static int run_flow(cnf_handle* h, int mode, const float* x, int rows_x, const float* ys,
                    const float* eps, long long B, float t0, float t1, float* logp, float* regs,
                    float* z_out, int z_rows, cnf_stats* stats) {
    int rc = check_call(h, mode, B);
    if (rc) return rc;
    CNF_CUDA(cudaSetDevice(h->device));
    const cnf_config& c = h->cfg;
    const int R = c.D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    h->cur_R = R;
    h->launches = 0;
    if ((rc = ensure_workspace(h, B, R))) return rc;
    IoPlan io = plan_io(c, B, rows_x);
    if ((rc = ensure_io(h, io.total))) return rc;
    CNF_CUDA(cudaEventRecord(h->ev0, h->stream));   // the device time of the call covers the batch-size exchange
    long long Bglobal = global_batch(h, B, &rc);
    if (rc) return rc;
    CnfBufs bf = make_bufs(h, B);
    if ((rc = load_problem(h, bf, io, mode, R, x, rows_x, ys, eps, B))) return rc;
    h->path_used = 1;
    CnfCtrl fin;
    if (h->vector_mode) {   // N4: every sample its own adaptive solve (no cross-sample coupling, no allreduce)
        if (h->nranks > 1) Bglobal = B;   // no cross-sample coupling: nothing is exchanged between ranks
        if (vector_narrow_ok(h, mode)) rc = vector_solve(h, bf, t0, t1, &fin);            // one launch, registers
        else rc = vector_solve_generic(h, bf, mode, R, t0, t1, &fin);                      // any net / mode / width
        if (rc) return rc;
    } else if ((rc = solve_device(h, bf, mode, R, Bglobal, t0, t1, false, &fin))) return rc;
    // epilogue writes in the caller's layout; host destinations go through the staging area
    const int nreg = mode == CNF_MODE_TRAIN ? 3 : 1;
    float* d_logp = logp ? (is_device_ptr(logp) ? logp : h->d_io + io.off_logp) : nullptr;
    float* d_regs = regs ? (is_device_ptr(regs) ? regs : h->d_io + io.off_regs) : nullptr;
    float* d_z = nullptr;
    const bool gen = z_rows < 0;   // generate: return rows 1:(-z_rows) of the final state
    if (gen) z_rows = -z_rows;
    if (z_out && !gen) d_z = is_device_ptr(z_out) ? z_out : h->d_io + io.off_z;
    CNF_CUDA(cudaMemsetAsync(h->d_lossacc, 0, sizeof(double), h->stream));
    if (logp || regs || d_z) launch_epilogue(h, bf, mode, d_logp, d_regs, d_z, Bglobal);
    if (z_out && gen) {
        float* dz = is_device_ptr(z_out) ? z_out : h->d_io + io.off_z;
        const float* fs = h->slots[fin.naccept & 1];
        launch_unpack_rows(h, fs, z_rows, B, bf.Bs, dz);
        if (dz != z_out)
            CNF_CUDA(cudaMemcpyAsync(z_out, dz, (size_t)z_rows * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    }
    if (logp && d_logp != logp)
        CNF_CUDA(cudaMemcpyAsync(logp, d_logp, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (regs && d_regs != regs)
        CNF_CUDA(cudaMemcpyAsync(regs, d_regs, (size_t)nreg * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    if (d_z && d_z != z_out)
        CNF_CUDA(cudaMemcpyAsync(z_out, d_z, (size_t)c.D * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CNF_CUDA(cudaEventRecord(h->ev3, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    CNF_CUDA(cudaGetLastError());
    if (h->vector_mode) fin.naccept = fin.iter;   // reported: the maxima over the samples
    fill_stats(h, fin, stats);
    return fin.retcode;
}

extern "C" int32_t cnfb200_logdensity(cnf_handle* h, int32_t mode, const float* xs, const float* ys,
                                      const float* eps, int64_t B, float t0, float t1, float* logp,
                                      float* regs, float* z_out, cnf_stats* stats) {
    if (!h) return CNF_ERR_BADARG;
    if (!xs) { h->err = "null xs"; return CNF_ERR_BADARG; }
    return run_flow(h, mode, xs, h->cfg.nvars, ys, eps, B, t0, t1, logp, regs, z_out, h->cfg.D, stats);
}

extern "C" int32_t cnfb200_generate(cnf_handle* h, int32_t mode, const float* z0, const float* ys,
                                    const float* eps, int64_t n, float t_from, float t_to,
                                    float* xs_out, cnf_stats* stats) {
    if (!h) return CNF_ERR_BADARG;
    if (!xs_out) { h->err = "null xs_out"; return CNF_ERR_BADARG; }
    if (!z0 && !h->rng_on) { h->err = "null z0 and no device RNG seeded (cnfb200_seed_rng)"; return CNF_ERR_BADARG; }
    return run_flow(h, mode, z0, h->cfg.D, ys, eps, n, t_from, t_to, nullptr, nullptr, xs_out,
                    -h->cfg.nvars, stats);
}

// ------------------------------------------------------------------ N1: device-side draws
extern "C" int32_t cnfb200_seed_rng(cnf_handle* h, uint64_t seed, uint64_t sample_offset) {
    if (!h) return CNF_ERR_BADARG;
    h->rng_on = true;
    h->rng_seed = seed;
    h->rng_offset = sample_offset;
    h->rng_draws = 0;
    return CNF_OK;
}

extern "C" int32_t cnfb200_last_eps(cnf_handle* h, float* eps_out, int64_t B) {
    if (!h || !eps_out) return CNF_ERR_BADARG;
    if (B < 1 || B != h->eps_B || !h->d_eps) { h->err = "last_eps: no eps of that batch size on the device"; return CNF_ERR_BADARG; }
    CNF_CUDA(cudaSetDevice(h->device));
    const cnf_config& c = h->cfg;
    const long long Bs = round_up(B, 256);
    if (is_device_ptr(eps_out)) {
        launch_unpack_rows(h, h->d_eps, c.D, B, Bs, eps_out);
    } else {
        int rc;
        if ((rc = ensure_io(h, (long long)c.D * B))) return rc;
        launch_unpack_rows(h, h->d_eps, c.D, B, Bs, h->d_io);
        CNF_CUDA(cudaMemcpyAsync(eps_out, h->d_io, (size_t)c.D * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    }
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    return CNF_OK;
}

// ------------------------------------------------------------------ loss + discrete adjoint
static int ensure_bw(cnf_handle* h, long long Bs) {
    const long long need = (long long)h->cfg.D * Bs;
    if (need <= h->cap_bw) return CNF_OK;
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    for (int j = 0; j < 13; ++j) {   // [12]: zbar scratch of the tensor-path reverse sweep; [13], [14]: allocated by the exact-trace tensor sweep only
        cudaFree(h->d_bw[j]);
        h->d_bw[j] = nullptr;
        CNF_CUDA(cudaMalloc(&h->d_bw[j], (size_t)need * sizeof(float)));
    }
    for (int j = 13; j < 15; ++j) { cudaFree(h->d_bw[j]); h->d_bw[j] = nullptr; }
    h->cap_bw13 = 0;
    h->cap_bw = need;
    return CNF_OK;
}

// the problem is loaded (slot 0, d_eps, d_ys): forward solve with the step tape, loss sum into
// h->d_lossacc (all ranks), and — want_grad — the reverse sweep: grad theta (summed over ranks) in
// h->d_grad, grad of the initial state in h->d_bw[11]
static int loss_grad_core(cnf_handle* h, const CnfBufs& bf, int mode, int R, long long Bglobal, float t0, float t1,
                          bool want_grad, CnfCtrl* fin) {
    int rc;
    h->path_used = 1;
    if ((rc = solve_device(h, bf, mode, R, Bglobal, t0, t1, want_grad, fin))) return rc;
    CNF_CUDA(cudaMemsetAsync(h->d_lossacc, 0, sizeof(double), h->stream));
    launch_epilogue(h, bf, mode, nullptr, nullptr, nullptr, Bglobal);
    if (h->nranks > 1 && (rc = cnf_nccl_allreduce_f64(h, h->d_lossacc, 1))) return rc;
    if (want_grad) {
        CNF_CUDA(cudaMemsetAsync(h->d_grad, 0, (size_t)h->nparams * sizeof(float), h->stream));
        launch_init_ubar(h, bf, h->d_bw[11], Bglobal, mode);
        if ((rc = backward_sweep(h, bf, mode, fin->naccept, Bglobal))) return rc;
        if (h->nranks > 1 && (rc = cnf_nccl_allreduce_f32(h, h->d_grad, h->nparams))) return rc;
    }
    return CNF_OK;
}

extern "C" int32_t cnfb200_loss_grad_mode(cnf_handle* h, int32_t mode, const float* xs, const float* ys,
                                          const float* eps, int64_t B, float t0, float t1, float* loss,
                                          float* grad_theta, float* grad_x, cnf_stats* stats) {
    int rc = check_call(h, mode, B);
    if (rc) return rc;
    if (!xs || !loss) { h->err = "null xs/loss"; return CNF_ERR_BADARG; }
    if (h->vector_mode && (grad_theta || grad_x)) {
        h->err = "VectorMode: the reverse sweep is built for MatrixMode only (use loss without gradients, or a MatrixMode flow)";
        return CNF_ERR_UNSUPPORTED;
    }
    CNF_CUDA(cudaSetDevice(h->device));
    const cnf_config& c = h->cfg;
    const int R = c.D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    const bool want_grad = grad_theta || grad_x;
    h->cur_R = R;
    h->launches = 0;
    if ((rc = ensure_workspace(h, B, R))) return rc;
    IoPlan io = plan_io(c, B, c.nvars);
    if ((rc = ensure_io(h, io.total))) return rc;
    CnfBufs bf = make_bufs(h, B);
    if (want_grad && (rc = ensure_bw(h, bf.Bs))) return rc;
    CNF_CUDA(cudaEventRecord(h->ev0, h->stream));
    long long Bglobal = global_batch(h, B, &rc);
    if (rc) return rc;
    if ((rc = load_problem(h, bf, io, mode, R, xs, c.nvars, ys, eps, B))) return rc;
    CnfCtrl fin;
    if ((rc = loss_grad_core(h, bf, mode, R, Bglobal, t0, t1, want_grad, &fin))) return rc;
    double lsum = 0.0;
    CNF_CUDA(cudaMemcpyAsync(&lsum, h->d_lossacc, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    if (want_grad) {
        if (grad_theta) {
            CNF_CUDA(cudaMemcpyAsync(grad_theta, h->d_grad, (size_t)h->nparams * sizeof(float),
                                     cudaMemcpyDefault, h->stream));
        }
        if (grad_x) {
            float* dg = is_device_ptr(grad_x) ? grad_x : h->d_io + io.off_z;
            launch_unpack_rows(h, h->d_bw[11], c.nvars, B, bf.Bs, dg);
            if (dg != grad_x)
                CNF_CUDA(cudaMemcpyAsync(grad_x, dg, (size_t)c.nvars * B * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
        }
    }
    CNF_CUDA(cudaEventRecord(h->ev3, h->stream));
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    CNF_CUDA(cudaGetLastError());
    *loss = (float)(lsum / (double)Bglobal);
    fill_stats(h, fin, stats);
    return fin.retcode;
}

extern "C" int32_t cnfb200_loss_grad(cnf_handle* h, const float* xs, const float* ys, const float* eps,
                                     int64_t B, float t0, float t1, float* loss, float* grad_theta,
                                     float* grad_x, cnf_stats* stats) {
    return cnfb200_loss_grad_mode(h, CNF_MODE_TRAIN, xs, ys, eps, B, t0, t1, loss, grad_theta, grad_x, stats);
}

// ------------------------------------------------------------------ FP32 FFMA peak probe
__global__ void k_ffma_peak(float* out, int iters) {
    float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
    float a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
    const float m = 0.999f, c = 1e-3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
            a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

extern "C" int32_t cnfb200_host_alloc(size_t bytes, void** out) {
    if (!out || bytes == 0) return CNF_ERR_BADARG;
    *out = nullptr;
    if (cudaHostAlloc(out, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); *out = nullptr; return CNF_ERR_OOM; }
    return CNF_OK;
}
extern "C" int32_t cnfb200_host_free(void* p) {
    if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return CNF_ERR_CUDA; }
    return CNF_OK;
}

extern "C" int32_t cnfb200_measure_fp32_peak(int32_t device, double* tflops_out) {
    if (!tflops_out) return CNF_ERR_BADARG;
    if (cudaSetDevice(device) != cudaSuccess) return CNF_ERR_CUDA;
    const int blocks = 148 * 8, threads = 256, iters = 4096;
    float* d = nullptr;
    if (cudaMalloc(&d, (size_t)blocks * threads * sizeof(float)) != cudaSuccess) return CNF_ERR_OOM;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a);
        k_ffma_peak<<<blocks, threads>>>(d, iters);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, a, b);
        double fl = 2.0 * 128.0 * iters * (double)blocks * threads;
        double tf = fl / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(d);
    *tflops_out = best;
    return cudaGetLastError() == cudaSuccess ? CNF_OK : CNF_ERR_CUDA;
}
