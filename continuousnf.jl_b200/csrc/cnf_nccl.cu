// cnf_nccl.cu — NCCL binding by dlopen (the library loads without NCCL for single-GPU use;
// in a torch process this resolves to the libnccl.so.2 torch already loaded).
#include "cnf_internal.h"
#include <dlfcn.h>
#include <stdlib.h>

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef int (*fn_getid)(ncclUniqueId_t*);
typedef int (*fn_init)(void**, int, ncclUniqueId_t, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_allgather)(const void*, void*, size_t, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);

static struct {
    void* lib = nullptr;
    fn_getid getid = nullptr;
    fn_init init = nullptr;
    fn_allreduce allreduce = nullptr;
    fn_allgather allgather = nullptr;
    fn_destroy destroy = nullptr;
    fn_errstr errstr = nullptr;
} g_nccl;

static bool nccl_load(std::string* err) {
    if (g_nccl.lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib) break;
    }
    if (!g_nccl.lib) { if (err) *err = std::string("dlopen libnccl failed: ") + dlerror(); return false; }
    g_nccl.getid = (fn_getid)dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.init = (fn_init)dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.allreduce = (fn_allreduce)dlsym(g_nccl.lib, "ncclAllReduce");
    g_nccl.allgather = (fn_allgather)dlsym(g_nccl.lib, "ncclAllGather");
    g_nccl.destroy = (fn_destroy)dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.errstr = (fn_errstr)dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.getid || !g_nccl.init || !g_nccl.allreduce) {
        if (err) *err = "libnccl: missing symbols";
        return false;
    }
    return true;
}

extern "C" int32_t cnfb200_comm_unique_id(char id_out[128]) {
    if (!id_out || !nccl_load(nullptr)) return CNF_ERR_NCCL;
    ncclUniqueId_t id;
    if (g_nccl.getid(&id) != 0) return CNF_ERR_NCCL;
    memcpy(id_out, id.internal, 128);
    return CNF_OK;
}


// ---- error-norm exchange over peer memory.  The per-attempt collective of this path is one double per
// rank; instead of an ncclAllReduce launch between the 7th stage kernel and the controller, the controller
// kernel itself stores its partial into every peer's mailbox (P2P stores over NVLink) and spins on its own
// mailbox.  Mailboxes are plain cudaMalloc memory shared through CUDA IPC; the handles travel by
// ncclAllGather.  Best effort: if any rank cannot set it up, all ranks keep the NCCL allreduce
// (the decision itself is an allreduce, so it is the same everywhere).  CNFB200_P2P=0 disables it.
void cnf_p2p_free(cnf_handle* h) {
    for (void* p : h->xpeer_maps) cudaIpcCloseMemHandle(p);
    h->xpeer_maps.clear();
    cudaFree(h->d_xslots); cudaFree(h->d_xpeers); cudaFree(h->d_xseq);
    h->d_xslots = nullptr; h->d_xpeers = nullptr; h->d_xseq = nullptr;
    h->p2p_ok = false;
}

static int cnf_p2p_setup(cnf_handle* h) {
    h->p2p_ok = false;
    const int n = h->nranks;
    if (n <= 1 || n > 32 || !g_nccl.allgather) return CNF_OK;
    const char* env = getenv("CNFB200_P2P");
    int fail = (env && env[0] == '0') ? 1 : 0;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    unsigned char* d_h = nullptr;     // [n + 1] handles: gathered, then mine
    int* d_flag = nullptr;
    std::vector<cudaIpcMemHandle_t> all((size_t)n);
    std::vector<CnfXSlot*> peers((size_t)n, nullptr);
    const size_t hs = sizeof(cudaIpcMemHandle_t);
    if (cudaMalloc(&d_h, hs * (n + 1)) != cudaSuccess || cudaMalloc(&d_flag, sizeof(int)) != cudaSuccess) {
        cudaFree(d_h); cudaFree(d_flag); cudaGetLastError();
        h->err = "p2p setup: cudaMalloc failed";
        return CNF_ERR_CUDA;   // nothing collective has happened yet on any rank that fails here
    }
    if (!fail) {
        if (cudaMalloc(&h->d_xslots, sizeof(CnfXSlot) * 2 * n) != cudaSuccess ||
            cudaMalloc(&h->d_xpeers, sizeof(CnfXSlot*) * n) != cudaSuccess ||
            cudaMalloc(&h->d_xseq, sizeof(unsigned long long)) != cudaSuccess ||
            cudaMemset(h->d_xslots, 0, sizeof(CnfXSlot) * 2 * n) != cudaSuccess ||
            cudaMemset(h->d_xseq, 0, sizeof(unsigned long long)) != cudaSuccess ||
            cudaIpcGetMemHandle(&mine, h->d_xslots) != cudaSuccess)
            fail = 1;
    }
    cudaGetLastError();
    // 1. everybody's handle (ncclInt8 = 0)
    cudaMemcpyAsync(d_h + hs * n, &mine, hs, cudaMemcpyHostToDevice, h->stream);
    int rc = g_nccl.allgather(d_h + hs * n, d_h, hs, 0, h->nccl_comm, h->stream);
    cudaMemcpyAsync(all.data(), d_h, hs * n, cudaMemcpyDeviceToHost, h->stream);
    if (rc != 0 || cudaStreamSynchronize(h->stream) != cudaSuccess) fail = 1;
    // 2. map the peers' mailboxes
    if (!fail) {
        for (int r = 0; r < n && !fail; ++r) {
            if (r == h->rank) { peers[r] = h->d_xslots; continue; }
            void* p = nullptr;
            if (cudaIpcOpenMemHandle(&p, all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { fail = 1; break; }
            h->xpeer_maps.push_back(p);
            peers[r] = (CnfXSlot*)p;
        }
        if (!fail && cudaMemcpyAsync(h->d_xpeers, peers.data(), sizeof(CnfXSlot*) * n, cudaMemcpyHostToDevice, h->stream) != cudaSuccess)
            fail = 1;
    }
    cudaGetLastError();
    // 3. the same decision on every rank: sum of the failure flags (ncclInt32 = 2, ncclSum = 0)
    int total = 1;
    cudaMemcpyAsync(d_flag, &fail, sizeof(int), cudaMemcpyHostToDevice, h->stream);
    rc = g_nccl.allreduce(d_flag, d_flag, 1, 2, 0, h->nccl_comm, h->stream);
    cudaMemcpyAsync(&total, d_flag, sizeof(int), cudaMemcpyDeviceToHost, h->stream);
    if (rc != 0 || cudaStreamSynchronize(h->stream) != cudaSuccess) total = 1;
    cudaFree(d_h); cudaFree(d_flag);
    cudaGetLastError();
    if (total == 0) h->p2p_ok = true;
    else cnf_p2p_free(h);
    return CNF_OK;
}

extern "C" int32_t cnfb200_comm_init(cnf_handle* h, int32_t nranks, int32_t rank, const char id[128]) {
    if (!h || !id || nranks < 1 || rank < 0 || rank >= nranks) return CNF_ERR_BADARG;
    if (!nccl_load(&h->err)) return CNF_ERR_NCCL;
    CNF_CUDA(cudaSetDevice(h->device));
    ncclUniqueId_t uid;
    memcpy(uid.internal, id, 128);
    int rc = g_nccl.init(&h->nccl_comm, nranks, uid, rank);
    if (rc != 0) {
        h->err = std::string("ncclCommInitRank: ") + (g_nccl.errstr ? g_nccl.errstr(rc) : "error");
        return CNF_ERR_NCCL;
    }
    h->nranks = nranks;
    h->rank = rank;
    h->bglobal_B = -1;
    return cnf_p2p_setup(h);
}

extern "C" int32_t cnfb200_comm_p2p(const cnf_handle* h) { return (h && h->p2p_ok) ? 1 : 0; }

// ncclDataType: float32 = 7, float64 = 8 ; ncclSum = 0
int cnf_nccl_allreduce_f64(cnf_handle* h, double* dptr, int n) {
    if (h->nranks <= 1) return CNF_OK;
    int rc = g_nccl.allreduce(dptr, dptr, (size_t)n, 8, 0, h->nccl_comm, h->stream);
    if (rc != 0) { h->err = std::string("ncclAllReduce: ") + (g_nccl.errstr ? g_nccl.errstr(rc) : "error"); return CNF_ERR_NCCL; }
    h->launches++;
    return CNF_OK;
}
int cnf_nccl_allreduce_f32(cnf_handle* h, float* dptr, long long n) {
    if (h->nranks <= 1) return CNF_OK;
    int rc = g_nccl.allreduce(dptr, dptr, (size_t)n, 7, 0, h->nccl_comm, h->stream);
    if (rc != 0) { h->err = std::string("ncclAllReduce: ") + (g_nccl.errstr ? g_nccl.errstr(rc) : "error"); return CNF_ERR_NCCL; }
    h->launches++;
    return CNF_OK;
}
void cnf_nccl_destroy(cnf_handle* h) {
    cnf_p2p_free(h);
    if (h->nccl_comm && g_nccl.destroy) g_nccl.destroy(h->nccl_comm);
    h->nccl_comm = nullptr;
}
