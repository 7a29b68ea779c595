// cnf_nccl.cu — NCCL binding by dlopen (the library loads without NCCL for single-GPU use;
// in a torch process this resolves to the libnccl.so.2 torch already loaded).
#include "cnf_internal.h"
#include <dlfcn.h>

typedef struct { char internal[128]; } ncclUniqueId_t;
typedef int (*fn_getid)(ncclUniqueId_t*);
typedef int (*fn_init)(void**, int, ncclUniqueId_t, int);
typedef int (*fn_allreduce)(const void*, void*, size_t, int, int, void*, cudaStream_t);
typedef int (*fn_destroy)(void*);
typedef const char* (*fn_errstr)(int);

static struct {
    void* lib = nullptr;
    fn_getid getid = nullptr;
    fn_init init = nullptr;
    fn_allreduce allreduce = nullptr;
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
    return CNF_OK;
}

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
    if (h->nccl_comm && g_nccl.destroy) g_nccl.destroy(h->nccl_comm);
    h->nccl_comm = nullptr;
}
