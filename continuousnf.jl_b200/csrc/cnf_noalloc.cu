// cnf_noalloc.cu — cnfb200_augmented_f_static: one augmented_f evaluation that allocates nothing.
//
// No counterpart in the reference: added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// cnfb200_augmented_f (cnf_api.cu) on the generic SIMT evaluation (cnf_stage_generic.cu): the same three steps — pack
// the caller's sample-major rows into SoA planes, evaluate a tile of samples per CTA, unpack the R output rows —
// but the SoA planes live in ONE file-scope device array instead of the handle's lazily grown workspaces, and the
// caller's pointers must already be device pointers (no staging copies).  The call makes no cudaMalloc / cudaFree and
// calls no allocating library routine; the parameters it reads were uploaded by cnfb200_set_params beforehand.
//
//  * the array is CNFB200_STATIC_SCRATCH_BYTES (16 MiB) and every GPU that loads the library holds it until exit;
//  * a call needs (2 D + C + R) * round_up(B, 256) floats of it and is refused with CNF_ERR_BADARG beyond that;
//  * every call (and every handle) on a device shares the array and it keeps the last call's contents: the pack
//    kernels rewrite every plane a call reads, padding columns included, so nothing has to start at zero — but two
//    calls that overlap on the GPU (two handles = two streams, or two host threads) race.  The entry point waits for
//    its own work before returning, so calls made one after another from one thread never overlap.
#include "cnf_internal.h"
#include <mutex>

static __device__ __align__(256) float g_static_scratch[CNFB200_STATIC_SCRATCH_BYTES / sizeof(float)];

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked up
// once per device, with that device current, and that pointer is what the kernels are given.
static float* g_static_ptr[64] = {};
static std::mutex g_static_mu;

static int static_scratch(cnf_handle* h, float** out) {
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (h->device < 0 || h->device >= 64) { h->err = "augmented_f_static: device ordinal out of range"; return CNF_ERR_BADARG; }
    if (!g_static_ptr[h->device]) {
        CNF_CUDA(cudaSetDevice(h->device));
        void* p = nullptr;
        CNF_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) { h->err = "augmented_f_static: cudaGetSymbolAddress returned a null address"; return CNF_ERR_CUDA; }
        g_static_ptr[h->device] = (float*)p;
    }
    *out = g_static_ptr[h->device];
    return CNF_OK;
}

// caller layout (sample-major, `ld` floats per sample) -> the first `rows` rows as SoA planes; padding columns are
// written too (zeros), so no plane depends on what the previous call left behind
__global__ void k_static_pack(const float* __restrict__ src, int ld, int rows, long long B, float* __restrict__ dst, long long Bs) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= Bs) return;
    for (int r = 0; r < rows; ++r) dst[(long long)r * Bs + b] = (b < B) ? src[b * ld + r] : 0.0f;
}
__global__ void k_static_unpack(const float* __restrict__ src, int rows, long long B, long long Bs, float* __restrict__ dst) {
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    for (int r = 0; r < rows; ++r) dst[b * rows + r] = src[(long long)r * Bs + b];
}

static bool static_is_device(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" int32_t cnfb200_augmented_f_static(cnf_handle* h, int32_t mode, const float* u, const float* ys,
                                              const float* eps, int64_t B, float* du) {
    if (!h) return CNF_ERR_BADARG;
    if (!h->have_params) { h->err = "set_params has not been called"; return CNF_ERR_BADARG; }
    if (mode != CNF_MODE_TEST && mode != CNF_MODE_TRAIN) { h->err = "bad mode"; return CNF_ERR_BADARG; }
    const cnf_config& c = h->cfg;
    const int R = c.D + (mode == CNF_MODE_TRAIN ? 3 : 1);
    const long long per_sample = 2LL * c.D + c.C + R;
    const long long cap = (long long)(CNFB200_STATIC_SCRATCH_BYTES / sizeof(float));
    if (B < 1) { h->err = "bad batch size"; return CNF_ERR_BADARG; }
    const long long Bmax = cap / per_sample / 256 * 256;
    if (B > Bmax) {
        h->err = "augmented_f_static: batch of " + std::to_string((long long)B) + " samples exceeds the static scratch (" +
                 std::to_string((long long)CNFB200_STATIC_SCRATCH_BYTES) + " bytes hold at most " + std::to_string(Bmax) +
                 " samples of this network and mode); use cnfb200_augmented_f";
        return CNF_ERR_BADARG;
    }
    if (!u || !du) { h->err = "null u/du"; return CNF_ERR_BADARG; }
    if (c.C > 0 && !ys) { h->err = "conditional flow: ys is required"; return CNF_ERR_BADARG; }
    if (mode == CNF_MODE_TRAIN && !eps) { h->err = "augmented_f_static: eps is required in TrainMode"; return CNF_ERR_BADARG; }
    CNF_CUDA(cudaSetDevice(h->device));
    if (!static_is_device(u) || !static_is_device(du) || (c.C > 0 && !static_is_device(ys)) ||
        (mode == CNF_MODE_TRAIN && !static_is_device(eps))) {
        h->err = "augmented_f_static: u, ys, eps and du must be device pointers (a host buffer would need a staging allocation)";
        return CNF_ERR_BADARG;
    }
    float* base = nullptr;
    int rc = static_scratch(h, &base);
    if (rc) return rc;
    const long long Bs = (B + 255) / 256 * 256;
    float* z = base;                         // D x Bs  stage input (the z rows of u)
    float* yp = z + (long long)c.D * Bs;     // C x Bs
    float* ep = yp + (long long)c.C * Bs;    // D x Bs
    float* out = ep + (long long)c.D * Bs;   // R x Bs
    const unsigned gp = (unsigned)((Bs + CNF_NT - 1) / CNF_NT);
    k_static_pack<<<gp, CNF_NT, 0, h->stream>>>(u, R, c.D, B, z, Bs);
    if (c.C > 0) k_static_pack<<<gp, CNF_NT, 0, h->stream>>>(ys, c.C, c.C, B, yp, Bs);
    if (mode == CNF_MODE_TRAIN) k_static_pack<<<gp, CNF_NT, 0, h->stream>>>(eps, c.D, c.D, B, ep, Bs);
    CnfBufs bf;
    bf.slots = nullptr;
    for (int j = 0; j < 7; ++j) bf.K[j] = nullptr;
    bf.S = nullptr;
    bf.eps = ep;
    bf.ys = yp;
    bf.Bs = Bs;
    bf.B = (int)B;
    if ((rc = launch_eval_generic(h, bf, mode, z, out))) return rc;   // explicit source / destination: no handle buffer is touched
    k_static_unpack<<<(unsigned)((B + CNF_NT - 1) / CNF_NT), CNF_NT, 0, h->stream>>>(out, R, B, Bs, du);
    CNF_CUDA(cudaStreamSynchronize(h->stream));
    CNF_CUDA(cudaGetLastError());
    return CNF_OK;
}
