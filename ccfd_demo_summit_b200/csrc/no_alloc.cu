// ccfd_predict_proba_noalloc: the latency (SPLIT) scoring for callers that must not allocate.
//
// Who this is for.  ccfd_predict_proba takes the SPLIT kernel's per-(tree,row) scratch from the
// stream-ordered allocator, and a session allocates its scratch when it is created.  Some hosts
// can do neither: a serving process that has handed all device memory to another allocator (a
// framework's caching pool, a co-tenant with a fixed budget), or a request thread on which an
// allocation's implicit device-wide synchronisation would stall every other stream.  For them the
// scratch lives in this library's own image instead: a fixed array in device global memory that
// comes into existence when a GPU loads the library and stays until the process exits.
//
// What that costs and what to watch:
//   * kNoAllocScratchBytes (16 MiB) of device memory on every GPU that loads the library, used or
//     not.  A request must satisfy n * n_trees <= CCFD_NOALLOC_MAX_TREE_ROWS; larger ones are
//     refused with CCFD_EINVAL (use ccfd_predict_proba or a session for those).
//   * One array per GPU, shared by every call on that GPU.  It still holds whatever the previous
//     call left in it.  Nothing has to be cleared: a request reads only the slots
//     [t * n + row] for t < n_trees, row < n, and its own first kernel writes every one of those
//     before its second kernel reads them.  Two requests that overlap on the GPU (different
//     streams, or different host threads) would write the same slots and corrupt each other:
//     keep calls on one GPU strictly one after another.
//   * The first call on each GPU looks the array's device address up (cudaGetSymbolAddress) and
//     keeps it.  Make that first call during start-up, before any CUDA stream capture begins.
//     With CUDA's lazy module loading that look-up is also what loads this library's device image
//     on the GPU, so "no allocation" holds from the second call on; the warm-up call is where the
//     16 MiB come into existence.  The kept address belongs to the GPU's primary context: after a
//     cudaDeviceReset() (or any other teardown of that context) it is stale and the process must
//     not call this entry point again.
#include <mutex>

#include "ccfd_internal.h"

namespace {

constexpr size_t kNoAllocScratchBytes = (size_t)CCFD_NOALLOC_MAX_TREE_ROWS * 16;  // (v0, v1) float64 per (tree,row)
constexpr int kMaxDevices = 64;

// In host code this name is only a handle for the runtime's symbol lookup, never an address a
// kernel could use; the device address is fetched once per GPU below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

std::mutex g_mu;
void* g_scratch_addr[kMaxDevices] = {};

// Device address of the scratch array on `device`, which must be the current device.
int scratch_address(int device, void** out) {
    if (device < 0 || device >= kMaxDevices) return ccfd_fail(CCFD_EUNSUPPORTED, "device index %d beyond %d", device, kMaxDevices);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return ccfd_cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!p) return ccfd_fail(CCFD_ECUDA, "cudaGetSymbolAddress returned a null address");
        g_scratch_addr[device] = p;
    }
    *out = g_scratch_addr[device];
    return CCFD_OK;
}

}  // namespace

static int ccfd_predict_proba_noalloc_impl(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, void* stream_) {
    if (!f) return ccfd_fail(CCFD_EINVAL, "null forest");
    if (n < 0) return ccfd_fail(CCFD_EINVAL, "n must be >= 0 (got %lld)", (long long)n);
    if (n == 0) return CCFD_OK;
    if (!X || !out) return ccfd_fail(CCFD_EINVAL, "X and out must be device pointers, not null");
    if (ld < f->dev.n_features) return ccfd_fail(CCFD_EINVAL, "row stride %lld is smaller than n_features %d", (long long)ld, f->dev.n_features);
    const int64_t tree_rows = n * (int64_t)f->dev.n_trees;
    if (tree_rows > (int64_t)CCFD_NOALLOC_MAX_TREE_ROWS)
        return ccfd_fail(CCFD_EINVAL,
                         "request of %lld rows x %d trees needs %lld scratch slots; the built-in scratch holds %lld "
                         "(CCFD_NOALLOC_MAX_TREE_ROWS): split the request or use ccfd_predict_proba / a session",
                         (long long)n, f->dev.n_trees, (long long)tree_rows, (long long)CCFD_NOALLOC_MAX_TREE_ROWS);
    DeviceGuard guard(f->device);
    void* scratch = nullptr;
    int rc = scratch_address(f->device, &scratch);
    if (rc != CCFD_OK) return rc;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    ScoreArgs a{X, n, ld, out, nullptr, 0.5, 0};
    rc = ccfd_launch_score(f, a, CCFD_KERNEL_SPLIT, scratch, stream);  // two launches, no allocation
    if (rc != CCFD_OK) return rc;
    CCFD_CUDA(cudaStreamSynchronize(stream));
    return CCFD_OK;
}

// ---- the C ABI: every body above runs inside ccfd_guarded (no C++ exception crosses the boundary)
extern "C" int ccfd_predict_proba_noalloc(ccfd_forest* f, const float* X, int64_t n, int64_t ld, double* out, void* stream_) {
    return ccfd_guarded("ccfd_predict_proba_noalloc", [&]() -> int { return ccfd_predict_proba_noalloc_impl(f, X, n, ld, out, stream_); });
}
