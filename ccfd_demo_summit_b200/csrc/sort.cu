// Row pre-sort: makes the 32 rows of a warp walk similar paths.
//
// Rows are independent (SK/ensemble/_forest.py:951-962 adds per row), so the ORDER in which a kernel
// takes them is free, and it matters: the GLOBAL kernel's node loads go through L1/L2, where a
// warp-wide load costs one wavefront per distinct 128-byte line and one sector per distinct 32 bytes.
// Rows in arrival order spread over ~22 distinct nodes per level (measured: ~20 sectors per load);
// rows that agree on the forest's most used split features share the upper levels of every tree.
// So, before a bulk GLOBAL launch: key = a few bits per feature for the features the forest splits
// on most (cut points = quantiles of the forest's own thresholds for that feature, so no statistics
// of the data are needed), counting sort of the row indices by key, and the kernel reads row
// perm[i] and writes result perm[i].  Bits cannot change: every row is still scored alone, by one
// lane, trees in order.  Measured on the B200 (scripts/sort_probe.py, 100 unbounded trees x 2 M rows,
// sort not timed): 8.11 -> 6.61 ms; the sort itself costs ~1 % of the launch it precedes.
//
// Three small kernels + a three-phase scan, all hand-written; scratch comes from the caller
// (stream-ordered allocator in api.cu).  The order of rows inside one key is whatever the atomics
// give: it does not affect any result.
#include <algorithm>
#include <cstring>
#include <vector>

#include "ccfd_internal.h"

namespace {

constexpr int kScanBlock = 1024;  // bins per scan block

__device__ __forceinline__ uint32_t key_bucket(const SortKey& sk, int k, float x) {
    // NaN compares false everywhere: bucket 0, like any other value it just has to be deterministic
    return (uint32_t)(x > sk.cut[k][0]) + (uint32_t)(x > sk.cut[k][1]) + (uint32_t)(x > sk.cut[k][2]);
}

// One warp = 32 consecutive rows.  The rows are read the way they lie in memory (one row per load
// instruction, lanes = consecutive features: full sectors), parked in shared memory with an odd pitch,
// and each lane then picks its own row's key features from there.  Wide rows (pitch beyond the shared-memory
// budget) fall back to strided loads.
constexpr int kKeyWarps = 8;
constexpr int kKeyMaxPitch = 47;  // 8 warps x 32 rows x 47 floats = 48,128 bytes: within the 48 KB a kernel gets without opting in
static_assert(kKeyWarps * 32 * kKeyMaxPitch * 4 <= 48 * 1024, "k_sort_keys' dynamic shared memory must fit the default limit");
__global__ void __launch_bounds__(kKeyWarps * 32) k_sort_keys(const float* __restrict__ X, int64_t n, int64_t ld, int F, SortKey sk, int shift,
                                                              uint32_t* __restrict__ keys, uint32_t* __restrict__ hist) {
    extern __shared__ float rows_s[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t r0 = ((int64_t)blockIdx.x * kKeyWarps + warp) * 32;
    if (r0 >= n) return;
    const int64_t i = r0 + lane;
    const int pitch = F + 1 + (F & 1);  // odd
    uint32_t key = 0;
    if (pitch <= kKeyMaxPitch) {
        float* mine = rows_s + warp * 32 * pitch;
        const int cnt = (int)min((int64_t)32, n - r0);
        for (int r = 0; r < cnt; ++r) {
            const float* src = X + (r0 + r) * ld;
            for (int c = lane; c < F; c += 32) mine[r * pitch + c] = __ldg(src + c);
        }
        __syncwarp();
        if (i < n) {
#pragma unroll 1
            for (int k = 0; k < sk.n_feat; ++k) key = (key << 2) | key_bucket(sk, k, mine[lane * pitch + sk.feat[k]]);
        }
    } else if (i < n) {
        const float* row = X + i * ld;
#pragma unroll 1
        for (int k = 0; k < sk.n_feat; ++k) key = (key << 2) | key_bucket(sk, k, __ldg(row + sk.feat[k]));
    }
    if (i >= n) return;
    key >>= shift;
    keys[i] = key;
    atomicAdd(hist + key, 1u);
}

// exclusive scan of `bins` counters, three phases
__global__ void __launch_bounds__(kScanBlock) k_scan_block_sums(const uint32_t* __restrict__ hist, int bins, uint32_t* __restrict__ sums) {
    __shared__ uint32_t warp_sum[32];
    const int i = blockIdx.x * kScanBlock + threadIdx.x;
    uint32_t v = i < bins ? hist[i] : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = warp_sum[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) sums[blockIdx.x] = v;
    }
}

// in-block exclusive scan helper: returns the exclusive prefix of v over the block, total in *total
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_tot[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += u;
    }
    if (lane == 31) warp_tot[w] = inc;
    __syncthreads();
    if (w == 0) {
        uint32_t t = warp_tot[lane];
        uint32_t ti = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t u = __shfl_up_sync(0xffffffffu, ti, o);
            if (lane >= o) ti += u;
        }
        warp_tot[lane] = ti - t;  // exclusive prefix of the warp totals
        if (lane == 31 && total) *total = ti;
    }
    __syncthreads();
    return inc - v + warp_tot[w];
}

// one block: exclusive scan of the block sums in place (<= kScanBlock * kScanBlock bins in all)
__global__ void __launch_bounds__(kScanBlock) k_scan_sums(uint32_t* __restrict__ sums, int n_blocks) {
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0u;
    __syncthreads();
    for (int base = 0; base < n_blocks; base += kScanBlock) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < n_blocks ? sums[i] : 0u;
        __shared__ uint32_t total;
        const uint32_t ex = block_exclusive_scan(v, &total);
        if (i < n_blocks) sums[i] = ex + carry;
        __syncthreads();
        if (threadIdx.x == 0) carry += total;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kScanBlock) k_scan_apply(uint32_t* __restrict__ hist, int bins, const uint32_t* __restrict__ sums) {
    const int i = blockIdx.x * kScanBlock + threadIdx.x;
    const uint32_t v = i < bins ? hist[i] : 0u;
    const uint32_t ex = block_exclusive_scan(v, nullptr);
    if (i < bins) hist[i] = ex + sums[blockIdx.x];
}

__global__ void __launch_bounds__(256) k_sort_scatter(const uint32_t* __restrict__ keys, int64_t n, uint32_t* __restrict__ offs,
                                                      uint32_t* __restrict__ perm) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t pos = atomicAdd(offs + keys[i], 1u);
    perm[pos] = (uint32_t)i;
}

int key_bits_for(int64_t n, int total_bits) {
    // about 8 rows per bin: more bins than that only costs histogram traffic
    int bits = 8;
    while (bits < total_bits && ((int64_t)1 << (bits + 3)) < n) ++bits;
    return std::min(bits, total_bits);
}

}  // namespace

// Cut points from the forest itself: the features with the most split nodes, most used first (they
// land in the key's top bits), each with three cuts at the quartiles of that feature's thresholds.
void ccfd_sort_key_from_desc(const ccfd_forest_desc* d, SortKey* sk) {
    memset(sk, 0, sizeof(*sk));
    const int F = d->n_features;
    const int64_t total = d->tree_offsets[d->n_trees];
    std::vector<std::vector<float>> thr((size_t)F);
    for (int64_t u = 0; u < total; ++u)
        if (d->left[u] >= 0 && d->feature[u] >= 0 && d->feature[u] < F) thr[(size_t)d->feature[u]].push_back(d->threshold[u]);
    std::vector<int> order((size_t)F);
    for (int f = 0; f < F; ++f) order[(size_t)f] = f;
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return thr[(size_t)a].size() > thr[(size_t)b].size(); });
    for (int k = 0; k < F && sk->n_feat < kSortMaxFeatures; ++k) {
        std::vector<float>& t = thr[(size_t)order[(size_t)k]];
        if (t.size() < 3) break;
        std::sort(t.begin(), t.end());
        const int j = sk->n_feat++;
        sk->feat[j] = order[(size_t)k];
        for (int q = 0; q < 3; ++q) sk->cut[j][q] = t[(q + 1) * t.size() / 4];
    }
}

int64_t ccfd_sort_scratch_bytes(const ccfd_forest* f, int64_t n) {
    const int bits = key_bits_for(n, 2 * f->sort_key.n_feat);
    const int64_t bins = (int64_t)1 << bits;
    const int64_t blocks = (bins + kScanBlock - 1) / kScanBlock;
    return 4 * (2 * n + bins + blocks) + 1024;
}

// keys | perm | hist | block sums, in that order inside `scratch`; returns the device pointer of perm
int ccfd_launch_row_sort(const ccfd_forest* f, const float* X, int64_t n, int64_t ld, void* scratch, const uint32_t** perm_out,
                         cudaStream_t stream) {
    if (n > 0xFFFFFFFFll) return ccfd_fail(CCFD_EINVAL, "too many rows for the row pre-sort");
    const int total_bits = 2 * f->sort_key.n_feat;
    const int bits = key_bits_for(n, total_bits);
    const int bins = 1 << bits;
    const int blocks = (bins + kScanBlock - 1) / kScanBlock;
    uint32_t* keys = static_cast<uint32_t*>(scratch);
    uint32_t* perm = keys + n;
    uint32_t* hist = perm + n;
    uint32_t* sums = hist + bins;
    CCFD_CUDA(cudaMemsetAsync(hist, 0, sizeof(uint32_t) * (size_t)bins, stream));
    const unsigned grid = (unsigned)((n + 255) / 256);
    const int F = f->dev.n_features, pitch = F + 1 + (F & 1);
    const size_t key_smem = pitch <= kKeyMaxPitch ? sizeof(float) * kKeyWarps * 32 * (size_t)pitch : 0;
    k_sort_keys<<<grid, kKeyWarps * 32, key_smem, stream>>>(X, n, ld, F, f->sort_key, total_bits - bits, keys, hist);
    k_scan_block_sums<<<blocks, kScanBlock, 0, stream>>>(hist, bins, sums);
    k_scan_sums<<<1, kScanBlock, 0, stream>>>(sums, blocks);
    k_scan_apply<<<blocks, kScanBlock, 0, stream>>>(hist, bins, sums);
    k_sort_scatter<<<grid, 256, 0, stream>>>(keys, n, hist, perm);
    CCFD_CUDA(cudaGetLastError());
    *perm_out = perm;
    return CCFD_OK;
}
