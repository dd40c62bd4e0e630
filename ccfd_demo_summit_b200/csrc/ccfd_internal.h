// Internal declarations shared by the library's translation units (not part of the ABI).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <exception>
#include <functional>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <system_error>
#include <vector>

#include "ccfd_rf.h"

// ----------------------------------------------------------------------------------------------
// Packed node, 8 bytes.  Trees are laid out with the two children of a node in adjacent slots, so
// one child reference serves both: next = child + (go_right ? one slot : 0).
//   internal : thr  = largest fp32 <= sklearn's fp64 threshold      (x <= thr  <=>  go left)
//   leaf     : thr  = bit pattern of the tree-local leaf-value index (a non-negative tiny float);
//              its feature is n_features -> it reads the all-zero dummy feature row, 0 <= thr always
//              holds and its child reference is itself, so a leaf steps to itself and fixed-trip
//              loops need no leaf test.
// Two encodings of `meta`, one per forest (TreeRec::group_flags bit 2, ForestDev::wide_nodes):
//   compact (every tree <= 96 KB: the shared-memory kernels' format)
//              meta = child_byte_offset << 16 | feature << 8 | missing_go_to_left
//              The child reference is a BYTE offset from the tree's first node, so the next node's
//              shared-memory address is base + (meta >> 16) (+ 8 for the right child): one shift-add
//              instead of shift, add, scale, add per level of every row.
//   wide    (forests with a larger tree: GLOBAL kernel only)
//              meta = child_slot << 9 | missing_go_to_left << 8 | feature
// Breadth-first order (compact) keeps each level contiguous: the top levels (<= 16 nodes = 128 B)
// are shared-memory bank-conflict free however the 32 rows of a warp spread over them.
// ----------------------------------------------------------------------------------------------
struct PackedNode {
    float thr;
    uint32_t meta;
};
static_assert(sizeof(PackedNode) == 8, "node must be 8 bytes");

constexpr int64_t kMaxNodesPerTree = (int64_t)1 << 23;
constexpr uint32_t kWideNodesFlag = 4u;  // TreeRec::group_flags

// Field access for the two encodings.  child() is in node slots.
template <bool WIDE>
struct NodeFmt;
template <>
struct NodeFmt<true> {
    static __host__ __device__ __forceinline__ uint32_t feature(uint32_t meta) { return meta & 0xFFu; }
    static __host__ __device__ __forceinline__ bool mgl(uint32_t meta) { return (meta & 0x100u) != 0u; }
    static __host__ __device__ __forceinline__ uint32_t child(uint32_t meta) { return meta >> 9; }
    static __host__ __device__ __forceinline__ uint32_t make(uint32_t child_slot, bool mgl, uint32_t feature) {
        return (child_slot << 9) | (mgl ? 0x100u : 0u) | feature;
    }
};
template <>
struct NodeFmt<false> {
    static __host__ __device__ __forceinline__ uint32_t feature(uint32_t meta) { return (meta >> 8) & 0xFFu; }
    static __host__ __device__ __forceinline__ bool mgl(uint32_t meta) { return (meta & 1u) != 0u; }
    static __host__ __device__ __forceinline__ uint32_t child(uint32_t meta) { return meta >> 19; }
    static __host__ __device__ __forceinline__ uint32_t child_bytes(uint32_t meta) { return meta >> 16; }
    static __host__ __device__ __forceinline__ uint32_t make(uint32_t child_slot, bool mgl, uint32_t feature) {
        return (child_slot << 19) | (feature << 8) | (mgl ? 1u : 0u);
    }
};
constexpr int64_t kMaxCompactSlots = 8192;  // 16-bit byte offsets

// Per-tree record read by the kernels (32 bytes, two 16-byte uniform loads).
struct TreeRec {
    uint32_t blob_off16;   // offset of the tree's blob in the forest buffer, in 16-byte units
    uint32_t blob_bytes;   // nodes + padding + leaf values, multiple of 16
    uint32_t values_off;   // byte offset of the leaf values inside the blob (multiple of 16)
    uint32_t depth;        // longest root-to-leaf path = trip count of the descent loop
    // Breadth-first order puts the deepest level last and that level holds leaves only, with
    // consecutive value indices: a row that ends at position p >= last_pos has value index
    // p - last_pos + last_vidx and need not read its leaf slot at all.
    uint32_t last_pos;
    uint32_t last_vidx;
    // filled when the forest is bound to a device (STREAM ring geometry), 0 in ccfd_forest_pack output
    uint32_t rel_off;      // byte offset of the blob inside its tree group (= inside a ring stage)
    uint32_t group_flags;  // bit 0: first tree of its group, bit 1: last tree of its group (both set at bind
                           // time), bit 2: wide node encoding (set by the packer)
};
static_assert(sizeof(TreeRec) == 32, "TreeRec must be 32 bytes");

// The top three levels of a tree in a fixed "virtual complete" order (root; its two children; the
// four grandchildren), copies of the packed nodes.  Where a node of the top is a leaf, its slot
// repeats downwards (a leaf steps to itself).
#ifndef CCFD_TOP_LEVELS
#define CCFD_TOP_LEVELS 3
#endif
constexpr int kTopLevels = CCFD_TOP_LEVELS;
constexpr int kTopNodes = (1 << kTopLevels) - 1;
struct TopNodes {
    PackedNode n[kTopNodes];
};

// Everything a consumer warp needs to know about a tree, passed to the STREAM kernel BY VALUE so
// that it is read through the constant bank: warp-uniform per-tree metadata fetched with ordinary
// loads costs a full 32-lane write-back in the same L1/shared-memory data pipe that bounds the
// kernel, the constant bank costs nothing there.
struct TreeParam {
    TopNodes top;          // 56 B (three levels)
    uint32_t rel_off;      // blob offset inside its stage
    uint32_t values_off;   // leaf values inside the blob
    uint32_t depth_flags;  // depth | first_of_group << 16 | last_of_group << 17
    uint32_t last_pos;
    uint32_t last_vidx;
    uint32_t pad;
};
static_assert(sizeof(TreeParam) == 8 * kTopNodes + 24, "TreeParam layout");
constexpr int kMaxParamTrees = 32000 / (int)sizeof(TreeParam);  // 400 * 80 B = 32,000 B of the 32,764 B kernel parameter space

// Consecutive trees whose blobs fit one shared-memory stage together: one bulk copy per group.
struct GroupRec {
    uint32_t first_tree;
    uint32_t n_trees;
    uint32_t blob_off16;   // of the first tree
    uint32_t bytes;        // sum of the group's blob_bytes (blobs are contiguous)
};
static_assert(sizeof(GroupRec) == 16, "GroupRec must be 16 bytes");

struct ForestDev {
    const uint8_t* blob;     // all tree blobs back to back, 16-byte aligned each
    const TreeRec* trees;    // [n_trees]
    const GroupRec* groups;  // [n_groups] STREAM kernel's unit of transfer (null if trees do not fit)
    int32_t n_trees;
    int32_t n_features;
    int32_t n_groups;
    uint32_t max_tree_bytes; // largest blob
    uint32_t max_depth;
    uint32_t stage_bytes;    // STREAM ring stage: >= largest blob and >= 128 rows of features
    uint32_t chunk_rows;     // rows of a row chunk that fit a stage, multiple of 32
    uint32_t wide_nodes;     // node encoding: 0 compact, 1 wide (see PackedNode)
    uint32_t blob_bytes;     // all tree blobs together (0 if beyond 4 GiB): what the forest-resident STREAM variant loads once
};

// Row pre-sort key (sort.cu): up to 11 features x 2 bits, most used split feature in the top bits.
constexpr int kSortMaxFeatures = 11;
struct SortKey {
    int32_t n_feat;
    int32_t feat[kSortMaxFeatures];
    float cut[kSortMaxFeatures][3];
};

struct ccfd_forest {
    int device = 0;
    ForestDev dev{};
    SortKey sort_key{};
    ccfd_forest_info info{};
    std::vector<TreeRec> host_trees;
    std::vector<TopNodes> host_top;
    std::vector<TreeParam> host_params;  // STREAM kernel parameter table (n_trees <= kMaxParamTrees)
    std::vector<TreeParam> host_params_res;  // the same for the forest-resident variant (absolute offsets)
    int sm_count = 0;
    int smem_optin = 0;
    // host-pointer path: staging owned by the forest, one host call at a time
    std::mutex host_mu;
    float* stage_x[2] = {nullptr, nullptr};
    double* stage_p[2] = {nullptr, nullptr};
    uint8_t* stage_d[2] = {nullptr, nullptr};
    cudaStream_t stage_stream[2] = {nullptr, nullptr};
    int64_t stage_rows = 0;
    // pageable-host path: pinned staging the rows are cast / copied into and the results come back to
    float* pin_x[2] = {nullptr, nullptr};
    double* pin_p[2] = {nullptr, nullptr};
    uint8_t* pin_d[2] = {nullptr, nullptr};
    int64_t pin_rows = 0;
};

// error plumbing -------------------------------------------------------------------------------
int ccfd_fail(int code, const char* fmt, ...);
int ccfd_cuda_fail(cudaError_t e, const char* what);
#define CCFD_CUDA(call)                                              \
    do {                                                             \
        cudaError_t _e = (call);                                     \
        if (_e != cudaSuccess) return ccfd_cuda_fail(_e, #call);     \
    } while (0)

// Every extern "C" body runs inside ccfd_guarded: the header promises "never throws", and a C++
// exception that crossed the C boundary (ctypes, cgo, JNI) would end the host process -- the Seldon
// model server -- instead of failing one request.  Allocation failures (std::bad_alloc from vector /
// map growth, std::system_error from thread creation) come back as CCFD_ENOMEM, anything else as
// CCFD_ECUDA with the exception's text in ccfd_last_error().
template <class F>
int ccfd_guarded(const char* where, F&& body) noexcept {
    try {
        return body();
    } catch (const std::bad_alloc&) {
        return ccfd_fail(CCFD_ENOMEM, "%s: out of host memory", where);
    } catch (const std::length_error& e) {
        return ccfd_fail(CCFD_ENOMEM, "%s: %s", where, e.what());
    } catch (const std::system_error& e) {
        return ccfd_fail(CCFD_ENOMEM, "%s: %s", where, e.what());
    } catch (const std::exception& e) {
        return ccfd_fail(CCFD_ECUDA, "%s: unexpected failure: %s", where, e.what());
    } catch (...) {
        return ccfd_fail(CCFD_ECUDA, "%s: unexpected failure", where);
    }
}

// Host worker pool shared by the multi-threaded host passes (ccfd_cast_rows, ccfd_parse_rows): threads
// are created once, on first use, not per call.  Runs fn(0..n_parts-1); the calling thread takes part
// in the work.  If the pool is busy with another caller, or its threads could not be created, the
// caller simply runs every part itself.
void ccfd_parallel_for(int n_parts, const std::function<void(int)>& fn);
int ccfd_host_threads();  // parts worth asking for (pool size + the caller), >= 1
// float64 / float32 rows -> dense float32 rows (dst == null: scan only) + "holds +-inf" (ccfd_cast_rows' body)
// values_per_part: how finely the work is cut for the host pool (one part = the calling thread only)
// streaming_stores: dst is read next by a DMA engine, not by the CPU (non-temporal stores, x86)
int ccfd_cast_span(const void* src, int32_t src_is_f64, int64_t n, int64_t n_cols, int64_t src_ld, float* dst, int32_t* has_inf,
                   int64_t values_per_part, bool streaming_stores);

// Is `dev`'s primary context alive in this process?  (true when the driver cannot be asked)
bool ccfd_primary_ctx_active(int dev);

// Makes the forest's device current for the calling thread and puts the previous one back on exit --
// unless the previous device has no context in this process: a thread that never chose a device
// reports device 0, and cudaSetDevice(0) would then CREATE a context there (~1 s and several hundred
// MB on a GPU this process does not use, e.g. every rank but 0 of a one-process-per-GPU job).
class DeviceGuard {
  public:
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev_) != cudaSuccess) prev_ = -1;
        if (prev_ != dev) cudaSetDevice(dev);
        dev_ = dev;
    }
    ~DeviceGuard() {
        if (prev_ >= 0 && prev_ != dev_ && ccfd_primary_ctx_active(prev_)) cudaSetDevice(prev_);
    }

  private:
    int prev_ = -1, dev_ = 0;
};

// api.cu
int ccfd_session_geometry(const ccfd_session* ss, int64_t* max_batch, int32_t* n_slots, int32_t* n_features);

// kernels.cu -------------------------------------------------------------------------------------
struct ScoreArgs {
    const float* X;     // device
    int64_t n;
    int64_t ld;
    double* out;        // device, may be null when only decisions are wanted
    uint8_t* decision;  // device, may be null
    double threshold;
    int ge;
    const uint32_t* perm = nullptr;  // device, optional (GLOBAL kernel): row i of the launch is row perm[i] of X / out / decision
};

// Pick + launch.  `kernel` is one of CCFD_KERNEL_*; scratch is used by SPLIT only
// (n * n_trees * 16 bytes) and may be null for the others.
int ccfd_launch_score(const ccfd_forest* f, const ScoreArgs& a, uint32_t kernel, void* scratch,
                      cudaStream_t stream);
uint32_t ccfd_pick_kernel(const ccfd_forest* f, int64_t n);
int64_t ccfd_split_scratch_bytes(const ccfd_forest* f, int64_t n);
int ccfd_kernels_init(ccfd_forest* f);

// sort.cu ----------------------------------------------------------------------------------------
void ccfd_sort_key_from_desc(const ccfd_forest_desc* d, SortKey* sk);
int64_t ccfd_sort_scratch_bytes(const ccfd_forest* f, int64_t n);
int ccfd_launch_row_sort(const ccfd_forest* f, const float* X, int64_t n, int64_t ld, void* scratch, const uint32_t** perm_out,
                         cudaStream_t stream);
bool ccfd_row_sort_wanted(const ccfd_forest* f, uint32_t kernel, int64_t n);
