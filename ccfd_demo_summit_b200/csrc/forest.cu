// Forest handle: validation, breadth-first re-layout, packing and upload (host code).
//
// Replaces the model-load step of the Seldon model image (reference
// deploy/model/modelfull.json:24): what scikit-learn keeps as 64-byte Node structs plus a
// [node, 1, class] float64 value array (SK/tree/_tree.pxd:15-25, SK/tree/_tree.pyx:656-765)
// becomes, per tree, one 16-byte-aligned blob of 8-byte nodes followed by 16-byte leaf values,
// ready to be moved into shared memory by a single bulk copy.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <new>

#include "ccfd_internal.h"

// ------------------------------------------------------------------------------------ errors
static thread_local char g_err[512] = "";

int ccfd_fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

bool ccfd_primary_ctx_active(int dev) {
    // cuDevicePrimaryCtxGetState through the runtime's driver entry-point lookup: no link-time dependency on
    // libcuda, so the library still loads (for ccfd_forest_pack and the symbol checks) on a box without a driver.
    using GetState = int (*)(int /*CUdevice*/, unsigned int* /*flags*/, int* /*active*/);
    static const GetState fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult st;
        if (cudaGetDriverEntryPoint("cuDevicePrimaryCtxGetState", &p, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess) p = nullptr;
        return reinterpret_cast<GetState>(p);
    }();
    if (!fn) return true;
    unsigned int flags = 0;
    int active = 0;
    if (fn(dev, &flags, &active) != 0) return true;
    return active != 0;
}

int ccfd_cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_err, sizeof(g_err), "CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    cudaGetLastError();  // clear the non-sticky error state
    return e == cudaErrorMemoryAllocation ? CCFD_ENOMEM : CCFD_ECUDA;
}

extern "C" const char* ccfd_last_error(void) { return g_err; }

extern "C" int ccfd_device_count(void) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) return ccfd_cuda_fail(e, "cudaGetDeviceCount");
    return n;
}

// ------------------------------------------------------------------------------------ packing
namespace {

struct PackedTree {
    std::vector<PackedNode> nodes;
    std::vector<double> values;  // 2 per leaf
    uint32_t depth = 0;
    uint32_t last_pos = 0;   // first position of the deepest level
    uint32_t last_vidx = 0;  // value index of the leaf at last_pos
    TopNodes top{};
};

// Orders the nodes of one tree (the two children of a node always adjacent) and packs them.
// `big` is the layout for trees that can never sit in shared memory (the GLOBAL kernel reads them
// from L2/HBM): equal leaf values share one table entry (an unbounded-depth forest has pure leaves
// only, so its 16-byte-per-leaf table collapses to two entries and the forest halves), and every
// leaf slot is kept because the value index no longer follows from the position.
int pack_tree(const ccfd_forest_desc* d, int t, bool big, PackedTree* out) {
    const int64_t lo = d->tree_offsets[t], hi = d->tree_offsets[t + 1];
    const int64_t cnt = hi - lo;
    if (cnt < 1) return ccfd_fail(CCFD_EINVAL, "tree %d has no nodes", t);
    if (cnt > kMaxNodesPerTree) return ccfd_fail(CCFD_EUNSUPPORTED, "tree %d has %lld nodes (max %lld)", t, (long long)cnt, (long long)kMaxNodesPerTree);
    const int F = d->n_features;

    // ---- phase 1: the order of the nodes.  Whatever the order, the two children of a node are adjacent.
    std::vector<int32_t> order;          // position -> source node
    std::vector<uint32_t> level;         // position -> depth
    std::vector<uint8_t> seen((size_t)cnt, 0);
    order.reserve((size_t)cnt);
    level.reserve((size_t)cnt);
    auto check_internal = [&](int32_t src) -> int {  // validates an internal node before its children are placed
        const int64_t u = lo + src;
        const int32_t l = d->left[u], r = d->right[u];
        if (l < 0 && r < 0) return 0;  // leaf
        if (l < 0 || r < 0) return ccfd_fail(CCFD_EINVAL, "tree %d node %d has exactly one child", t, src);
        if (l >= cnt || r >= cnt || l == r) return ccfd_fail(CCFD_EINVAL, "tree %d node %d has a child out of range", t, src);
        if (seen[l] || seen[r]) return ccfd_fail(CCFD_EINVAL, "tree %d is not a tree (node reached twice)", t);
        seen[l] = seen[r] = 1;
        return 1;
    };
    seen[0] = 1;
    if (!big) {
        // breadth-first: every level contiguous, the deepest level (leaves only) last
        order.push_back(0);
        level.push_back(0);
        for (size_t pos = 0; pos < order.size(); ++pos) {
            const int rc = check_internal(order[pos]);
            if (rc < 0) return rc;
            if (rc == 1) {
                const int64_t u = lo + order[pos];
                order.push_back(d->left[u]);
                order.push_back(d->right[u]);
                level.push_back(level[pos] + 1);
                level.push_back(level[pos] + 1);
            }
        }
    } else {
        // Blocks of three levels: a sibling pair, its (up to) two child pairs and four grandchild pairs,
        // 14 nodes = 112 bytes, stored together.  A root-to-leaf path then crosses one 128-byte line
        // (two at most) per three levels instead of one per level: fewer L2 sectors per path and three
        // times as many levels per L1 line.  Blocks themselves are laid out breadth-first, top first.
        struct BlockRoot {
            int32_t a, b;  // the sibling pair that heads the block (b < 0: the tree root, which stands alone)
            uint32_t depth;
        };
        std::vector<BlockRoot> blocks;
        blocks.push_back(BlockRoot{0, -1, 0u});
        std::vector<int32_t> cur, nxt;
        for (size_t bi = 0; bi < blocks.size(); ++bi) {
            const BlockRoot br = blocks[bi];
            cur.clear();
            cur.push_back(br.a);
            if (br.b >= 0) cur.push_back(br.b);
            for (int rel = 0; rel < 3; ++rel) {
                nxt.clear();
                for (int32_t src : cur) {
                    order.push_back(src);
                    level.push_back(br.depth + (uint32_t)rel);
                    const int rc = check_internal(src);
                    if (rc < 0) return rc;
                    if (rc == 1) {
                        nxt.push_back(d->left[lo + src]);
                        nxt.push_back(d->right[lo + src]);
                    }
                }
                if (rel == 2) {
                    for (size_t i = 0; i + 1 < nxt.size(); i += 2) blocks.push_back(BlockRoot{nxt[i], nxt[i + 1], br.depth + 3u});
                } else {
                    cur.swap(nxt);
                }
            }
        }
    }
    if (!big && (int64_t)order.size() > kMaxCompactSlots) return ccfd_fail(CCFD_EUNSUPPORTED, "internal: tree %d too large for the compact encoding", t);
    std::vector<uint32_t> pos_of((size_t)cnt, 0u);
    for (size_t pos = 0; pos < order.size(); ++pos) pos_of[(size_t)order[pos]] = (uint32_t)pos;

    // ---- phase 2: emit the packed nodes in that order
    out->nodes.clear();
    out->values.clear();
    out->depth = 0;
    std::map<std::pair<uint64_t, uint64_t>, uint32_t> seen_values;
    for (size_t pos = 0; pos < order.size(); ++pos) {
        const int64_t u = lo + order[pos];
        const int32_t l = d->left[u];
        PackedNode pn;
        if (l < 0) {
            uint32_t vidx = (uint32_t)(out->values.size() / 2);
            const double v0 = d->value[u * 2], v1 = d->value[u * 2 + 1];
            if (!std::isfinite(v0) || !std::isfinite(v1)) return ccfd_fail(CCFD_EINVAL, "tree %d node %d has a non-finite leaf value", t, order[pos]);
            bool fresh = true;
            if (big) {
                uint64_t b0, b1;
                memcpy(&b0, &v0, 8);
                memcpy(&b1, &v1, 8);
                auto ins = seen_values.emplace(std::make_pair(b0, b1), vidx);
                fresh = ins.second;
                vidx = ins.first->second;
            }
            if (fresh) {
                out->values.push_back(v0);
                out->values.push_back(v1);
            }
            memcpy(&pn.thr, &vidx, 4);
            pn.meta = big ? NodeFmt<true>::make((uint32_t)pos, false, (uint32_t)F) : NodeFmt<false>::make((uint32_t)pos, false, (uint32_t)F);
            if (level[pos] > out->depth) out->depth = level[pos];
        } else {
            const int32_t f = d->feature[u];
            if (f < 0 || f >= F) return ccfd_fail(CCFD_EINVAL, "tree %d node %d splits on feature %d (n_features %d)", t, order[pos], f, F);
            const float thr = d->threshold[u];
            if (std::isnan(thr)) return ccfd_fail(CCFD_EINVAL, "tree %d node %d has a NaN threshold", t, order[pos]);
            const uint32_t child = pos_of[(size_t)l];
            if (pos_of[(size_t)d->right[u]] != child + 1) return ccfd_fail(CCFD_EINVAL, "internal: siblings of tree %d node %d are not adjacent", t, order[pos]);
            pn.thr = thr;
            const bool mgl = d->missing_go_to_left[u] != 0;
            pn.meta = big ? NodeFmt<true>::make(child, mgl, (uint32_t)f) : NodeFmt<false>::make(child, mgl, (uint32_t)f);
        }
        out->nodes.push_back(pn);
    }
    // deepest level: contiguous tail of the breadth-first order, leaves only
    uint32_t lp = (uint32_t)order.size();
    while (lp > 0 && level[lp - 1] == out->depth) --lp;
    if (big) lp = (uint32_t)order.size();  // keep every slot: "pos >= last_pos" never holds
    out->last_pos = lp;
    out->last_vidx = 0;
    if (lp < order.size()) memcpy(&out->last_vidx, &out->nodes[lp].thr, 4);
    // top three levels in virtual complete order; a leaf repeats itself downwards
    auto feature_of = [big](const PackedNode& n) { return big ? NodeFmt<true>::feature(n.meta) : NodeFmt<false>::feature(n.meta); };
    auto is_leaf = [&](const PackedNode& n) { return feature_of(n) == (uint32_t)F; };
    TopNodes& tp = out->top;
    tp.n[0] = out->nodes[0];
    for (int parent = 0; parent < (kTopNodes - 1) / 2; ++parent) {
        const PackedNode& pn = tp.n[parent];
        const uint32_t c = big ? NodeFmt<true>::child(pn.meta) : NodeFmt<false>::child(pn.meta);
        tp.n[2 * parent + 1] = out->nodes[c];
        tp.n[2 * parent + 2] = is_leaf(pn) ? out->nodes[c] : out->nodes[c + 1];
    }
    return CCFD_OK;
}

}  // namespace

namespace {

struct PackedForest {
    std::vector<uint8_t> blob;
    std::vector<TreeRec> trees;
    std::vector<TopNodes> top;
    int64_t n_nodes = 0, n_leaves = 0;
    uint32_t max_depth = 0, max_bytes = 0;
};

int validate_desc(const ccfd_forest_desc* d) {
    if (!d) return ccfd_fail(CCFD_EINVAL, "null argument");
    if (d->reserved != 0) return ccfd_fail(CCFD_EINVAL, "ccfd_forest_desc.reserved must be 0");
    if (d->n_trees < 1) return ccfd_fail(CCFD_EINVAL, "n_trees must be >= 1 (got %d)", d->n_trees);
    if (d->n_features < 1 || d->n_features > 255) return ccfd_fail(CCFD_EINVAL, "n_features must be in 1..255 (got %d)", d->n_features);
    if (d->n_classes != 2) return ccfd_fail(CCFD_EUNSUPPORTED, "this build scores binary forests only (n_classes=%d)", d->n_classes);
    if (!d->tree_offsets || !d->feature || !d->threshold || !d->left || !d->right || !d->missing_go_to_left || !d->value)
        return ccfd_fail(CCFD_EINVAL, "ccfd_forest_desc has a null array");
    if (d->tree_offsets[0] != 0) return ccfd_fail(CCFD_EINVAL, "tree_offsets[0] must be 0");
    for (int t = 0; t < d->n_trees; ++t)
        if (d->tree_offsets[t + 1] <= d->tree_offsets[t]) return ccfd_fail(CCFD_EINVAL, "tree_offsets must increase (tree %d)", t);
    return CCFD_OK;
}

// Per tree: [nodes][pad to 16][leaf values], blobs back to back, each 16-byte aligned and a
// multiple of 16 bytes long: the unit a 1-D TMA bulk copy moves.
int pack_forest(const ccfd_forest_desc* d, PackedForest* pf) {
    int rc = validate_desc(d);
    if (rc != CCFD_OK) return rc;
    pf->trees.resize((size_t)d->n_trees);
    pf->top.resize((size_t)d->n_trees);
    PackedTree pt;
    // A tree whose blob exceeds this can never fit a shared-memory stage on any sm_100 part: if the
    // forest has one, lay the whole forest out for the GLOBAL kernel (second pass).
    constexpr size_t kBigTreeBytes = 96 * 1024;
    bool big = false;
    for (int t = 0; t < d->n_trees && !big; ++t) {
        const int64_t cnt = d->tree_offsets[t + 1] - d->tree_offsets[t];
        big = (size_t)cnt * 16 > kBigTreeBytes;  // ~8 B per slot + 16 B per leaf (half the nodes) = 16 B per node
    }
    for (int t = 0; t < d->n_trees; ++t) {
        rc = pack_tree(d, t, big, &pt);
        if (rc != CCFD_OK) return rc;
        // the deepest level's leaf slots are never read (TreeRec::last_pos), so they are not stored
        const size_t node_bytes = (size_t)pt.last_pos * sizeof(PackedNode);
        const size_t values_off = (node_bytes + 15) & ~(size_t)15;
        const size_t bytes = values_off + pt.values.size() * sizeof(double);
        const size_t off = pf->blob.size();
        if ((off + bytes) / 16 > 0xFFFFFFFFull) return ccfd_fail(CCFD_EUNSUPPORTED, "forest larger than 64 GiB");
        pf->blob.resize(off + bytes, 0);
        memcpy(pf->blob.data() + off, pt.nodes.data(), node_bytes);
        memcpy(pf->blob.data() + off + values_off, pt.values.data(), pt.values.size() * sizeof(double));
        TreeRec& tr = pf->trees[(size_t)t];
        tr.blob_off16 = (uint32_t)(off / 16);
        tr.blob_bytes = (uint32_t)bytes;
        tr.values_off = (uint32_t)values_off;
        tr.depth = pt.depth;
        tr.last_pos = pt.last_pos;
        tr.last_vidx = pt.last_vidx;
        tr.rel_off = 0;
        tr.group_flags = big ? kWideNodesFlag : 0u;
        pf->top[(size_t)t] = pt.top;
        pf->n_nodes += (int64_t)pt.nodes.size();
        pf->n_leaves += (int64_t)pt.values.size() / 2;
        if (pt.depth > pf->max_depth) pf->max_depth = pt.depth;
        if (bytes > pf->max_bytes) pf->max_bytes = (uint32_t)bytes;
    }
    return CCFD_OK;
}

}  // namespace

static int ccfd_forest_pack_impl(const ccfd_forest_desc* d, void* blob, int64_t blob_capacity, uint32_t* tree_recs, uint32_t* top_nodes,
                                int64_t* blob_bytes) {
    if (!blob_bytes) return ccfd_fail(CCFD_EINVAL, "null argument");
    PackedForest pf;
    int rc = pack_forest(d, &pf);
    if (rc != CCFD_OK) return rc;
    *blob_bytes = (int64_t)pf.blob.size();
    if (!blob && !tree_recs) return CCFD_OK;  // size query
    if (!blob || !tree_recs) return ccfd_fail(CCFD_EINVAL, "blob and tree_recs must both be given");
    if (blob_capacity < (int64_t)pf.blob.size()) return ccfd_fail(CCFD_EINVAL, "blob buffer too small (%lld < %lld)", (long long)blob_capacity, (long long)pf.blob.size());
    memcpy(blob, pf.blob.data(), pf.blob.size());
    memcpy(tree_recs, pf.trees.data(), sizeof(TreeRec) * pf.trees.size());
    if (top_nodes) memcpy(top_nodes, pf.top.data(), sizeof(TopNodes) * pf.top.size());
    return CCFD_OK;
}

static int ccfd_forest_create_impl(const ccfd_forest_desc* d, int device, ccfd_forest** out) {
    if (!d || !out) return ccfd_fail(CCFD_EINVAL, "null argument");
    *out = nullptr;
    PackedForest pf;
    int rc0 = pack_forest(d, &pf);
    if (rc0 != CCFD_OK) return rc0;

    int ndev = 0;
    CCFD_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return ccfd_fail(CCFD_EINVAL, "device %d out of range (%d visible)", device, ndev);

    ccfd_forest* f = new (std::nothrow) ccfd_forest();
    if (!f) return ccfd_fail(CCFD_ENOMEM, "out of host memory");
    f->device = device;
    f->host_trees = pf.trees;
    f->host_top = pf.top;
    ccfd_sort_key_from_desc(d, &f->sort_key);
    const std::vector<uint8_t>& blob = pf.blob;
    const int64_t n_nodes = pf.n_nodes, n_leaves = pf.n_leaves;
    const uint32_t max_depth = pf.max_depth, max_bytes = pf.max_bytes;

    DeviceGuard guard(device);
    uint8_t* dblob = nullptr;
    TreeRec* dtrees = nullptr;
    cudaError_t e = cudaMalloc(&dblob, blob.size());
    if (e == cudaSuccess) e = cudaMalloc(&dtrees, sizeof(TreeRec) * (size_t)d->n_trees);
    if (e == cudaSuccess) e = cudaMemcpy(dblob, blob.data(), blob.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(dtrees, f->host_trees.data(), sizeof(TreeRec) * (size_t)d->n_trees, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        cudaFree(dblob);
        cudaFree(dtrees);
        delete f;
        return ccfd_cuda_fail(e, "forest upload");
    }
    f->dev.blob = dblob;
    f->dev.trees = dtrees;
    f->dev.n_trees = d->n_trees;
    f->dev.n_features = d->n_features;
    f->dev.max_tree_bytes = max_bytes;
    f->dev.blob_bytes = blob.size() <= 0xFFFFFFF0ull ? (uint32_t)blob.size() : 0u;
    f->dev.max_depth = max_depth;

    f->info.n_trees = d->n_trees;
    f->info.n_features = d->n_features;
    f->info.n_classes = 2;
    f->info.max_depth = (int32_t)max_depth;
    f->info.n_nodes = n_nodes;
    f->info.n_leaves = n_leaves;
    f->info.model_bytes = (int64_t)blob.size() + (int64_t)sizeof(TreeRec) * d->n_trees;
    f->info.max_tree_bytes = max_bytes;
    f->info.device = device;

    int rc = ccfd_kernels_init(f);
    if (rc != CCFD_OK) {
        ccfd_forest_destroy(f);
        return rc;
    }
    *out = f;
    return CCFD_OK;
}

static int ccfd_forest_destroy_impl(ccfd_forest* f) {
    if (!f) return CCFD_OK;
    DeviceGuard guard(f->device);
    for (int i = 0; i < 2; ++i) {
        if (f->stage_stream[i]) cudaStreamSynchronize(f->stage_stream[i]);
        cudaFree(f->stage_x[i]);
        cudaFree(f->stage_p[i]);
        cudaFree(f->stage_d[i]);
        if (f->stage_stream[i]) cudaStreamDestroy(f->stage_stream[i]);
        cudaFreeHost(f->pin_x[i]);
        cudaFreeHost(f->pin_p[i]);
        cudaFreeHost(f->pin_d[i]);
    }
    cudaFree((void*)f->dev.blob);
    cudaFree((void*)f->dev.trees);
    cudaFree((void*)f->dev.groups);
    delete f;
    return CCFD_OK;
}

static int ccfd_forest_get_info_impl(const ccfd_forest* f, ccfd_forest_info* info) {
    if (!f || !info) return ccfd_fail(CCFD_EINVAL, "null argument");
    *info = f->info;
    return CCFD_OK;
}

// ---- the C ABI: every body above runs inside ccfd_guarded (no C++ exception crosses the boundary)
extern "C" int ccfd_forest_pack(const ccfd_forest_desc* d, void* blob, int64_t blob_capacity, uint32_t* tree_recs, uint32_t* top_nodes, int64_t* blob_bytes) {
    return ccfd_guarded("ccfd_forest_pack", [&]() -> int { return ccfd_forest_pack_impl(d, blob, blob_capacity, tree_recs, top_nodes, blob_bytes); });
}
extern "C" int ccfd_forest_create(const ccfd_forest_desc* d, int device, ccfd_forest** out) {
    return ccfd_guarded("ccfd_forest_create", [&]() -> int { return ccfd_forest_create_impl(d, device, out); });
}
extern "C" int ccfd_forest_destroy(ccfd_forest* f) {
    return ccfd_guarded("ccfd_forest_destroy", [&]() -> int { return ccfd_forest_destroy_impl(f); });
}
extern "C" int ccfd_forest_get_info(const ccfd_forest* f, ccfd_forest_info* info) {
    return ccfd_guarded("ccfd_forest_get_info", [&]() -> int { return ccfd_forest_get_info_impl(f, info); });
}
