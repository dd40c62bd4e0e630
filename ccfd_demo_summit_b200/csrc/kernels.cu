// sm_100a kernels of the scoring path: forest descent + leaf gather + ordered accumulate + divide,
// fused.  Replaces Tree._apply_dense (SK/tree/_tree.pyx:954-996), Tree.predict's value.take
// (SK/tree/_tree.pyx:939-945), _accumulate_prediction (SK/ensemble/_forest.py:704-717) and the
// final divide (SK/ensemble/_forest.py:961-962).
//
// Mapping, common to the three kernels: one lane = one row, so a row's float64 sum runs over
// trees 0..T-1 in order inside one thread exactly as scikit-learn's n_jobs=1 loop does, and
// results are bit-identical.  Rows live in shared memory transposed so that the data-dependent
// `x[feature]` fetch of 32 lanes always hits 32 different banks: [feature][row] with a row stride
// that is a multiple of 32, or, for up to 31 features, 64-row blocks of 32 feature rows on 8 KB
// boundaries, where the address is one and-or of the node's meta word (XStrided / XBlocked below).
// Nodes of forests that fit shared memory use the compact encoding of ccfd_internal.h: a level of a
// chain is six instructions (LDS.64 node, LOP3 x address, LDS x, FSETP, SEL base, LEA.HI next).
//
//   STREAM  throughput.  A persistent CTA per SM keeps 512*SPT rows resident; rows (as row-major
//           chunks) and the forest (as tree groups) arrive through one shared-memory ring fed by
//           1-D TMA bulk copies (cp.async.bulk + mbarrier complete_tx) from one producer warp;
//           16 consumer warps transpose their own rows and walk TPT trees x SPT rows per thread.
//           Per-tree metadata and the top three levels of every tree are kernel parameters
//           (constant bank).
//           Forests small enough to sit in shared memory next to the rows take the forest-resident variant
//           (RES): loaded once per CTA, the ring carries rows only, four trees x two rows per thread.
//   SPLIT   latency.  Grid = row tiles x tree chunks so a 1-1024 row request still covers the
//           chip; per-(tree,row) leaf values go to scratch and a second kernel adds them in tree
//           order.
//   GLOBAL  forests whose trees do not fit a shared-memory stage: nodes come from L1/L2/HBM.  Bulk launches
//           take their rows in the order of a counting sort on the forest's most used split features
//           (sort.cu) and then hold two CTAs per SM (latency-bound once a warp's rows share their paths).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "ccfd_internal.h"

namespace {

constexpr int kConsumerWarps = 16;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kStreamThreads = kConsumerThreads + 32;  // + one producer warp
constexpr int kMaxStages = 8;
constexpr int kTileThreads = 256;                      // SPLIT: one row per thread
// GLOBAL: one row per thread.  Bulk batches use one 768-row CTA per SM: all 24 warps of an SM then walk
// the same two trees, so L1 holds those trees' upper levels (100 unbounded trees x 2 M rows: 256-thread
// CTAs 8.97 ms, 512: 11.1, 768: 8.02, 1024: 8.11).  Batches too small to give every SM such a CTA use
// 256-row CTAs.
constexpr int kGlobalThreadsBulk = 768;
constexpr int kGlobalThreadsSmall = 256;
constexpr int kGlobalThreadsWide = 64;  // rows with so many features that 256 of them overflow shared memory
constexpr int kSplitStages = 2;

__host__ __device__ constexpr uint32_t round_up(uint32_t v, uint32_t a) { return (v + a - 1) / a * a; }

// ------------------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA: global -> shared, completion counted in bytes on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
// barrier + OR-reduction of a predicate over the participating threads
__device__ __forceinline__ bool named_bar_or(int id, int nthreads, bool pred) {
    uint32_t r;
    asm volatile(
        "{\n"
        ".reg .pred p, q;\n"
        "setp.ne.u32 q, %1, 0;\n"
        "bar.red.or.pred p, %2, %3, q;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(r)
        : "r"((uint32_t)pred), "r"(id), "r"(nthreads)
        : "memory");
    return r != 0;
}
// Shared-memory loads by 32-bit shared-window address.  The descent keeps node addresses in this form
// so that "next node" is one select and one shift-add (see descend).  volatile: ordered after the
// mbarrier wait that published the bytes.
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ double2 lds128d(uint32_t addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr));
    return v;
}
// streaming 128-bit load: rows are read once, keep them out of L1
__device__ __forceinline__ float4 ld_stream4(const float4* p) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// --------------------------------------------------------------------- row tile -> shared memory
// Copies `rows` rows starting at src into xs[feature * S + row] (transposed) with coalesced
// reads, zero-fills rows [rows, S) so the descent of idle lanes stays on finite values, and
// reports whether this thread saw a NaN.  The row stride S is a multiple of 32 on purpose: the
// bank of xs[f][row] is row % 32 whatever f is, so the descent's data-dependent reads (lane ==
// row, a different f per lane) never conflict.  The price is bank conflicts on these transposed
// stores, paid once per row instead of once per node visit; the STREAM kernel's main path avoids
// even that (see transpose_group).
template <int NT, int S, bool XB = false>
__device__ __forceinline__ bool load_tile(float* xs, const float* __restrict__ src, int rows, int64_t ld, int F, bool vec_ok, int tid) {
    auto at = [](int c, int r) { return XB ? ((r >> 6) * 2048 + c * 64 + (r & 63)) : (c * S + r); };  // see XStrided / XBlocked
    bool nan = false;
    const int total = rows * F;
    if (vec_ok) {
        const float4* src4 = reinterpret_cast<const float4*>(src);
        const int nvec = total >> 2;
#pragma unroll 4
        for (int v = tid; v < nvec; v += NT) {
            const float4 q = ld_stream4(src4 + v);
            const int e = v << 2;
            int r = e / F, c = e - r * F;
            const float qq[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                xs[at(c, r)] = qq[j];
                nan |= (qq[j] != qq[j]);
                if (++c == F) {
                    c = 0;
                    ++r;
                }
            }
        }
        for (int e = (nvec << 2) + tid; e < total; e += NT) {
            const int r = e / F, c = e - r * F;
            const float v = __ldg(src + e);
            xs[at(c, r)] = v;
            nan |= (v != v);
        }
    } else {
        for (int e = tid; e < total; e += NT) {
            const int r = e / F, c = e - r * F;
            const float v = __ldg(src + (int64_t)r * ld + c);
            xs[at(c, r)] = v;
            nan |= (v != v);
        }
    }
    if (rows < S) {
        const int pad = (S - rows) * F;
        for (int e = tid; e < pad; e += NT) {
            const int r = e / F, c = e - r * F;
            xs[at(c, rows + r)] = 0.0f;
        }
    }
    return nan;
}

// ------------------------------------------------------------------------------ the descent
// One step per level, no branch: x <= thr picks the left slot of the sibling pair, anything else
// (including NaN on the NaN-free path) the right one.  With NAN_AWARE a NaN follows the node's
// missing_go_to_left bit (SK/tree/_tree.pyx:984-988).  Leaves step to themselves.
template <bool NAN_AWARE, bool WIDE>
__device__ __forceinline__ bool goes_right(const uint2 nd, const float x) {
    bool right = !(x <= __uint_as_float(nd.x));
    if (NAN_AWARE) right = (x != x) ? !NodeFmt<WIDE>::mgl(nd.y) : right;
    return right;
}

// What a thread needs to walk one tree whose blob sits in shared memory (compact node encoding:
// child references are byte offsets from the first node).  Addresses are 32-bit shared-window
// addresses.
struct TreeView {
    uint32_t nodes;        // node slots [nodes, nodes + last_off): the deepest level's leaf slots are not stored
    uint32_t vals;         // (v0, v1) per leaf, 16 bytes each
    const uint2* top;      // seven top nodes in the constant bank (TOP kernels) or null
    int depth;
    uint32_t last_off;     // byte offset of the first slot of the deepest level
    uint32_t last_voff;    // byte offset, from vals, of that slot's leaf value
};

// Where a thread finds x[feature] of its rows.  Two layouts of the transposed rows:
//   strided  xs[feature * XS + row] (XS a multiple of 32): any feature count.
//   blocked  64-row blocks of 32 feature rows (8 KB, block-aligned in the shared window):
//            address = block | feature << 8 | row_in_block << 2.  The node's meta word carries the
//            feature at bits 8..15 already, so the x address is ONE and-or of meta with the thread's
//            own address bits.  Needs n_features + 1 <= 32.
// Both give bank = row % 32 whatever the feature: the data-dependent reads never conflict.
constexpr int kBlockRows = 64;
constexpr int kBlockFloats = 32 * kBlockRows;      // 2048 floats = 8 KB
constexpr uint32_t kBlockBytes = kBlockFloats * 4u;
constexpr int kMaxBlockedFeatures = 31;            // + the all-zero dummy row = 32

__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

template <int STRIDE>
struct XStrided {
    const float* xcol;  // &xs[row of this thread]
    int XS;
    __device__ __forceinline__ float get(uint32_t meta, int i) const { return xcol[NodeFmt<false>::feature(meta) * XS + i * STRIDE]; }
};
// STRIDE_ROWS: distance in rows between a thread's consecutive rows (a multiple of kBlockRows)
template <int STRIDE_ROWS>
struct XBlocked {
    uint32_t tb;  // shared address of x[feature 0] of this thread's first row; bits 8..12 are zero
    __device__ __forceinline__ float get(uint32_t meta, int i) const {
        return lds_f32(((meta & 0x1F00u) | tb) + (uint32_t)i * (uint32_t)(STRIDE_ROWS / kBlockRows) * kBlockBytes);
    }
};
// float index of (feature 0, row s) in the blocked layout
__device__ __forceinline__ int blocked_row(int s) { return (s >> 6) * kBlockFloats + (s & (kBlockRows - 1)); }

// NT trees x SPT rows = NT*SPT independent pointer chases per thread, interleaved so that the
// shared-memory latency of one is covered by the others.  With TOP the first three levels come
// from the kernel parameters (constant bank, warp-uniform loads: no shared-memory bandwidth) and
// the level-1 / level-2 nodes are fetched with a per-lane index.  Every tree runs exactly `depth` steps, so
// a row ends on its leaf; rows that end on the deepest level (nearly all rows of a depth-limited
// forest) get their value's place from the slot's place, the few others read their leaf slot.
// Per level and chain: node load, feature -> x address, x load, compare, select the base (left or
// right sibling), shift-add the child offset.
// Result: vaddr = shared address of the row's leaf value.
// TL: how many of the kTopLevels levels held in the parameter table are actually walked there (the rest
// come from the tree's own nodes in shared memory, whose top levels are bank-conflict free anyway).
template <int SPT, int NT, bool NAN_AWARE, bool TOP, class XA, int TL = kTopLevels>
__device__ __forceinline__ void descend(const TreeView (&tv)[NT], const XA xa, uint32_t (&vaddr)[NT][SPT]) {
    static_assert(TL >= 1 && TL <= kTopLevels, "top levels walked from the table");
    using Fmt = NodeFmt<false>;
    uint32_t at[NT][SPT];  // shared address of the current slot
    int d0 = 0;
    if (TOP) {
        // The top table is in virtual complete order (children of i at 2i+1, 2i+2; a leaf repeats itself
        // downwards), so a lane walks it with an index and fetches each node with a lane-indexed LDC: the
        // hardware replays the load once per distinct address (<= 2, <= 4 on levels 1, 2) in the constant
        // path, off the shared-memory pipe.  Measured against picking the nodes with selects: 7.13 ->
        // 7.03 ms (hard), 5.21 -> 4.79 ms (easy) per 12.5 M rows; four / five levels this way: 7.04 / 8.37
        // and 5.02 / 5.22 ms.
#pragma unroll
        for (int k = 0; k < NT; ++k) {
            const uint2* top = tv[k].top;
            const uint32_t nb = tv[k].nodes, nb8 = nb + 8u;
#pragma unroll
            for (int i = 0; i < SPT; ++i) {
                uint32_t idx = 0u;
                uint2 nd = top[0];
#pragma unroll
                for (int l = 0; l < TL - 1; ++l) {
                    idx = 2u * idx + 1u + (goes_right<NAN_AWARE, false>(nd, xa.get(nd.y, i)) ? 1u : 0u);
                    nd = top[idx];
                }
                const uint32_t base = goes_right<NAN_AWARE, false>(nd, xa.get(nd.y, i)) ? nb8 : nb;
                at[k][i] = base + Fmt::child_bytes(nd.y);
            }
        }
        d0 = TL;
    } else {
#pragma unroll
        for (int k = 0; k < NT; ++k)
#pragma unroll
            for (int i = 0; i < SPT; ++i) at[k][i] = tv[k].nodes;
    }
    int dmin = tv[0].depth, dmax = tv[0].depth;
#pragma unroll
    for (int k = 1; k < NT; ++k) {
        dmin = min(dmin, tv[k].depth);
        dmax = max(dmax, tv[k].depth);
    }
#pragma unroll 1  // measured: unroll 1 / 2 / 4 = 7.13 / 7.23 / 7.18 ms (hard), 5.21 / 5.32 / 5.36 ms (easy) per 12.5 M rows
    for (int d = d0; d < dmin; ++d) {
#pragma unroll
        for (int k = 0; k < NT; ++k) {
            const uint32_t nb = tv[k].nodes, nb8 = nb + 8u;
#pragma unroll
            for (int i = 0; i < SPT; ++i) {
                const uint2 nd = lds64(at[k][i]);
                const uint32_t base = goes_right<NAN_AWARE, false>(nd, xa.get(nd.y, i)) ? nb8 : nb;
                at[k][i] = base + Fmt::child_bytes(nd.y);
            }
        }
    }
    if (NT > 1) {
        for (int d = max(d0, dmin); d < dmax; ++d) {  // the deeper tree of the pair finishes alone
#pragma unroll
            for (int k = 0; k < NT; ++k)
                if (d < tv[k].depth) {
                    const uint32_t nb = tv[k].nodes, nb8 = nb + 8u;
#pragma unroll
                    for (int i = 0; i < SPT; ++i) {
                        const uint2 nd = lds64(at[k][i]);
                        const uint32_t base = goes_right<NAN_AWARE, false>(nd, xa.get(nd.y, i)) ? nb8 : nb;
                        at[k][i] = base + Fmt::child_bytes(nd.y);
                    }
                }
        }
    }
#pragma unroll
    for (int k = 0; k < NT; ++k) {
        const uint32_t last = tv[k].nodes + tv[k].last_off;
        const uint32_t vlast = tv[k].vals + tv[k].last_voff;
#pragma unroll
        for (int i = 0; i < SPT; ++i) {
            uint32_t v = ((at[k][i] - last) << 1) + vlast;  // 8-byte slot -> 16-byte value
            if (at[k][i] < last) v = tv[k].vals + (lds32(at[k][i]) << 4);
            vaddr[k][i] = v;
        }
    }
}

// =============================================================================== STREAM kernel
// smem: [0,128) mbarriers | xs (F+1 rows of S floats, last row = zeros) | ring of stages.
// The ring carries everything the consumers need, in order, batch after batch:
//   row chunks  (main path) the batch's rows, row-major, R rows per stage, straight from HBM by TMA;
//               each consumer warp transposes the 32-row groups it owns into xs and releases the stage
//   tree groups consecutive tree blobs that fit one stage (one bulk copy per group)
// so the main path has no block-wide barrier and no register-staged global load at all.

// One warp moves 32 rows (lane = row) from a row-major chunk in shared memory into its own columns
// of xs.  Row stride F floats: with 8-byte loads the 16 lanes of a half-warp hit 16 different bank
// pairs (15*r mod 16 is a permutation), and the transposed stores are lane-consecutive.
template <int FS>
__device__ __forceinline__ bool transpose_group(float* __restrict__ xcol, const float* __restrict__ rows_rm, int F, int lane) {
    bool nan = false;
    const float* rp = rows_rm + lane * F;
    if ((F & 1) == 0) {
        const float2* rp2 = reinterpret_cast<const float2*>(rp);
#pragma unroll 5
        for (int c2 = 0; c2 < (F >> 1); ++c2) {
            const float2 v = rp2[c2];
            xcol[(2 * c2) * FS] = v.x;
            xcol[(2 * c2 + 1) * FS] = v.y;
            nan |= (v.x != v.x) | (v.y != v.y);
        }
    } else {
        for (int c = 0; c < F; ++c) {
            const float v = rp[c];
            xcol[c * FS] = v;
            nan |= (v != v);
        }
    }
    return nan;
}

struct TopTable {
    TreeParam t[kMaxParamTrees];
};
struct NoTop {};

// Position of a consumer warp (or of the producer) in the ring.
struct RingCursor {
    int stage = 0;
    uint32_t phase = 0;
    __device__ __forceinline__ void advance(int n_stages) {
        if (++stage == n_stages) {
            stage = 0;
            phase ^= 1u;
        }
    }
};

// RES (forest-resident): a forest small enough to sit in shared memory NEXT to the rows (the Kaggle-like
// model: 100 trees, 60 KB) is loaded once per CTA and stays; the ring then carries row chunks only, and the
// tree loop loses its stage waits, stage releases and group bookkeeping.  Needs TOP (per-tree offsets come
// from the parameter table, here as absolute blob offsets).
template <int SPT, int TPT, bool TOP, bool XB, bool RES = false, int TL = kTopLevels>
__global__ void __launch_bounds__(kStreamThreads, 1)
k_stream(const ForestDev f, const ScoreArgs a, const int n_stages_planned, const int vec_ok, const typename std::conditional<TOP, TopTable, NoTop>::type top) {
    static_assert(!RES || TOP, "the resident variant takes its per-tree offsets from the parameter table");
    constexpr int S = kConsumerThreads * SPT;
    constexpr int XS = S;
    constexpr int FS = XB ? kBlockRows : XS;  // feature stride of xs in floats
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* empty = full + kMaxStages;  // [2 * kMaxStages] = 128 bytes; the resident forest's barrier borrows the last slot
    uint64_t* resbar = empty + kMaxStages - 1;
    const int F = f.n_features;
    const int T = f.n_trees;
    const uint32_t stage_bytes = f.stage_bytes;
    const uint32_t res_bytes = RES ? round_up(f.blob_bytes, 128u) : 0u;
    // blocked rows start on an 8 KB boundary of the shared window (XBlocked); what the padding
    // costs comes out of the ring
    uint32_t xs_off = 128u, xs_bytes = round_up((uint32_t)(F + 1) * XS * 4u, 128u);
    int n_stages = n_stages_planned;
    if (XB) {
        const uint32_t base = smem_u32(smem);
        xs_off = ((base + 128u + kBlockBytes - 1u) & ~(kBlockBytes - 1u)) - base;
        xs_bytes = (uint32_t)(S / kBlockRows) * kBlockBytes;
        uint32_t total;
        asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(total));
        n_stages = min(n_stages, (int)((total - xs_off - xs_bytes - res_bytes) / stage_bytes));
    }
    if (RES) n_stages = min(n_stages, kMaxStages - 1);  // the last barrier slot is resbar
    float* xs = reinterpret_cast<float*>(smem + xs_off);
    uint8_t* resident = smem + xs_off + xs_bytes;  // RES: the whole forest blob, loaded once
    uint8_t* ring = resident + res_bytes;
    const int tid = threadIdx.x;
    const int64_t n_batches = (a.n + S - 1) / S;
    const int R = (int)f.chunk_rows;  // rows per row chunk, a multiple of 32
    const int n_chunks = (S + R - 1) / R;

    if (tid == 0) {
        for (int s = 0; s < n_stages; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, kConsumerWarps);
        }
        if (RES) mbar_init(resbar, 1);
        mbar_fence_init();
    }
    for (int i = tid; i < S; i += kStreamThreads) xs[(XB ? blocked_row(i) : i) + F * FS] = 0.0f;
    __syncthreads();

    if (tid >= kConsumerThreads) {
        // ---- producer warp: one lane issues every bulk copy of this CTA
        if (tid == kConsumerThreads) {
            RingCursor cur;
            if (RES) {  // the forest, once: bulk copies of at most 32 KB on one barrier
                mbar_expect_tx(resbar, f.blob_bytes);
                for (uint32_t off = 0; off < f.blob_bytes; off += 32768u) {
                    const uint32_t bytes = min(32768u, f.blob_bytes - off);
                    bulk_g2s(resident + off, f.blob + off, bytes, resbar);
                }
            }
            for (int64_t b = blockIdx.x; b < n_batches; b += gridDim.x) {
                const int64_t row0 = b * S;
                const bool fast = vec_ok && (a.n - row0 >= (int64_t)S);
                if (fast) {
                    for (int c = 0; c < n_chunks; ++c) {
                        const int r_lo = c * R;
                        const int r_n = (S - r_lo) < R ? (S - r_lo) : R;
                        const uint32_t bytes = (uint32_t)r_n * (uint32_t)F * 4u;
                        mbar_wait(empty + cur.stage, cur.phase ^ 1u);
                        mbar_expect_tx(full + cur.stage, bytes);
                        bulk_g2s(ring + (size_t)cur.stage * stage_bytes, a.X + (row0 + r_lo) * (int64_t)F, bytes, full + cur.stage);
                        cur.advance(n_stages);
                    }
                }
                for (int g = 0; g < (RES ? 0 : f.n_groups); ++g) {
                    const int4 gr = __ldg(reinterpret_cast<const int4*>(f.groups + g));  // GroupRec
                    mbar_wait(empty + cur.stage, cur.phase ^ 1u);
                    mbar_expect_tx(full + cur.stage, (uint32_t)gr.w);
                    bulk_g2s(ring + (size_t)cur.stage * stage_bytes, f.blob + (size_t)(uint32_t)gr.z * 16u, (uint32_t)gr.w, full + cur.stage);
                    cur.advance(n_stages);
                }
            }
        }
        return;
    }

    // ---- consumers
    const int lane = tid & 31, warp = tid >> 5;
    const double Td = (double)T;
    typename std::conditional<XB, XBlocked<kConsumerThreads>, XStrided<kConsumerThreads>>::type xa;
    if constexpr (XB)
        xa.tb = smem_u32(xs) + (uint32_t)blocked_row(tid) * 4u;
    else {
        xa.xcol = xs + tid;
        xa.XS = XS;
    }
    RingCursor acq, rel;  // next stage to wait for / oldest stage this warp still holds
    if (RES) mbar_wait(resbar, 0u);
    for (int64_t b = blockIdx.x; b < n_batches; b += gridDim.x) {
        const int64_t row0 = b * S;
        const int rows = (int)((a.n - row0) < (int64_t)S ? (a.n - row0) : (int64_t)S);
        bool any_nan;
        if (vec_ok && rows == S) {
            // main path: each warp lifts the 32-row groups it owns out of the row chunks
            bool my_nan = false;
            const int k = R >> 5;  // groups per chunk
            for (int c = 0; c < n_chunks; ++c) {
                const int cnt = (((S - c * R) < R ? (S - c * R) : R)) >> 5;
                mbar_wait(full + acq.stage, acq.phase);
                const float* chunk = reinterpret_cast<const float*>(ring + (size_t)acq.stage * stage_bytes);
                for (int lg = (warp - c * k) & (kConsumerWarps - 1); lg < cnt; lg += kConsumerWarps) {
                    const int i = (c * k + lg) / kConsumerWarps;  // which of this thread's SPT rows
                    const int s = tid + i * kConsumerThreads;
                    my_nan |= transpose_group<FS>(xs + (XB ? blocked_row(s) : s), chunk + (size_t)lg * 32 * F, F, lane);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(empty + acq.stage);
                acq.advance(n_stages);
            }
            rel = acq;
            any_nan = __any_sync(0xffffffffu, my_nan);  // a warp only ever reads its own rows
        } else {
            // ragged last batch, strided or unaligned rows: cooperative load with block barriers
            named_bar_sync(1, kConsumerThreads);
            const bool my_nan = load_tile<kConsumerThreads, S, XB>(xs, a.X + row0 * a.ld, rows, a.ld, F, vec_ok != 0, tid);
            any_nan = named_bar_or(2, kConsumerThreads, my_nan);
        }

        double acc0[SPT], acc1[SPT];
#pragma unroll
        for (int i = 0; i < SPT; ++i) acc0[i] = acc1[i] = 0.0;

        // Trees are taken TPT at a time.  A tree's stage is waited for when its group starts and
        // handed back once the group's last tree is done, so a pair may straddle two stages.
        const uint8_t* gb = RES ? resident : nullptr;
        for (int t0 = 0; t0 < T; t0 += TPT) {
            TreeView tv[TPT];
            int last_of_group[TPT];
#pragma unroll
            for (int k = 0; k < TPT; ++k) {
                const bool valid = t0 + k < T;
                const int t = valid ? (t0 + k) : (T - 1);  // odd tail: the slot repeats the last tree and is ignored
                uint32_t rel_off, values_off, depth_flags;
                if constexpr (TOP) {
                    const TreeParam& tp = top.t[t];  // constant bank
                    rel_off = tp.rel_off;
                    values_off = tp.values_off;
                    depth_flags = tp.depth_flags;
                    tv[k].last_off = tp.last_pos << 3;
                    tv[k].last_voff = tp.last_vidx << 4;
                    tv[k].top = reinterpret_cast<const uint2*>(tp.top.n);
                } else {
                    const int4 raw = __ldg(reinterpret_cast<const int4*>(f.trees + t));      // TreeRec, first half
                    const int4 raw2 = __ldg(reinterpret_cast<const int4*>(f.trees + t) + 1);  // second half
                    rel_off = (uint32_t)raw2.z;
                    values_off = (uint32_t)raw.z;
                    depth_flags = (uint32_t)raw.w | (((uint32_t)raw2.w & 3u) << 16);
                    tv[k].last_off = (uint32_t)raw2.x << 3;
                    tv[k].last_voff = (uint32_t)raw2.y << 4;
                    tv[k].top = nullptr;
                }
                if (!RES && valid && (depth_flags & 0x10000u)) {  // first tree of a group: its stage must have landed
                    mbar_wait(full + acq.stage, acq.phase);
                    gb = ring + (size_t)acq.stage * stage_bytes;
                    acq.advance(n_stages);
                }
                const uint8_t* tb = gb + rel_off;
                tv[k].nodes = smem_u32(tb);
                tv[k].vals = tv[k].nodes + values_off;
                tv[k].depth = (int)(depth_flags & 0xFFFFu);
                last_of_group[k] = !RES && valid && (depth_flags & 0x20000u);
            }
            uint32_t voff[TPT][SPT];
            if (TPT == 1 || t0 + TPT <= T) {
                if (!any_nan)
                    descend<SPT, TPT, false, TOP, decltype(xa), TL>(tv, xa, voff);
                else
                    descend<SPT, TPT, true, TOP, decltype(xa), TL>(tv, xa, voff);
            } else {
                // the last, incomplete step: its trees one at a time
#pragma unroll
                for (int k = 0; k < TPT - 1; ++k) {
                    if (t0 + k < T) {
                        TreeView one[1] = {tv[k]};
                        uint32_t v1[1][SPT];
                        if (!any_nan)
                            descend<SPT, 1, false, TOP, decltype(xa), TL>(one, xa, v1);
                        else
                            descend<SPT, 1, true, TOP, decltype(xa), TL>(one, xa, v1);
#pragma unroll
                        for (int i = 0; i < SPT; ++i) voff[k][i] = v1[0][i];
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < TPT; ++k) {
                if (t0 + k < T) {
#pragma unroll
                    for (int i = 0; i < SPT; ++i) {
                        const double2 v = lds128d(voff[k][i]);
                        acc0[i] += v.x;  // tree order, plain adds: SK/ensemble/_forest.py:714
                        acc1[i] += v.y;
                    }
                }
            }
            // hand back every stage whose last tree was in this step (in ring order)
            if constexpr (!RES) {
                __syncwarp();
#pragma unroll
                for (int k = 0; k < TPT; ++k) {
                    if (last_of_group[k]) {
                        if (lane == 0) mbar_arrive(empty + rel.stage);
                        rel.advance(n_stages);
                    }
                }
            }
        }
#pragma unroll
        for (int i = 0; i < SPT; ++i) {
            const int s = tid + i * kConsumerThreads;
            if (s < rows) {
                const double p0 = acc0[i] / Td, p1 = acc1[i] / Td;  // true divide: SK/ensemble/_forest.py:962
                if (a.out) reinterpret_cast<double2*>(a.out)[row0 + s] = make_double2(p0, p1);
                if (a.decision) a.decision[row0 + s] = (uint8_t)(a.ge ? (p1 >= a.threshold) : (p1 > a.threshold));
            }
        }
    }
}

// =============================================================================== GLOBAL kernel
// One row tile per CTA; nodes via the read-only path (L1 -> L2 -> HBM).
// Branch-free like the shared-memory descent: a row that has reached the deepest level (whose leaf
// slots are not stored) is fed a synthetic self-looping leaf instead of a load.  Deep, ragged trees:
// the warp stops as soon as all its lanes sit on leaves.
struct GlobalTree {
    const uint2* nodes;
    const uint8_t* vals;
    int depth;
    uint32_t last_pos, last_vidx;
};

// NT trees walked together by each thread: NT independent chains of dependent L2/HBM loads in flight.
template <int NT, bool NAN_AWARE, bool WIDE>
__device__ __forceinline__ void descend_global(const GlobalTree (&gt)[NT], const float* __restrict__ xcol, int XS, uint32_t leaf_feat, uint32_t (&vidx)[NT]) {
    uint32_t pos[NT];
    int dmax = 0;
#pragma unroll
    for (int k = 0; k < NT; ++k) {
        pos[k] = 0u;
        dmax = max(dmax, gt[k].depth);
    }
    for (int d = 0; d < dmax; ++d) {
        bool all_leaf = true;
#pragma unroll
        for (int k = 0; k < NT; ++k) {
            using Fmt = NodeFmt<WIDE>;
            uint2 nd = make_uint2(0u, Fmt::make(pos[k], false, leaf_feat));  // synthetic self-looping leaf
            if (d < gt[k].depth && pos[k] < gt[k].last_pos) nd = __ldg(gt[k].nodes + pos[k]);
            const uint32_t feat = Fmt::feature(nd.y);
            pos[k] = Fmt::child(nd.y) + (goes_right<NAN_AWARE, WIDE>(nd, xcol[feat * XS]) ? 1u : 0u);
            all_leaf = all_leaf && (feat == leaf_feat);
        }
        if ((d & 3) == 3 && __all_sync(0xffffffffu, all_leaf)) break;
    }
#pragma unroll
    for (int k = 0; k < NT; ++k) {
        uint32_t v = pos[k] - gt[k].last_pos + gt[k].last_vidx;
        if (pos[k] < gt[k].last_pos) v = __ldg(gt[k].nodes + pos[k]).x;
        vidx[k] = v;
    }
}

// kGlobalTrees trees per thread at a time: two independent chains of dependent loads per thread.
// (Measured on unbounded-depth forests with the L1 carve-out below: 1 tree 12.1 ms, 2 trees 9.1 ms,
// 4 trees 10.8 ms per 2 M rows x 100 trees.)
// TWO_PER_SM (pre-sorted bulk launches): with the rows in key order a warp-wide node load touches ~5 sectors
// instead of ~20 and the kernel is bound by the latency of its dependent loads, not by L1 any more
// (profiles/r2_k_global_ncu_summary.txt: long_scoreboard 5.8 stall cycles per issued instruction), so it
// pays to give most of the SM to shared memory (85 % carve-out), hold TWO 768-row CTAs and double the
// chains in flight: 500 unbounded trees x 10 M rows 133.7 -> 101.5 ms.  For rows in arrival order the same
// setting is far worse (237 -> 368 ms: L1 is what keeps that case alive), hence a kernel of its own.
template <int kGlobalTrees, int kGlobalThreads, bool WIDE, bool TWO_PER_SM = false>
__global__ void __launch_bounds__(kGlobalThreads, TWO_PER_SM ? 2 : 1) k_global(const ForestDev f, const ScoreArgs a, const int vec_ok) {
    constexpr int S = kGlobalThreads, XS = S;
    extern __shared__ __align__(128) uint8_t smem[];
    float* xs = reinterpret_cast<float*>(smem);
    const int F = f.n_features, T = f.n_trees, tid = threadIdx.x;
    const int64_t row0 = (int64_t)blockIdx.x * S;
    const int rows = (int)((a.n - row0) < (int64_t)S ? (a.n - row0) : (int64_t)S);
    for (int i = tid; i < XS; i += kGlobalThreads) xs[F * XS + i] = 0.0f;
    bool my_nan;
    if (a.perm) {
        // pre-sorted launch (sort.cu): row r of this tile is row perm[row0 + r] of X.  A warp copies one row
        // at a time (lanes = consecutive features: one or two sectors per row), zero-fills the ragged tail.
        my_nan = false;
        const int lane = tid & 31, warp = tid >> 5;
        for (int r = warp; r < S; r += kGlobalThreads / 32) {
            const float* src = r < rows ? a.X + (int64_t)__ldg(a.perm + row0 + r) * a.ld : nullptr;
            for (int c = lane; c < F; c += 32) {
                const float v = src ? __ldg(src + c) : 0.0f;
                xs[c * XS + r] = v;
                my_nan |= (v != v);
            }
        }
    } else {
        my_nan = load_tile<kGlobalThreads, S>(xs, a.X + row0 * a.ld, rows, a.ld, F, vec_ok != 0, tid);
    }
    const bool any_nan = __syncthreads_or(my_nan) != 0;

    double acc0 = 0.0, acc1 = 0.0;
    for (int t0 = 0; t0 < T; t0 += kGlobalTrees) {
        GlobalTree gt[kGlobalTrees];
#pragma unroll
        for (int k = 0; k < kGlobalTrees; ++k) {
            const int t = min(t0 + k, T - 1);
            const int4 raw = __ldg(reinterpret_cast<const int4*>(f.trees + t));
            const int4 raw2 = __ldg(reinterpret_cast<const int4*>(f.trees + t) + 1);
            const uint8_t* tb = f.blob + (size_t)(uint32_t)raw.x * 16u;
            gt[k].nodes = reinterpret_cast<const uint2*>(tb);
            gt[k].vals = tb + (uint32_t)raw.z;
            gt[k].depth = (t0 + k < T) ? raw.w : 0;  // a slot past the last tree walks nothing and is ignored
            gt[k].last_pos = (uint32_t)raw2.x;
            gt[k].last_vidx = (uint32_t)raw2.y;
        }
        uint32_t vidx[kGlobalTrees];
        if (any_nan)
            descend_global<kGlobalTrees, true, WIDE>(gt, xs + tid, XS, (uint32_t)F, vidx);
        else
            descend_global<kGlobalTrees, false, WIDE>(gt, xs + tid, XS, (uint32_t)F, vidx);
#pragma unroll
        for (int k = 0; k < kGlobalTrees; ++k) {
            if (t0 + k < T) {
                const double2 v = __ldg(reinterpret_cast<const double2*>(gt[k].vals) + vidx[k]);
                acc0 += v.x;  // tree order
                acc1 += v.y;
            }
        }
    }
    if (tid < rows) {
        const double Td = (double)T;
        const double p0 = acc0 / Td, p1 = acc1 / Td;
        const int64_t dst = a.perm ? (int64_t)__ldg(a.perm + row0 + tid) : row0 + tid;
        if (a.out) reinterpret_cast<double2*>(a.out)[dst] = make_double2(p0, p1);
        if (a.decision) a.decision[dst] = (uint8_t)(a.ge ? (p1 >= a.threshold) : (p1 > a.threshold));
    }
}

// ================================================================================ SPLIT kernels
// grid.x = 256-row tiles, grid.y = tree chunks.  Each CTA brings its chunk's trees into a 2-stage
// shared-memory ring by TMA and writes scratch[t][row] = (v0, v1) of the leaf row lands in.
template <bool XB>
__global__ void __launch_bounds__(kTileThreads) k_split_descend(const ForestDev f, const ScoreArgs a, double2* __restrict__ scratch, const int trees_per_cta, const uint32_t stage_bytes, const int vec_ok) {
    constexpr int S = kTileThreads, XS = S;
    constexpr int FS = XB ? kBlockRows : XS;
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    const int F = f.n_features, tid = threadIdx.x;
    uint32_t xs_off = 128u, xs_bytes = round_up((uint32_t)(F + 1) * XS * 4u, 128u);
    if (XB) {  // blocked rows start on an 8 KB boundary of the shared window; the host sized for the worst case
        const uint32_t base = smem_u32(smem);
        xs_off = ((base + 128u + kBlockBytes - 1u) & ~(kBlockBytes - 1u)) - base;
        xs_bytes = (uint32_t)(S / kBlockRows) * kBlockBytes;
    }
    float* xs = reinterpret_cast<float*>(smem + xs_off);
    uint8_t* ring = smem + xs_off + xs_bytes;
    const int t_lo = blockIdx.y * trees_per_cta;
    const int t_hi = min(f.n_trees, t_lo + trees_per_cta);
    const int64_t row0 = (int64_t)blockIdx.x * S;
    const int rows = (int)((a.n - row0) < (int64_t)S ? (a.n - row0) : (int64_t)S);

    if (tid == 0) {
        for (int s = 0; s < kSplitStages; ++s) mbar_init(full + s, 1);
        mbar_fence_init();
        // prefetch the first trees while the tile loads (same thread inits and issues: ordered)
        for (int s = 0; s < kSplitStages && t_lo + s < t_hi; ++s) {
            const TreeRec tr = f.trees[t_lo + s];
            mbar_expect_tx(full + s, tr.blob_bytes);
            bulk_g2s(ring + (size_t)s * stage_bytes, f.blob + (size_t)tr.blob_off16 * 16u, tr.blob_bytes, full + s);
        }
    }
    for (int i = tid; i < S; i += kTileThreads) xs[(XB ? blocked_row(i) : i) + F * FS] = 0.0f;
    const bool my_nan = load_tile<kTileThreads, S, XB>(xs, a.X + row0 * a.ld, rows, a.ld, F, vec_ok != 0, tid);
    const bool any_nan = __syncthreads_or(my_nan) != 0;  // also publishes the mbarrier inits
    typename std::conditional<XB, XBlocked<kBlockRows>, XStrided<0>>::type xa;
    if constexpr (XB)
        xa.tb = smem_u32(xs) + (uint32_t)blocked_row(tid) * 4u;
    else {
        xa.xcol = xs + tid;
        xa.XS = XS;
    }

    int stage = 0;
    uint32_t phase = 0;
    for (int t = t_lo; t < t_hi; ++t) {
        const int4 raw = __ldg(reinterpret_cast<const int4*>(f.trees + t));
        const int4 raw2 = __ldg(reinterpret_cast<const int4*>(f.trees + t) + 1);
        mbar_wait(full + stage, phase);
        const uint8_t* tb = ring + (size_t)stage * stage_bytes;
        TreeView tv[1];
        tv[0].nodes = smem_u32(tb);
        tv[0].vals = tv[0].nodes + (uint32_t)raw.z;
        tv[0].top = nullptr;
        tv[0].depth = raw.w;
        tv[0].last_off = (uint32_t)raw2.x << 3;
        tv[0].last_voff = (uint32_t)raw2.y << 4;
        uint32_t voff[1][1];
        if (!any_nan)
            descend<1, 1, false, false>(tv, xa, voff);
        else
            descend<1, 1, true, false>(tv, xa, voff);
        const double2 v = lds128d(voff[0][0]);
        if (tid < rows) scratch[(int64_t)t * a.n + row0 + tid] = v;
        const int t_next = t + kSplitStages;
        if (t_next < t_hi) {
            __syncthreads();  // everyone is done reading this stage
            if (tid == 0) {
                const TreeRec tr = f.trees[t_next];
                mbar_expect_tx(full + stage, tr.blob_bytes);
                bulk_g2s(ring + (size_t)stage * stage_bytes, f.blob + (size_t)tr.blob_off16 * 16u, tr.blob_bytes, full + stage);
            }
        }
        if (++stage == kSplitStages) {
            stage = 0;
            phase ^= 1u;
        }
    }
}

// out[row] = (sum over t = 0..T-1, in that order, of scratch[t][row]) / T
// A CTA owns 32 rows.  All 128 threads first pull a [<=64 trees][32 rows] tile of scratch into
// shared memory with independent coalesced loads (one L2 round trip instead of one per tree), then
// the first warp adds each row's values in tree order.
constexpr int kReduceRows = 32;
constexpr int kReduceTrees = 64;
__global__ void __launch_bounds__(128) k_split_reduce(const double2* __restrict__ scratch, const int T, const ScoreArgs a) {
    __shared__ double2 tile[kReduceTrees][kReduceRows];
    const int64_t row0 = (int64_t)blockIdx.x * kReduceRows;
    const int rows = (int)((a.n - row0) < (int64_t)kReduceRows ? (a.n - row0) : (int64_t)kReduceRows);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double acc0 = 0.0, acc1 = 0.0;
    cudaGridDependencySynchronize();  // the descent kernel's scratch writes are complete and visible from here on
    for (int t0 = 0; t0 < T; t0 += kReduceTrees) {
        const int nt = min(kReduceTrees, T - t0);
        if (t0) __syncthreads();
        double2 v[kReduceTrees / 4];  // all of this thread's loads are issued before any is used
#pragma unroll
        for (int j = 0; j < kReduceTrees / 4; ++j) {
            const int t = w + 4 * j;
            v[j] = (t < nt && lane < rows) ? scratch[(int64_t)(t0 + t) * a.n + row0 + lane] : make_double2(0.0, 0.0);
        }
#pragma unroll
        for (int j = 0; j < kReduceTrees / 4; ++j) tile[w + 4 * j][lane] = v[j];
        __syncthreads();
        if (w == 0 && lane < rows) {
#pragma unroll 8
            for (int t = 0; t < nt; ++t) {
                const double2 v = tile[t][lane];
                acc0 += v.x;
                acc1 += v.y;
            }
        }
    }
    if (w == 0 && lane < rows) {
        const double Td = (double)T;
        const double p0 = acc0 / Td, p1 = acc1 / Td;
        if (a.out) reinterpret_cast<double2*>(a.out)[row0 + lane] = make_double2(p0, p1);
        if (a.decision) a.decision[row0 + lane] = (uint8_t)(a.ge ? (p1 >= a.threshold) : (p1 > a.threshold));
    }
}

// ------------------------------------------------------------------------------- host helpers
struct StreamPlan {
    bool ok = false;
    int n_stages = 0;
    uint32_t stage_bytes = 0;
    uint32_t smem_bytes = 0;
};

// Rows in the blocked layout (one and-or per x address) whenever the features and the dummy row fit
// a block's 32 feature rows.
bool stream_blocked(const ccfd_forest* f) {
    static const bool off = getenv("CCFD_NO_BLOCKED_ROWS") != nullptr;  // tuning knob for A/B runs
    return !off && f->dev.n_features <= kMaxBlockedFeatures;
}

StreamPlan plan_stream(const ccfd_forest* f, int spt) {
    StreamPlan p;
    const uint32_t S = kConsumerThreads * spt;
    const bool xb = stream_blocked(f);
    // blocked: the rows start on an 8 KB boundary of the shared window; the kernel finds out where that
    // is, the plan assumes the worst case and asks for all the shared memory there is
    const uint32_t xs_bytes = xb ? (S / kBlockRows) * kBlockBytes : round_up((uint32_t)(f->dev.n_features + 1) * S * 4u, 128u);
    const uint32_t head = xb ? 128u + kBlockBytes : 128u;
    p.stage_bytes = f->dev.stage_bytes;
    const int64_t budget = (int64_t)f->smem_optin - (int64_t)head - (int64_t)xs_bytes;
    if (budget < 2 * (int64_t)p.stage_bytes) return p;
    p.n_stages = (int)(budget / p.stage_bytes);
    if (p.n_stages > kMaxStages) p.n_stages = kMaxStages;
    p.smem_bytes = xb ? (uint32_t)f->smem_optin : 128 + xs_bytes + (uint32_t)p.n_stages * p.stage_bytes;
    p.ok = true;
    return p;
}

// Forest-resident STREAM variant: blocked rows, parameter table, and the whole blob fits beside the rows
// with at least two row-chunk stages left.
bool plan_resident(const ccfd_forest* f, int spt, StreamPlan* out) {
    const bool off = getenv("CCFD_NO_RESIDENT") != nullptr;  // knob for A/B runs and tests (read per launch)
    if (off || !stream_blocked(f) || f->host_params_res.empty() || f->dev.blob_bytes == 0) return false;
    const uint32_t S = kConsumerThreads * spt;
    const int64_t budget = (int64_t)f->smem_optin - 128 - kBlockBytes - (int64_t)(S / kBlockRows) * kBlockBytes - (int64_t)round_up(f->dev.blob_bytes, 128u);
    if (budget < 2 * (int64_t)f->dev.stage_bytes) return false;
    out->stage_bytes = f->dev.stage_bytes;
    out->n_stages = (int)std::min<int64_t>(budget / f->dev.stage_bytes, kMaxStages - 1);
    out->smem_bytes = (uint32_t)f->smem_optin;
    out->ok = true;
    return true;
}

// SPLIT tile: rows (blocked layout: worst-case padding to the next 8 KB boundary included)
uint32_t tile_xs_bytes(const ccfd_forest* f, bool blocked) {
    if (blocked) return kBlockBytes + (kTileThreads / kBlockRows) * kBlockBytes;
    return round_up((uint32_t)(f->dev.n_features + 1) * kTileThreads * 4u, 128u);
}

bool split_ok(const ccfd_forest* f, uint32_t* smem_bytes, bool blocked = false) {
    const uint32_t stage = round_up(f->dev.max_tree_bytes, 128u);
    const uint32_t need = 128 + tile_xs_bytes(f, blocked) + kSplitStages * stage;
    if (smem_bytes) *smem_bytes = need;
    // keep it small enough for 2 CTAs per SM so short requests overlap
    return need <= (uint32_t)f->smem_optin / 2;
}

}  // namespace

int ccfd_kernels_init(ccfd_forest* f) {
    DeviceGuard guard(f->device);
    CCFD_CUDA(cudaDeviceGetAttribute(&f->sm_count, cudaDevAttrMultiProcessorCount, f->device));
    CCFD_CUDA(cudaDeviceGetAttribute(&f->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, f->device));
    int major = 0;
    CCFD_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, f->device));
    if (major != 10) return ccfd_fail(CCFD_EUNSUPPORTED, "libccfd_rf is built for sm_100a (B200); device %d has compute capability major %d", f->device, major);
#define CCFD_STREAM_ATTRS(SPT_, TPT_)                                                                                                       \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin)); \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));  \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));  \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
#define CCFD_RES_ATTRS(SPT_, TPT_) \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, true, true, true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin)); \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, true, true, true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin)); \
    CCFD_CUDA(cudaFuncSetAttribute(k_stream<SPT_, TPT_, true, true, true, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
    CCFD_RES_ATTRS(1, 2)
    CCFD_RES_ATTRS(2, 2)
    CCFD_RES_ATTRS(1, 4)
    CCFD_RES_ATTRS(2, 4)
#undef CCFD_RES_ATTRS
    CCFD_STREAM_ATTRS(1, 1)
    CCFD_STREAM_ATTRS(2, 1)
    CCFD_STREAM_ATTRS(1, 2)
    CCFD_STREAM_ATTRS(2, 2)
#undef CCFD_STREAM_ATTRS
    CCFD_CUDA(cudaFuncSetAttribute(k_split_descend<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
    CCFD_CUDA(cudaFuncSetAttribute(k_split_descend<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
    {
        // GLOBAL walks nodes through L1/L2: leave most of the SM's 256 KB to L1 so that the upper levels
        // of the trees the resident CTAs are on stay in L1.  Measured on a 100-tree unbounded forest,
        // 2 M rows, 256-thread CTAs: default carve-out 25.0 ms, 40 % 9.1 ms, 60 % 12.1 ms.
        int carveout = 40;
        if (const char* e = getenv("CCFD_GLOBAL_CARVEOUT")) carveout = atoi(e);  // tuning knob
#define CCFD_GLOBAL_ATTRS(NT_, TH_)                                                                                                 \
    CCFD_CUDA(cudaFuncSetAttribute(k_global<NT_, TH_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));  \
    CCFD_CUDA(cudaFuncSetAttribute(k_global<NT_, TH_, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout));    \
    CCFD_CUDA(cudaFuncSetAttribute(k_global<NT_, TH_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));   \
    CCFD_CUDA(cudaFuncSetAttribute(k_global<NT_, TH_, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout));
        CCFD_GLOBAL_ATTRS(1, kGlobalThreadsBulk)
        CCFD_GLOBAL_ATTRS(4, kGlobalThreadsBulk)
        CCFD_GLOBAL_ATTRS(4, kGlobalThreadsSmall)
        CCFD_GLOBAL_ATTRS(4, kGlobalThreadsWide)
        CCFD_GLOBAL_ATTRS(2, kGlobalThreadsBulk)
        CCFD_GLOBAL_ATTRS(1, kGlobalThreadsSmall)
        CCFD_GLOBAL_ATTRS(2, kGlobalThreadsSmall)
        CCFD_GLOBAL_ATTRS(1, kGlobalThreadsWide)
        CCFD_GLOBAL_ATTRS(2, kGlobalThreadsWide)
#undef CCFD_GLOBAL_ATTRS
        int carveout_sorted = 85;
        if (const char* e = getenv("CCFD_GLOBAL_CARVEOUT_SORTED")) carveout_sorted = atoi(e);  // tuning knob
        CCFD_CUDA(cudaFuncSetAttribute(k_global<2, kGlobalThreadsBulk, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
        CCFD_CUDA(cudaFuncSetAttribute(k_global<2, kGlobalThreadsBulk, false, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout_sorted));
        CCFD_CUDA(cudaFuncSetAttribute(k_global<2, kGlobalThreadsBulk, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, f->smem_optin));
        CCFD_CUDA(cudaFuncSetAttribute(k_global<2, kGlobalThreadsBulk, true, true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout_sorted));
    }

    // STREAM ring geometry: a stage holds the largest tree and at least 128 rows of features.
    const uint32_t F = (uint32_t)f->dev.n_features;
    f->dev.stage_bytes = std::max(round_up(f->dev.max_tree_bytes, 128u), round_up(128u * F * 4u, 128u));
    f->dev.chunk_rows = f->dev.stage_bytes / (F * 4u) / 32u * 32u;
    f->dev.groups = nullptr;
    f->dev.n_groups = 0;
    f->dev.wide_nodes = (f->host_trees[0].group_flags & kWideNodesFlag) ? 1u : 0u;
    f->info.stream_kernel_ok = (!f->dev.wide_nodes && plan_stream(f, 1).ok && split_ok(f, nullptr)) ? 1 : 0;
    if (f->info.stream_kernel_ok) {
        std::vector<GroupRec> groups;
        for (int t = 0; t < f->dev.n_trees; ++t) {
            TreeRec& tr = f->host_trees[(size_t)t];
            if (tr.depth > 0xFFFFu) return ccfd_fail(CCFD_EUNSUPPORTED, "tree %d is %u levels deep", t, tr.depth);
            if (groups.empty() || groups.back().bytes + tr.blob_bytes > f->dev.stage_bytes) {
                if (!groups.empty()) f->host_trees[(size_t)t - 1].group_flags |= 2u;
                groups.push_back(GroupRec{(uint32_t)t, 0u, tr.blob_off16, 0u});
                tr.group_flags |= 1u;
            }
            tr.rel_off = (tr.blob_off16 - groups.back().blob_off16) * 16u;
            groups.back().n_trees += 1;
            groups.back().bytes += tr.blob_bytes;
        }
        f->host_trees.back().group_flags |= 2u;
        GroupRec* dg = nullptr;
        CCFD_CUDA(cudaMalloc(&dg, sizeof(GroupRec) * groups.size()));
        cudaError_t e = cudaMemcpy(dg, groups.data(), sizeof(GroupRec) * groups.size(), cudaMemcpyHostToDevice);
        if (e == cudaSuccess)  // the tree records now carry their place in the ring
            e = cudaMemcpy((void*)f->dev.trees, f->host_trees.data(), sizeof(TreeRec) * f->host_trees.size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) {
            cudaFree(dg);
            return ccfd_cuda_fail(e, "group table upload");
        }
        f->dev.groups = dg;
        f->dev.n_groups = (int32_t)groups.size();
        if (f->dev.n_trees <= kMaxParamTrees) {
            f->host_params.resize((size_t)f->dev.n_trees);
            for (int t = 0; t < f->dev.n_trees; ++t) {
                const TreeRec& tr = f->host_trees[(size_t)t];
                TreeParam& tp = f->host_params[(size_t)t];
                tp.top = f->host_top[(size_t)t];
                tp.rel_off = tr.rel_off;
                tp.values_off = tr.values_off;
                tp.depth_flags = tr.depth | ((tr.group_flags & 3u) << 16);
                tp.last_pos = tr.last_pos;
                tp.last_vidx = tr.last_vidx;
                tp.pad = 0;
            }
            // the same table for the forest-resident variant: absolute blob offsets, no group flags
            f->host_params_res = f->host_params;
            for (int t = 0; t < f->dev.n_trees; ++t) {
                f->host_params_res[(size_t)t].rel_off = f->host_trees[(size_t)t].blob_off16 * 16u;
                f->host_params_res[(size_t)t].depth_flags = f->host_trees[(size_t)t].depth;
            }
        }
    }
    return CCFD_OK;
}

// The row pre-sort (sort.cu) pays where node loads go through L1/L2 and the launch is long enough to hide
// it: bulk GLOBAL launches.  (For the shared-memory STREAM kernel it was built and measured, and does not
// pay: DESIGN.md section 5.)  CCFD_ROW_SORT=0 switches it off (A/B runs, tests), CCFD_ROW_SORT_MIN moves the
// threshold.
bool ccfd_row_sort_wanted(const ccfd_forest* f, uint32_t kernel, int64_t n) {
    if (kernel != CCFD_KERNEL_GLOBAL || f->sort_key.n_feat < 4 || n > 0xFFFFFFFFll) return false;
    int64_t min_rows = 262144;
    if (const char* e = getenv("CCFD_ROW_SORT")) {  // read per call: a knob for experiments, not a hot path
        if (atoi(e) == 0) return false;
    }
    if (const char* e = getenv("CCFD_ROW_SORT_MIN")) min_rows = atoll(e);
    return n >= std::max<int64_t>(min_rows, 1);
}

uint32_t ccfd_pick_kernel(const ccfd_forest* f, int64_t n) {
    if (!f->info.stream_kernel_ok) return CCFD_KERNEL_GLOBAL;
    // Measured crossover on the 100-tree forest: SPLIT 44 / 64 / 86 us at 16k / 32k / 48k rows, STREAM a
    // flat ~76 us until its CTAs take a second batch.  SPLIT's scratch is kept under 64 MiB.
    const bool small = n <= 32768 && ccfd_split_scratch_bytes(f, n) <= ((int64_t)64 << 20);
    return small ? CCFD_KERNEL_SPLIT : CCFD_KERNEL_STREAM;
}

int64_t ccfd_split_scratch_bytes(const ccfd_forest* f, int64_t n) { return n * (int64_t)f->dev.n_trees * (int64_t)sizeof(double2); }

int ccfd_launch_score(const ccfd_forest* f, const ScoreArgs& a, uint32_t kernel, void* scratch, cudaStream_t stream) {
    if (a.n <= 0) return CCFD_OK;
    if (a.out && (reinterpret_cast<uintptr_t>(a.out) & 15u)) return ccfd_fail(CCFD_EINVAL, "out must be 16-byte aligned");
    if (reinterpret_cast<uintptr_t>(a.X) & 3u) return ccfd_fail(CCFD_EINVAL, "X must be 4-byte aligned");
    const int vec_ok = (a.ld == f->dev.n_features) && ((reinterpret_cast<uintptr_t>(a.X) & 15u) == 0);
    if (kernel == CCFD_KERNEL_AUTO) kernel = ccfd_pick_kernel(f, a.n);
    if (a.perm && kernel != CCFD_KERNEL_GLOBAL) return ccfd_fail(CCFD_EINVAL, "internal: a row permutation is only understood by the GLOBAL kernel");

    if (kernel == CCFD_KERNEL_STREAM) {
        // two rows per thread (1,024 resident rows per CTA) as soon as 512-row batches would need
        // a second wave: a 1,024-row batch costs ~1.2x a 512-row one, not 2x (more chains in flight)
        int spt = (a.n > (int64_t)f->sm_count * kConsumerThreads) ? 2 : 1;
        {
            // small forests: resident variant (the ring carries rows only)
            StreamPlan rp;
            int rspt = spt;
            bool res = plan_resident(f, rspt, &rp);
            if (!res && rspt == 2) res = plan_resident(f, rspt = 1, &rp);
            if (res && f->info.stream_kernel_ok) {
                const int64_t S = (int64_t)kConsumerThreads * rspt;
                const int64_t n_batches = (a.n + S - 1) / S;
                const int grid = (int)(n_batches < f->sm_count ? n_batches : f->sm_count);
                int tpt = 4;
                if (const char* e = getenv("CCFD_STREAM_TPT")) tpt = atoi(e) == 2 ? 2 : 4;  // tuning knob
                TopTable top;
                memcpy(top.t, f->host_params_res.data(), sizeof(TreeParam) * f->host_params_res.size());
                int tl = kTopLevels;
                if (const char* e = getenv("CCFD_RES_TOP")) tl = std::min(std::max(atoi(e), 1), kTopLevels);  // tuning knob
#define CCFD_LAUNCH_RES(SPT_, TPT_)                                                                                                              \
    do {                                                                                                                                         \
        if (tl == 1) k_stream<SPT_, TPT_, true, true, true, 1><<<grid, kStreamThreads, rp.smem_bytes, stream>>>(f->dev, a, rp.n_stages, vec_ok, top);      \
        else if (tl == 2) k_stream<SPT_, TPT_, true, true, true, 2><<<grid, kStreamThreads, rp.smem_bytes, stream>>>(f->dev, a, rp.n_stages, vec_ok, top); \
        else k_stream<SPT_, TPT_, true, true, true, 3><<<grid, kStreamThreads, rp.smem_bytes, stream>>>(f->dev, a, rp.n_stages, vec_ok, top);              \
    } while (0)
                if (rspt == 2 && tpt == 4) CCFD_LAUNCH_RES(2, 4);
                else if (rspt == 2) CCFD_LAUNCH_RES(2, 2);
                else if (tpt == 4) CCFD_LAUNCH_RES(1, 4);
                else CCFD_LAUNCH_RES(1, 2);
#undef CCFD_LAUNCH_RES
                CCFD_CUDA(cudaGetLastError());
                return CCFD_OK;
            }
        }
        StreamPlan p = plan_stream(f, spt);
        if (!p.ok && spt == 2) p = plan_stream(f, spt = 1);
        if (!f->info.stream_kernel_ok) p.ok = false;  // no group table was built for this forest
        if (!p.ok) return ccfd_fail(CCFD_EUNSUPPORTED, "largest tree (%u bytes) does not fit the shared-memory ring of the STREAM kernel", f->dev.max_tree_bytes);
        const int64_t S = (int64_t)kConsumerThreads * spt;
        const int64_t n_batches = (a.n + S - 1) / S;
        const int grid = (int)(n_batches < f->sm_count ? n_batches : f->sm_count);
        // the 16-byte alignment of each batch's first row holds because S * F * 4 is a multiple of 16
        // Two trees at a time doubles the independent chains per thread; it needs a ring deep enough
        // to hold a pair and still prefetch (>= 4 single-tree stages, or stages that hold >= 2 trees).
        int tpt = (p.n_stages >= 4 || 2 * f->dev.n_groups <= f->dev.n_trees) ? 2 : 1;
        bool use_top = !f->host_params.empty();
        if (const char* e = getenv("CCFD_STREAM_TPT")) tpt = atoi(e) == 2 ? 2 : 1;   // tuning knobs for experiments
        if (const char* e = getenv("CCFD_STREAM_TOP")) use_top = use_top && atoi(e) != 0;
        const bool xb = stream_blocked(f);
#define CCFD_LAUNCH_STREAM_X(SPT_, TPT_, XB_)                                                                                            \
    do {                                                                                                                                 \
        if (use_top) {                                                                                                                   \
            static_assert(sizeof(TopTable) + 256 < 32764, "kernel parameter space");                                                     \
            TopTable top; /* per-tree metadata and top three levels travel as kernel parameters (constant bank) */                        \
            memcpy(top.t, f->host_params.data(), sizeof(TreeParam) * f->host_params.size());                                             \
            k_stream<SPT_, TPT_, true, XB_><<<grid, kStreamThreads, p.smem_bytes, stream>>>(f->dev, a, p.n_stages, vec_ok, top);          \
        } else {                                                                                                                         \
            k_stream<SPT_, TPT_, false, XB_><<<grid, kStreamThreads, p.smem_bytes, stream>>>(f->dev, a, p.n_stages, vec_ok, NoTop{});     \
        }                                                                                                                                \
    } while (0)
#define CCFD_LAUNCH_STREAM(SPT_, TPT_)                 \
    do {                                               \
        if (xb) CCFD_LAUNCH_STREAM_X(SPT_, TPT_, true); \
        else CCFD_LAUNCH_STREAM_X(SPT_, TPT_, false);   \
    } while (0)
        if (spt == 2 && tpt == 2) CCFD_LAUNCH_STREAM(2, 2);
        else if (spt == 2) CCFD_LAUNCH_STREAM(2, 1);
        else if (tpt == 2) CCFD_LAUNCH_STREAM(1, 2);
        else CCFD_LAUNCH_STREAM(1, 1);
#undef CCFD_LAUNCH_STREAM_X
#undef CCFD_LAUNCH_STREAM
    } else if (kernel == CCFD_KERNEL_SPLIT) {
        uint32_t smem_bytes = 0, smem_blocked = 0;
        if (!split_ok(f, &smem_bytes) || !f->info.stream_kernel_ok) return ccfd_fail(CCFD_EUNSUPPORTED, "largest tree (%u bytes) does not fit the shared-memory ring of the SPLIT kernel", f->dev.max_tree_bytes);
        if (!scratch) return ccfd_fail(CCFD_EINVAL, "SPLIT kernel needs scratch");
        const int64_t tiles = (a.n + kTileThreads - 1) / kTileThreads;
        if (tiles > 0x7FFFFFFFll) return ccfd_fail(CCFD_EINVAL, "too many rows for one launch");
        // aim at two CTAs per SM over the whole grid
        int64_t chunks = (2 * (int64_t)f->sm_count + tiles - 1) / tiles;
        if (chunks > f->dev.n_trees) chunks = f->dev.n_trees;
        if (chunks < 1) chunks = 1;
        const int trees_per_cta = (int)((f->dev.n_trees + chunks - 1) / chunks);
        chunks = (f->dev.n_trees + trees_per_cta - 1) / trees_per_cta;
        if (chunks > 65535) return ccfd_fail(CCFD_EINVAL, "too many tree chunks");
        const uint32_t stage_bytes = round_up(f->dev.max_tree_bytes, 128u);
        dim3 grid((unsigned)tiles, (unsigned)chunks);
        // The blocked row layout saves instructions on every level but its padding to an 8 KB boundary
        // costs the third resident CTA per SM: use it when the grid fits two CTAs per SM anyway
        // (1,024 rows: 23.6 -> 22.1 us; 16,384 rows, 320 CTAs: 39 us strided, 54 us blocked).
        if (stream_blocked(f) && tiles * chunks <= 2 * (int64_t)f->sm_count && split_ok(f, &smem_blocked, true))
            k_split_descend<true><<<grid, kTileThreads, smem_blocked, stream>>>(f->dev, a, static_cast<double2*>(scratch), trees_per_cta, stage_bytes, vec_ok);
        else
            k_split_descend<false><<<grid, kTileThreads, smem_bytes, stream>>>(f->dev, a, static_cast<double2*>(scratch), trees_per_cta, stage_bytes, vec_ok);
        {
            // Programmatic dependent launch: the reduce grid may be scheduled while the descent drains and
            // waits in cudaGridDependencySynchronize() for the descent's memory to be visible, which takes
            // the second launch's latency off the request's critical path.
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)((a.n + kReduceRows - 1) / kReduceRows));
            cfg.blockDim = dim3(128);
            cfg.dynamicSmemBytes = 0;
            cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            CCFD_CUDA(cudaLaunchKernelEx(&cfg, k_split_reduce, static_cast<const double2*>(scratch), (int)f->dev.n_trees, a));
        }
    } else if (kernel == CCFD_KERNEL_GLOBAL) {
        auto xs_bytes = [&](int threads) { return round_up((uint32_t)(f->dev.n_features + 1) * (uint32_t)threads * 4u, 128u); };
        int threads = a.n >= (int64_t)f->sm_count * kGlobalThreadsBulk ? kGlobalThreadsBulk : kGlobalThreadsSmall;
        if (threads == kGlobalThreadsBulk && xs_bytes(threads) > (uint32_t)f->smem_optin) threads = kGlobalThreadsSmall;
        if (threads == kGlobalThreadsSmall && xs_bytes(threads) > (uint32_t)f->smem_optin) threads = kGlobalThreadsWide;
        const int64_t tiles = (a.n + threads - 1) / threads;
        if (tiles > 0x7FFFFFFFll) return ccfd_fail(CCFD_EINVAL, "too many rows for one launch");
        const uint32_t gsmem = xs_bytes(threads);
        int nt = 2;
        if (const char* e = getenv("CCFD_GLOBAL_NT")) nt = atoi(e);  // tuning knob
        const bool wide = f->dev.wide_nodes != 0;
#define CCFD_LAUNCH_GLOBAL(TH_)                                                                        \
    do {                                                                                               \
        if (nt == 4 && wide) k_global<4, TH_, true><<<(unsigned)tiles, TH_, gsmem, stream>>>(f->dev, a, vec_ok);        \
        else if (nt == 2 && wide) k_global<2, TH_, true><<<(unsigned)tiles, TH_, gsmem, stream>>>(f->dev, a, vec_ok);        \
        else if (nt == 2) k_global<2, TH_, false><<<(unsigned)tiles, TH_, gsmem, stream>>>(f->dev, a, vec_ok);          \
        else if (wide) k_global<1, TH_, true><<<(unsigned)tiles, TH_, gsmem, stream>>>(f->dev, a, vec_ok);              \
        else k_global<1, TH_, false><<<(unsigned)tiles, TH_, gsmem, stream>>>(f->dev, a, vec_ok);                       \
    } while (0)
        const bool one_per_sm = getenv("CCFD_GLOBAL_ONE_PER_SM") != nullptr;  // tuning knob (read per launch): sorted launches as before
        if (a.perm && threads == kGlobalThreadsBulk && nt == 2 && !one_per_sm && 2 * gsmem + 4096 <= (uint32_t)f->smem_optin) {
            // pre-sorted bulk launch: two CTAs per SM (see k_global's comment)
            if (wide) k_global<2, kGlobalThreadsBulk, true, true><<<(unsigned)tiles, kGlobalThreadsBulk, gsmem, stream>>>(f->dev, a, vec_ok);
            else k_global<2, kGlobalThreadsBulk, false, true><<<(unsigned)tiles, kGlobalThreadsBulk, gsmem, stream>>>(f->dev, a, vec_ok);
        } else if (threads == kGlobalThreadsBulk) CCFD_LAUNCH_GLOBAL(kGlobalThreadsBulk);
        else if (threads == kGlobalThreadsSmall) CCFD_LAUNCH_GLOBAL(kGlobalThreadsSmall);
        else CCFD_LAUNCH_GLOBAL(kGlobalThreadsWide);
#undef CCFD_LAUNCH_GLOBAL
    } else {
        return ccfd_fail(CCFD_EINVAL, "unknown kernel selector 0x%x", kernel);
    }
    CCFD_CUDA(cudaGetLastError());
    return CCFD_OK;
}
