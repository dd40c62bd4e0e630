"""Test helper: numpy walk over the *packed* forest image exactly as the CUDA kernels walk it
(csrc/kernels.cu `goes_right` / `descend`): float32 compare, sibling-pair child slot, self-looping
leaves reading the all-zero dummy feature, fixed trip count = tree depth, float64 sums in tree
order, true divide.  Lets the CPU suite check converter + re-layout + bit packing without a GPU."""
import ctypes

import numpy as np

from ccfd_demo_summit_b200 import _lib


def _desc(fa):
    keep = [np.ascontiguousarray(x) for x in
            (fa.tree_offsets, fa.feature, fa.threshold, fa.left, fa.right, fa.missing_go_to_left, fa.value)]
    d = _lib.ForestDesc(fa.n_trees, fa.n_features, fa.n_classes, 0,
                        *[x.ctypes.data_as(ctypes.c_void_p) for x in keep])
    return d, keep


def pack(fa):
    L = _lib.lib()
    d, keep = _desc(fa)
    nbytes = ctypes.c_int64()
    _lib.check(L.ccfd_forest_pack(ctypes.byref(d), None, 0, None, None, ctypes.byref(nbytes)))
    blob = np.zeros(nbytes.value, dtype=np.uint8)
    recs = np.zeros((fa.n_trees, 8), dtype=np.uint32)
    top = np.zeros((fa.n_trees, 7, 2), dtype=np.uint32)
    _lib.check(L.ccfd_forest_pack(ctypes.byref(d), blob.ctypes.data_as(ctypes.c_void_p), blob.nbytes,
                                  recs.ctypes.data_as(ctypes.c_void_p), top.ctypes.data_as(ctypes.c_void_p),
                                  ctypes.byref(nbytes)))
    return blob, recs, top


class _Fmt:
    """Field access for the two node encodings (csrc/ccfd_internal.h NodeFmt): compact for forests
    whose trees fit shared memory, wide for the rest (TreeRec.group_flags bit 2)."""

    def __init__(self, wide):
        self.wide = bool(wide)

    def feature(self, meta):
        return (meta & 0xFF) if self.wide else ((meta >> 8) & 0xFF)

    def mgl(self, meta):
        return ((meta & 0x100) != 0) if self.wide else ((meta & 1) != 0)

    def child(self, meta):  # in node slots
        if self.wide:
            return (meta >> 9).astype(np.int64) if isinstance(meta, np.ndarray) else int(meta) >> 9
        off = meta >> 16
        assert np.all(off % 8 == 0), "compact child reference is a byte offset of an 8-byte slot"
        return (off // 8).astype(np.int64) if isinstance(off, np.ndarray) else int(off) // 8


def fmt_of(recs):
    wide = (recs[:, 7] & 4) != 0
    assert wide.all() or not wide.any(), "one encoding per forest"
    return _Fmt(wide.any())


def _right(fmt, thr_bits, meta, x):
    with np.errstate(invalid="ignore"):
        left = x <= thr_bits.view(np.float32)
    left = np.where(np.isnan(x), fmt.mgl(meta), left)
    return ~left


def _step(fmt, thr_bits, meta, x):
    return fmt.child(meta) + _right(fmt, thr_bits, meta, x).astype(np.int64)


def predict_proba(blob, recs, n_features, X32, top=None):
    """top=None: every level from the node array.  top given: the first three levels from the
    seven-node table in virtual complete order (children of i at 2i+1, 2i+2), as the STREAM kernel
    does with lane-indexed constant-bank loads (descend<TOP=true>)."""
    n = X32.shape[0]
    fmt = fmt_of(recs)
    xs = np.concatenate([X32, np.zeros((n, 1), np.float32)], axis=1)  # dummy feature row = 0
    acc = np.zeros((n, 2), dtype=np.float64)
    rows = np.arange(n)
    for t, (off16, nbytes, values_off, depth, last_pos, last_vidx, _r0, _r1) in enumerate(recs):
        tb = blob[int(off16) * 16: int(off16) * 16 + int(nbytes)]
        nodes = tb[: int(values_off)].view(np.uint32).reshape(-1, 2)
        vals = tb[int(values_off):].view(np.float64).reshape(-1, 2)
        thr_bits, meta = nodes[:, 0], nodes[:, 1]
        if top is None:
            pos, d0 = np.zeros(n, dtype=np.int64), 0
        else:
            tp = top[t]
            pick = lambda idx: (tp[idx, 0], tp[idx, 1])  # noqa: E731
            tb0, m0 = tp[0]
            r0 = _right(fmt, np.full(n, tb0, np.uint32), np.full(n, m0, np.uint32), xs[rows, fmt.feature(m0)])
            tb1, m1 = pick(np.where(r0, 2, 1))
            r1 = _right(fmt, tb1, m1, xs[rows, fmt.feature(m1)])
            tb2, m2 = pick(3 + 2 * r0.astype(np.int64) + r1.astype(np.int64))
            pos, d0 = _step(fmt, tb2, m2, xs[rows, fmt.feature(m2)]), 3
        last_pos = int(last_pos)
        assert nodes.shape[0] in (last_pos, last_pos + 1), "node array = slots above the deepest level (+ pad)"
        for _ in range(d0, int(depth)):
            assert np.all(pos < last_pos), "a stored slot is read on every step"
            m = meta[pos]
            pos = _step(fmt, thr_bits[pos], m, xs[rows, fmt.feature(m)])
        # rows on the deepest level compute their value index, the rest read their leaf slot
        shallow = pos < last_pos
        assert np.all(fmt.feature(meta[pos[shallow]]) == n_features), "descent did not end on a leaf"
        vidx = pos - last_pos + int(last_vidx)
        vidx[shallow] = thr_bits[pos[shallow]]
        assert vidx.min() >= 0 and vidx.max() < len(vals)
        acc += vals[vidx]
    acc /= len(recs)
    return acc
