"""Test helper: numpy walk over the *packed* forest image exactly as the CUDA kernels walk it
(csrc/kernels.cu `step` / `descend_smem`): float32 compare, sibling-pair child slot, self-looping
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
    _lib.check(L.ccfd_forest_pack(ctypes.byref(d), None, 0, None, ctypes.byref(nbytes)))
    blob = np.zeros(nbytes.value, dtype=np.uint8)
    recs = np.zeros((fa.n_trees, 4), dtype=np.uint32)
    _lib.check(L.ccfd_forest_pack(ctypes.byref(d), blob.ctypes.data_as(ctypes.c_void_p), blob.nbytes,
                                  recs.ctypes.data_as(ctypes.c_void_p), ctypes.byref(nbytes)))
    return blob, recs


def predict_proba(blob, recs, n_features, X32):
    n = X32.shape[0]
    xs = np.concatenate([X32, np.zeros((n, 1), np.float32)], axis=1)  # dummy feature row = 0
    acc = np.zeros((n, 2), dtype=np.float64)
    rows = np.arange(n)
    for off16, nbytes, values_off, depth in recs:
        tb = blob[int(off16) * 16: int(off16) * 16 + int(nbytes)]
        nodes = tb[: int(values_off)].view(np.uint32).reshape(-1, 2)
        vals = tb[int(values_off):].view(np.float64).reshape(-1, 2)
        thr_bits, meta = nodes[:, 0], nodes[:, 1]
        thr = thr_bits.view(np.float32)
        pos = np.zeros(n, dtype=np.int64)
        for _ in range(int(depth)):
            m = meta[pos]
            x = xs[rows, m & 0xFF]
            with np.errstate(invalid="ignore"):
                left = x <= thr[pos]
            left = np.where(np.isnan(x), (m & 0x100) != 0, left)
            pos = (m >> 9).astype(np.int64) + np.where(left, 0, 1)
        assert np.all((meta[pos] & 0xFF) == n_features), "descent did not end on a leaf"
        acc += vals[thr_bits[pos]]
    acc /= len(recs)
    return acc
