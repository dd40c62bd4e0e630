"""ORACLE (test infrastructure, never the product path).

CPU restatement of the arithmetic behind the demo's fraud model
``predict(X, feature_names)`` -> ``RandomForestClassifier.predict_proba``.
Only tests/, ``__graft_entry__.smoke()`` and bench.py's cpu_baseline leg may
import this package; the product (ccfd_demo_summit_b200) never does.

Where the algorithm lives: the reference repository deploys the model as the
opaque image ``nakfour/modelfull`` (reference deploy/model/modelfull.json:24) and
contains no source for it.  The arithmetic is scikit-learn's, a third-party
dependency the reference does not vendor and does not pin.  This file restates
scikit-learn 1.9.0 (the version installed in this image; SK/ below is its
site-packages directory):

  SK/ensemble/_forest.py:606-625   _validate_X_predict  (cast to float32, reject inf)
  SK/ensemble/_forest.py:921-967   ForestClassifier.predict_proba
  SK/ensemble/_forest.py:704-717   _accumulate_prediction (out += per-tree proba)
  SK/tree/_classes.py:1026-1055    DecisionTreeClassifier.predict_proba
  SK/tree/_tree.pyx:939-945        Tree.predict = value.take(apply(X))
  SK/tree/_tree.pyx:954-996        Tree._apply_dense (the descent, NaN routing)
  SK/tree/_tree.pxd:15-25          Node struct fields

PARITY PIN: the reference holds no golden vectors, tests or fixtures for this
path (it holds no code at all), so parity is *unpinned by the reference*.  The
restatement is instead pinned against scikit-learn 1.9.0 itself, run in this
container with n_jobs=1: tests/golden/*.npz were produced by
tests/golden/make_golden.py and tests/test_oracle.py checks this file and
rf_oracle.c bit-for-bit against them.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


@dataclass
class RawForest:
    """sklearn's own per-node arrays, concatenated over trees, nothing rounded."""

    tree_offsets: np.ndarray  # int64 [T+1]
    left: np.ndarray  # int64 [total]  tree-local, -1 at leaves (SK/tree/_tree.pxd:18)
    right: np.ndarray  # int64 [total]
    feature: np.ndarray  # int64 [total]  -2 at leaves
    threshold: np.ndarray  # float64 [total]
    missing_go_to_left: np.ndarray  # uint8 [total]
    value: np.ndarray  # float64 [total, n_classes] class fractions (SK/tree/_criterion.pyx:473-486)
    n_features: int
    n_classes: int

    @property
    def n_trees(self) -> int:
        return len(self.tree_offsets) - 1

    def save(self, prefix: str) -> dict:
        return {
            f"{prefix}tree_offsets": self.tree_offsets,
            f"{prefix}left": self.left,
            f"{prefix}right": self.right,
            f"{prefix}feature": self.feature,
            f"{prefix}threshold": self.threshold,
            f"{prefix}mgl": self.missing_go_to_left,
            f"{prefix}value": self.value,
            f"{prefix}shape": np.array([self.n_features, self.n_classes], dtype=np.int64),
        }

    @staticmethod
    def load(z, prefix: str) -> "RawForest":
        nf, nc = (int(v) for v in z[f"{prefix}shape"])
        return RawForest(
            z[f"{prefix}tree_offsets"], z[f"{prefix}left"], z[f"{prefix}right"], z[f"{prefix}feature"],
            z[f"{prefix}threshold"], z[f"{prefix}mgl"], z[f"{prefix}value"], nf, nc,
        )


def extract_raw(rf) -> RawForest:
    """Read a fitted sklearn forest's state (SK/tree/_tree.pyx:656-765 properties)."""
    offs, L, R, Fe, Th, M, V = [0], [], [], [], [], [], []
    for est in rf.estimators_:
        t = est.tree_
        L.append(np.asarray(t.children_left, dtype=np.int64))
        R.append(np.asarray(t.children_right, dtype=np.int64))
        Fe.append(np.asarray(t.feature, dtype=np.int64))
        Th.append(np.asarray(t.threshold, dtype=np.float64))
        M.append(np.asarray(t.missing_go_to_left, dtype=np.uint8))
        V.append(np.asarray(t.value, dtype=np.float64)[:, 0, : int(rf.n_classes_)])
        offs.append(offs[-1] + t.node_count)
    return RawForest(
        np.asarray(offs, dtype=np.int64), np.concatenate(L), np.concatenate(R), np.concatenate(Fe),
        np.concatenate(Th), np.concatenate(M), np.ascontiguousarray(np.concatenate(V)),
        int(rf.n_features_in_), int(rf.n_classes_),
    )


def validate_X(X, n_features: int) -> np.ndarray:
    """SK/ensemble/_forest.py:606-625: float32 C-array, 2-D, NaN allowed, inf rejected."""
    X = np.asarray(X)
    if X.ndim != 2:
        raise ValueError(f"Expected 2D array, got {X.ndim}D array instead")
    if X.shape[0] < 1:
        raise ValueError("Found array with 0 sample(s) while a minimum of 1 is required")
    if X.shape[1] != n_features:
        raise ValueError(f"X has {X.shape[1]} features, but the forest is expecting {n_features} features as input")
    with np.errstate(over="ignore"):
        X32 = np.ascontiguousarray(X, dtype=np.float32)
    if np.isinf(X32).any():
        raise ValueError("Input X contains infinity or a value too large for dtype('float32').")
    return X32


def tree_apply(f: RawForest, t: int, X32: np.ndarray) -> np.ndarray:
    """Leaf id per row for tree ``t`` (SK/tree/_tree.pyx:977-994), vectorised per level."""
    lo = int(f.tree_offsets[t])
    left, right = f.left[lo:], f.right[lo:]
    feat, thr, mgl = f.feature[lo:], f.threshold[lo:], f.missing_go_to_left[lo:]
    node = np.zeros(X32.shape[0], dtype=np.int64)
    idx = np.nonzero(left[node] != -1)[0]
    while idx.size:
        nd = node[idx]
        x = X32[idx, feat[nd]]
        # float32 feature against float64 threshold: C promotes x (SK/tree/_tree.pyx:989)
        go_left = np.where(np.isnan(x), mgl[nd] != 0, x.astype(np.float64) <= thr[nd])
        nxt = np.where(go_left, left[nd], right[nd])
        node[idx] = nxt
        idx = idx[left[nxt] != -1]
    return node


def predict_proba(f: RawForest, X) -> np.ndarray:
    """SK/ensemble/_forest.py:951-962 with n_jobs=1: zeros, += in tree order, true divide."""
    X32 = validate_X(X, f.n_features)
    out = np.zeros((X32.shape[0], f.n_classes), dtype=np.float64)
    for t in range(f.n_trees):
        leaf = tree_apply(f, t, X32) + int(f.tree_offsets[t])
        out += f.value.take(leaf, axis=0)
    out /= f.n_trees
    return out


def apply(f: RawForest, X) -> np.ndarray:
    """(n, T) tree-local leaf ids, like ``RandomForestClassifier.apply``."""
    X32 = validate_X(X, f.n_features)
    return np.stack([tree_apply(f, t, X32) for t in range(f.n_trees)], axis=1)


# --------------------------------------------------------------------------- C restatement
_LIB = None


def build_c(force: bool = False) -> str:
    """Compile rf_oracle.c with gcc into oracle/_build/ (git-ignored)."""
    out_dir = os.path.join(_HERE, "_build")
    so = os.path.join(out_dir, "librf_oracle.so")
    src = os.path.join(_HERE, "rf_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        os.makedirs(out_dir, exist_ok=True)
        # -ffp-contract=off: the accumulate must stay a plain add, as numpy's is.
        subprocess.check_call(
            ["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-shared", "-fPIC", "-o", so, src, "-lm"]
        )
    return so


def _lib():
    global _LIB
    if _LIB is None:
        lib = ctypes.CDLL(build_c())
        p = ctypes.c_void_p
        lib.rf_oracle_predict_proba.argtypes = [ctypes.c_int64, p, p, p, p, p, p, p, ctypes.c_int64,
                                                ctypes.c_int64, p, ctypes.c_int64, p]
        lib.rf_oracle_predict_proba.restype = ctypes.c_int
        _LIB = lib
    return _LIB


def predict_proba_c(f: RawForest, X) -> np.ndarray:
    """Same arithmetic through the plain-C loops in rf_oracle.c (fast enough for 284,807 rows)."""
    X32 = validate_X(X, f.n_features)
    out = np.empty((X32.shape[0], f.n_classes), dtype=np.float64)
    ptr = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    arrs = [np.ascontiguousarray(a) for a in
            (f.tree_offsets, f.left, f.right, f.feature, f.threshold, f.missing_go_to_left, f.value)]
    rc = _lib().rf_oracle_predict_proba(f.n_trees, *[ptr(a) for a in arrs], f.n_features, f.n_classes,
                                        ptr(X32), X32.shape[0], ptr(out))
    if rc != 0:
        raise RuntimeError(f"rf_oracle_predict_proba rc={rc}")
    return out
