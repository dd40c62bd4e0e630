"""Flatten a fitted sklearn RandomForestClassifier into plain structure-of-arrays.

Reads exactly what ``predict_proba`` reads from each ``estimators_[t].tree_``
(SK/tree/_tree.pyx:656-765: children_left/right, feature, threshold,
missing_go_to_left, value) and nothing else.  The one transformation is on the
thresholds: scikit-learn compares a float32 feature with a float64 threshold
(SK/tree/_tree.pyx:989).  For float32 ``x`` and float64 ``t``,
``x <= t``  <=>  ``x <= fl_down(t)`` where ``fl_down(t)`` is the largest float32
not above ``t`` (if x <= t then x, being a float32, cannot exceed the largest
float32 below-or-at t; the converse is immediate).  Rounding to *nearest* would
be wrong for about half of all fitted thresholds; rounding down is exact for
every input and halves the node size.  The breadth-first re-layout and bit
packing happen inside libccfd_rf (csrc/forest.cu).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class ForestArrays:
    tree_offsets: np.ndarray  # int64 [T+1]
    feature: np.ndarray  # int32 [total] (0 at leaves; ignored there)
    threshold: np.ndarray  # float32 [total] = fl_down(sklearn threshold)
    left: np.ndarray  # int32 [total], -1 at leaves
    right: np.ndarray  # int32 [total], -1 at leaves
    missing_go_to_left: np.ndarray  # uint8 [total]
    value: np.ndarray  # float64 [total, 2] class fractions
    n_features: int
    n_classes: int

    @property
    def n_trees(self) -> int:
        return len(self.tree_offsets) - 1

    @property
    def n_nodes(self) -> int:
        return int(self.tree_offsets[-1])

    def save(self, path: str) -> None:
        np.savez(path, tree_offsets=self.tree_offsets, feature=self.feature, threshold=self.threshold,
                 left=self.left, right=self.right, mgl=self.missing_go_to_left, value=self.value,
                 shape=np.array([self.n_features, self.n_classes], dtype=np.int64))

    @staticmethod
    def load(path: str) -> "ForestArrays":
        z = np.load(path)
        nf, nc = (int(v) for v in z["shape"])
        return ForestArrays(z["tree_offsets"], z["feature"], z["threshold"], z["left"], z["right"], z["mgl"],
                            z["value"], nf, nc)


def floor_to_float32(t) -> np.ndarray:
    """Largest float32 <= t, elementwise (``+inf`` stays ``+inf``)."""
    t = np.asarray(t, dtype=np.float64)
    with np.errstate(over="ignore"):
        t32 = t.astype(np.float32)  # round to nearest even
    too_big = t32.astype(np.float64) > t
    t32[too_big] = np.nextafter(t32[too_big], np.float32(-np.inf))
    return t32


def from_arrays(tree_offsets, left, right, feature, threshold64, missing_go_to_left, value, n_features,
                n_classes) -> ForestArrays:
    """Build from sklearn-shaped per-node arrays (tree-local child indices, fp64 thresholds)."""
    left = np.asarray(left)
    leaf = left < 0
    value = np.ascontiguousarray(np.asarray(value, dtype=np.float64).reshape(len(left), -1)[:, :n_classes])
    if value.shape[1] != n_classes:
        raise ValueError(f"value has {value.shape[1]} classes, expected {n_classes}")
    # sklearn >= 1.3 stores class *fractions* (SK/tree/_criterion.pyx:473-486); older versions
    # stored counts and normalised at predict time.  Read what is there, refuse what is not.
    s = value[leaf].sum(axis=1)
    if not np.allclose(s, 1.0, rtol=0, atol=1e-9):
        raise ValueError("tree_.value rows at leaves do not sum to 1: this converter expects class fractions "
                         "(scikit-learn >= 1.3)")
    feat = np.where(leaf, 0, np.asarray(feature)).astype(np.int32)
    thr = floor_to_float32(np.where(leaf, 0.0, np.asarray(threshold64, dtype=np.float64)))
    return ForestArrays(
        np.ascontiguousarray(tree_offsets, dtype=np.int64), np.ascontiguousarray(feat), np.ascontiguousarray(thr),
        np.ascontiguousarray(left, dtype=np.int32), np.ascontiguousarray(right, dtype=np.int32),
        np.ascontiguousarray(missing_go_to_left, dtype=np.uint8), value, int(n_features), int(n_classes),
    )


def from_sklearn(rf) -> ForestArrays:
    """RandomForestClassifier -> ForestArrays."""
    if not hasattr(rf, "estimators_"):
        raise ValueError("forest is not fitted")
    if getattr(rf, "n_outputs_", 1) != 1:
        raise ValueError("multi-output forests are not supported")
    n_classes = int(rf.n_classes_)
    offs, L, R, Fe, Th, M, V = [0], [], [], [], [], [], []
    for est in rf.estimators_:
        t = est.tree_
        L.append(t.children_left)
        R.append(t.children_right)
        Fe.append(t.feature)
        Th.append(t.threshold)
        M.append(t.missing_go_to_left)
        V.append(t.value[:, 0, :n_classes])
        offs.append(offs[-1] + int(t.node_count))
    return from_arrays(np.asarray(offs), np.concatenate(L), np.concatenate(R), np.concatenate(Fe),
                       np.concatenate(Th), np.concatenate(M), np.concatenate(V), rf.n_features_in_, n_classes)
