"""Generate tests/golden/*.npz and g_full_hashes.json from scikit-learn itself.

Run in the build container:  python tests/golden/make_golden.py
The arithmetic's home is scikit-learn (1.9.0 here); the reference repository has
no vectors for this path, so these files are the pin: seeded data -> seeded fit
-> ``predict_proba`` with n_jobs=1 (tree order 0..T-1, SK/ensemble/_forest.py:951-962)
-> stored next to the forest's own node arrays.  Nothing here touches the CUDA
product; tests compare the oracle *and* the CUDA path against these outputs.
"""
from __future__ import annotations

import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import sklearn  # noqa: E402
from sklearn.ensemble import RandomForestClassifier  # noqa: E402

from ccfd_demo_summit_b200 import synth  # noqa: E402
from oracle import rf_oracle  # noqa: E402


def _score(rf, X):
    rf.n_jobs = 1
    return rf.predict_proba(X), rf.apply(X).astype(np.int64)


def _save(name, rf, X):
    P, leaves = _score(rf, X)
    raw = rf_oracle.extract_raw(rf)
    np.savez_compressed(
        os.path.join(HERE, f"{name}.npz"), X=X, P=P, leaves=leaves,
        versions=np.array([sklearn.__version__, np.__version__]), **raw.save("f_"),
    )
    print(f"{name}: {rf.n_estimators} trees, nodes {raw.tree_offsets[-1]}, rows {len(X)}, "
          f"fraud@0.5 {(P[:, 1] > 0.5).sum()}")
    return raw, P


def adversarial_rows(raw, base_rows):
    """Rows sitting exactly on, and one fp32 step either side of, every split threshold."""
    rows = []
    internal = np.nonzero(raw.left != -1)[0]
    rng = np.random.default_rng(5)
    pick = rng.choice(internal, size=min(400, internal.size), replace=False)
    for k, nd in enumerate(pick):
        f, t = int(raw.feature[nd]), raw.threshold[nd]
        t32 = np.float32(t)
        if np.float64(t32) > t:  # largest fp32 <= t
            t32 = np.nextafter(t32, np.float32(-np.inf))
        for v in (t32, np.nextafter(t32, np.float32(np.inf)), np.nextafter(t32, np.float32(-np.inf))):
            r = base_rows[k % len(base_rows)].copy()
            r[f] = v
            rows.append(r)
    specials = [np.nan, np.finfo(np.float32).max, -np.finfo(np.float32).max, 0.0, -0.0,
                np.float32(1e-45), np.float32(-1e-45), np.finfo(np.float32).tiny]
    for s in specials:
        for f in range(raw.n_features):
            r = base_rows[(f * 7) % len(base_rows)].copy()
            r[f] = s
            rows.append(r)
        rows.append(np.full(raw.n_features, s, dtype=np.float32))
    return np.asarray(rows, dtype=np.float32)


def main():
    Xtr, ytr = synth.make_dataset(4000, seed=1, labels="hard")
    Xte = synth.make_features(1500, seed=11)

    rf = RandomForestClassifier(n_estimators=10, max_depth=6, random_state=0).fit(Xtr, ytr)
    raw, _ = _save("small_hard", rf, Xte)
    _save("adversarial", rf, adversarial_rows(raw, Xte))

    rf = RandomForestClassifier(n_estimators=6, max_depth=None, random_state=1).fit(Xtr, ytr)
    _save("deep", rf, Xte[:1000])

    rf = RandomForestClassifier(n_estimators=8, max_leaf_nodes=24, random_state=2).fit(Xtr, ytr)
    _save("bestfirst", rf, Xte[:1000])  # best-first builder: nodes are NOT in pre-order

    rf = RandomForestClassifier(n_estimators=5, max_depth=1, random_state=3).fit(Xtr, ytr)
    _save("stumps", rf, Xte[:300])
    rf = RandomForestClassifier(n_estimators=4, min_impurity_decrease=1e9, random_state=4).fit(Xtr, ytr)
    _save("single_node", rf, Xte[:300])

    # forest fitted WITH missing values, scored with missing values
    rng = np.random.default_rng(9)
    Xn = Xtr.copy()
    Xn[rng.uniform(size=Xn.shape) < 0.05] = np.nan
    rf = RandomForestClassifier(n_estimators=8, max_depth=7, random_state=5).fit(Xn, ytr)
    Xq = Xte[:1000].copy()
    Xq[rng.uniform(size=Xq.shape) < 0.08] = np.nan
    _save("nan_trained", rf, Xq)

    Xe, ye = synth.make_dataset(6000, seed=0, labels="easy")
    rf = RandomForestClassifier(n_estimators=12, max_depth=10, random_state=0).fit(Xe, ye)
    _save("small_easy", rf, synth.make_features(1000, seed=12))

    # Full-size goldens: too big to commit -> commit their digests.
    hashes = {"sklearn": sklearn.__version__, "numpy": np.__version__}
    for tag, seed, labels in (("G1_hard", 1, "hard"), ("G2_easy", 0, "easy")):
        X, y = synth.make_dataset(seed=seed, labels=labels)
        rf = synth.fit_forest(X, y, n_estimators=100, max_depth=10, random_state=0)
        rf.n_jobs = 1
        P = rf.predict_proba(X)
        nodes = [int(e.tree_.node_count) for e in rf.estimators_]
        hashes[tag] = {
            "rows": int(len(X)), "X_sha256": hashlib.sha256(X.tobytes()).hexdigest(),
            "P_sha256": hashlib.sha256(np.ascontiguousarray(P).tobytes()).hexdigest(),
            "nodes_min_mean_max": [min(nodes), float(np.mean(nodes)), max(nodes)],
            "fraud_at_0.5": int((P[:, 1] > 0.5).sum()),
            "P_head": P[:4].tolist(),
        }
        print(tag, hashes[tag]["P_sha256"][:16], hashes[tag]["nodes_min_mean_max"])
    with open(os.path.join(HERE, "g_full_hashes.json"), "w") as fh:
        json.dump(hashes, fh, indent=1)


G3_TREES, G3_ROWS, G3_SEED = 500, 1_000_000, 33


def g3_forest_and_rows():
    """SURVEY.md §8c G3 = BASELINE configs[2]'s forest: 500 trees of unbounded depth fitted on the
    284,807 hard-label rows, and a 1 M-row sample to score.  Shared with the GPU test, which refits
    the forest on the box (the fit is reproducible: one seed per tree) and compares digests."""
    X, y = synth.make_dataset(seed=1, labels="hard")
    rf = synth.fit_forest(X, y, n_estimators=G3_TREES, max_depth=None, random_state=0)
    return rf, synth.make_features(G3_ROWS, seed=G3_SEED)


def main_g3():
    """Too big to commit (107 MB packed, 16 MB of outputs): digests, the first 4,096 probability pairs
    and the per-tree node statistics go into g_full_hashes.json / g3_head.npy."""
    import time

    t0 = time.time()
    rf, Xq = g3_forest_and_rows()
    t1 = time.time()
    rf.n_jobs = 1  # tree order 0..T-1: the only bit-reproducible order
    P = rf.predict_proba(Xq)
    nodes = [int(e.tree_.node_count) for e in rf.estimators_]
    depth = [int(e.tree_.max_depth) for e in rf.estimators_]
    with open(os.path.join(HERE, "g_full_hashes.json")) as fh:
        hashes = json.load(fh)
    hashes["G3_unbounded"] = {
        "trees": G3_TREES, "rows": G3_ROWS, "rows_seed": G3_SEED,
        "X_sha256": hashlib.sha256(Xq.tobytes()).hexdigest(),
        "P_sha256": hashlib.sha256(np.ascontiguousarray(P).tobytes()).hexdigest(),
        "P_sha256_per_100k": [hashlib.sha256(np.ascontiguousarray(P[i:i + 100_000]).tobytes()).hexdigest()
                              for i in range(0, G3_ROWS, 100_000)],
        "nodes_min_mean_max": [min(nodes), float(np.mean(nodes)), max(nodes)], "n_nodes": int(sum(nodes)),
        "depth_min_max": [min(depth), max(depth)],
        "fraud_gt_0.5": int((P[:, 1] > 0.5).sum()), "fraud_ge_0.5": int((P[:, 1] >= 0.5).sum()),
        "sklearn": sklearn.__version__, "numpy": np.__version__,
    }
    np.save(os.path.join(HERE, "g3_head.npy"), P[:4096])
    with open(os.path.join(HERE, "g_full_hashes.json"), "w") as fh:
        json.dump(hashes, fh, indent=1)
    print(f"G3: fit {t1 - t0:.0f}s, predict {time.time() - t1:.0f}s, nodes {sum(nodes)}, depth {min(depth)}-{max(depth)}, "
          f"P {hashes['G3_unbounded']['P_sha256'][:16]}")


if __name__ == "__main__":
    main_g3() if "--g3" in sys.argv else main()
