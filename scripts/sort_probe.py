"""Does row coherence pay?  (development aid, GPU)

Feeds the STREAM and GLOBAL kernels the SAME rows in different orders and times the kernel alone
(CUDA events): arrival order, sorted by the leaves the rows reach in trees 0..k-1, and sorted by a
cheap key made of the most used split features quantised to a few bits.  The ordering itself is
done here with torch/numpy and is NOT timed: this prices the ceiling of a pre-sort before any of
it is built into the library (scripts/bank_sim.py predicts the shared-memory wavefronts of the
same orderings).

Usage: python scripts/sort_probe.py [--rows 4000000] [--unbounded-trees 100]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth  # noqa: E402


def time_kernel(forest, xs, kernel, iters=5, warm=2):
    n = xs.shape[0]
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(warm):
        forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), st, kernel, sync=False)
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), st, kernel, sync=False)
        b.record()
    torch.cuda.synchronize()
    ms = sorted(a.elapsed_time(b) for a, b in evs)
    return ms[len(ms) // 2], out


def feature_key(fa, X, k, bits):
    """k most used split features (weighted towards the top of the trees is not available here: plain
    node counts), each quantised at the quantiles of the forest's own thresholds for that feature."""
    internal = fa.left >= 0
    used = np.bincount(fa.feature[internal], minlength=fa.n_features)
    key = np.zeros(len(X), dtype=np.int64)
    for f in np.argsort(-used)[:k]:
        thr = np.sort(fa.threshold[internal & (fa.feature == f)])
        q = thr[(np.arange(1, 2 ** bits) * len(thr)) // (2 ** bits)]
        key = key * (2 ** bits) + np.searchsorted(q, X[:, f], side="left")
    return key


def probe(tag, rf, kernel, X, leaf_keys=(1, 2), feat_keys=((6, 3), (8, 3), (11, 2))):
    fa = converter.from_sklearn(rf)
    forest = model.Forest(fa)
    res = {"forest": tag, "kernel": kernel, "rows": len(X), **{k: forest.info[k] for k in ("n_nodes", "model_bytes", "max_depth")}}
    base_ms, base_out = time_kernel(forest, torch.from_numpy(X).cuda(), kernel)
    res["arrival_order_ms"] = round(base_ms, 3)
    orders = {}
    for k in leaf_keys:
        key = np.zeros(len(X), dtype=np.int64)
        for j in range(k):
            key = key * (1 << 20) + rf.estimators_[j].apply(X)
        orders[f"leaf{k}"] = np.argsort(key, kind="stable")
    for k, b in feat_keys:
        orders[f"feat{k}x{b}"] = np.argsort(feature_key(fa, X, k, b), kind="stable")
    for name, perm in orders.items():
        xs = torch.from_numpy(X[perm]).cuda()
        ms, out = time_kernel(forest, xs, kernel)
        same = torch.equal(out.view(torch.int64), base_out[torch.from_numpy(perm).cuda()].view(torch.int64))
        res[name + "_ms"] = round(ms, 3)
        res[name + "_speedup"] = round(base_ms / ms, 3)
        assert same, "rows are independent: a permutation must not change a bit"
    forest.close()
    print(json.dumps(res), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4_000_000)
    ap.add_argument("--unbounded-trees", type=int, default=100)
    ap.add_argument("--unbounded-rows", type=int, default=2_000_000)
    args = ap.parse_args()
    rng = np.random.default_rng(5)
    Xq = synth.make_features(args.rows, seed=77)
    Xq[:, 0] = rng.permutation(Xq[:, 0])  # the generator sorts Time; arrival order here is random
    X, y = synth.make_dataset(seed=1, labels="hard")
    t0 = time.time()
    rf = synth.fit_forest(X, y, n_estimators=100, max_depth=10, random_state=0)
    print(json.dumps({"fit_s": round(time.time() - t0, 1)}), flush=True)
    probe("hard 100 x depth 10", rf, "stream", Xq)
    Xe, ye = synth.make_dataset(seed=0, labels="easy")
    rf = synth.fit_forest(Xe, ye, n_estimators=100, max_depth=10, random_state=0)
    probe("easy 100 x depth 10", rf, "stream", Xq, leaf_keys=(2,), feat_keys=((6, 3),))
    if args.unbounded_trees > 0:
        rf = synth.fit_forest(X, y, n_estimators=args.unbounded_trees, max_depth=None, random_state=0)
        probe(f"hard {args.unbounded_trees} x unbounded", rf, "global", Xq[: args.unbounded_rows])


if __name__ == "__main__":
    main()
