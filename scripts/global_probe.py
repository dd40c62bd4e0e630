"""configs[2] tuning probe (development aid, GPU): fits the 500-tree unbounded forest once and times the
GLOBAL kernel on 10 M rows under different knobs (row pre-sort on/off, trees per thread), whole call
(sort included) with CUDA events.  Usage: python scripts/global_probe.py [--trees 500] [--rows 10000000]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench_configs import device_rows, time_device  # noqa: E402
from ccfd_demo_summit_b200 import converter, model, synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trees", type=int, default=500)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--variants", default="sort=1;sort=0;sort=1,one=1;sort=1,carves=100;sort=1,carves=70")
    args = ap.parse_args()
    X, y = synth.make_dataset(seed=1, labels="hard")
    t0 = time.time()
    rf = synth.fit_forest(X, y, n_estimators=args.trees, max_depth=None)
    fa = converter.from_sklearn(rf)
    forest = model.Forest(fa)
    print(json.dumps({"fit_s": round(time.time() - t0, 1), **forest.info}), flush=True)
    n = args.rows
    xs = device_rows(X, n)
    ref = None
    for var in args.variants.split(";"):
        kv = dict(p.split("=") for p in var.split(","))
        os.environ["CCFD_ROW_SORT"] = kv.get("sort", "1")
        for name, env in (("one", "CCFD_GLOBAL_ONE_PER_SM"),):
            if name in kv:
                os.environ[env] = kv[name]
            else:
                os.environ.pop(env, None)
        if "carve" in kv or "carves" in kv:  # carve-outs are set when a forest is bound to the device: a forest of its own
            os.environ["CCFD_GLOBAL_CARVEOUT"] = kv.get("carve", "40")             # arrival-order kernel
            os.environ["CCFD_GLOBAL_CARVEOUT_SORTED"] = kv.get("carves", "85")     # two-CTAs-per-SM kernel of sorted launches
            forest_v = model.Forest(fa)
        else:
            forest_v = forest
        if "nt" in kv:
            os.environ["CCFD_GLOBAL_NT"] = kv["nt"]
        else:
            os.environ.pop("CCFD_GLOBAL_NT", None)
        out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
        ms = time_device(forest_v, xs, n, out, "global", iters=3, warm=1)
        if forest_v is not forest:
            forest_v.close()
        same = True if ref is None else bool(torch.equal(out.view(torch.int64), ref.view(torch.int64)))
        if ref is None:
            ref = out
        print(json.dumps({"variant": var, "ms": round(ms, 2), "Mtx_s": round(n / ms / 1e3, 2), "same_bits_as_first": same}), flush=True)


if __name__ == "__main__":
    main()
