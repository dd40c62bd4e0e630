"""Kernel micro-benchmark (development aid): times each kernel on device-resident rows with CUDA
events.  Usage: python scripts/kbench.py [--labels hard|easy] [--sizes 1024,284807,4194304]"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth  # noqa: E402


def time_kernel(forest, xs, n, out, kernel, iters=5, warm=2):
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
    return sorted(a.elapsed_time(b) for a, b in evs)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--labels", default="hard")
    ap.add_argument("--sizes", default="1024,65536,284807,4194304")
    ap.add_argument("--trees", type=int, default=100)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--kernels", default="stream,global,split")
    ap.add_argument("--no-session", action="store_true")
    args = ap.parse_args()
    t0 = time.time()
    X, y = synth.make_dataset(seed=1 if args.labels == "hard" else 0, labels=args.labels)
    rf = synth.fit_forest(X, y, n_estimators=args.trees, max_depth=args.depth if args.depth > 0 else None)
    fa = converter.from_sklearn(rf)
    forest = model.Forest(fa)
    print(json.dumps({"fit_s": round(time.time() - t0, 1), **forest.info}))
    peak = 6532.5
    for n in [int(s) for s in args.sizes.split(",")]:
        big = X[:n] if n <= len(X) else synth.tile_rows(X, n)
        xs = torch.from_numpy(np.ascontiguousarray(big)).cuda()
        out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
        for kernel in args.kernels.split(","):
            if kernel == "split" and n > 70000:
                continue
            try:
                ms = time_kernel(forest, xs, n, out, kernel)
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"n": n, "kernel": kernel, "error": str(e)}))
                continue
            med = ms[len(ms) // 2]
            gbs = (136.0 * n + forest.info["model_bytes"]) / (med * 1e-3) / 1e9
            print(json.dumps({"n": n, "kernel": kernel, "ms_med": round(med, 4), "ms_min": round(ms[0], 4),
                              "Mtx_s": round(n / med / 1e3, 1), "GBs": round(gbs, 1), "frac": round(gbs / peak, 4)}))
    if args.no_session:
        return
    # latency path
    s = model.Session(forest, 1024, 4)
    s.x[0][:] = X[:1024]
    lat = []
    for i in range(2200):
        t = time.perf_counter()
        s.submit(0, 1024)
        s.wait(0)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(np.array(lat[200:]))
    print(json.dumps({"session_1024_us": {"p50": round(float(lat[len(lat) // 2]), 1), "p99": round(float(lat[int(len(lat) * .99)]), 1)}}))
    lat = []
    for i in range(2200):
        t = time.perf_counter()
        s.submit(1, 1)
        s.wait(1)
        lat.append((time.perf_counter() - t) * 1e6)
    lat = np.sort(np.array(lat[200:]))
    print(json.dumps({"session_1_us": {"p50": round(float(lat[len(lat) // 2]), 1), "p99": round(float(lat[int(len(lat) * .99)]), 1)}}))


if __name__ == "__main__":
    main()
