"""Secondary workloads of BASELINE.json (the headline is bench.py):

  config3  500-tree unbounded-depth forest on 10 M rows, one GPU, vs the memory roofline
  config5  mixed streaming load: concurrent single-transaction requests + back-to-back 65,536-row
           batches on the same GPU; under `python -m torch.distributed.run --nproc-per-node N` every
           GPU runs it at the same time and rank 0 prints the aggregate (configs[4] names 8 GPUs)
  easy     the 100-tree forest fitted on the Kaggle-like separable labels (tens of nodes per tree)

Each prints one JSON line.  CPU comparators are scikit-learn on all host cores, bounded samples.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ccfd_demo_summit_b200 import converter, model, synth  # noqa: E402

PEAK = 6532.5


def time_device(forest, xs, n, out, kernel="auto", iters=5, warm=3):
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
    return float(np.median([a.elapsed_time(b) for a, b in evs]))


def device_rows(X, n, seed=5):
    base = torch.from_numpy(X).cuda()
    g = torch.Generator(device="cuda").manual_seed(seed)
    xs = torch.empty((n, 30), dtype=torch.float32, device="cuda")
    CH = 1 << 21
    for r0 in range(0, n, CH):
        r1 = min(n, r0 + CH)
        blk = base[torch.arange(r0, r1, device="cuda") % base.shape[0]]
        blk += torch.randn(blk.shape, generator=g, device="cuda") * 1e-3 * blk.abs()
        xs[r0:r1] = blk
    return xs


def sklearn_tx_s(rf, X, budget_s=10.0):
    rf.n_jobs = os.cpu_count()
    rf.predict_proba(X[:2000])
    best, t_end, calls = float("inf"), time.perf_counter() + budget_s, 0
    while calls < 2 or time.perf_counter() < t_end:
        t = time.perf_counter()
        rf.predict_proba(X)
        best = min(best, time.perf_counter() - t)
        calls += 1
    return len(X) / best, calls


def config3(args):
    X, y = synth.make_dataset(seed=1, labels="hard")
    t0 = time.time()
    rf = synth.fit_forest(X, y, n_estimators=args.trees, max_depth=None)
    fit_s = time.time() - t0
    forest = model.Forest(converter.from_sklearn(rf))
    n = args.rows
    xs = device_rows(X, n)
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    ms = time_device(forest, xs, n, out)
    # parity spot check against scikit-learn (n_jobs=1 order)
    sl = slice(123_456, 123_456 + 4096)
    rf.n_jobs = 1
    ok = bool(np.array_equal(out[sl].cpu().numpy().view(np.uint64), rf.predict_proba(xs[sl].cpu().numpy()).view(np.uint64)))
    depths = [e.tree_.max_depth for e in rf.estimators_]
    cpu, calls = sklearn_tx_s(rf, X[:50_000])
    gbs = (136.0 * n + forest.info["model_bytes"]) / (ms * 1e-3) / 1e9
    print(json.dumps({"config": "configs[2]: unbounded-depth forest, 1 GPU", "n_trees": args.trees, "rows": n, "fit_s": round(fit_s, 1),
                      "n_nodes": forest.info["n_nodes"], "model_bytes": forest.info["model_bytes"], "depth_min_max": [min(depths), max(depths)],
                      "kernel": "k_global" if not forest.info["stream_kernel_ok"] else "k_stream", "ms": ms, "tx_s": n / ms * 1e3,
                      "tree_rows_s": n * args.trees / ms * 1e3, "algorithmic_GBs": gbs, "frac_of_measured_hbm": gbs / PEAK,
                      "bitwise_equal_sklearn_sample": ok,
                      "cpu_sklearn_tx_s": cpu, "cpu_cores": os.cpu_count(), "cpu_sample": f"50,000 rows, best of {calls}"}), flush=True)


def easy(args):
    X, y = synth.make_dataset(seed=0, labels="easy")
    rf = synth.fit_forest(X, y)
    forest = model.Forest(converter.from_sklearn(rf))
    n = args.rows
    xs = device_rows(X, n)
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    ms = time_device(forest, xs, n, out)
    cpu, calls = sklearn_tx_s(rf, X)
    gbs = (136.0 * n + forest.info["model_bytes"]) / (ms * 1e-3) / 1e9
    print(json.dumps({"config": "easy-label (Kaggle-like) 100-tree forest, 1 GPU", "rows": n, "n_nodes": forest.info["n_nodes"],
                      "model_bytes": forest.info["model_bytes"], "ms": ms, "tx_s": n / ms * 1e3, "algorithmic_GBs": gbs,
                      "frac_of_measured_hbm": gbs / PEAK, "cpu_sklearn_tx_s": cpu, "cpu_cores": os.cpu_count(),
                      "cpu_sample": f"284,807 rows, best of {calls}"}), flush=True)


def config5(args):
    # Under torch.distributed.run (one process per GPU) every rank runs the same mixed load on its own
    # GPU at the same time; rank 0 prints the aggregate.  gloo is only used to line the phases up and
    # to collect the per-GPU results: nothing is exchanged on the data path.
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")
    sync = (lambda: dist.barrier()) if dist else (lambda: None)
    X, y = synth.make_dataset(seed=1, labels="hard")
    rf = synth.fit_forest(X, y)
    forest = model.Forest(converter.from_sklearn(rf), device=local)
    K = args.clients
    sess = model.Session(forest, 64, K)
    nb = 65536
    xs = device_rows(X, nb)
    out = torch.empty((nb, 2), dtype=torch.float64, device="cuda")
    stop = threading.Event()
    batches = [0]

    def bulk():
        torch.cuda.set_device(local)  # the current device is per thread
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            while not stop.is_set():
                forest.predict_proba_device(xs.data_ptr(), nb, 30, out.data_ptr(), st.cuda_stream, "auto", sync=True)
                batches[0] += 1

    def client(k, lat, n_req):
        row = X[k:k + 1]
        for i in range(n_req):
            t = time.perf_counter()
            sess.x[k][:1] = row
            sess.submit(k, 1)
            sess.wait(k)
            lat.append(time.perf_counter() - t)

    def run_clients(n_req):
        lats = [[] for _ in range(K)]
        th = [threading.Thread(target=client, args=(k, lats[k], n_req)) for k in range(K)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        dt = time.perf_counter() - t0
        lat = np.sort(np.concatenate([np.array(l[n_req // 10:]) for l in lats])) * 1e6
        return {"p50_us": float(lat[len(lat) // 2]), "p99_us": float(lat[int(len(lat) * .99)]), "req_s": K * n_req / dt}

    sync()
    alone = run_clients(args.requests)
    sync()
    ms_alone = time_device(forest, xs, nb, out)
    sync()
    bt = threading.Thread(target=bulk)
    bt.start()
    time.sleep(0.2)
    sync()
    b0, t0 = batches[0], time.perf_counter()
    mixed = run_clients(args.requests)
    dt = time.perf_counter() - t0
    b1 = batches[0]
    sync()  # every GPU keeps its bulk stream busy until the slowest rank's clients are done
    stop.set()
    bt.join()
    mine = {"single_alone": alone, "single_under_bulk": mixed, "bulk_alone_tx_s": nb / ms_alone * 1e3,
            "bulk_under_requests_tx_s": (b1 - b0) * nb / dt}
    note = ("requests run on highest-priority streams through per-client session slots (CUDA graphs); "
            "bulk batches on a default-priority stream, host-synchronous per batch")
    if world == 1:
        print(json.dumps({"config": "configs[4] on one GPU: single-transaction requests + back-to-back 65,536-row batches",
                          "clients": K, "requests_per_client": args.requests, **mine, "note": note}), flush=True)
        return
    every = [None] * world
    dist.all_gather_object(every, mine)
    if rank == 0:
        agg = lambda k: {"p50_us_median_over_gpus": float(np.median([e[k]["p50_us"] for e in every])),  # noqa: E731
                         "p99_us_max_over_gpus": float(max(e[k]["p99_us"] for e in every)),
                         "req_s_total": float(sum(e[k]["req_s"] for e in every))}
        print(json.dumps({"config": f"configs[4] on {world} GPUs: one replica, one process, {K} request clients and one bulk stream per GPU",
                          "n_gpus": world, "clients_per_gpu": K, "requests_per_client": args.requests,
                          "single_alone": agg("single_alone"), "single_under_bulk": agg("single_under_bulk"),
                          "bulk_alone_tx_s_total": float(sum(e["bulk_alone_tx_s"] for e in every)),
                          "bulk_under_requests_tx_s_total": float(sum(e["bulk_under_requests_tx_s"] for e in every)),
                          "per_gpu": every, "note": note}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["config3", "config5", "easy"])
    ap.add_argument("--trees", type=int, default=500)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--clients", type=int, default=5)
    ap.add_argument("--requests", type=int, default=3000)
    a = ap.parse_args()
    {"config3": config3, "config5": config5, "easy": easy}[a.which](a)
