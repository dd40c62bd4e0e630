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


def run_config3(trees=500, rows=10_000_000, cpu_budget_s=8.0, steps=3):
    """BASELINE configs[2]: `trees` unbounded-depth trees fitted on the 284,807 hard-label rows, `rows`
    device-resident rows through ccfd_predict_proba(auto) -- the GLOBAL kernel behind the row pre-sort
    (timed with the sort) and, for comparison, in arrival order."""
    X, y = synth.make_dataset(seed=1, labels="hard")
    t0 = time.time()
    rf = synth.fit_forest(X, y, n_estimators=trees, max_depth=None)
    fit_s = time.time() - t0
    forest = model.Forest(converter.from_sklearn(rf))
    n = rows
    xs = device_rows(X, n)
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    kernel, presorted = forest.kernel_for(n)
    ms = time_device(forest, xs, n, out, iters=steps, warm=1)
    # parity against scikit-learn itself (n_jobs=1: the reproducible order) on a slice of the scored rows
    sl = slice(123_456, 123_456 + 4096)
    rf.n_jobs = 1
    want = rf.predict_proba(xs[sl].cpu().numpy())
    got = out[sl].cpu().numpy()
    parity = {"rows": 4096, "bitwise_equal": bool(np.array_equal(got.view(np.uint64), want.view(np.uint64))),
              "decisions_equal": bool(np.array_equal(got[:, 1] > 0.5, want[:, 1] > 0.5)), "against": "sklearn predict_proba(n_jobs=1)"}
    os.environ["CCFD_ROW_SORT"] = "0"
    try:
        ms_arrival = time_device(forest, xs, n, out, iters=2, warm=1)
    finally:
        del os.environ["CCFD_ROW_SORT"]
    depths = [e.tree_.max_depth for e in rf.estimators_]
    visits = float(np.mean([e.tree_.decision_path(X[:2000]).sum() / 2000 - 1 for e in rf.estimators_[:8]]))
    res = {"config": "configs[2]: unbounded-depth forest, 1 GPU, device-resident rows", "n_trees": trees, "rows": n, "fit_s": round(fit_s, 1),
           "n_nodes": forest.info["n_nodes"], "model_bytes": forest.info["model_bytes"], "depth_min_max": [min(depths), max(depths)],
           "mean_path_length": round(visits, 2), "kernel": "k_" + kernel, "rows_presorted": presorted, "ms": ms, "tx_s": n / ms * 1e3,
           "tree_rows_s": n * trees / ms * 1e3, "node_visits_s": n * trees * visits / ms * 1e3,
           "ms_arrival_order": ms_arrival, "tx_s_arrival_order": n / ms_arrival * 1e3,
           "algorithmic_GBs": (136.0 * n + forest.info["model_bytes"]) / (ms * 1e-3) / 1e9, "parity": parity}
    res["frac_of_measured_hbm"] = res["algorithmic_GBs"] / PEAK
    if cpu_budget_s > 0:
        cpu, calls = sklearn_tx_s(rf, X[:50_000], cpu_budget_s)
        res.update({"cpu_sklearn_tx_s": cpu, "cpu_cores": os.cpu_count(), "cpu_sample": f"50,000 rows, best of {calls}"})
    forest.close()
    return res


def config3(args):
    print(json.dumps(run_config3(args.trees, args.rows)), flush=True)


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


def run_mixed_load(forest, X, local=0, clients=5, requests=3000, sync=lambda: None):
    """BASELINE configs[4] on ONE GPU: `clients` threads of single-transaction requests (the KIE server's
    pool of 5, reference README.md:386-393, deploy/ccd-service.yaml:61-66) through per-client session slots
    while a bulk thread scores back-to-back 65,536-row batches on the same GPU.  `sync` lines the phases up
    across ranks when every GPU of the box runs this at the same time."""
    K = clients
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
    alone = run_clients(requests)
    sync()
    ms_alone = time_device(forest, xs, nb, out)
    sync()
    bt = threading.Thread(target=bulk)
    bt.start()
    time.sleep(0.2)
    sync()
    b0, t0 = batches[0], time.perf_counter()
    mixed = run_clients(requests)
    dt = time.perf_counter() - t0
    b1 = batches[0]
    sync()  # every GPU keeps its bulk stream busy until the slowest rank's clients are done
    stop.set()
    bt.join()
    sess.close()
    return {"single_alone": alone, "single_under_bulk": mixed, "bulk_alone_tx_s": nb / ms_alone * 1e3,
            "bulk_under_requests_tx_s": (b1 - b0) * nb / dt}


MIXED_NOTE = ("requests run on highest-priority streams through per-client session slots (CUDA graphs); "
              "bulk batches on a default-priority stream, host-synchronous per batch")


def aggregate_mixed(every):
    agg = lambda k: {"p50_us_median_over_gpus": float(np.median([e[k]["p50_us"] for e in every])),  # noqa: E731
                     "p99_us_max_over_gpus": float(max(e[k]["p99_us"] for e in every)),
                     "req_s_total": float(sum(e[k]["req_s"] for e in every))}
    return {"single_alone": agg("single_alone"), "single_under_bulk": agg("single_under_bulk"),
            "bulk_alone_tx_s_total": float(sum(e["bulk_alone_tx_s"] for e in every)),
            "bulk_under_requests_tx_s_total": float(sum(e["bulk_under_requests_tx_s"] for e in every))}


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
    mine = run_mixed_load(forest, X, local, K, args.requests, sync)
    if world == 1:
        print(json.dumps({"config": "configs[4] on one GPU: single-transaction requests + back-to-back 65,536-row batches",
                          "clients": K, "requests_per_client": args.requests, **mine, "note": MIXED_NOTE}), flush=True)
        return
    every = [None] * world
    dist.all_gather_object(every, mine)
    if rank == 0:
        print(json.dumps({"config": f"configs[4] on {world} GPUs: one replica, one process, {K} request clients and one bulk stream per GPU",
                          "n_gpus": world, "clients_per_gpu": K, "requests_per_client": args.requests, **aggregate_mixed(every),
                          "per_gpu": every, "note": MIXED_NOTE}), flush=True)
    dist.destroy_process_group()


def run_latency_1k(forest, X, batches=10_000, rate_records_s=4e6, burst=64):
    """BASELINE configs[1]: 1,024-transaction micro-batches.
    closed loop  one thread: copy 1,024 records into a pinned slot, submit, wait (what round 1 reported)
    replay       as specified: a native producer thread replays 120-byte records at
                 a fixed rate into the micro-batcher's pinned ring (csrc/batcher.cu), bursts of `burst` as a
                 consumer poll delivers them; a collector thread stamps each batch; latency = arrival of the
                 batch's last record -> results visible in host memory."""
    from ccfd_demo_summit_b200 import ingest

    sess = model.Session(forest, 1024, 4)
    rows = np.ascontiguousarray(X[:1024 * 16], dtype=np.float32)
    lat = []
    for i in range(batches + 300):
        r = rows[(i % 16) * 1024:(i % 16 + 1) * 1024]
        t = time.perf_counter()
        sess.x[i % 4][:] = r
        sess.submit(i % 4, 1024)
        sess.wait(i % 4)
        lat.append(time.perf_counter() - t)
    lat = np.sort(np.array(lat[300:])) * 1e6
    pct = lambda a: {"p50_us": float(a[len(a) // 2]), "p99_us": float(a[int(len(a) * 0.99)]), "p999_us": float(a[int(len(a) * 0.999)])}  # noqa: E731
    res = {"batch": 1024, "batches": len(lat), **pct(lat), "path": "closed loop: memcpy into pinned slot -> CUDA graph -> pinned result"}
    # replay through the native batcher
    lats = []
    mb = ingest.MicroBatcher(sess, batch=1024, on_result=lambda first, p, l: lats.append(l))
    stream = np.ascontiguousarray(np.tile(rows, (max(1, (batches + 200) // 16), 1)))
    t0 = time.perf_counter()
    errs = []

    def produce():
        try:
            mb.replay(stream, records_per_s=rate_records_s, burst=burst)
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    prod = threading.Thread(target=produce)
    prod.start()
    want, deadline = len(stream) // 1024, time.perf_counter() + 120.0
    while len(lats) < want and not errs and time.perf_counter() < deadline:
        mb.drain(50)  # the consumer side: one completed batch per call, in stream order
    prod.join()
    if errs:
        raise errs[0]
    mb.close()
    dt = time.perf_counter() - t0
    a = np.sort(np.array(lats[100:])) * 1e6
    res["replay"] = {"batches": len(a), **pct(a), "records_per_s_target": rate_records_s, "records_per_s_achieved": len(stream) / dt,
                     "burst_records": burst, "record_bytes": 120,
                     "path": "producer thread -> ccfd_batcher (pinned ring) -> CUDA graph per 1,024 records -> collector thread"}
    # the same ring unpaced: how many 120-byte records per second it takes when the producer never waits
    done = []
    mb = ingest.MicroBatcher(sess, batch=1024, on_result=lambda first, p, l: done.append(len(p)))
    t0 = time.perf_counter()
    prod = threading.Thread(target=lambda: mb.replay(stream, records_per_s=0.0, burst=1024))
    prod.start()
    deadline = time.perf_counter() + 60.0
    while sum(done) < len(stream) and time.perf_counter() < deadline:
        mb.drain(50)
    prod.join()
    dt = time.perf_counter() - t0
    mb.close()
    res["replay_unpaced"] = {"records_per_s": len(stream) / dt, "batches": len(done), "burst_records": 1024}
    sess.close()
    return res


def run_e2e_predict(fa, X, device=0, reps=8):
    """What the Seldon wrapper would see: the configs[0]-shaped 284,807 x 30 batch through
    CcfdForestModel.predict(X, names) from PAGEABLE numpy (float64 as decoded from JSON, and float32)."""
    m = model.CcfdForestModel(fa, device=device)
    res = {"rows": len(X), "api": "CcfdForestModel.predict(X, feature_names), pageable numpy in, numpy out, validation included"}
    for tag, arr in (("float32", np.ascontiguousarray(X, dtype=np.float32)), ("float64", X.astype(np.float64))):
        m.predict(arr, None)
        best = float("inf")
        for _ in range(reps):
            t = time.perf_counter()
            m.predict(arr, None)
            best = min(best, time.perf_counter() - t)
        res[tag] = {"ms": best * 1e3, "tx_s": len(X) / best}
    lat = []
    row = np.ascontiguousarray(X[:1], dtype=np.float64)
    for _ in range(2200):
        t = time.perf_counter()
        m.predict(row, None)
        lat.append(time.perf_counter() - t)
    lat = np.sort(np.array(lat[200:])) * 1e6
    res["single_transaction_us"] = {"p50": float(lat[len(lat) // 2]), "p99": float(lat[int(len(lat) * 0.99)])}
    m.close()
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("which", choices=["config3", "config5", "easy"])
    ap.add_argument("--trees", type=int, default=500)
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--clients", type=int, default=5)
    ap.add_argument("--requests", type=int, default=3000)
    a = ap.parse_args()
    {"config3": config3, "config5": config5, "easy": easy}[a.which](a)
