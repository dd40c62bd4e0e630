#!/usr/bin/env python
"""Headline benchmark: transactions/sec scored by the fraud forest's predict_proba.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      (scikit-learn on the host cores)

Workload = BASELINE.json configs[3] as written: ONE job of 100 M synthetic transactions (12 GB of
float32 rows, 1.6 GB of float64 results) scored by the demo's 100-tree, depth-10
RandomForestClassifier (fitted here on the seeded synthetic Kaggle-shaped set, SURVEY.md §8d "hard"
labels), rows sharded over the N GPUs in contiguous ranges, one forest replica per GPU, no collective
on the data path.  The job is fixed, so scaling is "strong": N=1 carries all 100 M rows (the model
node is `replicas: 1`, reference deploy/model/modelfull.json:46).  One step = one pass over the job.
`--mode weak --rows-per-gpu R` gives the fixed-rows-per-GPU run of round 1.

  value     whole-job tx/s with rows resident in HBM (one kernel launch per step per GPU, CUDA events)
  parity    4,096 of the scored rows per rank compared bit for bit with scikit-learn predict_proba(n_jobs=1)
  e2e       the same rows through the public host-buffer call (pinned host memory in, host memory
            out, copies inside the timed region); e2e_predict: the Seldon class from pageable numpy
  latency_1k configs[1]: 1,024-row micro-batches, closed loop and replayed by a producer thread at a fixed rate
  mixed_load configs[4]: single-transaction requests under back-to-back 65,536-row batches, every GPU
  config3   configs[2]: 500-tree unbounded-depth forest on 10 M rows (N=1 only)
  roofline  algorithmic bytes (136 B/tx + model bytes per launch) / kernel time vs measured HBM peak
  cpu_baseline  scikit-learn predict_proba, n_jobs = all host cores, same forest, bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "transactions/sec scored (RF predict_proba)"
UNIT = "tx/s"
ALGO_BYTES_PER_TX = 136  # 30 x fp32 in + 2 x fp64 out
TOTAL_ROWS = 100_000_000  # configs[3]
CPU_SAMPLE_ROWS = 2_848_070  # bounded CPU sample: ten configs[0] batches' worth of the job's own rows (~1 s per call on 16 cores)
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def build_forest(n_trees=100, max_depth=10):
    from ccfd_demo_summit_b200 import synth

    X, y = synth.make_dataset(seed=1, labels="hard")
    rf = synth.fit_forest(X, y, n_estimators=n_trees, max_depth=max_depth, random_state=0)
    return X, rf


def cpu_model_name():
    try:
        with open("/proc/cpuinfo") as fh:
            for line in fh:
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def time_sklearn(rf, X, budget_s=12.0, min_calls=3):
    """Best wall time of RandomForestClassifier.predict_proba through the Seldon-shaped CPU class."""
    from oracle.sklearn_seldon_model import SklearnSeldonModel

    m = SklearnSeldonModel(rf, n_jobs=os.cpu_count())
    m.predict(X[:4096], None)
    m.predict(X, None)  # warm-up at full size
    best, calls, t_end = float("inf"), 0, time.perf_counter() + budget_s
    while calls < min_calls or time.perf_counter() < t_end:
        t = time.perf_counter()
        m.predict(X, None)
        best = min(best, time.perf_counter() - t)
        calls += 1
    return best, calls


class ClockSampler:
    """SM clock and throttle reasons sampled through NVML (what nvidia-smi reads) every few
    milliseconds while the timed region runs; falls back to an nvidia-smi subprocess."""
    REASONS = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index, period_s=0.004):
        self.index, self.period_s = index, period_s
        self.sm, self.mx, self.bits, self.stop, self.th, self.nvml = [], None, 0, threading.Event(), None, None

    def _run(self):
        n = self.nvml
        while not self.stop.is_set():
            try:
                self.sm.append(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM))
                self.bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period_s)

    def __enter__(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.mx = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.nvml = pynvml
            self.th = threading.Thread(target=self._run, daemon=True)
            self.th.start()
        except Exception:  # noqa: BLE001
            self.nvml = None
        return self

    def __exit__(self, *a):
        self.stop.set()
        if self.th:
            self.th.join(timeout=2)

    def summary(self):
        if not self.sm:
            return self._smi_once()
        reasons = sorted(k for k, bit in self.REASONS.items() if self.bits & bit)
        return {"sm_mhz": float(np.median(self.sm)), "sm_max_mhz": float(self.mx), "reasons": reasons, "samples": len(self.sm),
                "source": "nvml, sampled during the timed region"}

    def _smi_once(self):
        try:
            out = subprocess.run(["nvidia-smi", f"--id={self.index}", "--query-gpu=clocks.sm,clocks.max.sm",
                                  "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=10).stdout
            sm, mx = (float(x) for x in out.strip().split(","))
            return {"sm_mhz": sm, "sm_max_mhz": mx, "reasons": [], "samples": 1, "source": "nvidia-smi after the timed region"}
        except Exception:  # noqa: BLE001
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": "unavailable"}


def job_rows(rank, n, base=None):
    """Rows [first, first + n) of the job: the 284,807-row synthetic block tiled, each row with its own
    small fp32 jitter so rows are distinct.  Generated on the device, seeded per rank."""
    import torch

    from ccfd_demo_summit_b200 import synth

    if base is None:
        base = torch.from_numpy(synth.make_features(seed=1)).cuda()
    g = torch.Generator(device="cuda").manual_seed(1234 + rank)
    xs = torch.empty((n, 30), dtype=torch.float32, device="cuda")
    CH = 1 << 21
    for r0 in range(0, n, CH):
        r1 = min(n, r0 + CH)
        idx = (torch.arange(r0, r1, device="cuda") + rank * 7919) % base.shape[0]
        blk = base[idx]
        blk += torch.randn(blk.shape, generator=g, device="cuda") * 1e-3 * blk.abs()
        xs[r0:r1] = blk
    return xs


def host_sample(n):
    """The first n rows of rank 0's shard, rebuilt on the host CPU arm's box without a GPU if need be."""
    import torch

    if torch.cuda.is_available():
        return job_rows(0, n).cpu().numpy()
    from ccfd_demo_summit_b200 import synth

    return synth.tile_rows(synth.make_features(seed=1), n)


def job_config(fa, world, mode, total_rows, n, kernel):
    """The `config` object of the JSON line: identical on both arms (ours and --impl reference)."""
    from ccfd_demo_summit_b200 import model

    nodes = int(fa.tree_offsets[-1])
    depth = int(max(_tree_depths(fa)))
    return {"workload": (f"configs[3]: one job of {total_rows} rows x 100-tree depth-10 forest (forest = configs[0]/[1]), HBM-resident, "
                         f"contiguous row shards over {world} GPU(s)") if mode == "strong" else
                        f"weak-scaling variant: {n} rows per GPU x 100-tree depth-10 forest, HBM-resident",
            "total_rows": total_rows, "rows_per_gpu": n, "n_trees": fa.n_trees, "max_depth": depth, "n_nodes": nodes,
            "model_bytes": model.packed_model_bytes(fa), "parallelism": f"rows sharded x{world}, forest replica per GPU, no collective",
            "l2": f"inputs ({n * 120 / 1e9:.1f} GB/GPU) exceed L2; no flush needed", "kernel": kernel}


def _tree_depths(fa):
    out = []
    for t in range(fa.n_trees):
        lo, hi = int(fa.tree_offsets[t]), int(fa.tree_offsets[t + 1])
        left, right = fa.left[lo:hi], fa.right[lo:hi]
        depth = np.zeros(hi - lo, dtype=np.int32)
        for i in range(hi - lo):  # sklearn stores parents before children
            if left[i] >= 0:
                depth[left[i]] = depth[right[i]] = depth[i] + 1
        out.append(int(depth.max()))
    return out


def job_shape(args, world, rank):
    from ccfd_demo_summit_b200 import sharding

    if args.mode == "strong":
        lo, hi = sharding.shard_range(args.total_rows, world, rank)
        return hi - lo, args.total_rows
    return args.rows_per_gpu, args.rows_per_gpu * world


def run_reference(args, rank, world):
    """scikit-learn's own CPU path (the arithmetic the reference image runs) on this box's cores, all of
    them, on a bounded sample of the job's rows per step."""
    if rank != 0:
        return
    X, rf = build_forest()
    from ccfd_demo_summit_b200 import converter
    from oracle.sklearn_seldon_model import SklearnSeldonModel

    n, total_rows = job_shape(args, world, 0)
    rows = host_sample(CPU_SAMPLE_ROWS)
    cores = os.cpu_count()
    m = SklearnSeldonModel(rf, n_jobs=cores)
    for _ in range(max(min(args.warmup, 2), 1)):
        m.predict(rows, None)
    steps = max(1, min(args.steps, 20))
    t0 = time.perf_counter()
    for _ in range(steps):
        m.predict(rows, None)
    dt = time.perf_counter() - t0
    value = steps * len(rows) / dt
    sample = (f"{len(rows)} of the job's 100 M rows per step ({steps} steps timed), sklearn {__import__('sklearn').__version__} "
              f"n_jobs={cores} through the Seldon-shaped class")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": args.warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": args.mode,
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": job_config(converter.from_sklearn(rf), world, args.mode, total_rows, n, args.kernel),
        "reference_note": "host CPU arm: each step scores a bounded sample (rows_per_step) of the job's rows; the rate is what is compared",
        "rows_per_step": len(rows),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample,
                         "cpu": cpu_model_name()},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="strong", choices=["strong", "weak"])
    ap.add_argument("--total-rows", type=int, default=TOTAL_ROWS)
    ap.add_argument("--rows-per-gpu", type=int, default=12_500_000, help="--mode weak only")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--no-mixed", action="store_true")
    ap.add_argument("--no-config3", action="store_true")
    ap.add_argument("--config3-trees", type=int, default=500)
    ap.add_argument("--kernel", default="auto")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    t_start, phases = time.perf_counter(), {}

    def mark(name, _last=[time.perf_counter()]):
        now = time.perf_counter()
        phases[name] = round(now - _last[0], 2)
        _last[0] = now

    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import bench_configs as extra
    from ccfd_demo_summit_b200 import converter, model, synth

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local_rank)
    if world > 1:
        # rank 0's stdout carries exactly one JSON line: NCCL prints its version banner to stdout when
        # NCCL_DEBUG=VERSION (as on the benchmark boxes), so stdout points at stderr while it initialises
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- model: rank 0 fits, everyone loads the flattened arrays
    tmp = os.path.join(tempfile.gettempdir(), f"ccfd_bench_forest_{os.environ.get('MASTER_PORT', '0')}.npz")
    X = rf = None
    if rank == 0:
        X, rf = build_forest()
        converter.from_sklearn(rf).save(tmp)
    barrier()
    fa = converter.ForestArrays.load(tmp)
    forest = model.Forest(fa, device=local_rank)
    info = forest.info
    mark("imports_fit_upload")

    # ---- rows: this rank's contiguous range of the job
    n, total_rows = job_shape(args, world, rank)
    xs = job_rows(rank, n)
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    kernel_name, presorted = forest.kernel_for(n) if args.kernel == "auto" else (args.kernel, False)

    def step():
        forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), stream, args.kernel, sync=False)

    mark("generate_rows")
    for _ in range(args.warmup):
        step()
    barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    with ClockSampler(local_rank) as clk:
        barrier()
        t0 = time.perf_counter()
        for a, b in evs:
            a.record()
            step()
            b.record()
        barrier()
        wall = time.perf_counter() - t0
    kernel_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = evs[0][0].elapsed_time(evs[-1][1])
    total_ms = max_over_ranks(total_ms)
    value = total_rows * args.steps / (total_ms * 1e-3)
    kavg_ms = float(np.mean(kernel_ms))
    clocks = clk.summary()
    checksum = float(out[:, 1].sum().item())

    # ---- parity inside the line: 4,096 of the rows just scored, per rank, against scikit-learn itself
    PAR = 4096
    p_lo = max(0, min(n - PAR, n // 2 + 12_345))
    xs_s, out_s = xs[p_lo:p_lo + PAR].contiguous(), out[p_lo:p_lo + PAR].contiguous()
    if world > 1:
        xg = [torch.empty_like(xs_s) for _ in range(world)]
        og = [torch.empty_like(out_s) for _ in range(world)]
        dist.all_gather(xg, xs_s)
        dist.all_gather(og, out_s)
        xs_s, out_s = torch.cat(xg), torch.cat(og)
    parity = None
    if rank == 0:
        rf.n_jobs = 1  # tree order 0..T-1: the bit-reproducible order (SK/ensemble/_forest.py:704-717)
        want = rf.predict_proba(xs_s.cpu().numpy())
        got = out_s.cpu().numpy()
        parity = {"rows": int(len(got)), "rows_per_rank": PAR, "bitwise_equal": bool(np.array_equal(got.view(np.uint64), want.view(np.uint64))),
                  "decisions_equal": bool(np.array_equal(got[:, 1] > 0.5, want[:, 1] > 0.5)), "max_abs_diff": float(np.abs(got - want).max()),
                  "against": f"scikit-learn {__import__('sklearn').__version__} predict_proba(n_jobs=1) on the same rows, this run"}

    mark("timed_steps_and_parity")
    # ---- roofline of the dominant (only) kernel of a step
    peak, peak_src = FALLBACK_HBM_GBS, "fallback"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peak, peak_src = float(json.load(fh)["hbm_gbs"]), "measured"
    except (OSError, KeyError, ValueError):
        pass
    algo_bytes = ALGO_BYTES_PER_TX * n + info["model_bytes"]
    achieved = algo_bytes / (kavg_ms * 1e-3) / 1e9
    # DRAM traffic and shared-memory wavefronts are NOT measured by this run (that needs ncu): they are the
    # per-row figures of the committed capture, scaled to this launch's rows, and are labelled as such.
    traffic = on_chip = traffic_src = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            tj = json.load(fh)
        if kernel_name == "stream":
            scale = n / float(tj["rows_per_launch"])
            traffic = int(tj["dram_bytes_per_launch"] * scale)
            traffic_src = f"committed ncu capture ({tj['round']}, {tj['rows_per_launch']} rows/launch) scaled by rows; not measured in this run"
            sm_hz = float(clocks["sm_mhz"]) * 1e6
            sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
            lsu, tma = tj["smem_lsu_wavefronts_per_launch"] * scale, tj["smem_tma_write_lines_per_launch"] * scale
            cycles = kavg_ms * 1e-3 * sm_hz * sms
            on_chip = {"resource": "shared-memory bank cycles (1 wavefront / clk / SM)", "lsu_read_frac": lsu / cycles,
                       "tma_write_frac": tma / cycles, "frac": (lsu + tma) / cycles,
                       "source": "wavefront counts: " + traffic_src + "; time and SM clock: this run",
                       "note": "~1.0 = banks saturated; the TMA share is an upper bound (a write may use banks a read wavefront leaves idle)"}
    except (OSError, KeyError, TypeError, ValueError):
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "kernel": "k_" + kernel_name,
                "rows_presorted": presorted, "kernel_ms": kavg_ms, "algorithmic_bytes_per_launch": algo_bytes,
                "node_visits_per_s": n * fa.n_trees * info["max_depth"] / (kavg_ms * 1e-3), "on_chip": on_chip}

    # ---- e2e: the same rows from pinned host memory through the host-buffer call
    e2e = None
    if not args.no_e2e:
        hx = torch.empty((n, 30), dtype=torch.float32, pin_memory=True)
        hx.copy_(xs)
        hp = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
        hxn, hpn = hx.numpy(), hp.numpy()
        e2e_steps = max(2, min(args.steps, 3))
        forest.predict_proba(hxn, args.kernel, out=hpn)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            forest.predict_proba(hxn, args.kernel, out=hpn)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        same = bool(np.array_equal(hpn[p_lo:p_lo + PAR].view(np.uint64), out[p_lo:p_lo + PAR].cpu().numpy().view(np.uint64)))
        e2e = {"value": total_rows * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": n * 120, "d2h_bytes_per_step": n * 16,
               "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3, "api": "Forest.predict_proba -> ccfd_predict_proba(host pointers, pinned)",
               "bytes_are": "per GPU", "parity_rows_bitwise_equal_to_device_run": same,
               "checksum_matches_device": bool(abs(float(hpn[:, 1].sum()) - checksum) <= 1e-6 * max(1.0, abs(checksum)))}
        # decisions only: 1 byte per row comes back instead of 16
        hd = torch.empty((n,), dtype=torch.uint8, pin_memory=True)
        hdn = hd.numpy()
        import ctypes

        from ccfd_demo_summit_b200 import _lib

        def decide_host():
            _lib.check(_lib.lib().ccfd_predict_decision(forest.handle, hxn.ctypes.data_as(ctypes.c_void_p), n, 30, None,
                                                        hdn.ctypes.data_as(ctypes.c_void_p), 0.5, 0, _lib.KERNELS[args.kernel], None))

        decide_host()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            decide_host()
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        e2e["decisions_only"] = {"value": total_rows * e2e_steps / dt, "unit": UNIT, "d2h_bytes_per_step": n,
                                 "equal_to_threshold_on_probabilities": bool(np.array_equal(hdn.astype(bool), hpn[:, 1] > 0.5))}
        del hx, hp, hd
    del xs, out
    torch.cuda.empty_cache()
    mark("e2e")

    # ---- configs[4]: requests under bulk load, on every rank's GPU at the same time
    mixed = None
    if not args.no_mixed:
        Xb = synth.make_features(seed=1)
        mine = extra.run_mixed_load(forest, Xb, local_rank, clients=5, requests=2000, sync=barrier)
        if world > 1:
            every = [None] * world
            dist.all_gather_object(every, mine)
            mixed = {"n_gpus": world, "clients_per_gpu": 5, **extra.aggregate_mixed(every), "note": extra.MIXED_NOTE} if rank == 0 else None
        else:
            mixed = {"n_gpus": 1, "clients_per_gpu": 5, **mine, "note": extra.MIXED_NOTE}

    mark("mixed_load")
    latency = e2e_predict = config3 = cpu = None
    if rank == 0:
        # ---- configs[1]: 1,024-row micro-batches, closed loop and replayed at a fixed rate
        if not args.no_latency:
            latency = extra.run_latency_1k(forest, X)
            # ---- the deployment's own e2e: the configs[0]-shaped batch through the Seldon class, pageable numpy
            e2e_predict = extra.run_e2e_predict(fa, X, device=local_rank)
        mark("latency_and_e2e_predict")
        # ---- configs[2]: 500 unbounded trees x 10 M rows (one GPU)
        if world == 1 and not args.no_config3:
            config3 = extra.run_config3(args.config3_trees, 10_000_000, cpu_budget_s=0.0)
        mark("config3")
        # ---- CPU baseline: scikit-learn on all host cores, bounded sample of the job's rows, N=1 only
        if world == 1 and not args.no_cpu_baseline:
            rows = host_sample(CPU_SAMPLE_ROWS)
            best, calls = time_sklearn(rf, rows, budget_s=6.0, min_calls=3)
            rf.n_jobs = 1  # the bit-exact order; also the faster CPU path for tiny requests (no joblib dispatch)
            t1 = time.perf_counter()
            rf.predict_proba(X)
            one_core = len(X) / (time.perf_counter() - t1)
            cpu = {"value": len(rows) / best, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "cpu": cpu_model_name(),
                   "value_n_jobs_1": one_core,
                   "sample": f"{len(rows)} of the job's rows, best of {calls} calls, scikit-learn {__import__('sklearn').__version__} "
                             f"predict_proba n_jobs={os.cpu_count()} through the Seldon-shaped class"}

        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": args.mode, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": job_config(fa, world, args.mode, total_rows, n, args.kernel),
            "parity": parity, "roofline": roofline, "e2e": e2e, "e2e_predict": e2e_predict, "latency_1k": latency, "mixed_load": mixed,
            "config3": config3, "cpu_baseline": cpu, "clocks": clocks,
            "gpu_launches": args.steps, "wall_s_timed": wall, "checksum_p_fraud": checksum,
        }
        mark("cpu_baseline")
        line["phases_s"] = {**phases, "total": round(time.perf_counter() - t_start, 1)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
