#!/usr/bin/env python
"""Headline benchmark: transactions/sec scored by the fraud forest's predict_proba.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      (scikit-learn on the host cores)

Workload (BASELINE.json): the demo's 100-tree, depth-10 RandomForestClassifier (configs[0]/[1]/[3]
forest, fitted here on the seeded synthetic Kaggle-shaped set, SURVEY.md §8d "hard" labels) scoring
12.5 M synthetic transactions per GPU (8 GPUs = configs[3]'s 100 M rows, row-sharded, one forest
replica per GPU, no collective on the data path).  One step = one pass over the rank's rows.

  value     whole-job tx/s with rows resident in HBM (one kernel launch per step, CUDA events)
  e2e       the same rows through the public host-buffer call (pinned host memory in, host memory
            out, copies inside the timed region)
  latency   configs[1]: p50/p99 of 1,024-row micro-batches through the pinned-slot CUDA-graph session
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
ROWS_PER_GPU = 12_500_000
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


def run_reference(args, rank, world):
    """scikit-learn's own CPU path (the arithmetic the reference image runs) on this box's cores."""
    if rank != 0:
        return
    X, rf = build_forest()
    from oracle.sklearn_seldon_model import SklearnSeldonModel

    cores = os.cpu_count()
    m = SklearnSeldonModel(rf, n_jobs=cores)
    for _ in range(max(args.warmup, 1)):
        m.predict(X, None)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        m.predict(X, None)
    dt = time.perf_counter() - t0
    value = args.steps * len(X) / dt
    sample = f"{len(X)} rows (configs[0] batch) per step, sklearn {__import__('sklearn').__version__} n_jobs={cores}"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "100-tree depth-10 forest, host CPU, bounded sample of the same rows", "rows_per_step": len(X)},
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
    ap.add_argument("--rows-per-gpu", type=int, default=ROWS_PER_GPU)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-latency", action="store_true")
    ap.add_argument("--kernel", default="auto")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

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

    # ---- rows: the 284,807-row block tiled with a per-row jitter, generated on the device
    n = args.rows_per_gpu
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
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream

    def step():
        forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), stream, args.kernel, sync=False)

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
    value = world * n * args.steps / (total_ms * 1e-3)
    kavg_ms = float(np.mean(kernel_ms))
    clocks = clk.summary()
    checksum = float(out[:, 1].sum().item())

    # ---- roofline of the dominant (only) kernel of a step
    peak, peak_src = FALLBACK_HBM_GBS, "fallback"
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            peak, peak_src = float(json.load(fh)["hbm_gbs"]), "measured"
    except (OSError, KeyError, ValueError):
        pass
    algo_bytes = ALGO_BYTES_PER_TX * n + info["model_bytes"]
    achieved = algo_bytes / (kavg_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            tj = json.load(fh)
        if tj.get("rows_per_launch") == n:
            traffic = tj.get("dram_bytes_per_launch")
    except (OSError, ValueError):
        pass
    # The resource that actually binds this kernel: shared-memory bank cycles.  Wavefront
    # counts per launch come from the committed ncu capture, the time and the SM clock from this run.
    on_chip = None
    try:
        sm_hz = float(clocks["sm_mhz"]) * 1e6
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        if traffic is not None and sm_hz > 0:
            lsu, tma = tj["smem_lsu_wavefronts_per_launch"], tj["smem_tma_write_lines_per_launch"]
            cycles = kavg_ms * 1e-3 * sm_hz * sms
            on_chip = {"resource": "shared-memory bank cycles (1 wavefront / clk / SM)", "lsu_read_frac": lsu / cycles,
                       "tma_write_frac": tma / cycles, "frac": (lsu + tma) / cycles, "source": "profiles/traffic.json (ncu) x this run's time and clock",
                       "note": "~1.0 = banks saturated; the TMA share is an upper bound (a write may use banks a read wavefront leaves idle)"}
    except (KeyError, TypeError, ValueError, NameError):
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel": "k_stream", "kernel_ms": kavg_ms,
                "algorithmic_bytes_per_launch": algo_bytes,
                "node_visits_per_s": n * fa.n_trees * info["max_depth"] / (kavg_ms * 1e-3), "on_chip": on_chip}

    # ---- e2e: same rows from pinned host memory through the host-buffer call
    e2e = None
    if not args.no_e2e:
        hx = torch.empty((n, 30), dtype=torch.float32, pin_memory=True)
        hx.copy_(xs)
        hp = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
        hxn, hpn = hx.numpy(), hp.numpy()
        e2e_steps = max(2, min(args.steps, 5))
        forest.predict_proba(hxn, args.kernel, out=hpn)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            forest.predict_proba(hxn, args.kernel, out=hpn)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": world * n * e2e_steps / dt, "unit": UNIT, "h2d_bytes_per_step": n * 120, "d2h_bytes_per_step": n * 16,
               "steps": e2e_steps, "ms_per_step": dt / e2e_steps * 1e3, "api": "Forest.predict_proba -> ccfd_predict_proba(host pointers)",
               "checksum_matches_device": bool(abs(float(hpn[:, 1].sum()) - checksum) <= 1e-6 * max(1.0, abs(checksum)))}
        del hx, hp

    # ---- configs[1]: 1,024-row micro-batches through the pinned-slot / CUDA-graph session
    latency = None
    if rank == 0 and not args.no_latency:
        sess = model.Session(forest, 1024, 4)
        rows = base[:1024 * 16].cpu().numpy()
        lat = []
        for i in range(10300):
            r = rows[(i % 16) * 1024:(i % 16 + 1) * 1024]
            t = time.perf_counter()
            sess.x[i % 4][:] = r           # producer writes the records into the pinned slot
            sess.submit(i % 4, 1024)
            sess.wait(i % 4)               # results visible on the host
            lat.append(time.perf_counter() - t)
        lat = np.sort(np.array(lat[300:])) * 1e6
        latency = {"batch": 1024, "batches": len(lat), "p50_us": float(lat[len(lat) // 2]), "p99_us": float(lat[int(len(lat) * 0.99)]),
                   "p999_us": float(lat[int(len(lat) * 0.999)]), "path": "ccfd_session (pinned slot -> CUDA graph -> pinned result)"}
        sess.close()

    # ---- CPU baseline: scikit-learn on all host cores, bounded sample, rank 0 at N=1 only
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        best, calls = time_sklearn(rf, X)
        rf.n_jobs = 1  # the bit-exact order; also the faster CPU path for tiny requests (no joblib dispatch)
        t1 = time.perf_counter()
        rf.predict_proba(X)
        one_core = len(X) / (time.perf_counter() - t1)
        cpu = {"value": len(X) / best, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "cpu": cpu_model_name(),
               "value_n_jobs_1": one_core,
               "sample": f"{len(X)} rows (configs[0] batch), best of {calls} calls, scikit-learn {__import__('sklearn').__version__} "
                         f"predict_proba n_jobs={os.cpu_count()} through the Seldon-shaped class"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "configs[3] shard: 100-tree depth-10 forest x 12.5M rows/GPU, HBM-resident (forest = configs[0]/[1])",
                       "rows_per_gpu": n, "n_trees": fa.n_trees, "max_depth": info["max_depth"], "n_nodes": info["n_nodes"],
                       "model_bytes": info["model_bytes"], "parallelism": f"rows sharded x{world}, forest replica per GPU, no collective",
                       "l2": "inputs (1.5 GB/GPU) exceed L2; no flush needed", "kernel": args.kernel},
            "roofline": roofline, "e2e": e2e, "latency_1k": latency, "cpu_baseline": cpu, "clocks": clocks,
            "gpu_launches": args.steps, "wall_s_timed": wall, "checksum_p_fraud": checksum,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
