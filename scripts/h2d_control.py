"""Control for the host-fed (e2e) scaling figure: bare host<->device copies, no scoring.

bench.py's `e2e` moves 120 B in and 16 B out per transaction through pinned buffers; on 8 GPUs it scales
far below the device-resident number.  Is that the library or the box?  This script runs nothing but
the copies: every rank (one process per GPU, like bench.py) streams the same pinned buffers to its GPU
(and the result-sized buffer back) with cudaMemcpyAsync on two streams, all ranks at once, and rank 0
prints per-GPU and aggregate GB/s.  If the aggregate stops growing with N the way e2e does, the limit
is the host's memory / PCIe fabric, not the scoring path.

  python scripts/h2d_control.py                                   (one GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 scripts/h2d_control.py
"""
import argparse
import json
import os
import time

import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=12_500_000)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo")
    n = args.rows
    CH = 1 << 20  # the library's chunk: 1 Mi rows
    hx = torch.empty((n, 30), dtype=torch.float32, pin_memory=True)
    hx.fill_(1.0)
    hp = torch.empty((n, 2), dtype=torch.float64, pin_memory=True)
    dx = [torch.empty((CH, 30), dtype=torch.float32, device="cuda") for _ in range(2)]
    dp = [torch.zeros((CH, 2), dtype=torch.float64, device="cuda") for _ in range(2)]
    st = [torch.cuda.Stream() for _ in range(2)]

    def one_pass(d2h=True):
        for k, r0 in enumerate(range(0, n, CH)):
            r1, s = min(n, r0 + CH), k & 1
            with torch.cuda.stream(st[s]):
                dx[s][: r1 - r0].copy_(hx[r0:r1], non_blocking=True)
                if d2h:
                    hp[r0:r1].copy_(dp[s][: r1 - r0], non_blocking=True)
        torch.cuda.synchronize()

    res = {}
    for tag, d2h in (("h2d_only", False), ("h2d_and_d2h", True)):
        one_pass(d2h)
        if dist:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            one_pass(d2h)
        if dist:
            dist.barrier()
        dt = (time.perf_counter() - t0) / args.reps
        res[tag] = {"ms": dt * 1e3, "h2d_GBs": n * 120 / dt / 1e9, "rows_per_s": n / dt}
    every = [res]
    if dist:
        every = [None] * world
        dist.all_gather_object(every, res)
    if rank == 0:
        out = {"n_gpus": world, "rows_per_gpu": n, "cpus": os.cpu_count()}
        for tag in ("h2d_only", "h2d_and_d2h"):
            out[tag] = {"aggregate_h2d_GBs": sum(e[tag]["h2d_GBs"] for e in every), "aggregate_rows_per_s": sum(e[tag]["rows_per_s"] for e in every),
                        "per_gpu_h2d_GBs": [round(e[tag]["h2d_GBs"], 1) for e in every]}
        print(json.dumps(out), flush=True)
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
