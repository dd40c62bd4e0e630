"""N>1 host logic on CPU: world_size-2 gloo run of the row-sharded scoring (no collective on the
data path; the gather is the optional result collection).  The scorer here is the oracle: this
test is about partitioning and reassembly, the CUDA parity tests cover the arithmetic."""
import os
import socket

import numpy as np
import pytest

from conftest import bits, load_golden
from ccfd_demo_summit_b200 import sharding


@pytest.mark.parametrize("n,world", [(0, 1), (1, 2), (31, 2), (32, 2), (33, 2), (1000, 3), (284807, 8), (100_000_000, 8)])
def test_shard_bounds_cover_and_align(n, world):
    b = sharding.shard_bounds(n, world)
    assert b[0] == 0 and b[-1] == n and np.all(np.diff(b) >= 0) and len(b) == world + 1
    assert np.all(b[:-1] % sharding.ALIGN_ROWS == 0)
    if n >= world * sharding.ALIGN_ROWS:
        sizes = np.diff(b)
        assert sizes.max() - sizes.min() <= 2 * sharding.ALIGN_ROWS


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist

    from oracle import rf_oracle

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    raw, X, P, _ = load_golden("small_hard")
    n = 1399  # ragged on purpose
    lo, hi = sharding.shard_range(n, world, rank)
    local = torch.from_numpy(rf_oracle.predict_proba_c(raw, X[lo:hi]) if hi > lo else np.zeros((0, 2)))
    full = sharding.gather_rows(local, n, world)
    ok = np.array_equal(bits(full.numpy()), bits(P[:n]))
    # timing reduction used by bench.py: max over ranks
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    q.put((rank, ok, float(t.item())))
    dist.destroy_process_group()


def test_two_rank_gloo_shards_reassemble_to_the_single_rank_result():
    import torch.multiprocessing as mp

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world = 2
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert [r[2] for r in res] == [2.0, 2.0]
