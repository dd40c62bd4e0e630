"""Shared-memory bank model of the STREAM kernel's gathers (development aid, CPU only).

Walks the PACKED forest image (ccfd_forest_pack) for a sample of rows exactly as the kernel does and
counts, per warp (32 consecutive rows) and per load instruction, the wavefronts the shared-memory
pipe needs: a 64-bit load is served one half-warp at a time, a 128-bit load one quarter-warp at a
time, and each such group costs max over banks of the number of DISTINCT addresses that fall into
that bank (same address = broadcast).  The model is checked against the ncu counters of the real
kernel (profiles/r1k_k_stream_ncu_summary.txt: 3.4-5.3 wavefronts per node fetch below the top,
9.87 per 16-byte leaf gather) and then used to price row orderings and layouts without a GPU.

Usage: python scripts/bank_sim.py [--labels hard|easy] [--rows 65536] [--order none|...]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from ccfd_demo_summit_b200 import converter, synth  # noqa: E402
import packed_emulator as pe  # noqa: E402


def group_wavefronts(addr, width, group):
    """addr: [n_warps, 32] byte addresses; width: bytes per lane; group: lanes served together.
    Returns [n_warps] wavefronts = sum over groups of max-per-bank distinct addresses."""
    nw = addr.shape[0]
    a = addr.reshape(nw, 32 // group, group)
    unit = a // width            # distinct addresses at access width
    bank = unit % (128 // width)  # bank group of that unit
    nb = 128 // width
    out = np.zeros(nw, dtype=np.int64)
    for g in range(32 // group):
        u, b = unit[:, g, :], bank[:, g, :]
        # count distinct units per bank: sort by (bank, unit)
        key = b.astype(np.int64) * (1 << 40) + u.astype(np.int64)
        key.sort(axis=1)
        first = np.ones_like(key, dtype=bool)
        first[:, 1:] = key[:, 1:] != key[:, :-1]
        kb = key >> 40
        mx = np.zeros(nw, dtype=np.int64)
        for bb in range(nb):
            mx = np.maximum(mx, ((kb == bb) & first).sum(axis=1))
        out += mx
    return out


def walk(blob, recs, top, X32, n_features, top_levels=3):
    """Yields per tree: list of per-level slot positions [n] (levels below the constant-bank top), leaf value index."""
    n = X32.shape[0]
    fmt = pe.fmt_of(recs)
    xs = np.concatenate([X32, np.zeros((n, 1), np.float32)], axis=1)
    rows = np.arange(n)
    for t, (off16, nbytes, values_off, depth, last_pos, last_vidx, _r0, _r1) in enumerate(recs):
        tb = blob[int(off16) * 16: int(off16) * 16 + int(nbytes)]
        nodes = tb[: int(values_off)].view(np.uint32).reshape(-1, 2)
        thr_bits, meta = nodes[:, 0], nodes[:, 1]
        pos = np.zeros(n, dtype=np.int64)
        levels = []
        last_pos = int(last_pos)
        for d in range(int(depth)):
            p = np.minimum(pos, last_pos - 1)
            if d >= top_levels:
                levels.append(pos.copy())
            m = meta[p]
            nxt = pe._step(fmt, thr_bits[p], m, xs[rows, fmt.feature(m)])
            pos = np.where(pos < last_pos, nxt, pos)
        shallow = pos < last_pos
        vidx = pos - last_pos + int(last_vidx)
        vidx[shallow] = thr_bits[pos[shallow]]
        yield t, levels, vidx, int(values_off), shallow


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--labels", default="hard")
    ap.add_argument("--rows", type=int, default=32768)
    ap.add_argument("--trees", type=int, default=100)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--order", default="none", help="none | leaf<k> (sort by leaf of trees 0..k-1) | proba | batch<B>-<key>")
    ap.add_argument("--eval-trees", type=int, default=100)
    ap.add_argument("--pool", type=int, default=0, help="sort a pool of this many rows, evaluate --rows of them (whole warps of the sorted order)")
    args = ap.parse_args()
    t0 = time.time()
    X, y = synth.make_dataset(seed=1 if args.labels == "hard" else 0, labels=args.labels)
    rf = synth.fit_forest(X, y, n_estimators=args.trees, max_depth=args.depth if args.depth > 0 else None)
    fa = converter.from_sklearn(rf)
    blob, recs, top = pe.pack(fa)
    print(f"fit+pack {time.time()-t0:.1f}s, blob {blob.nbytes} B")
    rng = np.random.default_rng(3)
    pool = max(args.pool, args.rows)
    Xs = synth.make_features(pool, seed=77)
    Xs[:, 0] = rng.permutation(Xs[:, 0])  # Time is sorted by the generator; arrival order here is random

    def sample(Xsorted):
        if len(Xsorted) == args.rows:
            return Xsorted
        w = rng.permutation(len(Xsorted) // 32)[: args.rows // 32]
        return Xsorted.reshape(-1, 32, Xsorted.shape[1])[np.sort(w)].reshape(-1, Xsorted.shape[1])

    for order in args.order.split(","):
        Xo = Xs
        if order.startswith("leaf"):
            k = int(order[4:])
            key = np.zeros(len(Xs), dtype=np.int64)
            for j in range(k):
                key = key * 4096 + rf.estimators_[j].apply(Xs)
            Xo = Xs[np.argsort(key, kind="stable")]
        elif order.startswith("feat"):
            # Morton-like key: the k most used split features, 3 bits each (quantiles)
            k = int(order[4:])
            used = np.bincount(fa.feature[fa.left >= 0], minlength=fa.n_features)
            key = np.zeros(len(Xs), dtype=np.int64)
            for f in np.argsort(-used)[:k]:
                q = np.quantile(Xs[:, f], np.linspace(0, 1, 9)[1:-1])
                key = key * 8 + np.searchsorted(q, Xs[:, f])
            Xo = Xs[np.argsort(key, kind="stable")]
        elif order == "proba":
            Xo = Xs[np.argsort(rf.predict_proba(Xs)[:, 1], kind="stable")]
        elif order.startswith("col"):
            # what the kernel can do without moving a row: inside a 1,024-row batch, slot s stays in its
            # shared-memory column s % 32 (the bank of x[f]); each column's 32 entries are ranked by key and
            # warp r takes the r-th of every column.  key: col<k> = path in trees 0..k-1, colf<k> = k features x 1 bit
            if order.startswith("colf"):
                k = int(order[4:])
                used = np.bincount(fa.feature[fa.left >= 0], minlength=fa.n_features)
                kk = np.zeros(len(Xs), dtype=np.int64)
                for f in np.argsort(-used)[:k]:
                    kk = kk * 2 + (Xs[:, f] > np.median(Xs[:, f]))
            else:
                k = int(order[3:])
                kk = np.zeros(len(Xs), dtype=np.int64)
                for j in range(k):
                    kk = kk * 4096 + rf.estimators_[j].apply(Xs)
            nb = len(Xs) // 1024
            idx = np.arange(nb * 1024).reshape(nb, 32, 32)          # [batch][entry][column]
            srt = np.argsort(kk[: nb * 1024].reshape(nb, 32, 32), axis=1, kind="stable")
            Xo = Xs[np.take_along_axis(idx, srt, axis=1).reshape(-1)]
        elif order.startswith("batch"):
            B, key = order[5:].split("-")
            B = int(B)
            if key == "proba":
                kk = rf.predict_proba(Xs)[:, 1]
            else:
                kk = rf.estimators_[0].apply(Xs)
            idx = np.arange(len(Xs)).reshape(-1, B)
            srt = np.argsort(kk.reshape(-1, B), axis=1, kind="stable")
            Xo = Xs[np.take_along_axis(idx, srt, axis=1).reshape(-1)]
        Xo = sample(Xo)
        node_wf, node_n, leaf_wf, leaf_split_wf, shallow_wf, n_t = 0.0, 0, 0.0, 0.0, 0.0, 0
        per_level = {}
        distinct_nodes = {}
        for t, levels, vidx, values_off, shallow in walk(blob, recs, top, Xo, fa.n_features):
            if t >= args.eval_trees:
                break
            for d, pos in enumerate(levels):
                wf = group_wavefronts((pos * 8).reshape(-1, 32), 8, 16)
                per_level.setdefault(d + 3, []).append(wf.mean())
                dn = np.array([len(np.unique(r)) for r in pos.reshape(-1, 32)[:256]]).mean()
                distinct_nodes.setdefault(d + 3, []).append(dn)
                node_wf += wf.mean()
                node_n += 1
            va = values_off + vidx * 16
            leaf_wf += group_wavefronts(va.reshape(-1, 32), 16, 8).mean()
            leaf_split_wf += 2 * group_wavefronts((vidx * 8).reshape(-1, 32), 8, 16).mean()
            n_t += 1
        print(f"order={order}: per warp-tree: node fetches {node_wf / n_t:.2f} wavefronts ({node_wf / max(node_n,1):.2f} each), "
              f"leaf LDS.128 {leaf_wf / n_t:.2f}, leaf as 2 x LDS.64 {leaf_split_wf / n_t:.2f}")
        print("   per level:", {d: round(float(np.mean(v)), 2) for d, v in per_level.items()})
        print("   distinct nodes per warp:", {d: round(float(np.mean(v)), 1) for d, v in distinct_nodes.items()})


if __name__ == "__main__":
    main()
