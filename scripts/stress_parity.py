"""One-off stress: many random forests x batch sizes x kernels against the C oracle (bit-exact)."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import random_forests as rfs
from ccfd_demo_summit_b200 import converter, model
from oracle import rf_oracle

def fa_of(raw):
    return converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                                 raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)

rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
n_cfg = int(sys.argv[2]) if len(sys.argv) > 2 else 200
bad = 0
t0 = time.time()
for c in range(n_cfg):
    seed = int(rng.integers(1 << 30))
    n_trees = int(rng.choice([1, 2, 3, 7, 16, 33, 100, 401, 450]))
    max_internal = int(rng.choice([1, 3, 20, 150, 700, 2500]))
    if n_trees >= 100:
        max_internal = min(max_internal, 150)
    F = int(rng.choice([1, 2, 3, 8, 30, 30, 30, 31, 64, 200]))
    skew = float(rng.choice([0.0, 0.3, 0.9, 0.99]))
    n = int(rng.choice([1, 31, 257, 513, 1024, 4097, 33000, 70001, 160000]))
    cfg = dict(seed=seed, n_trees=n_trees, max_internal=max_internal, F=F, skew=skew, n=n)
    raw = rfs.random_forest(seed, n_trees, max_internal, F, skew)
    X = rfs.random_rows(seed, raw, n)
    if rng.random() < 0.5:
        X = np.nan_to_num(X, nan=0.5)  # NaN-free batches take the other code path
    want = rf_oracle.predict_proba_c(raw, X)
    forest = model.Forest(fa_of(raw))
    # row stride and base alignment vary too: dense+aligned takes the TMA path, the rest the generic loader
    ld = F + int(rng.choice([0, 0, 1, 3, 8]))
    off = int(rng.choice([0, 0, 1, 2]))
    buf = torch.zeros(n * ld + 8, dtype=torch.float32, device="cuda")
    xs = buf[off:off + n * ld].view(n, ld)
    xs[:, :F] = torch.from_numpy(X).cuda()
    cfg.update(ld=ld, off=off)
    st = torch.cuda.current_stream().cuda_stream
    kernels = ["global"] + (["stream", "split"] if forest.info["stream_kernel_ok"] else [])
    for k in kernels:
        m = min(n, 20000) if k == "split" else n
        out = torch.full((m, 2), -1.0, dtype=torch.float64, device="cuda")
        try:
            forest.predict_proba_device(xs.data_ptr(), m, ld, out.data_ptr(), st, k)
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("ERROR", cfg, k, e, flush=True)
            continue
        got = out.cpu().numpy()
        if not np.array_equal(got.view(np.uint64), want[:m].view(np.uint64)):
            bad += 1
            print("MISMATCH", dict(seed=seed, n_trees=n_trees, max_internal=max_internal, F=F, skew=skew, n=n, kernel=k,
                                   rows_bad=int((got.view(np.uint64) != want[:m].view(np.uint64)).any(axis=1).sum())), flush=True)
    # host paths: pageable float32 / float64 matrices (pipelined staging), and the one-call session request
    host = []
    if n >= 4097:
        host.append(("host f32", lambda: forest.predict_proba(X)))
        host.append(("rows f64", lambda: forest.predict_rows(X.astype(np.float64))))
    elif forest.info["stream_kernel_ok"] or n <= 1024:
        def via_session():
            s = model.Session(forest, max(n, 32), 2)
            try:
                return s.score(int(rng.integers(2)), X.astype(np.float64) if rng.random() < 0.5 else X)
            finally:
                s.close()
        host.append(("session", via_session))
    for name, fn in host:
        try:
            got = fn()
        except Exception as e:  # noqa: BLE001
            bad += 1
            print("ERROR", cfg, name, e, flush=True)
            continue
        if not np.array_equal(got.view(np.uint64), want.view(np.uint64)):
            bad += 1
            print("MISMATCH", cfg, name, flush=True)
    forest.close()
print(f"{n_cfg} configurations, {bad} mismatches, {time.time() - t0:.0f} s")
