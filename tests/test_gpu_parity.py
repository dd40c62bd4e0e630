"""GPU parity: the CUDA path through the C ABI against scikit-learn's outputs (golden fixtures),
the oracle, and scikit-learn itself on the box.  Integer/bit work: float64 results must be
bit-identical (tolerance 0); north_star asks |dp| <= 1e-6 and exact decisions at 0.5."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN_DIR, bits, load_golden

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
KERNELS = ["auto", "stream", "split", "global"]


@pytest.fixture(scope="module")
def ccfd():
    import ccfd_demo_summit_b200 as pkg
    from ccfd_demo_summit_b200 import model  # raises if libccfd_rf.so is missing: no fallback

    assert torch.cuda.is_available()
    return pkg, model


def _fa(raw):
    from ccfd_demo_summit_b200 import converter

    return converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                                 raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)


def _score_device(forest, X32, kernel="auto", ld=None, misalign=False):
    n, F = X32.shape
    ld = F if ld is None else ld
    buf = torch.zeros(n * ld + 4, dtype=torch.float32, device="cuda")
    off = 1 if misalign else 0
    view = buf[off: off + n * ld].view(n, ld)
    view[:, :F] = torch.from_numpy(X32).cuda()
    out = torch.full((n, 2), -1.0, dtype=torch.float64, device="cuda")
    forest.predict_proba_device(view.data_ptr(), n, ld, out.data_ptr(), torch.cuda.current_stream().cuda_stream, kernel)
    return out.cpu().numpy()


@pytest.mark.parametrize("kernel", KERNELS)
def test_golden_bitwise_device_pointers(ccfd, golden, kernel):
    _, model = ccfd
    name, raw, X, P, _ = golden
    forest = model.Forest(_fa(raw))
    got = _score_device(forest, np.ascontiguousarray(X, dtype=np.float32), kernel)
    assert np.array_equal(bits(got), bits(P)), (name, kernel, np.abs(got - P).max())
    assert np.array_equal(got[:, 1] > 0.5, P[:, 1] > 0.5)
    forest.close()


@pytest.mark.parametrize("variant", ["streamed", "resident_tpt2", "resident_tpt4", "resident_top1", "resident_top2"])
def test_golden_bitwise_every_stream_variant(ccfd, golden, variant, monkeypatch):
    """The golden forests are small enough for the forest-resident STREAM variant, which `auto` and
    `stream` therefore take; the tree-streaming variant (what the 100-tree depth-10 forest runs) and both
    trees-per-thread settings of the resident one are pinned here explicitly, on small and on 40k-row batches."""
    _, model = ccfd
    name, raw, X, P, _ = golden
    if variant == "streamed":
        monkeypatch.setenv("CCFD_NO_RESIDENT", "1")
    elif "tpt" in variant:
        monkeypatch.setenv("CCFD_STREAM_TPT", variant[-1])
    else:
        monkeypatch.setenv("CCFD_RES_TOP", variant[-1])  # levels walked from the parameter table
    forest = model.Forest(_fa(raw))
    if forest.info["stream_kernel_ok"]:
        X32 = np.ascontiguousarray(X, dtype=np.float32)
        assert np.array_equal(bits(_score_device(forest, X32, "stream")), bits(P)), (name, variant)
        reps = 40_000 // len(X32) + 1
        big, want = np.tile(X32, (reps, 1))[:40_003], np.tile(P, (reps, 1))[:40_003]  # two rows per thread, ragged tail
        assert np.array_equal(bits(_score_device(forest, big, "stream")), bits(want)), (name, variant, "40k")
    forest.close()


@pytest.mark.parametrize("kernel", ["stream", "split", "global"])
def test_strided_and_misaligned_rows(ccfd, kernel):
    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    assert np.array_equal(bits(_score_device(forest, X, kernel, ld=37)), bits(P))
    assert np.array_equal(bits(_score_device(forest, X, kernel, misalign=True)), bits(P))
    forest.close()


@pytest.mark.parametrize("n", [1, 2, 31, 32, 33, 255, 256, 257, 511, 512, 513, 1023, 1024, 1025, 1500])
def test_ragged_batch_sizes_all_kernels(ccfd, n):
    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    for kernel in ("stream", "split", "global"):
        assert np.array_equal(bits(_score_device(forest, X[:n], kernel)), bits(P[:n])), (n, kernel)
    forest.close()


def test_host_pointer_path_and_oracle(ccfd, golden):
    from oracle import rf_oracle

    _, model = ccfd
    name, raw, X, P, _ = golden
    forest = model.Forest(_fa(raw))
    X32 = np.ascontiguousarray(X, dtype=np.float32)
    got = forest.predict_proba(X32)
    assert np.array_equal(bits(got), bits(P)), name
    assert np.array_equal(bits(got), bits(rf_oracle.predict_proba_c(raw, X32))), name
    forest.close()


def test_many_small_trees_take_the_table_free_stream_path(ccfd):
    """More trees than the kernel-parameter table holds (400): per-tree metadata comes from global
    memory instead of the constant bank.  450 trees = the 'small_easy' forest repeated."""
    from oracle import rf_oracle

    _, model = ccfd
    raw, X, _, _ = load_golden("small_easy")
    reps = 450 // raw.n_trees + 1
    sizes = np.diff(raw.tree_offsets)
    offs = np.concatenate([[0], np.cumsum(np.tile(sizes, reps))])[: 451]
    n_nodes = int(offs[-1])
    tile = lambda a: np.concatenate([a] * reps)[:n_nodes]  # noqa: E731
    big = rf_oracle.RawForest(offs.astype(np.int64), tile(raw.left), tile(raw.right), tile(raw.feature), tile(raw.threshold),
                              tile(raw.missing_go_to_left), np.concatenate([raw.value] * reps)[:n_nodes], raw.n_features, raw.n_classes)
    assert big.n_trees == 450
    rows = np.tile(X, (40, 1))  # 40k rows -> STREAM with two rows per thread
    want = rf_oracle.predict_proba_c(big, rows)
    forest = model.Forest(_fa(big))
    for kernel in ("stream", "split", "global"):
        got = _score_device(forest, rows if kernel != "split" else rows[:3000], kernel)
        assert np.array_equal(bits(got), bits(want[: len(got)])), kernel
    forest.close()


def test_unbounded_depth_forest_goes_through_the_global_kernel(ccfd):
    """configs[2] in small: trees far beyond a shared-memory stage, pure leaves, p == 0.5 ties occur."""
    from sklearn.ensemble import RandomForestClassifier

    from ccfd_demo_summit_b200 import _lib, converter, synth

    _, model = ccfd
    X, y = synth.make_dataset(60000, seed=1, labels="hard")
    rf = RandomForestClassifier(n_estimators=6, max_depth=None, random_state=0, n_jobs=-1).fit(X, y)
    rf.n_jobs = 1
    forest = model.Forest(converter.from_sklearn(rf))
    assert forest.info["stream_kernel_ok"] == 0
    Xq = synth.make_features(50000, seed=77)
    P = rf.predict_proba(Xq)
    got = _score_device(forest, Xq, "auto")
    assert np.array_equal(bits(got), bits(P))
    assert (P[:, 1] == 0.5).sum() > 0  # ties at the router's threshold exist: the operator matters
    for ge in (False, True):
        d = forest.decide(Xq, 0.5, ge=ge)
        assert np.array_equal(d, (P[:, 1] >= 0.5) if ge else (P[:, 1] > 0.5))
    got = forest.predict_proba(Xq[:300])  # host pointers, small: still GLOBAL
    assert np.array_equal(bits(got), bits(P[:300]))
    with pytest.raises(_lib.CcfdError):
        _score_device(forest, Xq[:100], "stream")
    forest.close()
    m = model.CcfdForestModel(converter.from_sklearn(rf), session_max_batch=256)  # session graphs capture GLOBAL
    assert np.array_equal(bits(m.predict(Xq[:100])), bits(P[:100]))
    assert np.array_equal(bits(m.predict(Xq[:3000])), bits(P[:3000]))
    m.close()


def test_row_presort_changes_no_bit(ccfd, monkeypatch):
    """Bulk GLOBAL launches take their rows in key order (csrc/sort.cu); rows are scored independently,
    so results land on the original rows and no bit changes.  The threshold is lowered so that small,
    ragged, strided and NaN-carrying batches go through the sort too."""
    _, model = ccfd
    monkeypatch.setenv("CCFD_ROW_SORT_MIN", "1")
    for name in ("deep", "small_hard", "nan_trained", "bestfirst"):
        raw, X, P, _ = load_golden(name)
        forest = model.Forest(_fa(raw))
        X32 = np.ascontiguousarray(X, dtype=np.float32)
        for n in (1, 33, 257, len(X32)):
            assert np.array_equal(bits(_score_device(forest, X32[:n], "global")), bits(P[:n])), (name, n)
        assert np.array_equal(bits(_score_device(forest, X32, "global", ld=41)), bits(P)), name
        assert np.array_equal(bits(_score_device(forest, X32, "global", misalign=True)), bits(P)), name
        d, p = forest.decide(X32, 0.5, ge=True, want_proba=True, kernel="global")
        assert np.array_equal(d, P[:, 1] >= 0.5) and np.array_equal(bits(p), bits(P)), name
        forest.close()
    # wide rows: the key kernel parks up to 46 features per row in shared memory and reads wider rows strided
    # (found by scripts/stress_parity.py: 64 features used to overflow the default shared-memory limit)
    import random_forests as rfs
    from oracle import rf_oracle

    for F in (46, 47, 64, 200):
        raw = rfs.random_forest(100 + F, 9, 150, F, 0.3)
        Xw = np.nan_to_num(rfs.random_rows(100 + F, raw, 3000), nan=0.25)
        forest = model.Forest(_fa(raw))
        assert np.array_equal(bits(_score_device(forest, Xw, "global")), bits(rf_oracle.predict_proba_c(raw, Xw))), F
        forest.close()


def test_fused_decision_matches_threshold_rule(ccfd):
    _, model = ccfd
    raw, X, P, _ = load_golden("deep")  # pure leaves: p is k/T, ties at the threshold really happen
    forest = model.Forest(_fa(raw))
    thr = float(np.median(P[:, 1][P[:, 1] > 0])) if (P[:, 1] > 0).any() else 0.5
    for t in (0.5, thr, 0.0, 1.0):
        for kernel in ("stream", "split", "global"):
            d_gt, p = forest.decide(X, t, ge=False, want_proba=True, kernel=kernel)
            d_ge = forest.decide(X, t, ge=True, kernel=kernel)
            assert np.array_equal(d_gt, P[:, 1] > t) and np.array_equal(d_ge, P[:, 1] >= t), (t, kernel)
            assert np.array_equal(bits(p), bits(P))
    forest.close()


def test_seldon_model_class_contract(ccfd):
    pkg, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    m = model.CcfdForestModel(_fa(raw), session_max_batch=256)
    names = pkg.synth.FEATURE_NAMES if hasattr(pkg, "synth") else None
    # wrapper passes names positionally; small request -> CUDA-graph session, large -> bulk path
    assert np.array_equal(bits(m.predict(X[:7], names)), bits(P[:7]))
    assert np.array_equal(bits(m.predict(X, names)), bits(P))
    assert np.array_equal(bits(m.predict(X[:256].astype(np.float64).tolist(), [])), bits(P[:256]))
    assert np.array_equal(bits(m.predict(np.asfortranarray(X[:100]))), bits(P[:100]))
    assert np.array_equal(bits(m.predict(X[:200:2])), bits(P[:200:2]))
    out = m.predict(X[:3])
    assert out.shape == (3, 2) and out.dtype == np.float64
    assert np.array_equal(m.decide(X), P[:, 1] > 0.5)
    keys = {g["key"] for g in m.metrics()}
    assert keys == {"proba_1", "Amount", "V17", "V10"}  # reference deploy/grafana/ModelPrediction.json:96-322
    for bad in (X[0], X[:0], X[:, :29], np.where(np.arange(30) == 3, np.inf, X[:2]),
                np.where(np.arange(30) == 3, 1e39, X[:2].astype(np.float64))):
        with pytest.raises(ValueError):
            m.predict(bad)
    m.close()


def test_bulk_requests_from_pageable_float64_and_float32_matrices(ccfd):
    """What the Seldon wrapper hands predict() for a bulk request: an ordinary numpy matrix, float64 after
    JSON decoding.  Cast, infinity check, copies and scoring are one pipelined native call
    (ccfd_predict_proba_rows); plain ccfd_predict_proba takes the same pipeline for pageable float32."""
    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    reps = 150  # 225,000 rows: four pipeline chunks, the last one ragged
    big, want = np.tile(np.ascontiguousarray(X, dtype=np.float32), (reps, 1)), np.tile(P, (reps, 1))
    m = model.CcfdForestModel(_fa(raw), session_max_batch=256)
    assert np.array_equal(bits(m.predict(big.astype(np.float64))), bits(want))
    assert np.array_equal(bits(m.predict(big)), bits(want))
    wide = np.zeros((len(big), 41), dtype=np.float64)
    wide[:, :30] = big
    assert np.array_equal(bits(m.predict(wide[:, :30])), bits(want))          # strided rows, read in place
    assert np.array_equal(bits(m.predict(np.asfortranarray(big[:70_000]))), bits(want[:70_000]))  # not row-major: numpy cast path
    assert np.array_equal(bits(m.forest.predict_proba(big)), bits(want))       # pageable float32 through ccfd_predict_proba
    assert np.array_equal(bits(m.forest.predict_rows(big[:9000].astype(np.float64), kernel="global")), bits(want[:9000]))
    for pos, val in ((3, np.inf), (150_000, -np.inf), (224_999, 1e39)):
        bad = big.astype(np.float64)
        bad[pos, 7] = val
        with pytest.raises(ValueError, match="infinity"):
            m.predict(bad)
    nan = big[:80_000].astype(np.float64)
    nan[::5, 2] = np.nan  # NaN is allowed
    from oracle import rf_oracle
    assert np.array_equal(bits(m.predict(nan)), bits(rf_oracle.predict_proba_c(raw, nan.astype(np.float32))))
    x_last, p_last = m._last
    assert x_last.dtype == np.float32 and p_last.shape == (2,)
    m.close()


def test_session_slots_and_buckets(ccfd):
    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    s = model.Session(forest, max_batch=1024, n_slots=3)
    for i, n in enumerate([1, 5, 31, 32, 33, 64, 65, 500, 1000, 1024, 3, 1024]):
        assert np.array_equal(bits(s.score(i % 3, X[:n])), bits(P[:n])), n
    # overlapped slots
    for k in range(3):
        s.x[k][:400] = X[k * 400: (k + 1) * 400]
        s.submit(k, 400)
    for k in range(3):
        s.wait(k)
        assert np.array_equal(bits(s.p[k][:400]), bits(P[k * 400: (k + 1) * 400]))
    with pytest.raises(ValueError):
        s.submit(0, 1025)
    from ccfd_demo_summit_b200 import _lib
    with pytest.raises(_lib.CcfdError):
        s.wait(0)  # nothing in flight
    s.close()
    forest.close()


def test_session_large_bucket_captures_the_stream_kernel_in_a_graph(ccfd):
    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    s = model.Session(forest, max_batch=40000, n_slots=1)  # buckets 32 .. 32768 run SPLIT, the 40000 one STREAM
    big, want = np.tile(X, (27, 1))[:40000], np.tile(P, (27, 1))[:40000]
    for n in (40000, 32769, 32768, 16385, 5000, 1):
        assert np.array_equal(bits(s.score(0, big[:n])), bits(want[:n])), n
    s.close()
    forest.close()


def test_concurrent_callers_on_their_own_streams(ccfd):
    """A forest is immutable and the device-pointer entry point re-entrant: several threads score
    different batches at once, each on its own stream (SPLIT takes its scratch per call)."""
    import threading

    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    big, want = np.tile(X, (40, 1)), np.tile(P, (40, 1))  # 60,000 rows
    results, errs = {}, []

    def worker(k):
        try:
            st = torch.cuda.Stream()
            n = [45000, 900, 33000, 64][k % 4] + k
            assert k + n <= len(big)
            xs = torch.from_numpy(big[k:k + n]).cuda()
            out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
            torch.cuda.synchronize()
            for _ in range(5):
                forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), st.cuda_stream, "auto")
            results[k] = (out.cpu().numpy(), n)
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    th = [threading.Thread(target=worker, args=(k,)) for k in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    for k, (got, n) in results.items():
        assert np.array_equal(bits(got), bits(want[k:k + n])), k
    forest.close()


def test_bad_arguments_come_back_as_error_codes_not_crashes(ccfd):
    import ctypes

    from ccfd_demo_summit_b200 import _lib

    _, model = ccfd
    raw, X, P, _ = load_golden("small_hard")
    forest = model.Forest(_fa(raw))
    L = _lib.lib()
    xs = torch.from_numpy(X[:64]).cuda()
    out = torch.empty((64, 2), dtype=torch.float64, device="cuda")
    xp, op = ctypes.c_void_p(xs.data_ptr()), ctypes.c_void_p(out.data_ptr())
    dev = _lib.DEVICE_PTRS
    assert L.ccfd_predict_proba(forest.handle, xp, 0, 30, op, dev, None) == _lib.OK          # empty batch: nothing to do
    assert L.ccfd_predict_proba(forest.handle, xp, -1, 30, op, dev, None) == _lib.EINVAL
    assert L.ccfd_predict_proba(forest.handle, xp, 64, 29, op, dev, None) == _lib.EINVAL     # stride below n_features
    assert L.ccfd_predict_proba(forest.handle, None, 64, 30, op, dev, None) == _lib.EINVAL
    assert L.ccfd_predict_proba(forest.handle, xp, 64, 30, None, dev, None) == _lib.EINVAL
    assert L.ccfd_predict_proba(None, xp, 64, 30, op, dev, None) == _lib.EINVAL
    assert L.ccfd_predict_proba(forest.handle, xp, 64, 30, op, dev | 0x8000, None) == _lib.EINVAL  # unknown flag
    assert L.ccfd_predict_proba(forest.handle, X[:64].ctypes.data_as(ctypes.c_void_p), 64, 30, op, _lib.ASYNC, None) == _lib.EINVAL
    assert L.ccfd_predict_proba(forest.handle, xp, 64, 30, ctypes.c_void_p(out.data_ptr() + 8), dev, None) == _lib.EINVAL  # misaligned out
    assert b"16-byte" in L.ccfd_last_error()
    assert L.ccfd_predict_decision(forest.handle, xp, 64, 30, op, None, 0.5, 0, dev, None) == _lib.EINVAL
    assert L.ccfd_predict_decision(forest.handle, xp, 64, 30, op, op, float("nan"), 0, dev, None) == _lib.EINVAL
    h = ctypes.c_void_p()
    assert L.ccfd_session_create(forest.handle, 0, 1, ctypes.byref(h)) == _lib.EINVAL
    assert L.ccfd_session_create(forest.handle, 64, 0, ctypes.byref(h)) == _lib.EINVAL
    d = _lib.ForestDesc()
    assert L.ccfd_forest_create(ctypes.byref(d), 0, ctypes.byref(h)) == _lib.EINVAL
    # still usable after all that
    forest.predict_proba_device(xs.data_ptr(), 64, 30, out.data_ptr(), 0, "auto")
    assert np.array_equal(bits(out.cpu().numpy()), bits(P[:64]))
    forest.close()
    with pytest.raises(ValueError):
        model.Forest(_fa(raw), device=99)


def test_nan_rows_follow_missing_go_to_left_in_large_batches(ccfd):
    _, model = ccfd
    from oracle import rf_oracle

    raw, X, P, _ = load_golden("nan_trained")
    forest = model.Forest(_fa(raw))
    rng = np.random.default_rng(3)
    big = np.tile(X, (40, 1))  # 40k rows, NaN only in some 512-row batches after shuffling a clean half
    clean = np.nan_to_num(big[: len(big) // 2], nan=0.25)
    big = np.concatenate([clean, big[len(big) // 2:]])
    rng.shuffle(big[len(big) // 2:])
    ref = rf_oracle.predict_proba_c(raw, big)
    for kernel in ("stream", "global"):
        assert np.array_equal(bits(_score_device(forest, big, kernel)), bits(ref)), kernel
    forest.close()


# ------------------------------------------------------------------ full-size forest (configs 1-4)
@pytest.fixture(scope="module")
def full_forest(ccfd):
    from ccfd_demo_summit_b200 import converter, synth

    _, model = ccfd
    X, y = synth.make_dataset(seed=1, labels="hard")
    rf = synth.fit_forest(X, y, n_estimators=100, max_depth=10, random_state=0)
    rf.n_jobs = 1
    return X, rf, model.Forest(converter.from_sklearn(rf))


def test_G1_full_batch_bitwise_vs_sklearn_on_this_box(full_forest):
    X, rf, forest = full_forest
    P = rf.predict_proba(X)  # n_jobs=1: the bit-reproducible order (SK/ensemble/_forest.py:704-717)
    with open(os.path.join(GOLDEN_DIR, "g_full_hashes.json")) as fh:
        g = json.load(fh)["G1_hard"]
    assert hashlib.sha256(X.tobytes()).hexdigest() == g["X_sha256"]
    assert hashlib.sha256(np.ascontiguousarray(P).tobytes()).hexdigest() == g["P_sha256"], "sklearn on this box differs from the committed golden"
    for kernel in ("stream", "global"):
        got = _score_device(forest, X, kernel)
        assert np.array_equal(bits(got), bits(P)), kernel
    got = forest.predict_proba(X)  # host pointers, chunked
    assert np.array_equal(bits(got), bits(P))
    assert int((got[:, 1] > 0.5).sum()) == g["fraud_at_0.5"]
    got = _score_device(forest, X[:5000], "split")
    assert np.array_equal(bits(got), bits(P[:5000]))


def test_G2_easy_label_forest_bitwise_vs_sklearn(ccfd):
    """Kaggle-like forest (0.17 % positives, separable): tens of nodes per tree, trees of very
    different depths, many trees per shared-memory stage."""
    from ccfd_demo_summit_b200 import converter, synth

    _, model = ccfd
    X, y = synth.make_dataset(seed=0, labels="easy")
    rf = synth.fit_forest(X, y, n_estimators=100, max_depth=10, random_state=0)
    rf.n_jobs = 1
    P = rf.predict_proba(X)
    with open(os.path.join(GOLDEN_DIR, "g_full_hashes.json")) as fh:
        g = json.load(fh)["G2_easy"]
    assert hashlib.sha256(np.ascontiguousarray(P).tobytes()).hexdigest() == g["P_sha256"]
    forest = model.Forest(converter.from_sklearn(rf))
    for kernel in ("stream", "global"):
        assert np.array_equal(bits(_score_device(forest, X, kernel)), bits(P)), kernel
    assert np.array_equal(bits(_score_device(forest, X[:4000], "split")), bits(P[:4000]))
    Xn = X[:50000].copy()
    Xn[::7, 3] = np.nan
    assert np.array_equal(bits(_score_device(forest, Xn, "stream")), bits(rf.predict_proba(Xn)))
    forest.close()


def test_G3_configs2_forest_500_unbounded_trees_1M_rows_vs_sklearn_digest(ccfd):
    """BASELINE configs[2] at full model size: 500 trees of unbounded depth (13.4 M
    nodes, depth 29-40, 107 MB packed, blob offsets in the millions) refitted here from the seeds, 1 M
    rows scored through `auto` (GLOBAL kernel, with and without the row pre-sort) from device pointers and
    through the chunked host-pointer path, compared with the digests of scikit-learn's own
    predict_proba(n_jobs=1) committed by tests/golden/make_golden.py --g3 (per 100k-row slice, so a
    mismatch names its slice), plus the first 4,096 probability pairs bit by bit."""
    import sys

    sys.path.insert(0, GOLDEN_DIR)
    import make_golden

    from ccfd_demo_summit_b200 import converter

    _, model = ccfd
    with open(os.path.join(GOLDEN_DIR, "g_full_hashes.json")) as fh:
        g = json.load(fh)["G3_unbounded"]
    head = np.load(os.path.join(GOLDEN_DIR, "g3_head.npy"))
    rf, Xq = make_golden.g3_forest_and_rows()
    assert hashlib.sha256(Xq.tobytes()).hexdigest() == g["X_sha256"]
    nodes = [int(e.tree_.node_count) for e in rf.estimators_]
    assert sum(nodes) == g["n_nodes"], "the refitted forest differs from the one the digest was taken from"
    forest = model.Forest(converter.from_sklearn(rf))
    del rf
    assert forest.info["stream_kernel_ok"] == 0 and forest.info["n_trees"] == g["trees"]

    def check(got, what):
        assert np.array_equal(bits(got[:4096]), bits(head)), what
        for i, want in enumerate(g["P_sha256_per_100k"]):
            sl = np.ascontiguousarray(got[i * 100_000:(i + 1) * 100_000])
            assert hashlib.sha256(sl.tobytes()).hexdigest() == want, (what, "slice", i)
        assert hashlib.sha256(np.ascontiguousarray(got).tobytes()).hexdigest() == g["P_sha256"], what
        assert int((got[:, 1] > 0.5).sum()) == g["fraud_gt_0.5"] and int((got[:, 1] >= 0.5).sum()) == g["fraud_ge_0.5"]

    check(_score_device(forest, Xq, "auto"), "device pointers, auto")
    check(forest.predict_proba(Xq), "host pointers, chunked")
    os.environ["CCFD_ROW_SORT"] = "0"
    try:
        check(_score_device(forest, Xq, "global"), "device pointers, arrival order")
    finally:
        del os.environ["CCFD_ROW_SORT"]
    d = forest.decide(Xq, 0.5, ge=True)
    assert int(d.sum()) == g["fraud_ge_0.5"]
    forest.close()


def test_session_request_overtakes_queued_bulk_batches(full_forest):
    """configs[4]: a single-transaction request (jBPM/Camel pattern, reference deploy/ccd-service.yaml:61-66)
    submitted BEHIND a queue of bulk batches completes before they do: session streams have the highest
    priority and the bulk kernel's CTAs yield between batches."""
    from ccfd_demo_summit_b200 import model

    X, rf, forest = full_forest
    n = 2_000_000
    xs = torch.from_numpy(np.tile(X, (8, 1))[:n]).cuda()
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    sess = model.Session(forest, 1024, 2)
    want = rf.predict_proba(X[:1])
    sess.score(0, X[:1])  # warm
    bulk = torch.cuda.Stream()
    done = torch.cuda.Event()
    torch.cuda.synchronize()
    with torch.cuda.stream(bulk):
        for _ in range(40):  # ~45 ms of queued bulk work
            forest.predict_proba_device(xs.data_ptr(), n, 30, out.data_ptr(), bulk.cuda_stream, "stream", sync=False)
        done.record(bulk)
    got = sess.score(0, X[:1])
    still_running = not done.query()
    torch.cuda.synchronize()
    assert np.array_equal(bits(got), bits(want))
    assert still_running, "the request waited for the whole bulk queue"
    sess.close()


def test_multi_gpu_row_sharding_in_one_process(full_forest):
    """One replica per GPU, contiguous row ranges, no exchange: the concatenation of the shards is the
    single-GPU result.  Uses every visible GPU (with one GPU it still exercises the sharded path)."""
    from ccfd_demo_summit_b200 import converter, model

    X, rf, forest = full_forest
    multi = model.MultiGpuForest(converter.from_sklearn(rf))
    assert len(multi.devices) == torch.cuda.device_count()
    rows = X[:200_003]
    want = forest.predict_proba(rows)
    got = multi.predict_proba(rows)
    assert np.array_equal(bits(got), bits(want))
    if len(multi.devices) > 1:
        two = model.MultiGpuForest(converter.from_sklearn(rf), devices=multi.devices[:2])
        assert np.array_equal(bits(two.predict_proba(rows[:77])), bits(want[:77]))  # a shard may be empty
        two.close()
    multi.close()


def test_large_batch_properties(full_forest):
    from ccfd_demo_summit_b200 import synth

    X, rf, forest = full_forest
    n = 3_000_000
    big = synth.tile_rows(X, n)
    xs = torch.from_numpy(big).cuda()
    out_s = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    out_g = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    forest.predict_proba_device(xs.data_ptr(), n, 30, out_s.data_ptr(), st, "stream")
    forest.predict_proba_device(xs.data_ptr(), n, 30, out_g.data_ptr(), st, "global")
    assert torch.equal(out_s.view(torch.int64), out_g.view(torch.int64))  # two kernels, same bits
    assert float((out_s.sum(dim=1) - 1).abs().max()) <= 1e-12
    # permutation invariance: rows are independent
    perm = torch.randperm(n, device="cuda", generator=torch.Generator(device="cuda").manual_seed(0))
    out_p = torch.empty_like(out_s)
    xp = xs[perm].contiguous()
    forest.predict_proba_device(xp.data_ptr(), n, 30, out_p.data_ptr(), st, "stream")
    assert torch.equal(out_p.view(torch.int64), out_s[perm].view(torch.int64))
    # spot-check a slice against sklearn
    sl = slice(1_234_567, 1_234_567 + 20_000)
    assert np.array_equal(bits(out_s[sl].cpu().numpy()), bits(rf.predict_proba(big[sl])))
    # decisions-only epilogue on the same rows
    dec = torch.empty(n, dtype=torch.uint8, device="cuda")
    forest.decide_device(xs.data_ptr(), n, 30, dec.data_ptr(), 0, 0.5, False, st, "stream")
    assert torch.equal(dec.bool(), out_s[:, 1] > 0.5)


@pytest.mark.gpu
def test_scoring_on_gpu1_from_a_fresh_thread_creates_no_context_on_gpu0():
    """A thread that never chose a device reports device 0; putting "the previous device" back after a
    call must not create a primary context there (every rank but 0 of a one-process-per-GPU job would
    pay ~1 s and several hundred MB on GPU 0).  Needs two GPUs; runs in its own process."""
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    code = r'''
import ctypes, threading, sys, numpy as np
sys.path.insert(0, %r); sys.path.insert(0, %r)
from conftest import load_golden
from ccfd_demo_summit_b200 import converter, model
raw, X, P, *_ = load_golden("small_hard")
fa = converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold, raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)
forest = model.Forest(fa, device=1)
got = []
t = threading.Thread(target=lambda: got.append(forest.predict_proba(X[:500])))
t.start(); t.join()
assert np.array_equal(np.ascontiguousarray(got[0]).view(np.uint64), np.ascontiguousarray(P[:500]).view(np.uint64))
cu = ctypes.CDLL("libcuda.so.1")
flags, active = ctypes.c_uint(), ctypes.c_int()
assert cu.cuInit(0) == 0 and cu.cuDevicePrimaryCtxGetState(0, ctypes.byref(flags), ctypes.byref(active)) == 0
print("active0", active.value)
''' % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert "active0 0" in r.stdout, r.stdout
