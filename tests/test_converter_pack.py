"""Converter + library re-layout, checked on the CPU against scikit-learn's outputs."""
import ctypes
import os

import numpy as np
import pytest

import packed_emulator
from conftest import bits, load_golden
from ccfd_demo_summit_b200 import _lib, converter


def _fa(raw):
    return converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                                 raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)


def test_floor_to_float32_is_largest_float32_not_above():
    rng = np.random.default_rng(0)
    a = rng.standard_normal(20000).astype(np.float32)
    b = np.nextafter(a, np.float32(np.inf))
    t = (a.astype(np.float64) + b.astype(np.float64)) / 2  # midpoints, like SK/tree/_splitter.pyx:459-461
    t = np.concatenate([t, a.astype(np.float64), [np.inf, 0.0, -0.0, 1e-46, -1e-46, 3.5e38]])
    f = converter.floor_to_float32(t)
    assert f.dtype == np.float32
    assert np.all(f.astype(np.float64) <= t)
    up = np.nextafter(f, np.float32(np.inf))
    finite = np.isfinite(t)
    assert np.all(up[finite].astype(np.float64) > t[finite])
    assert f[len(a) * 2] == np.inf


def test_floored_threshold_is_equivalent_for_every_float32_neighbour():
    raw, X, P, _ = load_golden("small_hard")
    t = raw.threshold[raw.left != -1]
    f = converter.floor_to_float32(t)
    for x in (f, np.nextafter(f, np.float32(np.inf)), np.nextafter(f, np.float32(-np.inf))):
        assert np.array_equal(x.astype(np.float64) <= t, x <= f)


def test_packed_walk_matches_sklearn_bitwise(golden):
    name, raw, X, P, _ = golden
    fa = _fa(raw)
    blob, recs, top = packed_emulator.pack(fa)
    X32 = np.ascontiguousarray(X, dtype=np.float32)
    got = packed_emulator.predict_proba(blob, recs, fa.n_features, X32)
    assert np.array_equal(bits(got), bits(P)), name
    got = packed_emulator.predict_proba(blob, recs, fa.n_features, X32, top=top)  # top levels from the 7-node table
    assert np.array_equal(bits(got), bits(P)), name


def test_pack_layout_invariants():
    raw, *_ = load_golden("bestfirst")  # best-first builder: source nodes are not in pre-order
    fa = _fa(raw)
    blob, recs, _ = packed_emulator.pack(fa)
    assert blob.nbytes % 16 == 0
    for t, (off16, nbytes, values_off, depth, last_pos, last_vidx, *_rest) in enumerate(recs):
        assert nbytes % 16 == 0 and values_off % 16 == 0
        n_src = int(raw.tree_offsets[t + 1] - raw.tree_offsets[t])
        n_leaves = (int(nbytes) - int(values_off)) // 16
        assert n_leaves == (n_src + 1) // 2
        # slots of the deepest level (leaves only) are not stored
        assert int(values_off) == (int(last_pos) * 8 + 15) // 16 * 16
        nodes = blob[int(off16) * 16: int(off16) * 16 + int(last_pos) * 8].view(np.uint32).reshape(-1, 2)
        fmt = packed_emulator.fmt_of(recs)
        assert not fmt.wide  # trees this small use the compact encoding: byte offsets, feature << 8, mgl bit 0
        internal = fmt.feature(nodes[:, 1]) != fa.n_features
        child = fmt.child(nodes[internal, 1])
        # siblings adjacent, children after parents, every slot referenced exactly once
        assert np.array_equal(np.sort(np.concatenate([child, child + 1])), np.arange(1, n_src))
        assert np.all(child > np.nonzero(internal)[0])
        assert np.array_equal(fmt.child(nodes[~internal, 1]), np.nonzero(~internal)[0])  # leaves point at themselves
        assert n_src - int(last_pos) == n_leaves - int(last_vidx)  # deepest level: consecutive value indices to the end


def test_from_sklearn_reads_a_fitted_forest():
    from sklearn.ensemble import RandomForestClassifier

    from ccfd_demo_summit_b200 import synth
    from oracle import rf_oracle

    X, y = synth.make_dataset(1500, seed=3)
    rf = RandomForestClassifier(n_estimators=5, max_depth=5, random_state=0).fit(X, y)
    fa = converter.from_sklearn(rf)
    raw = rf_oracle.extract_raw(rf)
    assert fa.n_trees == 5 and fa.n_features == 30 and fa.n_classes == 2
    assert np.array_equal(fa.tree_offsets, raw.tree_offsets)
    blob, recs, top = packed_emulator.pack(fa)
    rf.n_jobs = 1
    assert np.array_equal(bits(packed_emulator.predict_proba(blob, recs, 30, X[:500], top)), bits(rf.predict_proba(X[:500])))


def test_big_tree_layout_dedups_leaf_values_and_keeps_every_slot():
    from sklearn.ensemble import RandomForestClassifier

    from ccfd_demo_summit_b200 import synth

    X, y = synth.make_dataset(60000, seed=1, labels="hard")
    rf = RandomForestClassifier(n_estimators=3, max_depth=None, random_state=0, n_jobs=-1).fit(X, y)
    assert max(e.tree_.node_count for e in rf.estimators_) > 6144  # beyond any shared-memory stage
    fa = converter.from_sklearn(rf)
    blob, recs, top = packed_emulator.pack(fa)
    for t, (off16, nbytes, values_off, depth, last_pos, last_vidx, *_r) in enumerate(recs):
        n_src = rf.estimators_[t].tree_.node_count
        assert last_pos == n_src and values_off == (n_src * 8 + 15) // 16 * 16
        assert (nbytes - values_off) // 16 <= 4  # pure leaves: (1,0) and (0,1) only
    rf.n_jobs = 1
    got = packed_emulator.predict_proba(blob, recs, 30, X[:3000], top)
    assert np.array_equal(bits(got), bits(rf.predict_proba(X[:3000])))


@pytest.mark.parametrize("breakage", ["cycle", "feature", "one_child", "nan_thr", "offsets", "classes", "child_range"])
def test_malformed_forests_are_rejected_with_valueerror(breakage):
    raw, *_ = load_golden("stumps")
    fa = _fa(raw)
    if breakage == "cycle":
        fa.left[0] = 0
    elif breakage == "feature":
        fa.feature[0] = 30
    elif breakage == "one_child":
        fa.left[0] = -1
    elif breakage == "nan_thr":
        fa.threshold[0] = np.nan
    elif breakage == "offsets":
        fa.tree_offsets[1] = 0
    elif breakage == "classes":
        fa.n_classes = 3
    elif breakage == "child_range":
        fa.right[0] = 7
    expect = _lib.CcfdError if breakage == "classes" else ValueError
    with pytest.raises(expect):
        packed_emulator.pack(fa)


def test_library_exports_every_declared_symbol():
    import os
    import re

    header = open(os.path.join(os.path.dirname(_lib.PKG), "include", "ccfd_rf.h")).read()
    declared = set(re.findall(r"CCFD_API\s+[\w \*]+?\b(ccfd_\w+)\s*\(", header))
    bound = {name for name, _, _ in _lib.SYMBOLS}
    assert declared == bound, declared ^ bound
    L = _lib.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert isinstance(L.ccfd_last_error(), bytes)
    # constants mirrored in the binding
    m = re.search(r"#define CCFD_NOALLOC_MAX_TREE_ROWS \(1 << (\d+)\)", header)
    assert m and _lib.NOALLOC_MAX_TREE_ROWS == 1 << int(m.group(1))


def test_host_allocation_failure_comes_back_as_an_error_code():
    """include/ccfd_rf.h promises "never throws": a std::bad_alloc inside the packer must come back as
    CCFD_ENOMEM through the C ABI instead of ending the process.  Provoked for real, in a child process:
    a one-tree forest of 6 M nodes (the packer's working vectors need several hundred MB) under an
    address-space limit that leaves far less than that."""
    import subprocess
    import sys
    import textwrap

    code = textwrap.dedent(r'''
        import ctypes, resource, sys
        import numpy as np
        sys.path.insert(0, %r)
        from ccfd_demo_summit_b200 import _lib
        L = _lib.lib()
        n_int = 3_000_000                      # a comb: internal node i has the leaf 2i+1 and the internal node 2i+2
        n = 2 * n_int + 1
        left = np.full(n, -1, np.int32); right = np.full(n, -1, np.int32)
        idx = np.arange(n_int, dtype=np.int64) * 2
        left[idx] = (idx + 1).astype(np.int32); right[idx] = (idx + 2).astype(np.int32)
        keep = [np.array([0, n], np.int64), np.zeros(n, np.int32), np.zeros(n, np.float32), left, right,
                np.zeros(n, np.uint8), np.tile(np.array([0.5, 0.5]), (n, 1))]
        d = _lib.ForestDesc(1, 30, 2, 0, *[a.ctypes.data_as(ctypes.c_void_p) for a in keep])
        nbytes = ctypes.c_int64()
        with open("/proc/self/statm") as fh:
            vm_now = int(fh.read().split()[0]) * resource.getpagesize()
        resource.setrlimit(resource.RLIMIT_AS, (vm_now + (48 << 20), vm_now + (48 << 20)))
        rc = L.ccfd_forest_pack(ctypes.byref(d), None, 0, None, None, ctypes.byref(nbytes))
        print("rc", rc, L.ccfd_last_error().decode())
        sys.exit(0 if rc == _lib.ENOMEM else 3)
    ''') % os.path.dirname(_lib.PKG)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-1500:])
    assert "out of host memory" in r.stdout
