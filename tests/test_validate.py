"""validate_X (the class's mirror of ForestClassifier._validate_X_predict, SK/ensemble/_forest.py:606-625)
against numpy's own cast, through both its native threaded pass and its generic branch."""
import numpy as np
import pytest

from ccfd_demo_summit_b200 import model, synth


def _np_cast(a):
    with np.errstate(over="ignore"):
        return np.ascontiguousarray(a, dtype=np.float32)


@pytest.mark.parametrize("kind", ["f32", "f64", "f64_rows_strided", "f64_padded", "fortran", "int", "list", "small"])
def test_cast_is_bit_identical_to_numpy(kind):
    X = synth.make_features(5000, seed=4).astype(np.float64) * 1.00000013
    X[7, 3] = np.nan
    X[8, 4] = 1e-45       # float32 denormal range
    X[9, 5] = 3.4028234e38  # just inside float32
    X[10, 6] = -0.0
    a = {"f32": X.astype(np.float32), "f64": X, "f64_rows_strided": X[::3], "f64_padded": np.pad(X, ((0, 0), (0, 5)))[:, :30],
         "fortran": np.asfortranarray(X), "int": (X * 10).astype(np.int64)[:, :30], "list": X[:200].tolist(), "small": X[:5]}[kind]
    got = model.validate_X(a, 30)
    want = _np_cast(np.asarray(a))
    assert got.dtype == np.float32 and got.flags.c_contiguous
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows", [3, 6000])
def test_infinity_and_overflow_raise_like_sklearn(dtype, rows):
    X = synth.make_features(rows, seed=5).astype(dtype)
    for v in (np.inf, -np.inf) + ((1e39, -1e300) if dtype == np.float64 else ()):
        bad = X.copy()
        bad[rows - 1, 29] = v
        with pytest.raises(ValueError, match="infinity"):
            model.validate_X(bad, 30)
    ok = X.copy()
    ok[0, 0] = np.nan  # NaN is allowed: trees route it
    assert np.isnan(model.validate_X(ok, 30)[0, 0])


def test_shape_errors():
    X = synth.make_features(10, seed=6)
    for bad in (X[0], X[:0], X[:, :29], np.zeros((2, 3, 30))):
        with pytest.raises(ValueError):
            model.validate_X(bad, 30)
    with pytest.raises(ValueError):
        model.validate_X([["a"] * 30], 30)


def test_streaming_store_cast_is_bit_identical_too(monkeypatch):
    """The pageable-host path narrows rows into pinned staging with non-temporal stores (the copy engine
    reads them next); the same code is reachable through ccfd_cast_rows with CCFD_CAST_STREAM set.  Odd sizes,
    misaligned destinations, strided rows, both source types, infinity detection."""
    import ctypes

    from ccfd_demo_summit_b200 import _lib

    monkeypatch.setenv("CCFD_CAST_STREAM", "1")
    rng = np.random.default_rng(4)
    L = _lib.lib()
    for n, cols, ld, off in ((1, 30, 30, 0), (7, 30, 30, 1), (1001, 30, 37, 3), (20000, 30, 30, 0), (70001, 29, 29, 2), (40000, 31, 40, 1)):
        for dt in (np.float64, np.float32):
            src = (rng.standard_normal((n, ld)) * 10.0 ** rng.integers(-40, 39, (n, ld))).astype(dt)
            src[rng.uniform(size=src.shape) < 0.01] = np.nan
            with np.errstate(over="ignore", invalid="ignore"):
                want = src[:, :cols].astype(np.float32)
            buf = np.zeros(n * cols + 8, dtype=np.float32)
            dst = buf[off: off + n * cols].reshape(n, cols)
            bad = ctypes.c_int32(-1)
            rc = L.ccfd_cast_rows(src.ctypes.data_as(ctypes.c_void_p), int(dt == np.float64), n, cols, ld,
                                  dst.ctypes.data_as(ctypes.c_void_p), ctypes.byref(bad))
            assert rc == 0
            assert np.array_equal(dst.view(np.uint32), want.view(np.uint32)), (n, cols, ld, off, dt)
            assert bool(bad.value) == bool(np.isinf(want).any())
            assert buf[:off].sum() == 0 and buf[off + n * cols:].sum() == 0  # nothing written outside


def test_single_row_views_with_arbitrary_outer_stride_are_row_major():
    from ccfd_demo_summit_b200 import model

    x = np.arange(30, dtype=np.float32)
    for view in (x[None, :], np.broadcast_to(x, (1, 30)), x.reshape(1, 30)):
        assert model._row_major_float(view) and model._ld(view) == 30
        assert np.array_equal(model.validate_X(view, 30), x[None, :])
    wide = np.zeros((5, 41))
    assert model._row_major_float(wide[:, :30]) and model._ld(wide[:, :30]) == 41
    assert not model._row_major_float(np.asfortranarray(np.zeros((5, 30))))
    assert not model._row_major_float(np.zeros((5, 30), dtype=np.int64))
