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
