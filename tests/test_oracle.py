"""The oracle against scikit-learn's own outputs (tests/golden, made by make_golden.py)."""
import numpy as np
import pytest

from conftest import bits
from oracle import rf_oracle


def test_numpy_restatement_matches_sklearn_bitwise(golden):
    name, raw, X, P, leaves = golden
    got = rf_oracle.predict_proba(raw, X)
    assert got.shape == P.shape and got.dtype == np.float64
    assert np.array_equal(bits(got), bits(P)), name


def test_c_restatement_matches_sklearn_bitwise(golden):
    name, raw, X, P, leaves = golden
    got = rf_oracle.predict_proba_c(raw, X)
    assert np.array_equal(bits(got), bits(P)), name


def test_apply_matches_sklearn(golden):
    name, raw, X, P, leaves = golden
    assert np.array_equal(rf_oracle.apply(raw, X), leaves), name


def test_rows_sum_to_one(golden):
    # sklearn's own property test: SK/ensemble/tests/test_forest.py:278-293
    _, raw, X, P, _ = golden
    np.testing.assert_allclose(rf_oracle.predict_proba(raw, X).sum(axis=1), 1.0, rtol=0, atol=1e-12)


@pytest.mark.parametrize("bad", ["1d", "empty", "cols", "inf", "overflow"])
def test_input_validation_mirrors_sklearn(bad):
    # SK/ensemble/_forest.py:606-625 behaviour probed in SURVEY.md §8a A1
    from conftest import load_golden

    raw, X, _, _ = load_golden("small_hard")
    x = {"1d": X[0], "empty": X[:0], "cols": X[:, :29],
         "inf": np.where(np.arange(30) == 3, np.inf, X[:2]),
         "overflow": np.where(np.arange(30) == 3, 1e39, X[:2].astype(np.float64))}[bad]
    with pytest.raises(ValueError):
        rf_oracle.predict_proba(raw, x)


def test_accepts_lists_ints_fortran_order():
    # SK/ensemble/tests/test_forest.py:1117-1146 test_memory_layout
    from conftest import load_golden

    raw, X, P, _ = load_golden("small_hard")
    ref = rf_oracle.predict_proba(raw, X[:50])
    assert np.array_equal(rf_oracle.predict_proba(raw, np.asfortranarray(X[:50])), ref)
    assert np.array_equal(rf_oracle.predict_proba(raw, X[:50].astype(np.float64)), ref)
    assert np.array_equal(rf_oracle.predict_proba(raw, X[:50].tolist()), ref)
    assert np.array_equal(rf_oracle.predict_proba(raw, X[:100:2]), rf_oracle.predict_proba(raw, X[:100])[::2])
