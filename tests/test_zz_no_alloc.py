"""ccfd_predict_proba_noalloc, the entry point for hosts that may not allocate GPU memory.

Some deployments cannot let a scoring call touch cudaMalloc / cudaMallocAsync: the device memory
is already owned in full by another allocator, or the request thread cannot afford the device-wide
stall an allocation implies.  For them the library carries a fixed 16 MiB scratch array inside its
own image and this entry point scores with it.  The test drives it through the C ABI exactly as
such a host would (raw device pointers, no library-side allocation), two calls back to back on the
shared scratch - the second with fewer, different rows, so anything stale left by the first call
would show - and requires the same bits as the existing ccfd_predict_proba entry point (the
tolerance every parity test here uses: zero).  The file name sorts last so a -x run has already
exercised everything else; the test also passes alone in a fresh process.
"""
import ctypes

import numpy as np
import pytest

from conftest import bits, load_golden

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _forest():
    from ccfd_demo_summit_b200 import converter, model

    raw, X, P, _ = load_golden("small_hard")
    fa = converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                               raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)
    return model.Forest(fa), np.ascontiguousarray(X, dtype=np.float32), P


def _noalloc(forest, xs, n, out):
    from ccfd_demo_summit_b200 import _lib

    stream = torch.cuda.current_stream().cuda_stream
    rc = _lib.lib().ccfd_predict_proba_noalloc(forest.handle, ctypes.c_void_p(xs.data_ptr()), n, xs.shape[1],
                                               ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(stream))
    _lib.check(rc)


def test_noalloc_matches_existing_entry_point_on_two_calls_in_a_row():
    forest, X, P = _forest()
    stream = torch.cuda.current_stream().cuda_stream
    calls = [X[:1500], X[700:1300][::-1].copy()]  # second call: fewer rows, different rows, same scratch
    for rows in calls:
        n = len(rows)
        xs = torch.from_numpy(rows).cuda()
        want = torch.full((n, 2), -1.0, dtype=torch.float64, device="cuda")
        got = torch.full((n, 2), -2.0, dtype=torch.float64, device="cuda")
        forest.predict_proba_device(xs.data_ptr(), n, 30, want.data_ptr(), stream, "split")  # existing entry point
        torch.cuda.synchronize()
        _noalloc(forest, xs, n, got)
        assert torch.equal(got.view(torch.int64), want.view(torch.int64)), n
    # and both agree with scikit-learn's stored output for these rows
    assert np.array_equal(bits(got.cpu().numpy()), bits(P[700:1300][::-1]))
    forest.close()


def test_noalloc_rejects_requests_beyond_the_built_in_scratch():
    from ccfd_demo_summit_b200 import _lib

    forest, X, _ = _forest()
    limit = _lib.NOALLOC_MAX_TREE_ROWS  # CCFD_NOALLOC_MAX_TREE_ROWS of include/ccfd_rf.h (kept equal by a CPU test)
    n = limit // forest.n_trees + 1
    xs = torch.zeros((n, 30), dtype=torch.float32, device="cuda")
    out = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    with pytest.raises(ValueError, match="CCFD_NOALLOC_MAX_TREE_ROWS"):
        _noalloc(forest, xs, n, out)
    # the largest request that fits still works and matches the bulk path
    n = limit // forest.n_trees
    big = np.tile(X, (n // len(X) + 1, 1))[:n]
    xs = torch.from_numpy(big).cuda()
    want = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    got = torch.empty((n, 2), dtype=torch.float64, device="cuda")
    forest.predict_proba_device(xs.data_ptr(), n, 30, want.data_ptr(), torch.cuda.current_stream().cuda_stream, "stream")
    _noalloc(forest, xs, n, got)
    assert torch.equal(got.view(torch.int64), want.view(torch.int64))
    forest.close()


def test_noalloc_refuses_forests_whose_trees_do_not_fit_shared_memory():
    """The built-in scratch serves the shared-memory (SPLIT) kernels only: a forest with a tree far
    beyond a shared-memory stage is answered with CCFD_EUNSUPPORTED, not scored some other way."""
    from ccfd_demo_summit_b200 import _lib, converter, model

    rng = np.random.default_rng(0)
    n_int = 9000                          # one comb-shaped tree of 18,001 nodes: > 96 KB packed
    n = 2 * n_int + 1
    left = np.full(n, -1, np.int32)
    right = np.full(n, -1, np.int32)
    idx = np.arange(n_int) * 2
    left[idx], right[idx] = idx + 1, idx + 2
    val = np.tile(np.array([0.25, 0.75]), (n, 1))
    fa = converter.from_arrays(np.array([0, n]), left, right, rng.integers(0, 30, n), rng.normal(size=n), np.zeros(n, np.uint8),
                               val, 30, 2)
    forest = model.Forest(fa)
    assert forest.info["stream_kernel_ok"] == 0
    xs = torch.zeros((8, 30), dtype=torch.float32, device="cuda")
    out = torch.empty((8, 2), dtype=torch.float64, device="cuda")
    rc = _lib.lib().ccfd_predict_proba_noalloc(forest.handle, ctypes.c_void_p(xs.data_ptr()), 8, 30, ctypes.c_void_p(out.data_ptr()), None)
    assert rc == _lib.EUNSUPPORTED, _lib.lib().ccfd_last_error()
    forest.close()
