"""Property tests on random (unfitted) forests: the packed walk on the CPU and, with a GPU, every
CUDA kernel must reproduce the oracle bit for bit whatever the tree shapes and edge values."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import packed_emulator
import random_forests as rfs
from conftest import bits
from ccfd_demo_summit_b200 import converter
from oracle import rf_oracle


def _fa(raw):
    return converter.from_arrays(raw.tree_offsets, raw.left, raw.right, raw.feature, raw.threshold,
                                 raw.missing_go_to_left, raw.value, raw.n_features, raw.n_classes)


@settings(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck))
@given(seed=st.integers(0, 10**6), n_trees=st.integers(1, 12), max_internal=st.integers(1, 120),
       F=st.sampled_from([1, 2, 7, 30, 31, 64]), skew=st.floats(0.0, 0.95))
def test_packed_walk_equals_oracle_on_random_forests(seed, n_trees, max_internal, F, skew):
    raw = rfs.random_forest(seed, n_trees, max_internal, F, skew)
    X = rfs.random_rows(seed, raw, 300)
    want = rf_oracle.predict_proba_c(raw, X)
    assert np.array_equal(bits(want), bits(rf_oracle.predict_proba(raw, X)))  # the two restatements agree
    fa = _fa(raw)
    blob, recs, top = packed_emulator.pack(fa)
    assert np.array_equal(bits(packed_emulator.predict_proba(blob, recs, F, X)), bits(want))
    assert np.array_equal(bits(packed_emulator.predict_proba(blob, recs, F, X, top)), bits(want))


@pytest.mark.gpu
@pytest.mark.parametrize("seed,n_trees,max_internal,F,skew,n", [
    (1, 7, 200, 30, 0.3, 5000), (2, 3, 1500, 30, 0.0, 40000), (3, 40, 60, 30, 0.9, 20000), (4, 1, 1, 1, 0.0, 700),
    (5, 9, 300, 64, 0.5, 3000), (6, 12, 500, 7, 0.97, 17000), (7, 25, 40, 31, 0.2, 66000), (8, 2, 3000, 30, 0.1, 2049),
    (9, 5, 100, 200, 0.3, 900), (10, 3, 60, 255, 0.6, 120000)])  # rows too wide for the shared-memory kernels
def test_cuda_kernels_equal_oracle_on_random_forests(seed, n_trees, max_internal, F, skew, n):
    import torch

    from ccfd_demo_summit_b200 import model

    raw = rfs.random_forest(seed, n_trees, max_internal, F, skew)
    X = rfs.random_rows(seed, raw, n)
    want = rf_oracle.predict_proba_c(raw, X)
    forest = model.Forest(_fa(raw))
    xs = torch.from_numpy(X).cuda()
    st_ = torch.cuda.current_stream().cuda_stream
    kernels = ["global"] + (["stream", "split"] if forest.info["stream_kernel_ok"] else [])
    for kernel in kernels:
        m = min(n, 6000) if kernel == "split" else n
        out = torch.full((m, 2), -1.0, dtype=torch.float64, device="cuda")
        forest.predict_proba_device(xs.data_ptr(), m, F, out.data_ptr(), st_, kernel)
        assert np.array_equal(bits(out.cpu().numpy()), bits(want[:m])), kernel
    got = forest.predict_proba(X)  # host pointers, kernel picked by the library
    assert np.array_equal(bits(got), bits(want))
    forest.close()


@pytest.mark.gpu
@pytest.mark.parametrize("F", [31, 32])
@pytest.mark.parametrize("with_nan", [False, True])
def test_row_layout_boundary(F, with_nan):
    """31 features is the last count that takes the blocked shared-memory row layout (64-row blocks of
    32 feature rows, the x address one and-or of the node's meta word), 32 the first that takes the
    strided one.  Every way rows reach shared memory is exercised on both sides: TMA row chunks
    (dense, aligned), the cooperative loader (ragged last batch, strided or unaligned rows), one and
    two rows per thread, and the SPLIT tiles in both of their layouts."""
    import torch

    from ccfd_demo_summit_b200 import model

    raw = rfs.random_forest(100 + F, 9, 400, F, 0.4)
    n = 160_000 + 77  # two rows per thread in STREAM, with a ragged last batch
    X = rfs.random_rows(F, raw, n)
    if not with_nan:
        X = np.nan_to_num(X, nan=0.25)
    want = rf_oracle.predict_proba_c(raw, X)
    forest = model.Forest(_fa(raw))
    assert forest.info["stream_kernel_ok"]
    st_ = torch.cuda.current_stream().cuda_stream
    for ld, off in [(F, 0), (F + 3, 0), (F, 1)]:
        buf = torch.zeros(n * ld + 8, dtype=torch.float32, device="cuda")
        xs = buf[off:off + n * ld].view(n, ld)
        xs[:, :F] = torch.from_numpy(X).cuda()
        for kernel, m in [("stream", n), ("stream", 20_000), ("stream", 513), ("split", 1024), ("split", 6000), ("split", 1), ("global", 3000)]:
            out = torch.full((m, 2), -1.0, dtype=torch.float64, device="cuda")
            forest.predict_proba_device(xs.data_ptr(), m, ld, out.data_ptr(), st_, kernel)
            assert np.array_equal(bits(out.cpu().numpy()), bits(want[:m])), (kernel, m, ld, off)
    forest.close()
