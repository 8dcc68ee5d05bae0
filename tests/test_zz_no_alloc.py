"""prb_squared_l2_norm_noalloc -- the entry point for callers that must not allocate device memory.

Some users of this library ask for a gradient norm at moments when a device allocation is off the table: their
own memory pool already owns the GPU, they are inside a section that must not block on cudaMalloc/cudaFree, or
they are handling an out-of-memory condition.  For them the library offers a variant of squared_l2_norm whose
scratch is a fixed array inside the library image.  This test calls it through the C ABI and checks it against
the regular (allocating) entry point on the same inputs -- twice in a row, because every call reuses the same
scratch and must not depend on what the previous call left there -- and checks that inputs beyond the fixed
capacity are refused with a clear message instead of overrunning the array.

The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
It passes on its own in a fresh process as well (it needs nothing that other tests set up).
"""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

MAX_NUMEL = 268435456  # PRB_NOALLOC_MAX_NUMEL in include/paddle_ray_b200.h


@pytest.fixture(scope="module")
def prb():
    import paddle_ray_b200 as prb
    assert prb.device_count() > 0, "no CUDA device: the product has no CPU fallback"
    prb.set_device(0)
    return prb


def _noalloc(prb, x):
    out = prb.Tensor(1, x.dtype, x.device)
    a, b = x.prb, out.prb
    prb._lib.call("prb_squared_l2_norm_noalloc", C.byref(a), C.byref(b), prb.current_stream().handle)
    return out.numpy()


@pytest.mark.parametrize("dtype,tol", [(orc.F32, 1e-5), (orc.BF16, 8e-3), (orc.F16, 1e-3)])
def test_matches_the_allocating_entry_point_on_two_calls_in_a_row(prb, dtype, tol):
    rng = np.random.default_rng(42)
    # first a long input (many partial sums in the scratch), then a short one that must not see leftovers
    for n in ((1 << 22) + 12345, 777):
        x = orc.from_f32(rng.normal(0, 1e-2, n).astype(np.float32), dtype)
        t = prb.Tensor.from_numpy(x, dtype)
        got_bits = _noalloc(prb, t)
        want_bits = prb._C_ops.squared_l2_norm(t).numpy()
        # both entry points add the same partial sums in the same order: not close, EQUAL
        assert got_bits.tobytes() == want_bits.tobytes(), (n, got_bits, want_bits)
        got = orc.to_f32(got_bits, dtype)[0]
        _, want64 = orc.squared_l2_norm(x, dtype)
        np.testing.assert_allclose(got, want64, rtol=tol)
    # and the same input twice gives the same bits (nothing carried over between calls)
    assert _noalloc(prb, t).tobytes() == _noalloc(prb, t).tobytes()


def test_edge_sizes_and_unaligned_input(prb):
    rng = np.random.default_rng(1)
    for n in (0, 1, 8191, 8192, 8193):
        x = rng.normal(0, 1, n).astype(np.float32)
        t = prb.Tensor.from_numpy(x)
        got = _noalloc(prb, t)[0]
        np.testing.assert_allclose(got, float(np.sum(x.astype(np.float64) ** 2)), rtol=1e-5)
    big = prb.Tensor.from_numpy(rng.normal(0, 1, 20001).astype(np.float32))
    view = big.view(1, 20000)  # element-aligned only
    assert _noalloc(prb, view).tobytes() == prb._C_ops.squared_l2_norm(view).numpy().tobytes()


def test_inputs_beyond_the_fixed_capacity_are_refused(prb):
    from paddle_ray_b200._lib import PRBError, PRBTensor
    small = prb.Tensor.zeros(16, "float32")
    out = prb.Tensor(1, "float32")
    # the size check comes before anything touches memory, so a small buffer with an oversized count is safe here
    too_big = PRBTensor(small.data_ptr, MAX_NUMEL + 1, small.dtype, small.device)
    o = out.prb
    with pytest.raises(PRBError, match="fixed scratch covers at most %d" % MAX_NUMEL):
        prb._lib.call("prb_squared_l2_norm_noalloc", C.byref(too_big), C.byref(o), prb.current_stream().handle)
    host = prb.Tensor(16, "float32", device=-1)
    h = host.prb
    with pytest.raises(PRBError, match="on the device"):
        prb._lib.call("prb_squared_l2_norm_noalloc", C.byref(h), C.byref(o), prb.current_stream().handle)
