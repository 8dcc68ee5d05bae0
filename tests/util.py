"""Helpers shared by the GPU parity tests: seeded data in the BASELINE.md distributions, dtype plumbing."""
import numpy as np

from oracle import oracle as orc


def make_state(rng, n, dtype_param=orc.F32, dtype_grad=None, master=False, grad_scale=1.0):
    """params N(0,0.02), grads N(0,1e-2)*grad_scale, m1 N(0,1e-3), m2 U(0,1e-3)  (BASELINE.md section 3)."""
    dtype_grad = dtype_param if dtype_grad is None else dtype_grad
    p32 = rng.normal(0, 0.02, n).astype(np.float32)
    g32 = (rng.normal(0, 1e-2, n) * grad_scale).astype(np.float32)
    st = {
        "param": orc.from_f32(p32, dtype_param),
        "grad": orc.from_f32(g32, dtype_grad),
        "mom1": rng.normal(0, 1e-3, n).astype(np.float32),
        "mom2": rng.uniform(0, 1e-3, n).astype(np.float32),
        "master": None,
    }
    if master:
        st["master"] = orc.to_f32(st["param"], dtype_param)  # master = fp32 cast of the param (optimizer.py:640-666)
    return st


def copy_state(st):
    return {k: (None if v is None else v.copy()) for k, v in st.items()}


def bits_equal(a, b):
    """Bit-exact comparison that treats NaNs with identical payloads as equal."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    assert a.shape == b.shape and a.dtype == b.dtype, (a.shape, b.shape, a.dtype, b.dtype)
    va = a.view(np.uint8)
    vb = b.view(np.uint8)
    if not np.array_equal(va, vb):
        idx = np.flatnonzero(a.reshape(-1) != b.reshape(-1))
        raise AssertionError("bit mismatch at %d of %d elements, first at %s: %r vs %r" % (
            idx.size, a.size, idx[:5], a.reshape(-1)[idx[:5]], b.reshape(-1)[idx[:5]]))


def ulp_diff_bf16(a_bits, b_bits):
    """|a-b| in units of bf16 ulps for same-sign finite bit patterns (monotone integer mapping)."""
    def key(u):
        u = u.astype(np.int32)
        return np.where(u & 0x8000, 0x8000 - (u & 0x7FFF), 0x8000 + (u & 0x7FFF))
    return np.abs(key(a_bits) - key(b_bits))
