"""Kernel-level sweeps on one GPU (not the driver's bench): GB/s of each stage vs the measured HBM peak."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paddle_ray_b200 as prb  # noqa: E402
from paddle_ray_b200 import _C_ops  # noqa: E402

PEAK = 6541.8
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def timeit(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    s, e = prb.Event(), prb.Event()
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    e.synchronize()
    return s.elapsed_ms(e) / iters


def main():
    prb.set_device(0)
    print("device:", prb.core.device_info())
    T = prb.Tensor
    lr = T.from_numpy(np.array([1e-3], np.float32))
    b1 = T.from_numpy(np.array([0.9 ** 10], np.float32))
    b2 = T.from_numpy(np.array([0.999 ** 10], np.float32))
    for logn in [int(x) for x in os.environ.get('PRB_MB_SIZES', '20,22,24,26,28').split(',')]:
        n = 1 << logn
        p, g, m1 = (T.from_numpy(np.random.default_rng(i).normal(0, 0.02, n).astype(np.float32)) for i in range(3))
        m2 = T.from_numpy(np.random.default_rng(3).uniform(0, 1e-3, n).astype(np.float32))
        ms = timeit(lambda: _C_ops.adamw_(p, g, lr, m1, m2, b1, b2, None, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True, False, 1000, False, True))
        gbs = n * 28 / ms / 1e6
        print(f"adamw fp32 n=2^{logn}: {ms*1e3:9.1f} us  {gbs:8.1f} GB/s  {gbs/PEAK:.3f} of measured peak")
        ms = timeit(lambda: _C_ops.scale_(g, 0.125, 0.0, False))
        gbs = n * 8 / ms / 1e6
        print(f"scale fp32 n=2^{logn}: {ms*1e3:9.1f} us  {gbs:8.1f} GB/s  {gbs/PEAK:.3f}")
        x = _C_ops.squared_l2_norm(g)
        ms = timeit(lambda: _C_ops.squared_l2_norm(g))
        gbs = n * 4 / ms / 1e6
        print(f"sqnorm fp32 n=2^{logn}: {ms*1e3:9.1f} us  {gbs:8.1f} GB/s  {gbs/PEAK:.3f}")
        del p, g, m1, m2
    # bf16 multi-precision
    for logn in (24, 27):
        n = 1 << logn
        r = np.random.default_rng(7)
        mp = T.from_numpy(r.normal(0, 0.02, n).astype(np.float32))
        p = mp.astype("bfloat16")
        g = T.from_numpy(r.normal(0, 1e-2, n).astype(np.float32)).astype("bfloat16")
        m1 = T.from_numpy(r.normal(0, 1e-3, n).astype(np.float32))
        m2 = T.from_numpy(r.uniform(0, 1e-3, n).astype(np.float32))
        ms = timeit(lambda: _C_ops.adamw_(p, g, lr, m1, m2, b1, b2, mp, None, 0.9, 0.999, 1e-8, 1.0, 0.01, True, False, 1000, True, True))
        gbs = n * 28 / ms / 1e6
        print(f"adamw bf16+master n=2^{logn}: {ms*1e3:9.1f} us  {gbs:8.1f} GB/s  {gbs/PEAK:.3f}")


if __name__ == "__main__":
    main()
