"""Bucket allreduce alone, through the reducer, for both backends (PRB_ALLREDUCE=nccl|nvlink).

    PRB_ALLREDUCE=nvlink python tools/allreduce_bench.py --gpus 2
"""
import argparse
import os
import socket
import subprocess
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker(dtype):
    import paddle_ray_b200 as prb
    g = prb.distributed.init_parallel_env()
    n = g.nranks
    es = 4 if dtype == "float32" else 2
    for mb in (1, 8, 25, 102, 256):
        numel = mb * (1 << 20) // es
        p = prb.Tensor.zeros(numel, dtype, name="w")
        p.grad = prb.Tensor.from_numpy(np.full(numel, 1.0 + g.rank, np.float32)).astype(dtype) if dtype != "float32" \
            else prb.Tensor.from_numpy(np.full(numel, 1.0 + g.rank, np.float32))
        dp = prb.DataParallel([p], comm_buffer_size=4096, last_comm_buffer_size=4096)
        red = dp._reducer
        s = g.comm_stream
        dp.prepare_for_backward()
        dp.mark_all_ready()
        prb.synchronize()
        want = sum(1.0 + r for r in range(n)) / n
        got = p.grad.astype("float32").numpy() if dtype != "float32" else p.grad.numpy()
        assert np.all(got == np.float32(want)), (got[:4], want)
        for _ in range(5):
            red.allreduce_group(0, s)
        s.synchronize()
        g.barrier()
        e0, e1 = prb.Event(), prb.Event()
        iters = 20
        e0.record(s)
        for _ in range(iters):
            red.allreduce_group(0, s)
        e1.record(s)
        e1.synchronize()
        ms = e0.elapsed_ms(e1) / iters
        assert not red.comm_error()
        if g.rank == 0:
            alg = mb * (1 << 20) / (ms * 1e-3) / 1e9
            print("%-6s allreduce %4d MiB %s n=%d  %8.1f us  algbw %7.1f GB/s  busbw %7.1f GB/s" %
                  (red.allreduce_backend, mb, dtype, n, ms * 1e3, alg, alg * 2 * (n - 1) / n), flush=True)
        g.barrier()
        del dp, red, p


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--dtype", default="float32")
    args = ap.parse_args()
    if "WORLD_SIZE" in os.environ:
        worker(args.dtype)
    else:
        sk = socket.socket()
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
        sk.close()
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--gpus", str(args.gpus), "--dtype", args.dtype],
                                  env={**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(args.gpus),
                                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)}) for r in range(args.gpus)]
        sys.exit(max(p.wait() for p in procs))
