"""Allreduce sweep through the product's communicator (N ranks on one box): algbw / busbw per message size.

    python tools/comm_bench.py --gpus 2        (self-launches the ranks; honours NCCL_* and PRB_NCCL_LIB)
"""
import argparse
import os
import socket
import subprocess
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def worker():
    import paddle_ray_b200 as prb
    g = prb.distributed.init_parallel_env()
    n = g.nranks
    for mb in (1, 8, 25, 102, 256, 1024):
        numel = mb * (1 << 20) // 4
        t = prb.Tensor.zeros(numel, "float32")
        s = g.comm_stream
        for _ in range(5):
            g.all_reduce(t, stream=s)
        s.synchronize()
        g.barrier()
        e0, e1 = prb.Event(), prb.Event()
        iters = 20
        e0.record(s)
        for _ in range(iters):
            g.all_reduce(t, stream=s)
        e1.record(s)
        e1.synchronize()
        ms = e0.elapsed_ms(e1) / iters
        if g.rank == 0:
            alg = mb * (1 << 20) / (ms * 1e-3) / 1e9
            print("allreduce %5d MiB fp32  n=%d  %8.1f us  algbw %7.1f GB/s  busbw %7.1f GB/s" %
                  (mb, n, ms * 1e3, alg, alg * 2 * (n - 1) / n), flush=True)
        del t
    g.barrier()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    args = ap.parse_args()
    if "WORLD_SIZE" in os.environ:
        worker()
    else:
        sk = socket.socket()
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
        sk.close()
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--gpus", str(args.gpus)],
                                  env={**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(args.gpus),
                                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)}) for r in range(args.gpus)]
        sys.exit(max(p.wait() for p in procs))
