"""What the exchange kernel costs the calc stream, per grid cap: un-instrumented step time with a synthetic backward
(spin kernels between bucket-ready points) against the same schedule on a single-rank communicator, plus the
exchange alone per bucket size.  One launch sweeps every grid cap (prb_reducer_set_exchange_ctas is a run-time knob);
the memory-synchronisation domain is fixed per process (PRB_NVL_DOMAIN=default|remote).

    python tools/exchange_sweep.py --gpus 8 --t-bwd-us 730 [--workload resnet50]
"""
import argparse
import ctypes as C
import os
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker(args):
    import numpy as np

    import bench
    import paddle_ray_b200 as prb
    group = prb.distributed.init_parallel_env()
    rank = group.rank
    w = bench.build_workload(prb, args.workload, rank)
    dp, opt = w["dp"], w["opt"]
    _, tail = bench.make_step(w)
    stream = prb.current_stream()
    groups = list(reversed(dp.group_indices))
    idx = [(C.c_int64 * len(g))(*sorted(g, reverse=True)) for g in groups]
    per = int(args.t_bwd_us / len(groups))
    sync_t = prb.Tensor.zeros(1, "float32")

    def rank_max(x):
        t = prb.Tensor.from_numpy(np.array([x], np.float32))
        group.all_reduce(t, "max")
        return float(t.numpy()[0])

    def timed(dp_x, steps=30):
        def one():
            dp_x.prepare_for_backward()
            for arr in idx:
                prb._lib.call("prb_spin", per, stream.handle)
                dp_x._reducer.mark_vars_ready(arr)
            tail()
        for _ in range(4):
            one()
        group.barrier()
        prb.synchronize()
        a, b = prb.Event(), prb.Event()
        group.all_reduce(sync_t)
        a.record()
        for _ in range(steps):
            one()
        b.record()
        b.synchronize()
        return rank_max(a.elapsed_ms(b)) / steps * 1e3

    # baseline: same schedule, single-rank communicator
    h = C.c_void_p()
    prb._lib.call("prb_comm_init_single", group.device, C.byref(h))
    solo = prb.distributed.Group(h.value, 0, 1, group.device)
    dp_solo = prb.DataParallel(w["params"], comm_buffer_size=25, last_comm_buffer_size=1, group=solo)
    dp_solo.consume_from_buckets(True)
    dp_solo.pipelined = True
    base = timed(dp_solo)
    dp.consume_from_buckets(True)
    dom = os.environ.get("PRB_NVL_DOMAIN", "remote")
    if rank == 0:
        print("workload %s N=%d backend %s domain %s T_bwd %.0f us: no-exchange schedule %.1f us/step" %
              (args.workload, group.nranks, dp._reducer.allreduce_backend, dom, args.t_bwd_us, base), flush=True)
    red = dp._reducer
    cs = group.comm_stream
    for ctas in [int(x) for x in args.ctas.split(",")]:
        prb._lib.call("prb_reducer_set_exchange_ctas", red.handle, ctas)
        t = timed(dp)
        # the exchanges alone, back to back
        for _ in range(3):
            for k in range(len(groups)):
                red.allreduce_group(k, cs)
        cs.synchronize()
        group.barrier()
        e0, e1 = prb.Event(), prb.Event()
        group.all_reduce(sync_t, stream=cs)
        e0.record(cs)
        for _ in range(20):
            for k in range(len(groups)):
                red.allreduce_group(k, cs)
        e1.record(cs)
        e1.synchronize()
        comm = rank_max(e0.elapsed_ms(e1)) / 20 * 1e3
        if rank == 0:
            print("  ctas %3d: step %.1f us  exposed %.1f us  exchanges alone %.1f us  overlap %.3f" %
                  (ctas if ctas else 148, t, t - base, comm, 1.0 - max(0.0, t - base) / comm), flush=True)
    assert not red.comm_error()
    group.barrier()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--t-bwd-us", type=float, default=730.0)
    ap.add_argument("--workload", default="resnet50")
    ap.add_argument("--ctas", default="0,64,32,16")
    args = ap.parse_args()
    if "WORLD_SIZE" in os.environ:
        worker(args)
    else:
        sk = socket.socket()
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
        sk.close()
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__)] + sys.argv[1:],
                                  env={**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(args.gpus),
                                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)}) for r in range(args.gpus)]
        sys.exit(max(p.wait() for p in procs))
