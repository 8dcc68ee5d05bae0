"""Timeline of ONE data-parallel step with a synthetic backward (spin kernels), from CUDA events:
when each bucket is packed, when its allreduce is complete, when its optimizer launch starts and ends.

    python tools/overlap_timeline.py --gpus 2 --mult 2

Times are microseconds after the start of the instrumented step, per rank.  The step is embedded between
un-instrumented steps so the queues are in steady state.  The un-instrumented average ("plain") is printed first.

CAUTION (profiles/r1_experiments.md): the timing events perturb what they measure.  A timing event recorded on
the calc stream while an allreduce kernel is in flight was seen to hold the calc stream's next kernel back until
that allreduce had finished (40-70 us, every other bucket); the same schedule without events does not show it
("plain" 678 us vs 825 us instrumented, N=2, ResNet-50, T_bwd 460 us).  Use the timeline for the ORDER of things
and for kernel durations, the plain number for the step time.
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
    import bench
    import paddle_ray_b200 as prb
    group = prb.distributed.init_parallel_env()
    w = bench.build_workload(prb, "resnet50", group.rank)
    dp, opt = w["dp"], w["opt"]
    stream = prb.current_stream()
    side = prb.Stream()
    groups = list(reversed(dp.group_indices))
    idx = [(C.c_int64 * len(g))(*g) for g in groups]
    ng = len(groups)
    per = int(args.t_bwd_us / ng)

    def one(events=None):
        dp.prepare_for_backward()
        if events is not None:
            events["start"].record()
        for k, arr in enumerate(idx):
            prb._lib.call("prb_spin", per, stream.handle)
            if events is not None:
                events["spin%d" % k].record()
            dp._reducer.mark_vars_ready(arr)
            if events is not None:
                events["pack%d" % k].record()
                if group.nranks > 1:
                    dp.wait_group(k, side)
                    events["ar%d" % k].record(side)
        if events is not None:
            opt._launch_events = [(events["adam%d_a" % k], events["adam%d_b" % k]) for k in reversed(range(ng))]
            opt._launch_events_used = []
        opt.step()
        if events is not None:
            opt._launch_events = None
            events["end"].record()

    names = ["start", "end"] + ["%s%d" % (n, k) for k in range(ng) for n in ("spin", "pack", "ar")] + \
        ["adam%d_%s" % (k, s) for k in range(ng) for s in "ab"]
    # the same schedule without any instrumentation: 20 steps between two events (what bench.py times)
    for rep in range(2):
        for _ in range(3):
            one()
        group.barrier()
        prb.synchronize()
        a, b = prb.Event(), prb.Event()
        a.record()
        for _ in range(20):
            one()
        b.record()
        b.synchronize()
        print("rank %d plain: %.1f us per step" % (group.rank, a.elapsed_ms(b) * 1e3 / 20), flush=True)
    for rep in range(args.reps):
        ev = {n: prb.Event() for n in names}
        for _ in range(6):
            one()
        one(ev)
        for _ in range(3):
            one()
        prb.synchronize()
        group.barrier()
        used = len(opt._launch_events_used)
        rows = []
        for n in names:
            if n == "start" or (n.startswith("ar") and group.nranks == 1):
                continue
            if n.startswith("adam") and int(n[4]) >= used:
                continue
            try:
                rows.append((ev["start"].elapsed_ms(ev[n]) * 1e3, n))
            except Exception:
                pass
        rows.sort()
        for r in range(group.nranks):
            if r == group.rank:
                print("rank %d rep %d (spin %d us per bucket, %d optimizer launches): " % (r, rep, per, used) +
                      "  ".join("%s %.0f" % (n, t) for t, n in rows), flush=True)
            group.barrier()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--t-bwd-us", type=float, default=460.0)
    ap.add_argument("--reps", type=int, default=3, help="instrumented steps to print (0 = plain timing only)")
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
