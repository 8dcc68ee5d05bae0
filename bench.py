#!/usr/bin/env python
"""Benchmark of the data-parallel step tail (BASELINE.json metric: AdamW+grad-sync step GB/s & ms/step).

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload bert_base] [--no-configs]

One "step" = one pass of the hot path over one batch of synthetic gradients, all four north-star stages:
    prepare_for_backward -> every gradient marked ready (reverse parameter order) ->
    per bucket: pack + 1/nranks scale (+ AMP inf/nan check + squared-norm partials in the same kernel) ->
    bucket allreduce on the comm stream (N > 1; there the check + partials ride in the exchange kernel) ->
    norm finalize -> one fused multi-tensor AdamW launch that reads the reduced gradients from the bucket arena.
Headline workload = BASELINE.json configs[2]: BERT-base parameter shapes (206 tensors, 110,106,428 elements),
AMP bf16 with fp32 master weights, global-norm clip, paddle.DataParallel + paddle.optimizer.AdamW + GradScaler.
The other configs (ResNet-50 fp32 = configs[1], GPT-2 medium = configs[3], the flat sweep = configs[4], and the
"gradients produced in the arena" mode) are measured with few steps and nested under `configs` of the same line.

`value` = whole-job algorithmic GB/s with inputs resident in HBM; `e2e` = the same metric through the public Python
API with the gradients coming from pinned HOST memory every step (H2D per bucket inside the timed region) and the
step's metric (global grad norm, found_inf, loss scale) read back.  Algorithmic bytes per element:
update r14+w14 (bf16 + master) or r16+w12 (fp32) = 28, pack r s + w s; the check + norm ride in the pack, so a
bf16 AMP+clip step is 32 B/elem (SURVEY 8d's floor with a separate norm read is 34), fp32 without AMP/clip 36.

The oracle (oracle/) is imported only by the cpu_baseline / --impl reference legs and by the N>1 parity preamble
(as the checker of a tiny step, outside every timed region).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import bench_configs  # noqa: E402

METRIC = "adamw_grad_sync_step_throughput"
UNIT = "GB/s"

WORKLOADS = {
    # name: (shape list, param dtype, grad dtype, amp, clip, multi_precision, human description)
    "resnet50": ("resnet50", "float32", "float32", False, False, False,
                 "ResNet-50 param shapes (161 tensors, 25,557,032 elems), DataParallel + AdamW fp32"),
    "resnet18": ("resnet18", "float32", "float32", False, False, False,
                 "ResNet-18 param shapes (62 tensors, 11,689,512 elems), DataParallel + AdamW fp32"),
    "bert_base": ("bert_base", "bfloat16", "bfloat16", True, True, True,
                  "BERT-base shapes (206 tensors, 110,106,428 elems), AMP bf16 + fp32 master, global-norm clip"),
    "gpt2_medium": ("gpt2_medium", "bfloat16", "bfloat16", True, True, True,
                    "GPT-2 medium shapes (292 tensors, 354,823,168 elems), multi_precision bf16, 25 MB buckets"),
}
DEFAULT_WORKLOAD = "bert_base"
# configs[4]: the flat sweep (name = flat<log2 n | 2e9>x<tensors>[r = ragged sizes][:bf16])
FLAT_SWEEP = ["flat20x1", "flat22x1", "flat24x1", "flat26x1", "flat28x1", "flat30x1", "flat2e9x1", "flat26x1000r",
              "flat26x1:bf16", "flat30x1:bf16", "flat26x1000r:bf16"]


def workload_spec(name):
    """-> (shape list name, param dtype, grad dtype, amp, clip, multi_precision, description)"""
    if name in WORKLOADS:
        return WORKLOADS[name]
    base, _, dt = name.partition(":")
    if dt == "bf16":
        return (base, "bfloat16", "bfloat16", False, False, True, "flat sweep %s, bf16 + fp32 master, fused AdamW" % base)
    return (base, "float32", "float32", False, False, False, "flat sweep %s, fp32, fused AdamW" % base)


def bytes_per_elem(pdt, gdt, amp, clip, arena, fused_pass1=True):
    """Algorithmic HBM bytes per gradient element of one step."""
    s = 4 if gdt == "float32" else 2
    upd = 28 + (2 if (pdt != "float32" and gdt == "float32") else 0)   # fp32 main-grad next to a 16-bit param: 30
    pack = 0 if arena else 2 * s
    norm_read = s if ((amp or clip) and (arena or not fused_pass1)) else 0
    return upd + pack + norm_read


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# --------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region
# --------------------------------------------------------------------------------------------
class ClockSampler:
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTE = {"sw_power_cap": 0x4}

    def __init__(self, device_index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[device_index]) if vis and vis.split(",")[device_index].isdigit() else device_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.h = None

    def _sample(self):
        if self.h is None:
            return
        try:
            self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            for name, bit in {**self.BAD, **self.NOTE}.items():
                if mask & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def start(self):
        def run():
            while not self._stop.is_set():
                self._sample()
                time.sleep(0.002)
        self._thread = threading.Thread(target=run, daemon=True)
        self._thread.start()

    def stop(self):
        self._sample()
        self._stop.set()
        if self._thread:
            self._thread.join(1.0)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unsampled"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------
# workload construction (product API only)
# --------------------------------------------------------------------------------------------
_TILE = 1 << 22


def _fill(prb, t, rng, mean_scale, uniform=False):
    """Synthetic values.  Tensors above 16 M elements are filled by tiling one random
    4 M-element block on the device (host RNG for 2e9 elements would take minutes; throughput does not care)."""
    n = t.numel
    nb = n if n <= (1 << 24) else _TILE
    blk = (rng.random(nb, dtype=np.float32) if uniform else rng.standard_normal(nb, dtype=np.float32)) * np.float32(mean_scale)
    first = t if nb == n else t.view(0, nb)
    if t.dtype == prb._lib.FLOAT32:
        first.copy_from_numpy(blk)
    else:
        first.copy_(prb.Tensor.from_numpy(blk).astype(t.dtype))
        prb.current_stream().synchronize()
    es = prb._lib.DTYPE_SIZE[t.dtype]
    done = nb
    while done < n:
        m = min(done, n - done)
        prb._lib.call("prb_memcpy_d2d", t.data_ptr + done * es, t.data_ptr, m * es, prb.current_stream().handle)
        done += m
    return blk if nb == n else None


def build_workload(prb, name, rank, arena=False, host_mirror=False, producer_pass1=True, sharded=False):
    shapes_name, pdt, gdt, amp, clip, mp, desc = workload_spec(name)
    numels = bench_configs.param_numels(shapes_name)
    rng = np.random.default_rng(1234 + rank)
    gsz = 4 if gdt == "float32" else 2
    params = []
    for i, n in enumerate(numels):
        p = prb.Tensor(n, pdt, name="w_%d" % i)
        _fill(prb, p, rng, 0.02)
        params.append(p)
    grad_arena = host_arena = None
    offs, total = [], 0
    if not arena:
        # gradients live in ONE device arena (256-byte aligned views, parameter order), mirrored by one pinned host
        # arena, so the e2e leg moves each bucket's gradients with a single H2D copy
        align = 256 // gsz
        for n in numels:
            offs.append(total)
            total += (n + align - 1) // align * align
        grad_arena = prb.Tensor.zeros(total, gdt)
        for p, off in zip(params, offs):
            p.grad = grad_arena.view(off, p.numel)
    grad_clip = prb.nn.ClipGradByGlobalNorm(1.0) if clip else None
    opt = prb.optimizer.AdamW(learning_rate=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, parameters=params,
                              weight_decay=0.01, grad_clip=grad_clip, multi_precision=mp)
    # the reduced gradients are consumed from the bucket arena; nothing reads an unscaled copy back
    opt._write_back_unscaled_grads = False
    opt._use_producer_pass1 = producer_pass1
    scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15) if amp else None
    dp = prb.DataParallel(params, comm_buffer_size=25, last_comm_buffer_size=1, consume_from_buckets=True,
                          grads_in_arena=arena)
    inner = opt
    if sharded:
        # optimizer state and update sharded over the ranks: reduce-scatter -> update of the owned arena ranges ->
        # gather of the parameters (bit-identical parameters, 28/N instead of 28 B/elem in the update pass)
        opt = prb.distributed.GroupShardedOptimizerStage2(params, inner, dp.group, reduce_overlap=(sharded == "bucket"))
    inner._create_accumulators()
    # moments start away from zero (m1 N(0,1e-3), m2 U(0,1e-3), step 10)
    for p in inner._parameter_list:
        acc = inner._accumulators[p.name]
        _fill(prb, acc["moment1"], rng, 1e-3)
        _fill(prb, acc["moment2"], rng, 1e-3, uniform=True)
    inner._beta1_pow.copy_from_numpy(np.array([0.9 ** 10], np.float32))
    inner._beta2_pow.copy_from_numpy(np.array([0.999 ** 10], np.float32))
    for p in params:  # grads ~ N(0, 1e-2), times the loss scale for AMP runs
        _fill(prb, p.grad, rng, 1e-2 * (2.0 ** 15 if amp else 1.0))
    if host_mirror and not arena:
        host_arena = prb.PinnedBuffer(total, np.float32 if gsz == 4 else np.uint16)
        prb._lib.call("prb_memcpy_d2h", host_arena.ptr, grad_arena.data_ptr, grad_arena.nbytes, prb.current_stream().handle)
    prb.current_stream().synchronize()
    return {"name": name, "params": params, "opt": opt, "scaler": scaler, "dp": dp, "numels": numels, "desc": desc,
            "grad_arena": grad_arena, "host_arena": host_arena, "offs": offs, "amp": amp, "clip": clip, "pdt": pdt,
            "gdt": gdt, "arena": arena, "elems": int(sum(numels)), "buckets": len(dp.group_indices), "gsz": gsz,
            "bpe": bytes_per_elem(pdt, gdt, amp, clip, arena, producer_pass1), "clip_obj": grad_clip,
            "sharded": sharded}


def make_step(w):
    dp, opt, scaler = w["dp"], w["opt"], w["scaler"]

    def tail():
        if scaler is not None:
            scaler.step(opt)
            scaler.update()
        else:
            opt.step()

    def step():
        dp.prepare_for_backward()
        dp.mark_all_ready()
        tail()

    return step, tail


KERNEL_KINDS = 6
# algorithmic bytes per element each kernel kind moves (s = gradient element size, upd = 28 / 30)
def kernel_bytes(kind_name, w, nranks):
    s = w["gsz"]
    upd = 28 + (2 if (w["pdt"] != "float32" and w["gdt"] == "float32") else 0)
    return {"map_chunks_kernel": 2 * s, "grad_pass_kernel": s, "grad_pass_kernel<pack>": 2 * s,
            "adam_chunks_kernel": upd, "norm_finalize_kernel": 0,
            # exchange, per rank: reads its slice from n ranks, writes it to n ranks (HBM + NVLink together)
            "nvl_exchange_kernel": 2 * s}[kind_name]


def read_kernel_timing(prb, w, nranks, steps, peak):
    """Per-kernel rooflines from the CUDA events the library recorded around every launch (prb_kernel_timing)."""
    lib = prb._lib.lib
    out = {}
    for k in range(KERNEL_KINDS):
        ms, n = C.c_double(), C.c_int64()
        prb._lib.call("prb_kernel_timing_read", k, C.byref(ms), C.byref(n))
        if n.value == 0:
            continue
        name = lib.prb_kernel_kind_name(k).decode()
        per_step = ms.value / steps
        bpe = kernel_bytes(name, w, nranks)
        e = {"launches_per_step": n.value / steps, "ms_per_step": per_step}
        if bpe:
            elems = w["elems"]
            if w.get("sharded") and nranks > 1 and name == "adam_chunks_kernel":
                elems = elems / nranks   # each rank updates the arena ranges it owns only
            e["algorithmic_bytes_per_step"] = bpe * elems
            e["achieved_gbs"] = bpe * elems / (per_step * 1e-3) / 1e9
            e["frac_of_measured_hbm_peak"] = e["achieved_gbs"] / peak
        out[name] = e
    return out


def measure_workload(prb, group, w, steps, warmup, peak, instrumented=True, sample_clocks=False):
    """K timed steps between two CUDA events on the launching stream (barrier + synchronize on both sides, start
    events lined up on the device, max over ranks), then the same K steps once more with the library's per-kernel
    event timing switched on (for the per-kernel rooflines; the instrumented step time is reported next to the
    clean one so that the cost of the instrumentation is visible)."""
    lib = prb._lib.lib
    nranks = group.nranks
    step, tail = make_step(w)
    stream = prb.current_stream()

    def rank_max(x: float) -> float:
        if nranks == 1:
            return x
        t = prb.Tensor.from_numpy(np.array([x], np.float32))
        group.all_reduce(t, "max")
        return float(t.numpy()[0])

    sync_t = prb.Tensor.zeros(1, "float32")

    def device_barrier():
        """A one-element allreduce on the calc stream, queued right before a start event: the host barrier +
        synchronize the contract asks for leaves the ranks' hosts up to a millisecond apart, and max-over-ranks
        would charge that skew to the steps (the early rank waits for the late one inside its first allreduce)."""
        if nranks > 1:
            group.all_reduce(sync_t)

    for _ in range(max(warmup, 3)):   # also builds every cached plan and arms the producer-side pass 1
        step()
    stream.synchronize()

    def timed(k):
        group.barrier()
        prb.synchronize()
        ev0, ev1 = prb.Event(), prb.Event()
        device_barrier()
        ev0.record()
        for _ in range(k):
            step()
        ev1.record()
        ev1.synchronize()
        group.barrier()
        prb.synchronize()
        return rank_max(ev0.elapsed_ms(ev1)) / k

    sampler = ClockSampler(group.device) if sample_clocks else None
    launches0 = lib.prb_launch_count()
    if sampler:
        sampler.start()
    ms_per_step = timed(steps)
    clocks = sampler.stop() if sampler else None
    launches = lib.prb_launch_count() - launches0
    out = {"ms_per_step": ms_per_step, "value": w["elems"] * w["bpe"] * nranks / (ms_per_step * 1e-3) / 1e9,
           "bytes_per_elem": w["bpe"], "launches_per_step": launches / steps, "clocks": clocks, "gpu_launches": int(launches),
           "rank_max": rank_max, "device_barrier": device_barrier}
    if instrumented:
        prb._lib.call("prb_kernel_timing", 1)
        out["ms_per_step_instrumented"] = timed(steps)
        prb._lib.call("prb_kernel_timing", 0)
        out["kernels"] = read_kernel_timing(prb, w, nranks, steps, peak)
    return out


def nested_result(m, peak, nranks):
    """The part of a measurement that goes under `configs` of the JSON line."""
    r = {"ms_per_step": m["ms_per_step"], "value": m["value"], "bytes_per_elem": m["bytes_per_elem"],
         "pct_of_measured_hbm_peak_per_gpu": 100.0 * m["value"] / nranks / peak,
         "launches_per_step": m["launches_per_step"]}
    if "kernels" in m:
        r["kernels"] = m["kernels"]
        adam = m["kernels"].get("adam_chunks_kernel")
        if adam:
            r["roofline"] = {"bound": "hbm", "kernel": "adam_chunks_kernel", "achieved": adam["achieved_gbs"],
                             "peak": peak, "unit": "GB/s", "frac": adam["achieved_gbs"] / peak, "traffic": None}
    return r


# --------------------------------------------------------------------------------------------
# CPU baseline: the reference's CPU path restated in oracle/ (kind "port"), all host threads
# --------------------------------------------------------------------------------------------
def cpu_reference_run(workload, nranks, budget_s, threads, min_steps=2, max_steps=10 ** 9, warmup=1, dist=None, rank=0):
    """Per bucket: concat -> scale(1/nranks) -> allreduce (gloo, when nranks > 1) -> split; then AdamW per tensor.
    ref: reducer.cc:1037-1070 on CPUPlace, process_group_gloo.cc:279-343, cpu/adamw_kernel.cc:96-189.  fp32 only
    (the CPU kernel registers float/double only), so bf16 workloads are timed with fp32 storage of the same shapes;
    AMP unscale and the clip are left out (the CPU arm would only get slower)."""
    from oracle import oracle as orc
    shapes_name, pdt, gdt, amp, clip, mp, desc = workload_spec(workload)
    bpe = bytes_per_elem(pdt, gdt, amp, clip, False)
    numels = bench_configs.param_numels(shapes_name)
    rng = np.random.default_rng(1234 + rank)
    P = [rng.standard_normal(n, dtype=np.float32) * np.float32(0.02) for n in numels]
    G = [rng.standard_normal(n, dtype=np.float32) * np.float32(1e-2) for n in numels]
    M1 = [rng.standard_normal(n, dtype=np.float32) * np.float32(1e-3) for n in numels]
    M2 = [rng.random(n, dtype=np.float32) * np.float32(1e-3) for n in numels]
    b1p, b2p = np.float32(0.9 ** 10), np.float32(0.999 ** 10)
    elems = int(sum(numels))
    # the reference's buckets: 1 MiB for the last parameters, 25 MiB otherwise, launched in reverse order
    # (grouped by the workload's OWN gradient dtype, so that both arms run the same buckets; the CPU stores them as fp32)
    gcode = orc.F32 if gdt == "float32" else orc.BF16
    groups = orc.assign_group_by_size(numels, [gcode] * len(numels), [False] * len(numels), [1 << 20, 25 << 20])
    groups = list(reversed(groups))
    buckets = [np.empty(sum(numels[v] for v in g), np.float32) for g in groups]  # a caching allocator reuses them
    grads = [np.empty(n, np.float32) for n in numels]
    torch = None
    if dist is not None:
        import torch

    def one_step():
        for g, bucket in zip(groups, buckets):
            orc.concat([G[v] for v in g], out=bucket)                  # memcpy-loop concat
            orc.cpu_scale_f32(bucket, np.float32(1.0 / nranks), threads)
            if dist is not None:
                dist.all_reduce(torch.from_numpy(bucket))              # gloo, in place, sum
            orc.split(bucket, [numels[v] for v in g], outs=[grads[v] for v in g])   # memcpy-loop split
        for p, g, m1, m2 in zip(P, grads, M1, M2):
            orc.cpu_adamw(p, g, m1, m2, lr=1e-3, lr_ratio=1.0, beta1=0.9, beta2=0.999, epsilon=1e-8, coeff=0.01,
                          beta1_pow=b1p, beta2_pow=b2p, nthreads=threads)

    for _ in range(max(1, warmup)):  # untimed warm-up (page faults, thread pool, gloo connections)
        one_step()
    if dist is not None:
        dist.barrier()
    t0 = time.perf_counter()
    steps = 0
    while steps < max_steps:
        go = steps < min_steps or time.perf_counter() - t0 < budget_s
        if dist is not None:  # every rank must take the same decision
            flag = torch.tensor([1 if go else 0])
            dist.broadcast(flag, 0)
            go = bool(flag.item())
        if not go:
            break
        one_step()
        steps += 1
    if dist is not None:
        dist.barrier()
    dt = time.perf_counter() - t0
    return {"gbs": elems * bpe * nranks * steps / dt / 1e9, "ms_per_step": dt / steps * 1e3, "steps": steps,
            "threads": threads, "elems": elems, "seconds": dt, "bpe": bpe, "desc": desc, "buckets": len(groups)}


def reference_worker(args):
    """One of N local CPU processes of `--impl reference --gpus N` (N > 1): gloo over 127.0.0.1."""
    rank, world = int(os.environ["PRB_REF_RANK"]), int(os.environ["PRB_REF_WORLD"])
    threads = int(os.environ["PRB_REF_THREADS"])
    import torch
    import torch.distributed as dist
    torch.set_num_threads(1)
    import datetime
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % os.environ["PRB_REF_PORT"], rank=rank,
                            world_size=world, timeout=datetime.timedelta(seconds=120))
    r = cpu_reference_run(args.workload, world, budget_s=120.0, threads=threads, min_steps=1, max_steps=args.steps,
                          warmup=args.warmup, dist=dist, rank=rank)
    if rank == 0:
        print("PRB_REF_RESULT " + json.dumps(r), flush=True)
    dist.destroy_process_group()
    return 0


def config_block(desc, elems, tensors, buckets, pdt, gdt, amp, clip, bpe, nranks):
    """The `config` object: identical (keys and values) in both arms for the same workload and N -- the driver compares
    them.  Arm-specific facts (which allreduce, percentages of peak, the CPU path) go to `details` instead."""
    return {"workload": desc, "elements_per_gpu": elems, "tensors": tensors, "buckets": buckets, "bucket_mb": [25, 1],
            "param_dtype": pdt, "grad_dtype": gdt, "amp": amp, "clip": clip, "algorithmic_bytes_per_elem": bpe,
            "parallelism": "dp%d" % nranks,
            "l2": "working set %.0f MB per step > 126 MB L2 (inputs larger than L2, no flush)" % (elems * (bpe + 8) / 1e6)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # only rank 0 runs the CPU arm
    # each "step" of this arm = one full CPU pass over the workload: W untimed warm-up steps, then K timed ones; a time
    # budget cuts K short where K full CPU passes would take more than two minutes (the line reports the steps timed)
    cores = host_cores()
    n = args.gpus
    if n == 1:
        r = cpu_reference_run(args.workload, 1, budget_s=120.0, threads=cores, min_steps=1, max_steps=args.steps,
                              warmup=args.warmup)
        allreduce = "none (one process)"
    else:
        import socket
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
        s.close()
        per = max(1, cores // n)
        procs = []
        for q in range(n):
            # a clean environment: under torch.distributed.run the workers would otherwise inherit
            # TORCHELASTIC_USE_AGENT_STORE and try to JOIN the launcher's store instead of creating their own
            drop = ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "OMP_NUM_THREADS", "GROUP_RANK",
                    "GROUP_WORLD_SIZE", "LOCAL_WORLD_SIZE", "ROLE_RANK", "ROLE_NAME", "ROLE_WORLD_SIZE")
            env = {k: v for k, v in os.environ.items() if k not in drop and not k.startswith("TORCHELASTIC_")}
            env.update({"PRB_REF_RANK": str(q), "PRB_REF_WORLD": str(n), "PRB_REF_PORT": str(port),
                        "PRB_REF_THREADS": str(per), "OMP_NUM_THREADS": str(per)})
            procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__), "--impl", "reference-worker",
                                           "--workload", args.workload, "--steps", str(args.steps), "--warmup",
                                           str(args.warmup), "--gpus", str(n)], env=env, stdout=subprocess.PIPE))
        try:
            outs = [p.communicate(timeout=600)[0].decode() for p in procs]
        except subprocess.TimeoutExpired:
            for p in procs:
                p.kill()
            print(json.dumps({"impl": "reference", "unavailable": "the gloo workers of the CPU arm did not finish in 600 s"}),
                  flush=True)
            return 0
        if any(p.returncode for p in procs):
            print(json.dumps({"impl": "reference", "unavailable": "a gloo worker of the CPU arm failed"}), flush=True)
            return 0
        line = [ln for ln in outs[0].splitlines() if ln.startswith("PRB_REF_RESULT ")][0]
        r = json.loads(line[len("PRB_REF_RESULT "):])
        r["threads"] = per * n
        allreduce = "gloo (torch wheel), %d local processes x %d threads" % (n, per)
    shapes_name, pdt, gdt, amp, clip, mp, desc = workload_spec(args.workload)
    cfg = config_block(desc, r["elems"], len(bench_configs.param_numels(shapes_name)), r["buckets"], pdt, gdt, amp, clip,
                       r["bpe"], n)
    details = {"allreduce": allreduce,
               "path": ("CPU: per bucket concat -> scale(1/nranks) -> allreduce -> split, then AdamW per tensor (oracle "
                        "port of the reference's CPU kernels in fp32 storage; Paddle itself cannot be built here)")}
    line = {
        "impl": "reference", "metric": METRIC, "value": r["gbs"], "unit": UNIT, "n_gpus": n, "steps": r["steps"],
        "warmup": max(1, args.warmup), "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "details": details,
        "cpu_baseline": {"value": r["gbs"], "unit": UNIT, "cores": r["threads"], "kind": "port",
                         "sample": "%d full steps of the workload in %.1f s" % (r["steps"], r["seconds"])},
        "e2e": {"value": r["gbs"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------
# N > 1 parity preamble (untimed, < 1 s): the GPU-test box has one GPU, so the multi-rank path proves itself here
# --------------------------------------------------------------------------------------------
def parity_preamble(prb, group, arena=False, save_dir=None, sharded=False):
    """A tiny AMP bf16 + master + clip job through the same API on all ranks: one clean step, one step with an Inf
    injected on ONE rank (every rank must skip and halve its loss scale), two more clean steps (even steps mark all
    gradients ready in one call, odd steps hook by hook).  After each step
    parameters, masters, moments and the scaler state must be bit-equal across ranks (checksums compared through
    the communicator); with one or two ranks (at most one addition, so the order cannot matter) they must also equal the oracle's
    composition of the reference's stages bit for bit.  Returns "ok" or "fail: <what>".
    sharded: the same job through GroupShardedOptimizerStage2 (reduce-scatter, update of the owned ranges only,
    gather of the parameters): the parameters must still be those bits on every rank."""
    from oracle import oracle as orc
    rank, n = group.rank, group.nranks
    sizes = [64, 3 * 3 * 64 * 16, 1000, 7, 4096, 4097, 250000, 1, 300000]
    try:
        params, host = [], []
        for i, sz in enumerate(sizes):
            p0 = orc.from_f32(np.random.default_rng(100 + i).normal(0, 0.02, sz).astype(np.float32), orc.BF16)
            p = prb.Tensor.from_numpy(p0, orc.BF16, name="pp_%d" % i)
            if not arena:
                p.grad = prb.Tensor.zeros(sz, orc.BF16)
            params.append(p)
            host.append({"param": p0.copy(), "mom1": np.zeros(sz, np.float32), "mom2": np.zeros(sz, np.float32),
                         "master": orc.to_f32(p0, orc.BF16)})
        clip = prb.nn.ClipGradByGlobalNorm(1.0)
        opt = prb.optimizer.AdamW(learning_rate=1e-3, parameters=params, weight_decay=0.01, grad_clip=clip,
                                  multi_precision=True)
        scaler = prb.amp.GradScaler(init_loss_scaling=2.0 ** 15, decr_every_n_nan_or_inf=1)
        dp = prb.DataParallel(params, comm_buffer_size=0.4, last_comm_buffer_size=0.01, consume_from_buckets=True,
                              grads_in_arena=arena)
        inner = opt
        if sharded:
            opt = prb.distributed.GroupShardedOptimizerStage2(params, inner, group, reduce_overlap=(sharded == "bucket"))
        tag = ("shb" if sharded == "bucket" else "shd") if sharded else "amp"
        scale, b1p, b2p = 2.0 ** 15, np.float32(0.9), np.float32(0.999)
        for step in range(4):
            per_rank = [[orc.from_f32((np.random.default_rng(1000 * r + 10 * step + i).normal(0, 1e-2, sz) * 2.0 ** 15 * 30)
                                      .astype(np.float32), orc.BF16) for i, sz in enumerate(sizes)] for r in range(n)]
            if step == 1:
                per_rank[n - 1][4][17] = orc.from_f32(np.array([np.inf], np.float32), orc.BF16)[0]   # one rank only
            for p, g in zip(params, per_rank[rank]):
                p.grad.copy_from_numpy(g)
            dp.prepare_for_backward()
            if step % 2 == 0:
                dp.mark_all_ready()                       # one call: one pack + one exchange for all buckets
            else:
                for i in reversed(range(len(params))):    # hooks one by one: one pack + one exchange per bucket
                    dp.mark_var_ready(i)
            scaler.step(opt)
            scaler.update()
            prb.current_stream().synchronize()
            if save_dir is not None:
                for i, p in enumerate(params):
                    np.save(os.path.join(save_dir, "%s%d_param_%d_%d_r%d.npy" % (tag, int(arena), step, i, rank)), p.numpy())
                    if not sharded:
                        np.save(os.path.join(save_dir, "amp%d_m2_%d_%d_r%d.npy" % (int(arena), step, i, rank)),
                                opt._accumulators[p.name]["moment2"].numpy())
            found = bool(scaler._found_inf.numpy()[0])
            if found != (step == 1):
                return "fail: step %d found_inf=%s on rank %d" % (step, found, rank)
            # cross-rank bit equality through a checksum
            h = 0
            for p in params:
                for t in ((p,) if sharded else (p, opt._master_weights[p.name], opt._accumulators[p.name]["moment1"],
                                                opt._accumulators[p.name]["moment2"])):
                    a = t.numpy()
                    h = (h * 1000003 + int(np.bitwise_xor.reduce(a.view(np.uint16 if a.dtype.itemsize == 2 else np.uint32)
                                                                 .astype(np.uint64) * np.arange(1, a.size + 1, dtype=np.uint64)))) % (1 << 52)
            h = (h * 1000003 + int(scaler._scale.numpy().view(np.uint32)[0])) % (1 << 52)
            t = prb.Tensor.from_numpy(np.array([h, -h], np.float64))
            group.all_reduce(t, "max")
            mx, mn = t.numpy()
            if mx != -mn:
                return "fail: step %d state differs across ranks" % step
            if n <= 2 and dp._reducer.allreduce_backend != "nvls":   # (the NVSwitch's own adder is not bit-pinned)
                red = []
                for i in range(len(sizes)):
                    s = sum(orc.to_f32(orc.scale(per_rank[r][i], orc.BF16, 1.0 / n), orc.BF16).astype(np.float64) for r in range(n))
                    red.append(orc.from_f32(s.astype(np.float32), orc.BF16))
                if orc.check_finite_and_unscale(red, orc.BF16, scale) != found:
                    return "fail: step %d oracle found_inf differs" % step
                if found:
                    scale /= 2
                else:
                    c = float(clip.last_clip.numpy()[0])
                    _, c_or = orc.global_norm_clip([g.copy() for g in red], [orc.BF16] * len(red), 1.0)
                    if abs(c - c_or) > 4e-3 * abs(c_or):
                        return "fail: step %d clip factor %g vs oracle %g" % (step, c, c_or)
                    for hst, g, p in zip(host, red, params):
                        orc.clip_mul(g, orc.BF16, c)
                        orc.adam(hst["param"], g, hst["mom1"], hst["mom2"], hst["master"], dtype_param=orc.BF16,
                                 dtype_grad=orc.BF16, lr=1e-3, beta1=0.9, beta2=0.999, epsilon=1e-8, coeff=0.01,
                                 use_decay=True, beta1_pow=b1p, beta2_pow=b2p)
                        if p.numpy().tobytes() != hst["param"].tobytes():
                            return "fail: step %d %s differs from the oracle" % (step, p.name)
                        if sharded and p.name in opt._param2range:   # state exists for the owned range only
                            off, ln = opt._param2range[p.name]
                            if inner._master_weights[p.name].numpy().tobytes() != hst["master"][off:off + ln].tobytes() or \
                                    inner._accumulators[p.name]["moment2"].numpy().tobytes() != hst["mom2"][off:off + ln].tobytes():
                                return "fail: step %d %s: sharded state differs from the oracle" % (step, p.name)
                        elif not sharded and opt._master_weights[p.name].numpy().tobytes() != hst["master"].tobytes():
                            return "fail: step %d %s differs from the oracle" % (step, p.name)
                    b1p, b2p = np.float32(0.9) * b1p, np.float32(0.999) * b2p
        if float(scaler._scale.numpy()[0]) != 2.0 ** 14:
            return "fail: loss scale %g" % float(scaler._scale.numpy()[0])
        if dp._reducer.comm_error():
            return "fail: exchange kernel reported an error"
        if sharded and n > 1 and dp._reducer.allreduce_backend == "nvlink" and not dp._sharded_armed:
            return "fail: the reduce-scatter was never switched on"
        return "ok"
    except Exception as e:  # noqa: BLE001
        return "fail: %s: %s" % (type(e).__name__, str(e)[:200])


# --------------------------------------------------------------------------------------------
# e2e: gradients arrive from pinned host memory every step, the step's metric is read back every step
# --------------------------------------------------------------------------------------------
def measure_e2e(prb, group, w, steps, rank_max):
    """Public API, host buffers: per bucket (launch order) one H2D copy of that bucket's gradients on a copy stream,
    an event, and mark_vars_ready for the bucket's parameters on the calc stream -- so the copy of bucket k+1 runs
    under the pack of bucket k, and the step costs max(copy, kernels) + the update, not their sum.  Read back: what
    a training loop logs every step -- global gradient norm, found_inf, loss scale (9 bytes)."""
    dp, opt, scaler = w["dp"], w["opt"], w["scaler"]
    stream = prb.current_stream()
    copy_stream = prb.Stream()
    groups = list(reversed(dp.group_indices))   # launch order
    gsz = w["gsz"]
    plan = []
    for g in groups:
        lo, hi = min(g), max(g)
        assert sorted(g) == list(range(lo, hi + 1))   # a bucket is a contiguous parameter range
        off0 = w["offs"][lo] * gsz
        off1 = (w["offs"][hi] + w["numels"][hi]) * gsz
        plan.append((off0, off1 - off0, (C.c_int64 * len(g))(*sorted(g, reverse=True)), prb.Event()))
    h2d = sum(p[1] for p in plan)
    res_host = prb.PinnedBuffer(4, np.float32)
    reads = []
    if w["clip_obj"] is not None:
        reads.append((w["clip_obj"].last_global_norm.data_ptr, 4))
    if scaler is not None:
        reads += [(scaler._found_inf.data_ptr, 1), (scaler._scale.data_ptr, 4)]
    if not reads:
        reads = [(opt._beta1_pow.data_ptr, 4), (opt._beta2_pow.data_ptr, 4)]
    d2h = sum(n for _, n in reads)
    done = prb.Event()
    _, tail = make_step(w)

    def e2e_step():
        copy_stream.wait_event(done)   # the previous step's packs have consumed the gradient arena
        dp.prepare_for_backward()
        for off, nbytes, vars_, ev in plan:
            prb._lib.call("prb_memcpy_h2d", w["grad_arena"].data_ptr + off, w["host_arena"].ptr + off, nbytes,
                          copy_stream.handle)
            ev.record(copy_stream)
            stream.wait_event(ev)
            dp._reducer.mark_vars_ready(vars_)
        done.record(stream)
        tail()
        at = 0
        for ptr, n in reads:
            prb._lib.call("prb_memcpy_d2h", res_host.ptr + at, ptr, n, stream.handle)
            at += 4
        stream.synchronize()  # the caller consumes the result on the host

    done.record(stream)
    for _ in range(3):
        e2e_step()
    group.barrier()
    prb.synchronize()
    k = max(3, min(steps, 20))
    t0 = time.perf_counter()
    for _ in range(k):
        e2e_step()
    prb.synchronize()
    ms = rank_max((time.perf_counter() - t0) * 1e3) / k
    return {"value": w["elems"] * w["bpe"] * group.nranks / (ms * 1e-3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d,
            "d2h_bytes_per_step": d2h, "ms_per_step": ms, "steps": k,
            "read_back": "global grad norm + found_inf + loss scale (what a training loop logs per step)",
            "h2d": "one copy per bucket on a copy stream, overlapped with the packs of earlier buckets"}


# --------------------------------------------------------------------------------------------
# allreduce / backward overlap (N > 1)
# --------------------------------------------------------------------------------------------
def measure_overlap(prb, w, group, rank_max, device_barrier, steps=40):
    """A synthetic backward (spin kernels on the calc stream between bucket-ready points, total T_bwd) produces the
    buckets over time; exposed comm = step time with the real communicator minus step time of the SAME schedule on
    a single-rank communicator (no exchange); overlap = 1 - exposed / total_comm, total_comm = the buckets'
    exchanges alone, back to back."""
    stream = prb.current_stream()
    dp = w["dp"]
    _, tail = make_step(w)
    groups = list(reversed(dp.group_indices))  # launch order
    idx = [(C.c_int64 * len(g))(*sorted(g, reverse=True)) for g in groups]
    red = dp._reducer
    ng = len(groups)
    cs = group.comm_stream
    prb.synchronize()
    for _ in range(3):
        for k in range(ng):
            red.allreduce_group(k, cs)
    cs.synchronize()
    lineup = prb.Tensor.zeros(1, "float32")
    group.barrier()
    e0, e1 = prb.Event(), prb.Event()
    group.all_reduce(lineup, stream=cs)   # device-side line-up, as in device_barrier()
    e0.record(cs)
    for _ in range(steps):
        for k in range(ng):
            red.allreduce_group(k, cs)
    e1.record(cs)
    e1.synchronize()
    total_comm = rank_max(e0.elapsed_ms(e1)) / steps
    # the LAST bucket of a backward has nothing left to hide under (its gradients are the last ones produced, and the
    # update cannot start before it is reduced): its exchange bounds the overlap any schedule can reach
    group.barrier()
    group.all_reduce(lineup, stream=cs)
    e0.record(cs)
    for _ in range(steps):
        red.allreduce_group(ng - 1, cs)
    e1.record(cs)
    e1.synchronize()
    last_comm = rank_max(e0.elapsed_ms(e1)) / steps

    host_ms = []

    def run(dp_x, t_bwd_us):
        per = int(t_bwd_us / len(groups))

        def one():
            dp_x.prepare_for_backward()
            for arr in idx:
                prb._lib.call("prb_spin", per, stream.handle)       # backward computing this bucket's gradients
                dp_x._reducer.mark_vars_ready(arr)
            tail()
        for _ in range(4):
            one()
        group.barrier()
        prb.synchronize()
        a, b = prb.Event(), prb.Event()
        t0 = time.perf_counter()
        device_barrier()
        a.record()
        for _ in range(steps):
            one()
        b.record()
        host_ms.append((time.perf_counter() - t0) * 1e3 / steps)  # the host's share: time to ISSUE a step
        b.synchronize()
        return rank_max(a.elapsed_ms(b)) / steps

    # the same schedule without communication: a single-rank communicator over the same gradients
    h = C.c_void_p()
    prb._lib.call("prb_comm_init_single", group.device, C.byref(h))
    solo_group = prb.distributed.Group(h.value, 0, 1, group.device)
    dp_solo = prb.DataParallel(w["params"], comm_buffer_size=25, last_comm_buffer_size=1, group=solo_group)
    out = {"total_comm_ms": total_comm, "last_bucket_comm_ms": last_comm, "buckets": ng,
           "overlap_cap_estimate": 1.0 - last_comm / total_comm, "cases": []}
    for mult in (1.0, 2.0, 4.0):
        t_bwd_us = mult * total_comm * 1e3
        dp.consume_from_buckets(True)
        t_comm = run(dp, t_bwd_us)
        dp_solo.consume_from_buckets(True)      # re-points the parameters at the solo arena
        dp_solo.pipelined = True                # per-bucket optimizer launches where the real schedule has them
        t_same = run(dp_solo, t_bwd_us)
        exposed = max(0.0, t_comm - t_same)
        out["cases"].append({"t_bwd_ms": t_bwd_us / 1e3, "step_ms": t_comm, "step_no_comm_ms": t_same,
                             "exposed_comm_ms": exposed, "overlap_frac": 1.0 - exposed / total_comm,
                             "host_issue_ms": host_ms[-2]})
    dp.consume_from_buckets(True)
    out["overlap_frac"] = out["cases"][1]["overlap_frac"]
    out["overlap_frac_of_cap"] = out["overlap_frac"] / out["overlap_cap_estimate"] if out["overlap_cap_estimate"] > 0 else None
    out["definition"] = ("1 - exposed_comm/total_comm; exposed_comm = step - the same schedule (per-bucket pack, same "
                         "optimizer launches) on a single-rank communicator; synthetic backward = spin kernels between "
                         "bucket-ready points; start events lined up by a device-side allreduce; headline at "
                         "T_bwd = 2 x total_comm.  overlap_cap_estimate = 1 - (exchange of the last bucket)/total_comm: the last "
                         "bucket's exchange can hide under nothing (BERT-base: the 47 MB word-embedding gradient is a "
                         "bucket of its own by the reference's assignment rule, reducer.cc:45-156)")
    return out


def measure_exposed_tail(prb, group, w, t_bwd_us, rank_max, device_barrier, steps=20):
    """What a training job pays AFTER its backward pass ends: a synthetic backward (spin kernels on the calc stream,
    t_bwd_us in total, one slice per bucket) marks the buckets ready one by one, then the optimizer steps; returned
    is step time minus the backward time.  The replicated step hides all exchanges but the last under the backward;
    the sharded step hides its reduce-scatters only with the per-bucket layout."""
    stream = prb.current_stream()
    dp = w["dp"]
    _, tail = make_step(w)
    groups = list(reversed(dp.group_indices))  # launch order
    idx = [(C.c_int64 * len(g))(*sorted(g, reverse=True)) for g in groups]
    per = max(1, int(t_bwd_us / len(groups)))

    def one():
        dp.prepare_for_backward()
        for arr in idx:
            prb._lib.call("prb_spin", per, stream.handle)
            dp._reducer.mark_vars_ready(arr)
        tail()
    for _ in range(4):
        one()
    group.barrier()
    prb.synchronize()
    a, b = prb.Event(), prb.Event()
    device_barrier()
    a.record()
    for _ in range(steps):
        one()
    b.record()
    b.synchronize()
    step_ms = rank_max(a.elapsed_ms(b)) / steps
    return {"t_bwd_ms": per * len(groups) / 1e3, "step_ms": step_ms, "exposed_tail_ms": step_ms - per * len(groups) / 1e3}


def measure_sparse_adam(prb, peak, iters=20):
    """adam_ with a SelectedRows gradient on BERT-base's word-embedding table (30522 x 768, fp32): 8192 looked-up rows
    per step with duplicates (a batch of token ids), lazy_mode.  One call = host sort + dedup of the indices, one
    H2D copy of the index tables, the merge kernel, the update kernel, the beta-pow kernel; timed with CUDA events
    around `iters` calls (so the host part counts when it is slower than the device part).  Algorithmic bytes:
    merge r nrows x w x 4 + w u x w x 4, update 28 B per element of the u unique rows."""
    height, width, nrows = 30522, 768, 8192
    rng = np.random.default_rng(99)
    T = prb.Tensor
    p = T.from_numpy(rng.normal(0, 0.02, height * width).astype(np.float32))
    m1, m2 = T.zeros(height * width, "float32"), T.zeros(height * width, "float32")
    lr = T.from_numpy(np.array([1e-3], np.float32))
    b1p, b2p = T.from_numpy(np.array([0.9], np.float32)), T.from_numpy(np.array([0.999], np.float32))
    rows = rng.integers(0, height, nrows)
    uniq = int(np.unique(rows).size)
    g = prb.SelectedRows(rows, height, T.from_numpy(rng.normal(0, 1e-2, nrows * width).astype(np.float32)))

    def call():
        prb._C_ops.adam_(p, g, lr, m1, m2, b1p, b2p, None, None, 0.9, 0.999, 1e-8, True, 1000, False, False)

    for _ in range(3):
        call()
    prb.current_stream().synchronize()
    e0, e1 = prb.Event(), prb.Event()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(iters):
        call()
    e1.record()
    host_ms = (time.perf_counter() - t0) * 1e3 / iters
    e1.synchronize()
    ms = e0.elapsed_ms(e1) / iters
    nbytes = nrows * width * 4 + uniq * width * 4 + 28 * uniq * width
    return {"workload": "adam_ (SelectedRows grad, lazy_mode) on a 30522 x 768 fp32 embedding table, %d rows per step, "
                        "%d unique" % (nrows, uniq),
            "ms_per_step": ms, "host_issue_ms_per_step": host_ms, "algorithmic_bytes_per_step": nbytes,
            "value": nbytes / (ms * 1e-3) / 1e9, "pct_of_measured_hbm_peak_per_gpu": 100.0 * nbytes / (ms * 1e-3) / 1e9 / peak,
            "launches_per_step": 3.0, "steps": iters,
            "dense_equivalent_bytes": 28 * height * width}


# --------------------------------------------------------------------------------------------
# the GPU arm
# --------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import gc

    import paddle_ray_b200 as prb  # fails loudly if the CUDA library is missing

    if prb.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    # NCCL prints its version banner on stdout; the driver wants ONE JSON line there
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        group = prb.distributed.init_parallel_env()
        if group.nranks > 1:
            group.barrier()   # the first collective is where the banner appears
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)
    rank, nranks = group.rank, group.nranks
    if nranks != args.gpus:
        raise SystemExit("bench.py: --gpus %d but WORLD_SIZE=%d (launch N>1 through torch.distributed.run)" % (args.gpus, nranks))
    peak, peak_kind = measured_peak()
    t_start = time.perf_counter()

    parity = parity_preamble(prb, group) if nranks > 1 else None
    parity_sharded = parity_preamble(prb, group, sharded=True) if nranks > 1 else None
    parity_sharded_pb = parity_preamble(prb, group, sharded="bucket") if nranks > 1 else None
    gc.collect()

    # ---- headline workload
    w = build_workload(prb, args.workload, rank, host_mirror=True)
    m = measure_workload(prb, group, w, args.steps, args.warmup, peak, sample_clocks=True)
    rank_max, device_barrier = m["rank_max"], m["device_barrier"]
    step, _ = make_step(w)

    # host cost of issuing a step (diagnostic): 10 steps issued into an empty queue, clock stopped before the sync
    prb.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        step()
    host_issue_ms = (time.perf_counter() - t0) * 1e3 / 10
    prb.synchronize()

    # roofline of the dominant kernel: algorithmic bytes per launch / its average launch duration (CUDA events
    # recorded by the library around every launch of the instrumented K steps)
    kern = m["kernels"]
    adam = kern["adam_chunks_kernel"]
    per_launch_ms = adam["ms_per_step"] / adam["launches_per_step"]
    per_launch_bytes = adam["algorithmic_bytes_per_step"] / adam["launches_per_step"]
    achieved = per_launch_bytes / (per_launch_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "kernel": "adam_chunks_kernel",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peak_kind == "measured" else "profiling-recipe fallback",
                "algorithmic_bytes_per_launch": per_launch_bytes, "avg_launch_ms": per_launch_ms,
                "frac_of_nominal_8TBps": achieved / 8000.0,
                "share_of_step": adam["ms_per_step"] / m["ms_per_step_instrumented"]}
    try:  # dram bytes per launch from the committed ncu capture of THIS workload, when there is one
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            roofline["traffic"] = json.load(f).get(args.workload, {}).get("adam_chunks_kernel")
    except Exception:
        pass

    e2e = measure_e2e(prb, group, w, args.steps, rank_max)

    overlap = None
    t_bwd_tail_us = 0.0
    if nranks > 1 and not args.no_overlap:
        overlap = measure_overlap(prb, w, group, rank_max, device_barrier)
        # the same synthetic backward (2 x the exchanges' total) in front of the replicated and the sharded steps
        t_bwd_tail_us = overlap["cases"][1]["t_bwd_ms"] * 1e3
        try:
            overlap["after_backward"] = measure_exposed_tail(prb, group, w, t_bwd_tail_us, rank_max, device_barrier)
        except Exception as e:  # noqa: BLE001  (an extra: never lose the line over it)
            overlap["after_backward"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:160])}

    # CPU baseline (rank 0, N = 1 only): bounded sample of the same workload on the host cores
    cpu_baseline = None
    if rank == 0 and nranks == 1 and not args.no_cpu_baseline:
        r = cpu_reference_run(args.workload, 1, budget_s=12.0, threads=host_cores())
        cpu_baseline = {"value": r["gbs"], "unit": UNIT, "cores": r["threads"], "kind": "port",
                        "ms_per_step": r["ms_per_step"],
                        "sample": "%d full steps of the same workload (%d elems, fp32) in %.1f s" % (r["steps"], r["elems"], r["seconds"])}

    config = config_block(w["desc"], w["elems"], len(w["numels"]), w["buckets"], w["pdt"], w["gdt"], w["amp"], w["clip"],
                          w["bpe"], nranks)
    details = {
        "allreduce": w["dp"]._reducer.allreduce_backend,
        "pct_of_measured_hbm_peak_per_gpu": 100.0 * m["value"] / nranks / peak,
        "pct_of_nominal_8TBps_per_gpu": 100.0 * m["value"] / nranks / 8000.0,
        "value_at_survey_34B_floor": (m["value"] * 34 / w["bpe"]) if (w["amp"] and w["gsz"] == 2 and not w["arena"]) else None,
        "stages": "pack+scale+inf-check+norm-partials (1 kernel) -> exchange (N>1) -> norm finalize -> AdamW (1 kernel)",
    }
    del w, step, m["rank_max"], m["device_barrier"]
    gc.collect()

    # ---- the other BASELINE configs, few steps each, nested in the same line
    nested = {}
    if not args.no_configs:
        ksteps = max(5, min(args.steps, 20))
        todo = [(n, False, False) for n in ("resnet50", "gpt2_medium", "bert_base") if n != args.workload]
        if nranks > 1:
            # the sharded step (GroupShardedOptimizerStage2): same parameters bit for bit, update pass 1/N the size
            todo = [(args.workload, False, True), (args.workload, False, "bucket")] + todo + \
                [("gpt2_medium", False, True), ("resnet50", False, True)]
        todo += [(args.workload, True, False), ("resnet50", True, False)]   # gradients produced in the arena (no pack)
        if nranks == 1:
            todo += [(n, False, False) for n in FLAT_SWEEP]
        else:
            todo += [("flat26x1", False, False), ("flat26x1:bf16", False, False)]
        for name, arena, shd in todo:
            label = name + ("+arena" if arena else "") + (("+sharded_per_bucket" if shd == "bucket" else "+sharded") if shd else "")
            if time.perf_counter() - t_start > args.config_budget_s:
                nested["_skipped_for_time"] = nested.get("_skipped_for_time", []) + [label]
                continue
            try:
                wn = build_workload(prb, name, rank, arena=arena, sharded=shd)
                mn = measure_workload(prb, group, wn, ksteps, 3, peak)
                r = nested_result(mn, peak, nranks)
                if nranks > 1 and name == "resnet50" and not arena and not shd and not args.no_overlap:
                    # round 1's overlap workload (5 buckets, the last one small), for comparison with its figures
                    r["overlap"] = measure_overlap(prb, wn, group, mn["rank_max"], mn["device_barrier"], steps=20)
                if nranks > 1 and name == args.workload and shd and t_bwd_tail_us:
                    r["after_backward"] = measure_exposed_tail(prb, group, wn, t_bwd_tail_us, mn["rank_max"],
                                                               mn["device_barrier"])
                r["workload"] = wn["desc"] + (" [gradients produced in the bucket arena: no pack]" if arena else "") + \
                    (" [sharded step: reduce-scatter -> update of the owned 1/%d -> parameter gather%s]"
                     % (nranks, "; ownership and reduce-scatter per bucket" if shd == "bucket" else "") if shd else "")
                r["steps"] = ksteps
                nested[label] = r
                del wn, mn
            except Exception as e:  # noqa: BLE001  (e.g. out of memory on a smaller GPU): report, keep the line
                nested[label] = {"error": "%s: %s" % (type(e).__name__, str(e)[:160])}
            gc.collect()

    if not args.no_configs and nranks == 1:
        try:
            nested["sparse_adam:bert_embedding"] = measure_sparse_adam(prb, peak)
        except Exception as e:  # noqa: BLE001
            nested["sparse_adam:bert_embedding"] = {"error": "%s: %s" % (type(e).__name__, str(e)[:160])}

    if rank == 0:
        line = {
            "metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": nranks, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config, "details": details,
            "clocks": m["clocks"], "e2e": e2e,
            "gpu_launches": m["gpu_launches"], "host_issue_ms_per_step": host_issue_ms, "roofline": roofline,
            "kernels": kern, "ms_per_step_instrumented": m["ms_per_step_instrumented"],
        }
        if cpu_baseline is not None:
            line["cpu_baseline"] = cpu_baseline
        if overlap is not None:
            line["overlap"] = overlap
        if parity is not None:
            line["parity_n"] = parity
            line["parity_n_sharded"] = parity_sharded
            line["parity_n_sharded_per_bucket"] = parity_sharded_pb
        if nested:
            line["configs"] = nested
        print(json.dumps(line), flush=True)
    group.barrier()
    return 0


def self_launch(args):
    """`python bench.py --gpus N` without a launcher: start N ranks of this script on 127.0.0.1."""
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(args.gpus):
        env = {**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(args.gpus),
               "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "PRB_SELF_LAUNCHED": "1"}
        procs.append(subprocess.Popen([sys.executable, os.path.abspath(__file__)] + sys.argv[1:], env=env))
    rc = 0
    for p in procs:
        rc = rc or p.wait()
    return rc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "reference-worker"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the nested measurements of the other configs")
    ap.add_argument("--config-budget-s", type=float, default=150.0,
                    help="no further nested config is started once the run is this many seconds old")
    args = ap.parse_args()
    if args.impl == "reference-worker":
        return reference_worker(args)
    if args.impl == "reference":
        return run_reference_arm(args)
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        return self_launch(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
