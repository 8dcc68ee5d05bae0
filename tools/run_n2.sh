#!/bin/bash
# Multi-GPU check: multi-rank parity tests, then the bench line (run under gpurun --gpus N; N = $2, default 2)
export PRB_NVL_TIMEOUT_MS=${PRB_NVL_TIMEOUT_MS:-20000}
out=gpurun_out/${1:-n2}
N=${2:-2}
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > $out/pytest_multi.log 2>&1
tail -25 $out/pytest_multi.log
timeout 400 python bench.py --gpus $N --steps 50 > $out/bench_n$N.json 2> $out/bench_n$N.err
tail -c 1500 $out/bench_n$N.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("$out/bench_n$N.json") if l.startswith("{")][-1])
except Exception as e:
    print("no bench line:", e); raise SystemExit
print(d["ms_per_step"], d["value"], d["ms_per_step_instrumented"], d["gpu_launches"], d.get("parity_n"), d.get("parity_n_sharded"), d.get("parity_n_sharded_per_bucket")); print("after_backward", d["overlap"].get("after_backward"), {k:v.get("after_backward") for k,v in d.get("configs",{}).items() if isinstance(v,dict) and "after_backward" in v})
print({k:(round(v["ms_per_step"],4), v["launches_per_step"]) for k,v in d["kernels"].items()})
print(d["e2e"]["ms_per_step"])
print(d["overlap"]["total_comm_ms"], [(round(c["t_bwd_ms"],3),round(c["step_ms"],3),round(c["step_no_comm_ms"],3),round(c["overlap_frac"],3)) for c in d["overlap"]["cases"]])
for k,v in d.get("configs",{}).items(): print(k, {a:b for a,b in v.items() if a not in ("kernels","workload","roofline","overlap")}, {kk:round(vv["ms_per_step"],4) for kk,vv in v.get("kernels",{}).items()})
PY
