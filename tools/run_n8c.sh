#!/bin/bash
# 8-GPU box: multi-rank parity at N = 8 (library exchange, incl. sparse groups and the sharded step), bench lines at N = 8 and 4
export PRB_NVL_TIMEOUT_MS=${PRB_NVL_TIMEOUT_MS:-20000}
out=gpurun_out/${1:-n8c}
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q -k "8-nvlink" > $out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_multi.log
for n in 8 4; do
  timeout 300 python bench.py --gpus $n --steps 50 > $out/bench_n$n.json 2> $out/bench_n$n.err; echo "bench N=$n rc=$?"; tail -c 300 $out/bench_n$n.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$out/bench_n$n.json") if l.startswith("{")][-1])
    print("N=$n", d["ms_per_step"], d["value"], d.get("parity_n"), d.get("parity_n_sharded"), {k:round(v["ms_per_step"],4) for k,v in d["kernels"].items()})
    o=d["overlap"]; print("  overlap", o["total_comm_ms"], o["overlap_cap_estimate"], [(round(c["t_bwd_ms"],3),round(c["step_ms"],3),round(c["step_no_comm_ms"],3),round(c["overlap_frac"],3)) for c in o["cases"]])
    for k,v in d.get("configs",{}).items():
        if isinstance(v, dict):
            print("  ",k, v.get("ms_per_step"), v.get("value"), v.get("error"), {kk:round(vv["ms_per_step"],4) for kk,vv in v.get("kernels",{}).items()})
            if "overlap" in v: print("     overlap", v["overlap"]["total_comm_ms"], [round(c["overlap_frac"],3) for c in v["overlap"]["cases"]])
        else: print("  ", k, v)
except Exception as e:
    print("no bench line:", e)
PY
done
