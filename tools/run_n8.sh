#!/bin/bash
# One 8-GPU box: multi-rank parity at N = 8 / 4, the bench lines at N = 8 and 4, the exchange sweep (grid cap x
# memory-synchronisation domain) with a 730 us synthetic backward, the exchange alone per bucket size.
export PRB_NVL_TIMEOUT_MS=${PRB_NVL_TIMEOUT_MS:-20000}
out=gpurun_out/${1:-n8}
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_multi.py -x -q -k "8-nvlink or 4-nvlink or 8-nccl or 8-nvls" > $out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest_multi.log
for n in 8 4; do
  timeout 300 python bench.py --gpus $n --steps 50 > $out/bench_n$n.json 2> $out/bench_n$n.err; echo "bench N=$n rc=$?"; tail -c 400 $out/bench_n$n.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$out/bench_n$n.json") if l.startswith("{")][-1])
    print("N=$n", d["ms_per_step"], d["value"], d.get("parity_n"), {k:round(v["ms_per_step"],4) for k,v in d["kernels"].items()})
    print("  overlap", d["overlap"]["total_comm_ms"], [(round(c["t_bwd_ms"],3),round(c["step_ms"],3),round(c["step_no_comm_ms"],3),round(c["overlap_frac"],3)) for c in d["overlap"]["cases"]])
    for k,v in d.get("configs",{}).items(): print("  ",k, v.get("ms_per_step"), v.get("value"), {kk:round(vv["ms_per_step"],4) for kk,vv in v.get("kernels",{}).items()})
except Exception as e:
    print("no bench line:", e)
PY
done
for dom in remote default; do
  for wl in resnet50 bert_base; do
    PRB_NVL_DOMAIN=$dom timeout 200 python tools/exchange_sweep.py --gpus 8 --t-bwd-us 730 --workload $wl 2>&1 | grep -v "^NCCL" | tee -a $out/sweep_n8.log
  done
done
PRB_ALLREDUCE=nccl timeout 200 python tools/exchange_sweep.py --gpus 8 --t-bwd-us 730 --workload resnet50 --ctas 0 2>&1 | grep -v "^NCCL" | tee -a $out/sweep_n8.log
timeout 120 python tools/allreduce_bench.py --gpus 8 2>&1 | grep allreduce | tee -a $out/allreduce_n8.log
timeout 120 python tools/allreduce_bench.py --gpus 8 --dtype bfloat16 2>&1 | grep allreduce | tee -a $out/allreduce_n8.log
