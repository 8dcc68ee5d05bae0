#!/bin/bash
# 1-GPU check: all GPU tests, the default bench line, the ncu launch list of a short bench run, one full ncu capture
out=gpurun_out/${1:-n1}
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest.log
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -c 600 $out/bench.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("$out/bench.json") if l.startswith("{")][-1])
except Exception as e:
    print("no bench line:", e); raise SystemExit
print(d["ms_per_step"], d["value"], d["ms_per_step_instrumented"], d["gpu_launches"])
print({k:(round(v["ms_per_step"],4), v["launches_per_step"]) for k,v in d["kernels"].items()})
print(d["roofline"]["frac"], d["e2e"]["ms_per_step"], d.get("cpu_baseline",{}).get("value"))
for k,v in d.get("configs",{}).items(): print(k, {a:b for a,b in v.items() if a not in ("kernels","workload","roofline","overlap")}, {kk:round(vv["ms_per_step"],4) for kk,vv in v.get("kernels",{}).items()})
PY
if [ "$2" = "ncu" ]; then
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"grad_pass_kernel|adam_chunks_kernel|norm_finalize_kernel|map_chunks_kernel|update_loss_scaling" -c 200 --csv --log-file $out/launches.csv \
      python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline > $out/ncu_list.log 2>&1; echo "ncu list rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"grad_pass_kernel|adam_chunks_kernel|norm_finalize_kernel" \
      --launch-skip 12 --launch-count 3 -o $out/bert_full python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline > $out/ncu_full.log 2>&1; echo "ncu full rc=$?"
  ncu -i $out/bert_full.ncu-rep --page raw --csv > $out/bert_full_raw.csv 2>/dev/null
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"map_chunks_kernel|adam_chunks_kernel" \
      --launch-skip 8 --launch-count 2 -o $out/resnet_full python bench.py --workload resnet50 --steps 5 --warmup 3 --no-configs --no-cpu-baseline > $out/ncu_full_resnet.log 2>&1; echo "ncu resnet rc=$?"
  ncu -i $out/resnet_full.ncu-rep --page raw --csv > $out/resnet_full_raw.csv 2>/dev/null
  rm -f $out/*.ncu-rep
fi
