#!/bin/bash
# 1-GPU check, final: all GPU tests, smoke(), the default bench line, an ncu capture of the sparse-gradient kernels
out=gpurun_out/${1:-n1b}
mkdir -p $out
timeout 600 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; tail -c 300 $out/bench.err
python - <<PY
import json
d=json.loads([l for l in open("$out/bench.json") if l.startswith("{")][-1])
print(d["ms_per_step"], d["value"], d["roofline"]["frac"], d["e2e"]["ms_per_step"], d.get("cpu_baseline",{}).get("value"))
print(d["configs"].get("sparse_adam:bert_embedding"))
PY
timeout 100 python tools/sparse_bench.py > $out/sparse.json 2>&1; cat $out/sparse.json
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"sparse_adam_kernel|merge_rows_kernel" --launch-skip 6 --launch-count 4 \
    -o $out/sparse_full python tools/sparse_bench.py > $out/ncu_sparse.log 2>&1; echo "ncu rc=$?"
ncu -i $out/sparse_full.ncu-rep --page raw --csv > $out/sparse_full_raw.csv 2>/dev/null
rm -f $out/*.ncu-rep
