#!/bin/bash
# One 8-GPU box: multi-rank parity tests, allreduce latency A/B of the done barrier, bench lines at N = 8, 4, 2.
mkdir -p gpurun_out/final
O=gpurun_out/final
timeout 400 python -m pytest tests/test_gpu_multi.py -q -x > $O/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_multi.log
for pr in 0 1; do
  PRB_NVL_PARALLEL_RELEASE=$pr timeout 90 python tools/allreduce_bench.py --gpus 8 2>&1 | grep allreduce | sed "s/^/par=$pr /" | tee -a $O/allreduce_n8.log
done
PRB_ALLREDUCE=nvls timeout 90 python tools/allreduce_bench.py --gpus 8 2>&1 | grep allreduce | tee -a $O/allreduce_n8.log
for pr in 0 1; do
  if [ $pr = 1 ]; then L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517"; else L="python"; fi
  PRB_NVL_PARALLEL_RELEASE=$pr timeout 200 $L bench.py --gpus 8 --steps 100 --warmup 10 2>/dev/null | grep '^{' > $O/bench_n8_par$pr.json
  python -c "
import json; d=json.load(open('$O/bench_n8_par$pr.json')); o=d['overlap']
print('N=8 par=$pr ms/step %.4f value %.0f comm %.4f overlap %.3f / single-launch %.3f' % (d['ms_per_step'], d['value'], o['total_comm_ms'], o['overlap_frac'], o['overlap_frac_vs_single_launch']))"
done
for n in 4 2; do
  timeout 150 python bench.py --gpus $n --steps 100 --warmup 10 2>/dev/null | grep '^{' > $O/bench_n$n.json
  python -c "
import json; d=json.load(open('$O/bench_n$n.json')); o=d['overlap']
print('N=$n ms/step %.4f value %.0f comm %.4f overlap %.3f / single-launch %.3f' % (d['ms_per_step'], d['value'], o['total_comm_ms'], o['overlap_frac'], o['overlap_frac_vs_single_launch']))"
done
timeout 200 python bench.py --gpus 8 --steps 30 --warmup 5 --workload gpt2_medium 2>/dev/null | grep '^{' > $O/bench_gpt2_n8.json
python -c "
import json; d=json.load(open('$O/bench_gpt2_n8.json')); print('gpt2 N=8 ms/step %.4f value %.0f' % (d['ms_per_step'], d['value']))"
