export PRB_NVL_TIMEOUT_MS=20000
mkdir -p gpurun_out/r2g
for dom in remote default; do
  echo "== domain $dom"
  PRB_NVL_DOMAIN=$dom timeout 120 python tools/allreduce_bench.py --gpus 2 2>&1 | grep allreduce
  PRB_NVL_DOMAIN=$dom timeout 200 python tools/exchange_sweep.py --gpus 2 --t-bwd-us 460 --workload resnet50 --ctas 0,32 2>&1 | grep -v "^NCCL"
done
PRB_NVL_DOMAIN=default timeout 200 python bench.py --gpus 2 --steps 50 --workload resnet50 --no-configs 2>/dev/null | grep "^{" > gpurun_out/r2g/bench_r50_default.json
python -c "
import json; d=json.load(open('gpurun_out/r2g/bench_r50_default.json')); print('resnet50 N=2 default domain', d['ms_per_step'], {k:round(v['ms_per_step'],4) for k,v in d['kernels'].items()}, d['overlap']['total_comm_ms'], [round(c['overlap_frac'],3) for c in d['overlap']['cases']])"
