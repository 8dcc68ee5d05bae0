#!/bin/bash
# 2-GPU box: multi-rank tests with the final defaults, smoke(), and both arms of the bench the way the driver runs them
export PRB_NVL_TIMEOUT_MS=${PRB_NVL_TIMEOUT_MS:-20000}
out=gpurun_out/${1:-verify2}
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > $out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_multi.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus 2 --steps 5 --warmup 2 > $out/ref_n2.json 2> $out/ref_n2.err; echo "reference arm rc=$?"; cut -c1-700 $out/ref_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 20 --warmup 5 > $out/ours_n2.json 2> $out/ours_n2.err; echo "our arm rc=$?"; grep -c "^{" $out/ours_n2.json; head -c 300 $out/ours_n2.json; tail -3 $out/ours_n2.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > $out/ref_n1.json 2>$out/ref_n1.err; echo "reference N=1 rc=$?"; cut -c1-400 $out/ref_n1.json
