"""adam_ with a SelectedRows gradient alone (the bench's `sparse_adam:bert_embedding` entry), for ncu captures.

    python tools/sparse_bench.py
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import bench  # noqa: E402
import paddle_ray_b200 as prb  # noqa: E402

if __name__ == "__main__":
    prb.set_device(0)
    peak, _ = bench.measured_peak()
    print(json.dumps(bench.measure_sparse_adam(prb, peak)))
