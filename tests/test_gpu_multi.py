"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): NCCL bucket allreduce through DataParallel."""
import glob
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _wait_all(procs, timeout_s):
    """Wait for every rank; as soon as ONE rank fails the others (now stuck in a collective it left) are killed, so a
    failed assertion on one rank is reported at once instead of after the whole time limit."""
    import time
    t0 = time.time()
    while time.time() - t0 < timeout_s:
        codes = [p.poll() for p in procs]
        if all(c is not None for c in codes):
            break
        if any(c not in (None, 0) for c in codes):
            time.sleep(2.0)   # let the others print what they have
            break
        time.sleep(0.2)
    for p in procs:
        if p.poll() is None:
            p.kill()
    return [p.communicate()[0].decode() for p in procs]


def _ngpus():
    import paddle_ray_b200 as prb
    return prb.device_count()


@pytest.mark.parametrize("backend", ["nccl", "nvlink", "nvls"])
@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_dataparallel_grads_equal_across_ranks(tmp_path, nranks, backend):
    if _ngpus() < nranks:
        pytest.skip("needs %d GPUs" % nranks)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(nranks):
        env = {**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nranks), "MASTER_ADDR": "127.0.0.1",
               "MASTER_PORT": str(port), "PRB_ALLREDUCE": backend}
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), str(tmp_path)],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = _wait_all(procs, 600)
    own = [r for r, p in enumerate(procs) if p.returncode not in (0, -9)]   # ranks that failed by themselves first
    for r in own + [r for r in range(nranks) if r not in own]:
        assert procs[r].returncode == 0, "rank %d failed (rc %s):\n%s" % (r, procs[r].returncode, outs[r][-3000:])
    # grads and updated params are bit-equal across ranks (ref parallel_dygraph_gradient_check_in_eager_mode.py:139-150)
    files = sorted(glob.glob(os.path.join(str(tmp_path), "*_r0.npy")))
    assert files
    for f0 in files:
        a = np.load(f0)
        for r in range(1, nranks):
            b = np.load(f0.replace("_r0.npy", "_r%d.npy" % r))
            assert a.tobytes() == b.tobytes(), os.path.basename(f0)


def test_late_rank_makes_the_exchange_fail_loudly(tmp_path):
    """a rank that is later than the exchange's time limit must not make its peers reduce garbage."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    procs = []
    for r in range(2):
        env = {**os.environ, "RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": "2", "MASTER_ADDR": "127.0.0.1",
               "MASTER_PORT": str(port), "PRB_ALLREDUCE": "nvlink", "PRB_NVL_TIMEOUT_MS": "1500", "PRB_TEST_LATE_S": "4"}
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), str(tmp_path), "timeout"],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=120)[0].decode() for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and ("timeout ok" in o or "timeout skipped" in o), "rank %d:\n%s" % (r, o[-3000:])
