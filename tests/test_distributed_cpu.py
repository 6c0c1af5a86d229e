"""Multi-process host-side logic on CPU (gloo, 2 ranks): client placement, global sampling, distributed aggregation."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 3])
def test_split_job_equals_single_process_job(world):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29640 + world),
           os.path.join(ROOT, "tests", "helpers", "gloo_fl_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)
    assert out.returncode == 0 and "GLOO_FL_OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
