"""Multi-GPU paths, launched with torchrun on one node (needs >= 2 GPUs)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _torchrun(script, nproc, port, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), script]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)


def test_fused_cross_gpu_aggregation_matches_nccl_baseline():
    n = min(torch.cuda.device_count(), 8)
    res = _torchrun(os.path.join(ROOT, "tests", "helpers", "multigpu_agg_worker.py"), n, 29617)
    assert "MULTIGPU_AGG_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_sharded_statevector_matches_single_gpu_engine():
    n = torch.cuda.device_count()
    n = 8 if n >= 8 else (4 if n >= 4 else 2)
    res = _torchrun(os.path.join(ROOT, "tests", "helpers", "multigpu_sharded_worker.py"), n, 29618)
    assert "MULTIGPU_SHARDED_OK" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]
