"""Two-GPU test of the front end fused with its all-gather (config 4 of BASELINE.json): needs at least two GPUs on
the box, skipped otherwise (the single-GPU suite covers the kernel's multi-destination stores in
test_gpu_processors.py::test_run_gather_stores_every_row_to_all_destinations)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def test_fused_gather_equals_nccl_all_gather():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(here, "fused_gather_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    assert "fused gather ok on 2 ranks" in res.stdout
