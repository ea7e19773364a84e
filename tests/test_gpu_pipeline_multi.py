"""Multi-GPU parity of the layer-per-GPU pipeline inside the suite: on a box with >= 2 GPUs, spawn one rank per GPU
(tools/check_pipeline.py under torch.distributed.run) and require that the peer-memory transport, the NCCL
transport and the single-GPU step give the same weights and losses (bit-identical between transports).  Skipped on
a one-GPU box; `bench.py` additionally prints loss bits and a weight checksum at every N for the driver's records."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_pipeline_transports_match_single_gpu():
    n = 2 if torch.cuda.device_count() < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", "29731", os.path.join(REPO, "tools", "check_pipeline.py")]
    r = subprocess.run(cmd, cwd=REPO, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "PIPELINE CHECK PASS" in r.stdout, r.stdout[-3000:]
