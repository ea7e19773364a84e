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


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_eval_driver_over_the_pipeline(tmp_path):
    """`python -m ddnn_b200.infer --gpus 2`: the reference client's CLI and log lines (run_grpc_inference.py:162-252)
    with the chain replaced by the forward-only GPU pipeline; the accuracy line must equal the oracle's argmax count
    on the same exported model, ragged last batch included."""
    import json
    import re

    import numpy as np

    import oracle
    from ddnn_b200.config import make_config, make_params, synthetic_batch
    dims, acts = [784, 128, 64, 10], ["relu", "relu", "softmax"]
    cfg = make_config(dims, acts, seed=1, layer_distribution=[2, 1])
    x, y = synthetic_batch(300, 784, 10, seed=9)
    ref = oracle.process_matrix(make_params(dims, acts, seed=1), x, 784)
    want = int((ref.argmax(axis=1) == y).sum())
    (tmp_path / "model.json").write_text(json.dumps(cfg))
    (tmp_path / "inputs.json").write_text(json.dumps(
        {"examples": [{"input": r.tolist(), "label": int(l)} for r, l in zip(x, y)]}))
    cmd = [sys.executable, "-m", "ddnn_b200.infer", "--gpus", "2", "--config", str(tmp_path / "model.json"),
           "--inputs", str(tmp_path / "inputs.json"), "--batch-size", "128", "--micro-batches", "3"]
    r = subprocess.run(cmd, cwd=REPO, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600,
                       env=dict(os.environ, PYTHONPATH=REPO, DDNN_MASTER_PORT="29743"))
    assert r.returncode == 0, r.stdout[-3000:]
    m = re.search(r"Correct predictions: (\d+) out of (\d+)", r.stdout)
    assert m and int(m.group(2)) == 300 and int(m.group(1)) == want, r.stdout[-3000:]
    assert "Batch 3/3 completed" in r.stdout


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_data_parallel_replicas_match_the_full_batch_step():
    """tools/check_replicas.py: Trainer.train_step_replicas (dW / db accumulators all-reduced over NCCL, one update)
    on 2 shards of a batch = the single-GPU step on the whole batch, and the replicas stay bit-identical."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29735", os.path.join(REPO, "tools", "check_replicas.py")]
    r = subprocess.run(cmd, cwd=REPO, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0 and "REPLICA CHECK PASS" in r.stdout, r.stdout[-3000:]
