"""Batched inference / evaluation driver -- the GPU counterpart of
/root/reference/src/run_grpc_inference.py (same CLI shape, same batch order, same log lines).

    python -m ddnn_b200.infer [input_index] --config model.json --inputs examples.json
                              [--batch-size N] [--dtype f32|bf16] [--timeout S]

The reference client sends each batch through the gRPC chain on port 5101; here the batch goes
through the layers on the GPU (``Layer.process_matrix``), or -- ``--gpus N`` -- through the layer-per-GPU
pipeline forward-only: one process per GPU, stages from the model's ``layer_distribution`` (even split when it
does not name N stages), each batch relayed in ``--micro-batches`` chunks over the peer transport
(``PipelineStage.infer``), rank 0 printing the reference's log lines.  ``--gpus N`` re-launches this module under
``torch.distributed.run`` unless it already runs there.  ``--port`` is accepted and ignored
(there is no server); ``--config`` is new because no launcher has loaded the model beforehand.
Batches are sequential, unshuffled, ``ceil(N/bs)`` of them with the ragged tail kept
(run_grpc_inference.py:198-205); accuracy = count(argmax == label) (:191-194, :207-210).
"""
from __future__ import annotations

import argparse
import logging
import math
import os
import subprocess
import sys
import time
from typing import List, Optional, Tuple

import numpy as np

from .config import layers_to_params, load_example_inputs, load_model_config

DEFAULT_GRPC_TIMEOUT = 10  # run_grpc_inference.py:28 (kept for CLI compatibility)


def evaluate(config, examples: List[dict], batch_size: Optional[int] = None, input_index: Optional[int] = None,
             dtype: Optional[str] = None, device=None) -> Tuple[int, int, float, List[float]]:
    """Returns (correct, num_examples, total_seconds, per_batch_seconds) with the reference's
    semantics: ``input_index`` selects one example but the reported denominator stays the size of the
    file (run_grpc_inference.py:173-177, :215)."""
    import torch

    from .layer import Layer
    cfg = load_model_config(config)
    params = layers_to_params(cfg["layers"])
    num_examples = len(examples)
    if input_index is not None:
        if not 0 <= input_index < len(examples):
            raise IndexError(f"Input index {input_index} is out of range (0-{len(examples) - 1}).")
        examples = [examples[input_index]]
    in_dim = int(np.asarray(params[0]["weights"]).shape[0])
    layer = Layer.from_params(params, in_dim, container_name="infer", dtype=dtype or cfg.get("dtype", "f32"),
                              device=device, is_output=True)
    correct = 0
    per_batch = []
    total_start = time.monotonic()
    n = len(examples)
    if batch_size is None or batch_size >= num_examples:
        slices = [(0, n)]
    else:
        slices = [(b * batch_size, min(b * batch_size + batch_size, n)) for b in range(math.ceil(n / batch_size))]
    for b, (s, e) in enumerate(slices):
        batch = examples[s:e]
        x = np.array([ex.get("input") for ex in batch], dtype=np.float64)
        labels = [ex.get("label") for ex in batch]
        t0 = time.monotonic()
        out = layer.process_matrix(x)
        torch.cuda.synchronize(layer.device)
        per_batch.append(time.monotonic() - t0)
        for idx, row in enumerate(out):
            pred = int(np.argmax(row)) if row.size > 0 else None
            if labels[idx] == pred:
                correct += 1
        if len(slices) == 1:
            logging.info(f"Batch inference completed in {per_batch[-1]:.4f} seconds for {num_examples} examples.")
        else:
            logging.info(f"Batch {b + 1}/{len(slices)} completed in {per_batch[-1]:.4f} seconds ({e - s} examples).")
    return correct, num_examples, time.monotonic() - total_start, per_batch


def evaluate_pipeline(config, examples: List[dict], batch_size: Optional[int] = None,
                      input_index: Optional[int] = None, dtype: Optional[str] = None, micro_batches: int = 4,
                      transport: str = "peer") -> Tuple[int, int, float, List[float]]:
    """``evaluate`` over the GPUs of this process group (call on every rank under torch.distributed.run): stage r =
    the layers ``layer_distribution`` gives container r (run_grpc_fcnn.py:176-252), the batch enters on rank 0 and
    the last rank counts argmax hits; every rank returns the same (correct, num_examples, seconds, per_batch)."""
    import torch
    import torch.distributed as dist

    from .layer import Layer
    from .pipeline import PipelineStage, stage_assignment
    rank, world = dist.get_rank(), dist.get_world_size()
    device = torch.device("cuda", torch.cuda.current_device())
    cfg = load_model_config(config)
    params = layers_to_params(cfg["layers"])
    num_examples = len(examples)
    if input_index is not None:
        if not 0 <= input_index < len(examples):
            raise IndexError(f"Input index {input_index} is out of range (0-{len(examples) - 1}).")
        examples = [examples[input_index]]
    lo, hi = stage_assignment(len(params), world, cfg.get("layer_distribution"))[rank]
    dims_in = int(np.asarray(params[lo]["weights"]).shape[0])
    layer = Layer.from_params(params[lo:hi], dims_in, container_name=f"fcnn_layer_container_{rank}",
                              dtype=dtype or cfg.get("dtype", "f32"), device=device,
                              input_activation=(params[lo - 1]["activation"] if lo > 0 else "linear"),
                              is_output=(hi == len(params)))
    stage = PipelineStage(layer, rank, world, transport=transport)
    n = len(examples)
    if batch_size is None or batch_size >= num_examples:
        slices = [(0, n)]
    else:
        slices = [(b * batch_size, min(b * batch_size + batch_size, n)) for b in range(math.ceil(n / batch_size))]
    correct = torch.zeros((1,), dtype=torch.int64, device=device)
    per_batch = []
    total_start = time.monotonic()
    for b, (s, e) in enumerate(slices):
        batch = examples[s:e]
        rows = e - s
        x = None
        if rank == 0:
            x = torch.as_tensor(np.array([ex.get("input") for ex in batch], dtype=np.float64),
                                dtype=torch.float32).to(device)
        t0 = time.monotonic()
        out = stage.infer(x, rows, min(micro_batches, rows))
        if rank == world - 1:
            labels = torch.as_tensor([(-1 if ex.get("label") is None else int(ex.get("label"))) for ex in batch],
                                     device=device)
            correct += (out.argmax(dim=1) == labels).sum()
        torch.cuda.synchronize(device)
        dist.barrier(device_ids=[device.index])
        per_batch.append(time.monotonic() - t0)
        if rank == 0:
            if len(slices) == 1:
                logging.info(f"Batch inference completed in {per_batch[-1]:.4f} seconds for {num_examples} examples.")
            else:
                logging.info(f"Batch {b + 1}/{len(slices)} completed in {per_batch[-1]:.4f} seconds ({rows} examples).")
    dist.broadcast(correct, src=world - 1)
    return int(correct.item()), num_examples, time.monotonic() - total_start, per_batch


def main(argv=None) -> int:
    logging.basicConfig(level=logging.INFO, format="%(asctime)s - %(levelname)s - %(message)s")
    p = argparse.ArgumentParser(description="Run inference on a docker-dist-nn model on a B200.")
    p.add_argument("input_index", type=int, nargs="?", help="Index of the input example to use (default: all examples).")
    p.add_argument("--inputs", type=str, required=True, help="Path to the example inputs JSON file.")
    p.add_argument("--config", type=str, required=True, help="Path to the model JSON file (reference schema).")
    p.add_argument("--port", type=int, default=5101, help="Ignored (kept for CLI compatibility with the gRPC client).")
    p.add_argument("--timeout", type=float, default=DEFAULT_GRPC_TIMEOUT, help="Ignored (no RPC).")
    p.add_argument("--batch-size", type=int, default=None, help="Number of examples per batch (default: all).")
    p.add_argument("--dtype", type=str, default=None, choices=["f32", "bf16"])
    p.add_argument("--gpus", type=int, default=1, help="Run the layer-per-GPU pipeline over this many GPUs.")
    p.add_argument("--micro-batches", type=int, default=4, help="Chunks a batch is relayed in (pipeline only).")
    args = p.parse_args(argv)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # one process per GPU: start over under the launcher (rendezvous on 127.0.0.1, like the reference's client)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("DDNN_MASTER_PORT", "29741"),
               "-m", "ddnn_b200.infer"] + list(sys.argv[1:] if argv is None else argv)
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))  # where the ddnn_b200 import shim lives
        env = dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
        return subprocess.call(cmd, env=env)
    try:
        examples = load_example_inputs(args.inputs)
        if not examples:
            logging.error("No examples loaded, cannot run inference.")
            return 1
    except Exception as e:  # noqa: BLE001
        logging.error(f"An unexpected error occurred loading inputs: {e}")
        return 1
    rank = 0
    try:
        if world > 1:
            import datetime

            import torch
            import torch.distributed as dist
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
            dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=120))
            rank = dist.get_rank()
            try:
                correct, n, total, _ = evaluate_pipeline(args.config, examples, args.batch_size, args.input_index,
                                                         args.dtype, args.micro_batches)
            finally:
                dist.destroy_process_group()
        else:
            correct, n, total, _ = evaluate(args.config, examples, args.batch_size, args.input_index, args.dtype)
    except IndexError as e:
        logging.error(str(e))
        return 1
    if rank == 0:
        logging.info(f"Total inference time: {total:.4f} seconds")
        logging.info(f"Correct predictions: {correct} out of {n}")
        logging.info("Inference process completed.")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
