"""Batched inference / evaluation driver -- the GPU counterpart of
/root/reference/src/run_grpc_inference.py (same CLI shape, same batch order, same log lines).

    python -m ddnn_b200.infer [input_index] --config model.json --inputs examples.json
                              [--batch-size N] [--dtype f32|bf16] [--timeout S]

The reference client sends each batch through the gRPC chain on port 5101; here the batch goes
through the layers on the GPU (``Layer.process_matrix``).  ``--port`` is accepted and ignored
(there is no server); ``--config`` is new because no launcher has loaded the model beforehand.
Batches are sequential, unshuffled, ``ceil(N/bs)`` of them with the ragged tail kept
(run_grpc_inference.py:198-205); accuracy = count(argmax == label) (:191-194, :207-210).
"""
from __future__ import annotations

import argparse
import logging
import math
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
    args = p.parse_args(argv)
    try:
        examples = load_example_inputs(args.inputs)
        if not examples:
            logging.error("No examples loaded, cannot run inference.")
            return 1
    except Exception as e:  # noqa: BLE001
        logging.error(f"An unexpected error occurred loading inputs: {e}")
        return 1
    try:
        correct, n, total, _ = evaluate(args.config, examples, args.batch_size, args.input_index, args.dtype)
    except IndexError as e:
        logging.error(str(e))
        return 1
    logging.info(f"Total inference time: {total:.4f} seconds")
    logging.info(f"Correct predictions: {correct} out of {n}")
    logging.info("Inference process completed.")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
