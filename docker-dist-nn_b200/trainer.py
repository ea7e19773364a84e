"""Single-GPU training driver: all layers of a model on one B200 as one ``Layer`` node.

``train_step`` = forward -> fused softmax-CE -> dX chain -> dW GEMMs with the SGD update in
their epilogue.  For the launch-bound small nets (MNIST shapes: ~10 kernels of a few
microseconds) the whole step can be captured once into a CUDA graph (``capture``) and
replayed; ``train_step_host`` is the end-to-end call a user makes with HOST data: pinned
staging buffers, H2D copy of the batch, the step, D2H read of the loss.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .config import batch_slices, examples_to_arrays, layers_to_params, load_model_config
from .layer import Layer


class Trainer:
    def __init__(self, model, dtype: Optional[str] = None, device=None, input_dim: Optional[int] = None):
        """``model``: a config dict / path in the reference schema, or a list of
        {"weights" (K,N), "biases", "activation"} dicts."""
        if isinstance(model, (list, tuple)):
            params, cfg = list(model), {}
        else:
            cfg = load_model_config(model)
            params = layers_to_params(cfg["layers"])
        self.cfg = cfg
        self.lr = float(cfg.get("learning_rate", 0.01))
        dtype = dtype or cfg.get("dtype", "f32")
        in_dim = input_dim if input_dim is not None else int(np.asarray(params[0]["weights"]).shape[0])
        self.layer = Layer.from_params(params, in_dim, container_name="trainer", dtype=dtype, device=device,
                                       is_output=True)
        self.device = self.layer.device
        self.in_dim = in_dim
        self.num_classes = int(self.layer.layer_params[-1]["weights"].shape[1])
        self._graph = None
        self._static = None
        self._pinned = None

    # ------------------------------------------------------------------ device-resident step
    def train_step(self, x: torch.Tensor, labels: torch.Tensor, lr: Optional[float] = None) -> torch.Tensor:
        """x (B, in) CUDA fp32 or the compute dtype, labels (B,) int32 CUDA.  Returns the
        device loss scalar (mean CE before the update); does not synchronise."""
        L = self.layer
        L.begin_batch(x.shape[0])
        L.forward_train(x, 0)
        loss = L.loss_and_grad(labels, 0)
        L.backward(0, x.shape[0], need_input_grad=False)
        L.step(self.lr if lr is None else lr, x.shape[0])
        return loss

    def capture(self, rows: int, lr: Optional[float] = None, x_dtype: torch.dtype = torch.float32) -> None:
        """Captures one ``train_step`` for a fixed batch size into a CUDA graph."""
        dev = self.device
        sx = torch.zeros((rows, self.in_dim), dtype=x_dtype, device=dev)
        sy = torch.zeros((rows,), dtype=torch.int32, device=dev)
        self.layer.begin_batch(rows)
        # warm-up on a side stream (sets kernel attributes, builds TMA descriptors) with lr = 0
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(s):
            for _ in range(2):
                self.train_step(sx, sy, 0.0)
        torch.cuda.current_stream(dev).wait_stream(s)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            loss = self.train_step(sx, sy, lr)
        self._graph, self._static = g, (sx, sy, loss)

    def replay(self) -> torch.Tensor:
        self._graph.replay()
        return self._static[2]

    @property
    def static_inputs(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return self._static[0], self._static[1]

    # ------------------------------------------------------------------ end-to-end host call
    def train_step_host(self, x_host: torch.Tensor, y_host: torch.Tensor, lr: Optional[float] = None) -> float:
        """x_host (B, in) fp32 and y_host (B,) int32 in (pinned) HOST memory.  Copies the batch
        to the device, runs the step (graph replay when captured for this size) and returns
        the loss as a Python float (a device->host read, so this call synchronises)."""
        rows = x_host.shape[0]
        if self._graph is not None and self._static[0].shape[0] == rows and lr is None:
            sx, sy, loss = self._static
            sx.copy_(x_host, non_blocking=True)
            sy.copy_(y_host, non_blocking=True)
            self._graph.replay()
        else:
            x = x_host.to(self.device, non_blocking=True)
            y = y_host.to(self.device, non_blocking=True)
            loss = self.train_step(x, y, lr)
        return float(loss.item())

    def fit(self, examples, epochs: Optional[int] = None, batch_size: Optional[int] = None,
            lr: Optional[float] = None) -> List[float]:
        """Trains on ``{"examples": [{"input", "label"}]}`` data in the reference's batch order
        (sequential, unshuffled, ragged last batch kept: run_grpc_inference.py:198-205)."""
        if isinstance(examples, tuple):
            x, y = examples
        else:
            x, y = examples_to_arrays(examples)
        epochs = int(self.cfg.get("epochs", 1)) if epochs is None else epochs
        batch_size = self.cfg.get("batch_size") if batch_size is None else batch_size
        xs = torch.as_tensor(np.asarray(x), dtype=torch.float32).pin_memory()
        ys = torch.as_tensor(np.asarray(y), dtype=torch.int32).pin_memory()
        losses = []
        for _ in range(epochs):
            for (a, b) in batch_slices(xs.shape[0], batch_size):
                losses.append(self.train_step_host(xs[a:b], ys[a:b], lr))
        return losses

    # ------------------------------------------------------------------ inference / export
    def predict(self, x):
        return self.layer.process_matrix(x)

    def export_params(self) -> List[Dict[str, Any]]:
        return self.layer.export_params()
