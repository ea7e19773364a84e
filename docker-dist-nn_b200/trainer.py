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
        w0 = params[0]["weights"]  # numpy / nested list / torch tensor on any device
        in_dim = input_dim if input_dim is not None else int(w0.shape[0] if hasattr(w0, "shape") else len(w0))
        self.layer = Layer.from_params(params, in_dim, container_name="trainer", dtype=dtype, device=device,
                                       is_output=True)
        self.device = self.layer.device
        self.in_dim = in_dim
        self.num_classes = int(self.layer.layer_params[-1]["weights"].shape[1])
        self._side = torch.cuda.Stream(device=self.device)  # upper-layer dW GEMMs run beside the dX chain
        self._graph = None
        self._static = None
        self._static_gen = -1
        self._pinned = None

    # ------------------------------------------------------------------ device-resident step
    def train_step(self, x: torch.Tensor, labels: torch.Tensor, lr: Optional[float] = None) -> torch.Tensor:
        """x (B, in) CUDA fp32 or the compute dtype, labels (B,) int32 CUDA.  Returns the
        device loss scalar (mean CE before the update); does not synchronise."""
        L = self.layer
        L.begin_batch(x.shape[0])
        lr = self.lr if lr is None else lr
        if L.can_fuse_tail2():
            # small fp32 nets: the last two layers (forward, loss, backward, partial gradients) are one kernel
            L.forward_train(x, 0, skip_last=2)
            loss = L.tail2_step(labels, lr, x.shape[0], need_input_grad=False, side=self._side)
            L.backward_and_step(lr, x.shape[0], side=self._side, top=len(L.layer_params) - 3)
            torch.cuda.current_stream(self.device).wait_stream(self._side)
            return loss
        if L.can_fuse_output_layer():
            # skinny output layer: logits, softmax-CE, dX, dW/db and the SGD update in one pass over its input
            L.forward_train(x, 0, skip_last=True)
            loss = L.output_layer_step(labels, lr, x.shape[0], need_input_grad=False, side=self._side)
            L.backward_and_step(lr, x.shape[0], side=self._side, top=len(L.layer_params) - 2)
            torch.cuda.current_stream(self.device).wait_stream(self._side)  # loss + output-layer update are in
            return loss
        L.forward_train(x, 0)
        loss = L.loss_and_grad(labels, 0)
        L.backward_and_step(lr, x.shape[0], side=self._side)
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
        self._graph, self._static, self._static_gen = g, (sx, sy, loss), self.layer._gen

    def replay(self) -> torch.Tensor:
        if self._static_gen != self.layer._gen:
            raise RuntimeError("the batch buffers were reallocated after capture(); capture again")
        self._graph.replay()
        return self._static[2]

    @property
    def static_inputs(self) -> Tuple[torch.Tensor, torch.Tensor]:
        return self._static[0], self._static[1]

    def train_step_replicas(self, x: torch.Tensor, labels: torch.Tensor, lr: Optional[float] = None, group=None
                            ) -> torch.Tensor:
        """Data-parallel step (the reference has no replicated parameters): every rank of
        ``group`` holds a full replica and calls this with ITS rows of the global batch.  Each replica runs
        forward / softmax-CE / dX on its rows with the loss scaled by the GLOBAL row count, accumulates dW / db in
        fp32 (``Layer.accumulate_grads``), the accumulators are summed over the ranks (one all-reduce per layer:
        NCCL on GPUs), and every rank applies the same update -- the result is the full-batch step of the
        concatenated batch up to fp32 summation order.  Returns the global mean loss (device scalar)."""
        import torch.distributed as dist
        L = self.layer
        lr = self.lr if lr is None else lr
        rows = x.shape[0]
        total = torch.tensor([rows], dtype=torch.int64, device=self.device)
        dist.all_reduce(total, group=group)
        denom = int(total.item())
        L.begin_batch(rows)
        L.forward_train(x, 0)
        loss = L.loss_and_grad(labels, 0, denom=denom)
        L.backward(0, rows, need_input_grad=False)
        L.accumulate_grads(0, rows, first=True)
        b = L._bufs
        for gw, gb in zip(b["gw"], b["gb"]):
            dist.all_reduce(gw, group=group)
            dist.all_reduce(gb, group=group)
        dist.all_reduce(loss, group=group)
        L.apply_update(lr)
        return loss

    # ------------------------------------------------------------------ end-to-end host calls
    def train_step_host(self, x_host: torch.Tensor, y_host: torch.Tensor, lr: Optional[float] = None) -> float:
        """One synchronous step from HOST data: x_host (B, in) fp32 and y_host (B,) int32 in
        (pinned) host memory -> H2D copy -> step (graph replay when captured for this size) ->
        loss read back.  Returns the loss as a Python float, so this call synchronises."""
        rows = x_host.shape[0]
        if (self._graph is not None and self._static[0].shape[0] == rows and lr is None
                and self._static_gen == self.layer._gen):
            sx, sy, loss = self._static
            sx.copy_(x_host, non_blocking=True)
            sy.copy_(y_host, non_blocking=True)
            self._graph.replay()
        else:
            x = x_host.to(self.device, non_blocking=True)
            y = y_host.to(self.device, non_blocking=True)
            loss = self.train_step(x, y, lr)
        return float(loss.item())

    def _stream_setup(self, rows: int, lr: Optional[float], x_dtype: torch.dtype = torch.float32) -> None:
        """Two device staging slots (in the host batches' dtype), one captured step per slot, a copy stream and
        the events that order (H2D of slot s) -> (step on slot s) -> (next H2D into slot s)."""
        dev = self.device
        self.layer.begin_batch(rows)
        key = (rows, lr, self.layer._gen, x_dtype)  # a reallocation of the batch buffers invalidates the graphs
        if self._pinned is not None and self._pinned["key"] == key:
            return
        sx = [torch.zeros((rows, self.in_dim), dtype=x_dtype, device=dev) for _ in range(2)]
        sy = [torch.zeros((rows,), dtype=torch.int32, device=dev) for _ in range(2)]
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):  # warm-up outside capture (kernel attributes, TMA descriptors), lr = 0
            self.train_step(sx[0], sy[0], 0.0)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        graphs, loss = [], None
        for s_ in range(2):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                loss = self.train_step(sx[s_], sy[s_], lr)
            graphs.append(g)
        self._pinned = {"key": key, "sx": sx, "sy": sy, "graphs": graphs, "loss": loss,
                        "copy": torch.cuda.Stream(device=dev),
                        "h2d": [torch.cuda.Event() for _ in range(2)],
                        "done": [torch.cuda.Event() for _ in range(2)]}

    def fit_stream(self, batches, lr: Optional[float] = None) -> List[float]:
        """Trains on a sequence of HOST batches ``(x_pinned (B, in), y_pinned int32 (B,))`` in order; ``x`` is fp32
        or already in the compute dtype (a bf16 model fed bf16 rows moves half the bytes over PCIe and the step
        reads the staging slot in place; fp32 rows are cast on the device).  The H2D copy of batch i+1 runs on a copy stream while batch i computes
        (double-buffered staging slots, one captured CUDA graph per slot); each step's loss is
        copied back to pinned host memory asynchronously and all losses are returned after one
        final synchronisation.  Batches whose row count differs from the first (a ragged tail)
        run through ``train_step_host``."""
        batches = list(batches)
        if not batches:
            return []
        dev = self.device
        rows = batches[0][0].shape[0]
        self._stream_setup(rows, lr, batches[0][0].dtype)
        st = self._pinned
        losses_host = torch.empty((len(batches),), dtype=torch.float32).pin_memory()
        tail = {}
        cur = torch.cuda.current_stream(dev)
        for ev in st["done"]:
            ev.record(cur)
        for i, (hx, hy) in enumerate(batches):
            if hx.shape[0] != rows:
                tail[i] = self.train_step_host(hx, hy, lr)
                continue
            s_ = i & 1
            with torch.cuda.stream(st["copy"]):
                st["copy"].wait_event(st["done"][s_])       # slot free: its previous step has finished
                st["sx"][s_].copy_(hx, non_blocking=True)
                st["sy"][s_].copy_(hy, non_blocking=True)
                st["h2d"][s_].record(st["copy"])
            cur.wait_event(st["h2d"][s_])
            st["graphs"][s_].replay()
            losses_host[i:i + 1].copy_(st["loss"], non_blocking=True)
            st["done"][s_].record(cur)
        torch.cuda.synchronize(dev)
        out = losses_host.tolist()
        for i, v in tail.items():
            out[i] = v
        return out

    def fit(self, examples, epochs: Optional[int] = None, batch_size: Optional[int] = None,
            lr: Optional[float] = None) -> List[float]:
        """Trains on ``{"examples": [{"input", "label"}]}`` data (or an ``(x, y)`` pair of
        arrays) in the reference's batch order: sequential, unshuffled, ragged last batch kept
        (run_grpc_inference.py:198-205).  Returns the per-step losses."""
        if isinstance(examples, tuple):
            x, y = examples
        else:
            x, y = examples_to_arrays(examples)
        epochs = int(self.cfg.get("epochs", 1)) if epochs is None else epochs
        batch_size = self.cfg.get("batch_size") if batch_size is None else batch_size
        xs = torch.as_tensor(np.asarray(x), dtype=torch.float32).pin_memory()
        ys = torch.as_tensor(np.asarray(y), dtype=torch.int32).pin_memory()
        slices = batch_slices(xs.shape[0], batch_size)
        losses: List[float] = []
        for _ in range(epochs):
            losses += self.fit_stream([(xs[a:b], ys[a:b]) for (a, b) in slices], lr)
        return losses

    # ------------------------------------------------------------------ inference / export
    def predict(self, x):
        return self.layer.process_matrix(x)

    def export_params(self) -> List[Dict[str, Any]]:
        return self.layer.export_params()

    def export_config(self) -> Dict[str, Any]:
        """The trained model in the reference's JSON schema (one neuron per output column with
        ``weights``/``bias``/``activation``, ``type`` hidden|output, ``nodes``: the exporter layout of
        Centralized_MNIST_Experimentation.ipynb:462-506), so the unmodified reference chain can
        serve it; training keys and ``layer_distribution`` are carried over."""
        from .config import params_to_layers
        cfg = {"layers": params_to_layers(self.export_params())}
        for k in ("layer_distribution", "learning_rate", "epochs", "batch_size", "seed", "micro_batches", "dtype"):
            if k in self.cfg:
                cfg[k] = self.cfg[k]
        return cfg

    def save(self, path: str) -> None:
        import json
        with open(path, "w") as f:
            json.dump(self.export_config(), f)
