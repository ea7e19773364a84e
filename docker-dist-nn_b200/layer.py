"""``Layer`` -- the node/layer API of /root/reference/src/grpc_node.py:16-97 on a B200.

Same constructor contract (environment variables CONTAINER_NAME, LISTEN_PORT,
EXPECTED_INPUT_DIM, NEURONS_CONFIG | NEURONS_FILE_CONFIG, NEXT_NODES), same attributes
(``container_name, listen_port, expected_input_dim, layer_params, next_nodes``), same
methods (``activate``, ``process_matrix``) with the same error text; the arithmetic runs
in libddnn.so.  ``layer_params[i]["weights"]`` stays (K, N) like grpc_node.py:51 but is a
CUDA fp32 tensor (the master copy the SGD epilogue updates in place).

Added for training (no reference counterpart): ``begin_batch``,
``forward_train``, ``loss_and_grad``, ``backward``, ``step``, ``gradients``.
"""
from __future__ import annotations

import json
import os
import sys
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, ops
from .config import node_config_to_params

_TORCH_DT = {"f32": torch.float32, "fp32": torch.float32, "float32": torch.float32,
             "bf16": torch.bfloat16, "bfloat16": torch.bfloat16}


def _device(device) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("docker-dist-nn_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device(device)


class _NullSpan:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


_NULL_SPAN = _NullSpan()


class _Span:
    def __init__(self, sink, name, device):
        self.sink, self.name, self.device = sink, name, device

    def __enter__(self):
        self.e0 = torch.cuda.Event(enable_timing=True)
        self.e0.record(torch.cuda.current_stream(self.device))
        return self

    def __exit__(self, *a):
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record(torch.cuda.current_stream(self.device))
        self.sink.append((self.name, self.e0, e1))
        return False


class Layer:
    supports_signals = True  # forward_train / backward take wait= / signal= (flags folded into the GEMM kernels)

    def __init__(self, dtype: Optional[str] = None, device=None):
        # grpc_node.py:17-58 -- identical environment contract and failure behaviour
        self.container_name = os.environ.get("CONTAINER_NAME", "layer")
        self.listen_port = int(os.environ.get("LISTEN_PORT", "8000"))
        self.expected_input_dim = int(os.environ.get("EXPECTED_INPUT_DIM", "0"))
        neurons_config_data = None
        if "NEURONS_CONFIG" in os.environ:
            neurons_config_data = json.loads(os.environ["NEURONS_CONFIG"])
            print(f"({self.container_name}) Loaded neurons config from env var.")
        elif "NEURONS_FILE_CONFIG" in os.environ:
            config_path = os.environ["NEURONS_FILE_CONFIG"]
            try:
                with open(config_path, "r") as f:
                    neurons_config_data = json.load(f)
                print(f"({self.container_name}) Loaded neurons config from file: {config_path}")
            except FileNotFoundError:
                print(f"({self.container_name}) ERROR: Neurons config file not found at {config_path}")
                sys.exit(1)
            except json.JSONDecodeError:
                print(f"({self.container_name}) ERROR: Failed to decode JSON from {config_path}")
                sys.exit(1)
        else:
            print(f"({self.container_name}) ERROR: No neuron configuration provided "
                  "(NEURONS_CONFIG or NEURONS_FILE_CONFIG env var).")
            sys.exit(1)
        self.next_nodes = json.loads(os.environ.get("NEXT_NODES", "[]"))
        self._init(node_config_to_params(neurons_config_data or {}),
                   dtype or os.environ.get("DDNN_DTYPE", "f32"), device, "linear", None)
        print(f"({self.container_name}) Initialized {len(self.layer_params)} layer(s).")

    # -- programmatic constructor (benchmarks build multi-GiB layers from a seed, not JSON) --
    @classmethod
    def from_params(cls, layer_params: Sequence[Dict[str, Any]], expected_input_dim: int,
                    container_name: str = "layer", next_nodes: Optional[list] = None, dtype: str = "f32",
                    device=None, input_activation: str = "linear", is_output: Optional[bool] = None,
                    listen_port: int = 8000) -> "Layer":
        self = cls.__new__(cls)
        self.container_name = container_name
        self.listen_port = listen_port
        self.expected_input_dim = int(expected_input_dim)
        self.next_nodes = list(next_nodes or [])
        self._init(layer_params, dtype, device, input_activation, is_output)
        return self

    def _init(self, params, dtype, device, input_activation, is_output) -> None:
        self.device = _device(device)
        self.dtype = "bf16" if _TORCH_DT[dtype] == torch.bfloat16 else "f32"
        self._cdt = _TORCH_DT[dtype]
        # activation of the layer feeding this node (its derivative is fused into our dX)
        self.input_activation = input_activation
        # the last node of the chain owns the loss (grpc_node.py:143: no next_nodes)
        self.is_output = (not self.next_nodes) if is_output is None else bool(is_output)
        self.layer_params: List[Dict[str, Any]] = []
        self._shadow: List[Optional[torch.Tensor]] = []
        for p in params:
            w = self._to_dev(p["weights"])
            b = self._to_dev(p["biases"])
            if w.dim() != 2 or b.numel() != w.shape[1]:
                raise ValueError("weights must be (in, out) with one bias per output")
            self.layer_params.append({"weights": w, "biases": b, "activation": p.get("activation", "linear")})
            self._shadow.append(ops.cast(w, torch.bfloat16) if self._cdt == torch.bfloat16 else None)
        self._rows = 0
        self._bufs = None
        self._ws = None
        self._ws_cache: Dict[int, Any] = {}
        self._chunk_ws_cache: Dict[int, int] = {}
        self._ws_chunk: Optional[torch.Tensor] = None
        self._a0_ext: Optional[torch.Tensor] = None  # caller's batch read in place as layer 1's operand
        self._gen = 0  # bumped whenever the batch buffers are reallocated (captured graphs go stale)
        self._probe = None  # bench.py: list collecting (name, start_event, end_event) per entry-point call

    def _to_dev(self, t) -> torch.Tensor:
        """numpy / list / torch (any device) -> contiguous fp32 tensor on this node's GPU."""
        if isinstance(t, torch.Tensor):
            return t.detach().to(device=self.device, dtype=torch.float32).contiguous()
        return torch.as_tensor(np.asarray(t), dtype=torch.float32).contiguous().to(self.device)

    def _span(self, name: str):
        """CUDA-event bracket around one entry-point call while a probe list is attached
        (in-situ kernel timing for bench.py); a no-op context otherwise."""
        return _Span(self._probe, name, self.device) if self._probe is not None else _NULL_SPAN

    @property
    def wire_dtype(self) -> torch.dtype:
        """Element type of the activations / gradients exchanged between stages."""
        return self._cdt

    # ------------------------------------------------------------------ inference
    def activate(self, values, func: str):
        """grpc_node.py:62-73 on the GPU.  numpy in -> numpy (float64) out; CUDA tensor in ->
        CUDA fp32 tensor out.  Unknown ``func`` strings are the identity."""
        is_np = isinstance(values, np.ndarray)
        v = torch.as_tensor(values, dtype=torch.float32).to(self.device) if is_np else values.float()
        # any rank, like numpy's ufuncs; softmax runs over the last axis (grpc_node.py:70-71, axis=-1)
        shape = v.shape
        v2 = v.reshape(-1, shape[-1] if v.dim() else 1).contiguous()
        out = ops.activate(v2, func) if v2.numel() else v2.clone()
        out = out.reshape(shape)
        return out.double().cpu().numpy() if is_np else out

    def _w(self, i: int) -> torch.Tensor:
        return self._shadow[i] if self._cdt == torch.bfloat16 else self.layer_params[i]["weights"]

    def _check_dim(self, i: int, got: int, expected: int) -> None:
        if got != expected or got != self.layer_params[i]["weights"].shape[0]:
            exp = expected if got != expected else self.layer_params[i]["weights"].shape[0]
            raise ValueError(f"({self.container_name}) Layer {i + 1}: expected input dim {exp}, got {got}")

    def process_matrix(self, input_matrix):
        """grpc_node.py:75-97: chain ``act(X.W + b)`` over this node's layers.  Accepts a 2-D
        numpy array (returns float64 numpy, like the reference) or a CUDA tensor (returns a
        CUDA fp32 tensor).  Raises the reference's ValueError on a width mismatch."""
        is_np = isinstance(input_matrix, np.ndarray)
        cur = (torch.as_tensor(input_matrix, dtype=torch.float32).to(self.device) if is_np else input_matrix)
        if cur.dim() != 2:
            raise ValueError(f"({self.container_name}) input must be 2-D (rows, features)")
        cur = cur.contiguous()
        if cur.dtype != self._cdt:
            cur = ops.cast(cur, self._cdt) if cur.numel() else cur.to(self._cdt)
        expected = self.expected_input_dim
        n_layers = len(self.layer_params)
        for i, p in enumerate(self.layer_params):
            self._check_dim(i, cur.shape[1], expected)
            act = p["activation"]
            last = i == n_layers - 1
            if cur.shape[0] == 0:  # empty request -> empty result (grpc_node.py:108-111)
                cur = torch.empty((0, p["weights"].shape[1]), dtype=self._cdt, device=self.device)
            elif act == "softmax" or (last and self.is_output):
                out = ops.dense_forward(cur, self._w(i), p["biases"], act, out_dtype=torch.float32)
                cur = out if last else (ops.cast(out, self._cdt) if self._cdt != torch.float32 else out)
            else:
                cur = ops.dense_forward(cur, self._w(i), p["biases"], act)
            expected = cur.shape[1]
        out = cur if cur.dtype == torch.float32 else (ops.cast(cur.contiguous(), torch.float32) if cur.numel()
                                                      else cur.float())
        return out.double().cpu().numpy() if is_np else out

    # ------------------------------------------------------------------ training
    def begin_batch(self, rows: int) -> None:
        """(Re)allocates the per-batch buffers: layer inputs A[l], pre-softmax logits Z of
        an output node (fp32), one dZ per local layer, the gradient handed upstream, the
        split-K workspace.  Buffers are reused while ``rows`` does not grow."""
        if self._bufs is not None and rows <= self._bufs["cap"]:
            # split plans are not monotone in the batch (a ragged last batch can need MORE partials than the
            # full one), so the scratch is re-checked for this row count and grown when it falls short
            ws_bytes, ws2 = self._ws_need(rows)
            if ws_bytes > self._ws.numel() or ws2 > self._ws_side.numel():
                dev = self.device
                self._ws = torch.empty((max(ws_bytes, self._ws.numel()),), dtype=torch.uint8, device=dev)
                self._ws_side = torch.empty((max(ws2, self._ws_side.numel()),), dtype=torch.uint8, device=dev)
                self._gen += 1
            self._rows = rows
            return
        dev, cdt = self.device, self._cdt
        lp = self.layer_params
        a = [torch.empty((rows, lp[0]["weights"].shape[0]), dtype=cdt, device=dev)]
        dz = []
        for i, p in enumerate(lp):
            n = p["weights"].shape[1]
            last = i == len(lp) - 1
            a.append(torch.empty((rows, n), dtype=torch.float32 if (last and self.is_output) else cdt, device=dev))
            dz.append(torch.empty((rows, n), dtype=cdt, device=dev))
        ws_bytes, ws2 = self._ws_need(rows)
        self._bufs = {
            "cap": rows, "a": a, "dz": dz,
            "dx": torch.empty((rows, lp[0]["weights"].shape[0]), dtype=cdt, device=dev),
            "row_loss": torch.empty((rows,), dtype=torch.float32, device=dev),
            "loss": torch.zeros((1,), dtype=torch.float32, device=dev),
        }
        self._ws = torch.empty((max(ws_bytes, 256),), dtype=torch.uint8, device=dev)
        self._ws_side = torch.empty((max(ws2, 256),), dtype=torch.uint8, device=dev)
        self._rows = rows
        self._gen += 1

    def _ws_need(self, rows: int):
        """Bytes of split-K scratch a step over ``rows`` rows needs: (main stream, side stream).  The side
        workspace serves the dW GEMMs of the upper layers, which run beside dW_1 in ``backward_and_step`` and
        must not share its scratch, and the deferred reductions of the fused output layer / two-layer tail."""
        hit = self._ws_cache.get(rows)
        if hit is not None:
            return hit
        lp, cdt = self.layer_params, self._cdt
        dw = [ops.dw_workspace_bytes(rows, p["weights"].shape[0], p["weights"].shape[1], cdt) for p in lp]
        out_b = ops.output_layer_workspace_bytes(rows, lp[-1]["weights"].shape[0], lp[-1]["weights"].shape[1], cdt)
        tail_b = 0
        if len(lp) >= 2 and self.is_output:
            tail_b = ops.tail2_workspace_bytes(rows, lp[-2]["weights"].shape[0], lp[-2]["weights"].shape[1],
                                               lp[-1]["weights"].shape[1], cdt)
        self._ws_cache[rows] = (max(max(dw), out_b, tail_b), max(max(dw[1:], default=0), out_b, tail_b))
        return self._ws_cache[rows]

    def input_buffer(self, row0: int, rows: int) -> torch.Tensor:
        """Rows [row0, row0+rows) of this node's input activation buffer (receive target)."""
        return self._bufs["a"][0][row0:row0 + rows]

    def _a(self, i: int) -> torch.Tensor:
        """Input of local layer i for the whole batch: the caller's own tensor for layer 1 when it was
        adopted (``set_input``), else the node's buffer."""
        if i == 0 and self._a0_ext is not None:
            return self._a0_ext
        return self._bufs["a"][i]

    def set_input(self, x: Optional[torch.Tensor]) -> bool:
        """Makes layer 1 (forward and dW_1, its only readers) read the batch straight from the caller's
        tensor instead of a copy in ``input_buffer``: saves one read + one write of the batch per step.
        Taken when ``x`` has the compute dtype, is contiguous and 16-byte aligned; the caller keeps it
        unchanged until the step's weight update has run.  Returns whether it was adopted; ``None`` (or a
        tensor that does not qualify) goes back to the internal buffer."""
        ok = (x is not None and x.dtype == self._cdt and x.dim() == 2 and x.is_contiguous()
              and x.data_ptr() % 16 == 0 and x.device == self.device)
        if ok and x.shape[1] != self.expected_input_dim:
            raise ValueError(f"({self.container_name}) Layer 1: expected input dim {self.expected_input_dim}, "
                             f"got {x.shape[1]}")
        self._a0_ext = x if ok else None
        return ok

    def grad_buffer(self, row0: int, rows: int) -> torch.Tensor:
        """Rows of dZ of the LAST local layer (receive target for the downstream gradient)."""
        return self._bufs["dz"][-1][row0:row0 + rows]

    def output_buffer(self, row0: int, rows: int) -> torch.Tensor:
        return self._bufs["a"][-1][row0:row0 + rows]

    def forward_train(self, x: Optional[torch.Tensor], row0: int = 0, rows: Optional[int] = None,
                      out_target: Optional[torch.Tensor] = None, skip_last: int = 0,
                      wait=None, signal=None, infer: bool = False) -> torch.Tensor:
        """Forward of rows [row0, row0+rows) keeping what backward needs.  ``x`` None means the
        rows are already in ``input_buffer``.  An output node stops at the fp32 logits Z
        (softmax is fused into ``loss_and_grad``).  ``out_target`` (rows, N_last) redirects the
        LAST local layer's output, e.g. to the next stage's input buffer mapped from a peer GPU:
        the GEMM epilogue then stores over NVLink and no local copy of that activation exists
        (this node never reads its own last output).  ``wait = (flag_address, value)`` makes the FIRST local
        layer's kernel wait for that flag before it loads its input rows, ``signal = (flag_address, value,
        counter_address)`` makes the LAST local layer's kernel publish that flag when its output is complete
        (``ddnn_signals``): the pipeline's arrival / departure flags without a launch of their own.
        ``infer=True`` is the forward-only relay of the reference chain (grpc_node.py:75-97) on the same buffers:
        an output node applies its real activation (softmax) and any activation is allowed."""
        a = self._bufs["a"]
        if x is not None:
            rows = x.shape[0]
            if x.shape[1] != self.expected_input_dim:
                raise ValueError(f"({self.container_name}) Layer 1: expected input dim {self.expected_input_dim}, "
                                 f"got {x.shape[1]}")
            if row0 == 0 and rows == self._rows and self.set_input(x):
                pass  # the whole batch in the compute dtype: read in place, no staging copy
            else:
                self._a0_ext = None
                dst = a[0][row0:row0 + rows]
                if x.dtype == self._cdt:
                    dst.copy_(x)
                else:  # the cast is the one pass that moves the batch into the node
                    ops.cast(x.contiguous(), self._cdt, out=dst)
        n_layers = len(self.layer_params)
        for i, p in enumerate(self.layer_params):
            last = i == n_layers - 1
            if i >= n_layers - int(skip_last):  # the fused output-layer (1) / two-layer tail (2) step runs these
                break
            act = p["activation"]
            if infer:
                pass
            elif last and self.is_output:
                if act in ("relu", "sigmoid"):
                    raise ValueError(f"output activation {act!r} has no softmax-CE loss")
                act = "linear"  # logits; softmax happens inside the fused CE kernel
            elif act == "softmax":
                raise ValueError("hidden softmax layers are not supported in training (row Jacobian)")
            k_, n_ = p["weights"].shape
            dst = out_target if (last and out_target is not None) else a[i + 1][row0:row0 + rows]
            sig = None
            if (wait is not None and i == 0) or (signal is not None and last):
                w_, s_ = (wait if i == 0 else None), (signal if last else None)
                sig = ops.make_signals(w_[0] if w_ else None, w_[1] if w_ else 0, s_[0] if s_ else None,
                                       s_[1] if s_ else 0, s_[2] if s_ else None)
            with self._span(f"forward L{i + 1} {k_}->{n_}"):
                ops.dense_forward(self._a(i)[row0:row0 + rows], self._w(i), p["biases"], act, out=dst, signals=sig)
        return out_target if out_target is not None else a[-1][row0:row0 + rows]

    def loss_and_grad(self, labels: torch.Tensor, row0: int = 0, denom: Optional[int] = None,
                      accumulate: bool = False) -> torch.Tensor:
        """Softmax-CE on the logits of rows [row0, row0+len(labels)); fills dZ of the last
        layer with (p - onehot)/denom and (adds to) the device loss scalar.  ``denom`` =
        global batch size (default: the rows of this call)."""
        if not self.is_output:
            raise RuntimeError("loss_and_grad is only defined on the output node")
        rows = labels.numel()
        denom = rows if denom is None else denom
        b = self._bufs
        with self._span(f"softmax-CE N={b['a'][-1].shape[1]}"):
            ops.softmax_ce(b["a"][-1][row0:row0 + rows], labels, b["dz"][-1][row0:row0 + rows],
                           b["row_loss"][row0:row0 + rows], b["loss"], 1.0 / float(denom), accumulate=accumulate)
        return b["loss"]

    def backward(self, row0: int = 0, rows: Optional[int] = None, need_input_grad: bool = True,
                 dx_target: Optional[torch.Tensor] = None, wait=None, signal=None) -> Optional[torch.Tensor]:
        """dX chain for rows [row0, row0+rows): dZ[l-1] = (dZ[l].W_l^T) * f'_{l-1}(A[l]).  Returns
        dZ for the node feeding this one (already multiplied by that node's activation
        derivative, ``input_activation``) or None when ``need_input_grad`` is False.  ``wait`` / ``signal`` as in
        ``forward_train``: the first dX kernel of the chain (the one that reads the gradient received from
        downstream) waits, the dX that produces the upstream gradient publishes."""
        rows = self._rows - row0 if rows is None else rows
        b = self._bufs
        sl = slice(row0, row0 + rows)
        top = len(self.layer_params) - 1
        for i in range(top, 0, -1):
            k_, n_ = self.layer_params[i]["weights"].shape
            sig = ops.make_signals(wait[0], wait[1]) if (wait is not None and i == top) else None
            with self._span(f"dX L{i + 1} {n_}->{k_}"):
                ops.dense_backward_dx(b["dz"][i][sl], self._w(i), b["a"][i][sl],
                                      self.layer_params[i - 1]["activation"], out=b["dz"][i - 1][sl], signals=sig)
        if not need_input_grad:
            return None
        k_, n_ = self.layer_params[0]["weights"].shape
        dst = dx_target if dx_target is not None else b["dx"][sl]  # dx_target: the upstream stage's dZ buffer (peer)
        w_ = wait if (wait is not None and top == 0) else None
        sig = None
        if w_ is not None or signal is not None:
            sig = ops.make_signals(w_[0] if w_ else None, w_[1] if w_ else 0, signal[0] if signal else None,
                                   signal[1] if signal else 0, signal[2] if signal else None)
        with self._span(f"dX L1 {n_}->{k_}"):
            ops.dense_backward_dx(b["dz"][0][sl], self._w(0), self._a(0)[sl], self.input_activation, out=dst,
                                  signals=sig)
        return dst

    def step(self, lr: float, rows: Optional[int] = None, chunk_rows: Optional[int] = None) -> None:
        """dW = A^T.dZ over the first ``rows`` buffered rows with the SGD update fused into the
        GEMM epilogue: W -= lr*dW, b -= lr*db, bf16 shadow refreshed.  With ``chunk_rows`` (< rows) the
        reduction is cut into row chunks accumulated in fp32 (``accumulate_grads``) and applied once."""
        rows = self._rows if rows is None else rows
        if chunk_rows and chunk_rows < rows:
            for r0 in range(0, rows, chunk_rows):
                self.accumulate_grads(r0, min(rows, r0 + chunk_rows), first=(r0 == 0))
            self.apply_update(lr)
            return
        b = self._bufs
        for i, p in enumerate(self.layer_params):
            k_, n_ = p["weights"].shape
            with self._span(f"dW+SGD L{i + 1} {k_}x{n_}"):
                ops.dense_backward_dw_sgd(self._a(i)[:rows], b["dz"][i][:rows], p["weights"], p["biases"], lr,
                                          w_shadow=self._shadow[i], workspace=self._ws)

    def accumulate_grads(self, row0: int, row1: int, first: bool) -> None:
        """Adds the weight / bias gradients of rows [row0, row1) of every local layer to the node's fp32
        accumulators (``first`` overwrites them); weights stay untouched until ``apply_update``.  A pipeline stage
        calls this for the micro-batches it has finished while it waits for its neighbours; a very long batch is
        cut into such chunks so that the operands of each GEMM stay in L2.  The chunk boundaries are part of the
        arithmetic (fp32 rounding of the partial sums)."""
        b = self._bufs
        if "gw" not in b:
            b["gw"] = [torch.empty_like(p["weights"]) for p in self.layer_params]
            b["gb"] = [torch.empty_like(p["biases"]) for p in self.layer_params]
        sl = slice(row0, row1)
        need = self._chunk_ws_cache.get(row1 - row0)
        if need is None:  # a chunk's split plan can need more scratch than the full batch's (plans are not monotone)
            need = max(ops.dw_workspace_bytes(row1 - row0, p["weights"].shape[0], p["weights"].shape[1], self._cdt)
                       for p in self.layer_params)
            self._chunk_ws_cache[row1 - row0] = need
        ws = self._ws
        if need > ws.numel():  # own buffer: the batch buffers (and what peers / captured graphs hold of them) stay put
            if self._ws_chunk is None or self._ws_chunk.numel() < need:
                self._ws_chunk = torch.empty((need,), dtype=torch.uint8, device=self.device)
            ws = self._ws_chunk
        for i, p in enumerate(self.layer_params):
            k_, n_ = p["weights"].shape
            with self._span(f"dW chunk L{i + 1} {k_}x{n_}"):
                ops.dense_backward_dw_accumulate(self._a(i)[sl], b["dz"][i][sl], b["gw"][i], b["gb"][i], first,
                                                 workspace=ws)

    def apply_update(self, lr: float) -> None:
        """W -= lr * dW, b -= lr * db from the accumulators of ``accumulate_grads``; bf16 shadow refreshed."""
        b = self._bufs
        for i, p in enumerate(self.layer_params):
            k_, n_ = p["weights"].shape
            with self._span(f"SGD apply L{i + 1} {k_}x{n_}"):
                ops.sgd_apply(p["weights"], p["biases"], b["gw"][i], b["gb"][i], lr, w_shadow=self._shadow[i])

    def backward_and_step(self, lr: float, rows: Optional[int] = None, side: Optional[torch.cuda.Stream] = None,
                          top: Optional[int] = None) -> None:
        """backward() + step() for a single-node model with the weight updates of the upper layers
        running on ``side`` beside the dX chain: dW_l starts as soon as dX_l (the last reader of the
        old W_l) has finished, so only  dX_L .. dX_2 -> dW_1  stays on the critical path.  Works
        eagerly and under CUDA-graph capture (the cross-stream events become graph edges)."""
        rows = self._rows if rows is None else rows
        top = len(self.layer_params) - 1 if top is None else top  # highest layer still to be processed
        if top < 0:
            return
        if side is None:
            side = torch.cuda.current_stream(self.device)
        b = self._bufs
        main = torch.cuda.current_stream(self.device)
        sl = slice(0, rows)
        for i in range(top, 0, -1):
            p = self.layer_params[i]
            k_, n_ = p["weights"].shape
            with self._span(f"dX L{i + 1} {n_}->{k_}"):
                ops.dense_backward_dx(b["dz"][i][sl], self._w(i), b["a"][i][sl],
                                      self.layer_params[i - 1]["activation"], out=b["dz"][i - 1][sl])
            ev = torch.cuda.Event()
            ev.record(main)
            side.wait_event(ev)
            with torch.cuda.stream(side):
                ops.dense_backward_dw_sgd(b["a"][i][sl], b["dz"][i][sl], p["weights"], p["biases"], lr,
                                          w_shadow=self._shadow[i], workspace=self._ws_side)
        p = self.layer_params[0]
        k_, n_ = p["weights"].shape
        with self._span(f"dW+SGD L1 {k_}x{n_}"):
            ops.dense_backward_dw_sgd(self._a(0)[sl], b["dz"][0][sl], p["weights"], p["biases"], lr,
                                      w_shadow=self._shadow[0], workspace=self._ws)
        main.wait_stream(side)

    def can_fuse_output_layer(self) -> bool:
        """True when the last local layer is the network's skinny output layer (N <= 16, K <= 256,
        softmax / linear) that ``output_layer_step`` can run as one kernel."""
        if not self.is_output or self._bufs is None:
            return False
        p = self.layer_params[-1]
        if p["activation"] in ("relu", "sigmoid"):
            return False
        prev_act = self.layer_params[-2]["activation"] if len(self.layer_params) > 1 else self.input_activation
        if prev_act == "softmax":
            return False
        k, n = p["weights"].shape
        return ops.output_layer_supported(self._a(len(self.layer_params) - 1), k, n)

    def output_layer_step(self, labels: torch.Tensor, lr: float, rows: Optional[int] = None,
                          denom: Optional[int] = None, need_input_grad: bool = True,
                          side: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """Forward + softmax-CE + dX + dW/db + SGD of the output layer in one pass over its input
        (``ddnn_output_layer_train``).  Call after ``forward_train(..., skip_last=True)``; fills dZ of the
        layer below (already multiplied by that layer's activation derivative) and the device loss.

        With ``side`` the loss reduction and the split-K reduction + update run on that stream
        (``ddnn_output_layer_finish``): the dX chain below only needs dZprev, so they leave the critical
        path.  The caller joins ``side`` before it reads the loss or runs the next step
        (``backward_and_step`` does)."""
        rows = self._rows if rows is None else rows
        b = self._bufs
        i = len(self.layer_params) - 1
        p = self.layer_params[i]
        prev_act = self.layer_params[i - 1]["activation"] if i > 0 else self.input_activation
        dz_prev = None
        if i > 0:
            dz_prev = b["dz"][i - 1][:rows]
        elif need_input_grad:
            dz_prev = b["dx"][:rows]
        k_, n_ = p["weights"].shape
        inv = 1.0 / float(rows if denom is None else denom)
        main = torch.cuda.current_stream(self.device)
        if side is None or side == main:
            with self._span(f"output layer fused {k_}->{n_}"):
                ops.output_layer_train(self._a(i)[:rows], self._w(i), labels, b["row_loss"][:rows], b["loss"], inv,
                                       prev_act, self._ws, w_master=p["weights"], bias=p["biases"], lr=lr,
                                       w_shadow=self._shadow[i], dz=b["dz"][i][:rows], dz_prev=dz_prev)
            return b["loss"]
        with self._span(f"output layer fused {k_}->{n_}"):
            ops.output_layer_train(self._a(i)[:rows], self._w(i), labels, b["row_loss"][:rows], None, inv, prev_act,
                                   self._ws_side, bias=p["biases"], dz=b["dz"][i][:rows], dz_prev=dz_prev, defer=True)
        ev = torch.cuda.Event()
        ev.record(main)
        side.wait_event(ev)
        with torch.cuda.stream(side):
            ops.output_layer_finish(rows, k_, n_, self._cdt, b["row_loss"][:rows], b["loss"], inv, self._ws_side,
                                    w_master=p["weights"], bias=p["biases"], lr=lr, w_shadow=self._shadow[i])
        return b["loss"]

    def can_fuse_tail2(self) -> bool:
        """True when the last two local layers (hidden K1 -> N2, output N2 -> N3 with softmax / linear) fit the
        fused fp32 tail kernel (``ddnn_tail2_train``: K1 <= 128, N2 <= 64, N3 <= 16)."""
        lp = self.layer_params
        if not self.is_output or self._bufs is None or len(lp) < 2 or self._cdt != torch.float32:
            return False
        if lp[-1]["activation"] in ("relu", "sigmoid") or lp[-2]["activation"] == "softmax":
            return False
        act1 = lp[-3]["activation"] if len(lp) > 2 else self.input_activation
        if act1 == "softmax":
            return False
        k1, n2 = lp[-2]["weights"].shape
        return ops.tail2_workspace_bytes(self._bufs["cap"], k1, n2, lp[-1]["weights"].shape[1], self._cdt) > 0

    def tail2_step(self, labels: torch.Tensor, lr: float, rows: Optional[int] = None, denom: Optional[int] = None,
                   need_input_grad: bool = True, side: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """Forward, softmax-CE, backward and SGD of the last TWO layers in one kernel plus its reducers.  Call
        after ``forward_train(..., skip_last=2)``; fills dZ of the layer below them.  With ``side`` the loss
        reduction, both split-K reductions and the updates run on that stream (the caller joins it)."""
        main = torch.cuda.current_stream(self.device)
        deferred = side is not None and side != main
        ws = self._ws_side if deferred else self._ws
        self.tail2_kernel(labels, rows, denom, need_input_grad, ws)
        if deferred:
            ev = torch.cuda.Event()
            ev.record(main)
            side.wait_event(ev)
        with torch.cuda.stream(side if deferred else main):
            self.tail2_finish(lr, rows, denom, ws)
        return self._bufs["loss"]

    def tail2_kernel(self, labels: torch.Tensor, rows: Optional[int] = None, denom: Optional[int] = None,
                     need_input_grad: bool = True, ws: Optional[torch.Tensor] = None) -> None:
        """First half of ``tail2_step``: the fused kernel (row losses, dZ below the tail, split partials in ``ws``)."""
        rows = self._rows if rows is None else rows
        b, lp = self._bufs, self.layer_params
        i2 = len(lp) - 2
        p2, p3 = lp[i2], lp[i2 + 1]
        act1 = lp[i2 - 1]["activation"] if i2 > 0 else self.input_activation
        dz1 = b["dz"][i2 - 1][:rows] if i2 > 0 else (b["dx"][:rows] if need_input_grad else None)
        k1, n2 = p2["weights"].shape
        with self._span(f"tail fused {k1}->{n2}->{p3['weights'].shape[1]}"):
            ops.tail2_train(self._a(i2)[:rows], p2["weights"], p2["biases"], p3["weights"], p3["biases"], labels,
                            b["row_loss"][:rows], 1.0 / float(rows if denom is None else denom), act1,
                            p2["activation"], self._ws if ws is None else ws, dz1=dz1)

    def tail2_finish(self, lr: float, rows: Optional[int] = None, denom: Optional[int] = None,
                     ws: Optional[torch.Tensor] = None) -> None:
        """Second half on the current stream: loss reduction, both split-K reductions, SGD."""
        rows = self._rows if rows is None else rows
        b, lp = self._bufs, self.layer_params
        p2, p3 = lp[-2], lp[-1]
        ops.tail2_finish(rows, b["row_loss"][:rows], b["loss"], 1.0 / float(rows if denom is None else denom),
                         self._ws if ws is None else ws, p2["weights"], p2["biases"], p3["weights"], p3["biases"], lr=lr)

    def gradients(self, rows: Optional[int] = None):
        """[(dW, db)] per local layer from the buffered activations/gradients (no update)."""
        rows = self._rows if rows is None else rows
        b = self._bufs
        return [ops.dense_backward_dw(self._a(i)[:rows], b["dz"][i][:rows], workspace=self._ws)
                for i in range(len(self.layer_params))]

    def zero_loss(self) -> None:
        self._bufs["loss"].zero_()

    # ------------------------------------------------------------------ export
    def export_params(self) -> List[Dict[str, Any]]:
        """float64 numpy copies in the oracle/reference layout (for JSON export and parity)."""
        return [{"weights": p["weights"].double().cpu().numpy(), "biases": p["biases"].double().cpu().numpy(),
                 "activation": p["activation"]} for p in self.layer_params]
