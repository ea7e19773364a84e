"""Layer-per-GPU pipeline: replaces the container-per-layer gRPC relay chain
(/root/reference/src/grpc_node.py:120-135, run_grpc_fcnn.py:176-252) with one process per
GPU and point-to-point NCCL send/recv of activations (forward) and gradients (backward)
between adjacent stages over NVLink.

Stage s owns the consecutive layers ``layer_distribution`` assigns to container s -- the
reference's partition rule.  One training step is GPipe-shaped:

    for each micro-batch m:  recv A (from s-1) -> forward -> send A (to s+1)
    for each micro-batch m:  recv dZ (from s+1) | softmax-CE on the last stage
                             -> dX chain -> send dZ (to s-1)
    one dW GEMM per layer over the WHOLE batch with the SGD update in its epilogue

Deferring dW to the end keeps full-batch semantics exactly (no micro-batch sees a
half-updated W, no dW accumulator) and takes dW out of the pipeline bubble: only the
forward and dX phases pay the (S-1) fill/drain slots.

The stage object is duck-typed (``begin_batch, input_buffer, grad_buffer, forward_train,
loss_and_grad, backward, step``) so the schedule itself is tested on CPU with gloo.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .config import even_distribution, micro_batch_slices, stage_layer_ranges

_TYPESTR = {torch.float32: ("<f4", torch.float32), torch.bfloat16: ("<u2", torch.bfloat16), torch.int32: ("<i4", torch.int32)}


class _RawCudaBuffer:
    """Minimal __cuda_array_interface__ holder: lets torch wrap a raw (peer-mapped) device pointer."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, shape, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    """Zero-copy torch view of ``shape``/``dtype`` over a raw device pointer (plumbing only: the
    tensor is used to hand the pointer to libddnn)."""
    typestr, dt = _TYPESTR[dtype]
    t = torch.as_tensor(_RawCudaBuffer(ptr, shape, typestr), device=device)
    return t.view(dt) if t.dtype != dt else t


def stage_assignment(num_layers: int, world_size: int, distribution: Optional[Sequence[int]] = None
                     ) -> List[Tuple[int, int]]:
    """[first, last) layer range per rank.  A ``layer_distribution`` with exactly
    ``world_size`` non-empty stages is honoured; otherwise layers are split evenly."""
    if distribution is not None:
        ranges = stage_layer_ranges(num_layers, distribution)
        if len(ranges) == world_size:
            return ranges
    return stage_layer_ranges(num_layers, even_distribution(num_layers, world_size))


class PipelineStage:
    """``transport="nccl"``: dist.send/recv between stages (baseline; also what the CPU/gloo test
    drives).  ``transport="peer"``: the neighbours' receive buffers are mapped with CUDA IPC and the
    GEMM epilogues write into them directly over NVLink; arrival is signalled with 4-byte flags and
    stream memory waits -- no NCCL kernel, no copy and no host synchronisation in the data path."""

    def __init__(self, layer, rank: int, world_size: int, group=None, transport: str = "nccl"):
        self.layer = layer
        self.rank = rank
        self.world = world_size
        self.group = group
        self.first = rank == 0
        self.last = rank == world_size - 1
        self.transport = transport if world_size > 1 else "nccl"
        self._peer = None
        self._t = 0  # step counter: flags carry it, so they never need clearing

    # ------------------------------------------------------------------ peer transport set-up
    def _peer_setup(self, rows: int, n_micro: int) -> None:
        """Collective over all ranks: export this stage's receive buffers and flags, import the
        neighbours'.  Re-done whenever the batch buffers were reallocated."""
        from . import _lib
        import ctypes as C
        L = self.layer
        key = (rows, n_micro, L._gen)
        if self._peer is not None and self._peer["key"] == key:
            return
        dev = L.device
        # flags: [0, M) activation of micro-batch m has landed (written by rank-1);
        #        [M, 2M) gradient of micro-batch m has landed (written by rank+1);
        #        2M: rank+1 has consumed my activations of step t; 2M+1: rank-1 has consumed my gradients
        flags = torch.zeros((2 * n_micro + 2,), dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

        def export(t: torch.Tensor):
            h = (C.c_ubyte * 64)()
            off = C.c_int64(0)
            _lib.check(_lib.lib.ddnn_ipc_export(t.data_ptr(), C.cast(h, C.c_void_p), C.byref(off)))
            return (bytes(h), int(off.value), tuple(t.shape), str(t.dtype))

        mine = {"a0": export(L._bufs["a"][0]), "dz": export(L._bufs["dz"][-1]), "flags": export(flags)}
        everyone = [None] * self.world
        dist.all_gather_object(everyone, mine, group=self.group)

        def imp(desc):
            handle, off, shape, dt = desc
            out = C.c_void_p(0)
            buf = (C.c_ubyte * 64).from_buffer_copy(handle)
            _lib.check(_lib.lib.ddnn_ipc_import(C.cast(buf, C.c_void_p), off, C.byref(out)))
            return tensor_from_ptr(out.value, shape, getattr(torch, dt.split(".")[1]), dev)

        peer = {"key": key, "flags": flags, "M": n_micro}
        if not self.last:
            nxt = everyone[self.rank + 1]
            peer["next_a0"], peer["next_flags"] = imp(nxt["a0"]), imp(nxt["flags"])
        if not self.first:
            prv = everyone[self.rank - 1]
            peer["prev_dz"], peer["prev_flags"] = imp(prv["dz"]), imp(prv["flags"])
        self._peer = peer
        self._t = 0
        dist.barrier(group=self.group)  # nobody starts writing before every mapping exists

    def _flag_write(self, flags: torch.Tensor, idx: int, value: int) -> None:
        from . import _lib
        _lib.check(_lib.lib.ddnn_flag_write(flags.data_ptr() + 4 * idx, value,
                                            torch.cuda.current_stream(self.layer.device).cuda_stream))

    def _flag_wait(self, idx: int, value: int) -> None:
        from . import _lib
        _lib.check(_lib.lib.ddnn_flag_wait(self._peer["flags"].data_ptr() + 4 * idx, value,
                                           torch.cuda.current_stream(self.layer.device).cuda_stream))

    def _train_step_peer(self, x, labels, rows: int, micro_batches: int, lr: float):
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        M = len(slices)
        self._peer_setup(rows, M)
        P = self._peer
        self._t += 1
        t = self._t
        with torch.cuda.device(L.device):
            # ---- forward: the last local layer's epilogue stores into the next GPU's input buffer ----
            in_place = self.first and self._adopt_input(x)
            for m, (r0, r1) in enumerate(slices):
                n = r1 - r0
                if not self.first:
                    self._flag_wait(m, t)                    # micro-batch m of step t has landed here
                if not self.last and m == 0 and t > 1:
                    self._flag_wait(2 * M, t - 1)            # rank+1 is done with my step t-1 activations
                tgt = None if self.last else P["next_a0"][r0:r1]
                L.forward_train(x[r0:r1] if (self.first and not in_place) else None, r0, n, out_target=tgt)
                if not self.last:
                    self._flag_write(P["next_flags"], m, t)
            # ---- backward: the dX epilogue stores dZ into the previous GPU's gradient buffer ----
            loss = None
            for m, (r0, r1) in enumerate(slices):
                n = r1 - r0
                if self.last:
                    loss = L.loss_and_grad(labels[r0:r1], r0, denom=rows, accumulate=m > 0)
                else:
                    self._flag_wait(M + m, t)
                if not self.first and m == 0 and t > 1:
                    self._flag_wait(2 * M + 1, t - 1)        # rank-1 is done with my step t-1 gradients
                tgt = None if self.first else P["prev_dz"][r0:r1]
                L.backward(r0, n, need_input_grad=not self.first, dx_target=tgt)
                if not self.first:
                    self._flag_write(P["prev_flags"], M + m, t)
            # ---- update, then hand the receive buffers back to the neighbours ----
            L.step(lr, rows)
            if not self.first:
                self._flag_write(P["prev_flags"], 2 * M, t)      # consumed rank-1's activations
            if not self.last:
                self._flag_write(P["next_flags"], 2 * M + 1, t)  # consumed rank+1's gradients
        return loss

    def _adopt_input(self, x) -> bool:
        """First stage: let layer 1 read the batch where the caller keeps it (no staging copy) when the
        stage offers that (``Layer.set_input``) and the tensor qualifies."""
        setter = getattr(self.layer, "set_input", None)
        return bool(setter(x)) if setter is not None else False

    def _send(self, t: torch.Tensor, dst: int) -> None:
        dist.send(t, dst=dst, group=self.group)

    def _recv(self, t: torch.Tensor, src: int) -> None:
        dist.recv(t, src=src, group=self.group)

    def train_step(self, x: Optional[torch.Tensor], labels: Optional[torch.Tensor], rows: int,
                   micro_batches: int, lr: float):
        """One GPipe step over ``rows`` examples.  ``x`` is needed on the first stage,
        ``labels`` (int32) on the last.  Returns the loss tensor on the last stage, else None."""
        if self.transport == "peer":
            return self._train_step_peer(x, labels, rows, micro_batches, lr)
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        # ---- forward: fill ----
        in_place = self.first and self._adopt_input(x)
        for (r0, r1) in slices:
            n = r1 - r0
            if self.first:
                out = L.forward_train(None, r0, n) if in_place else L.forward_train(x[r0:r1], r0)
            else:
                self._recv(L.input_buffer(r0, n), self.rank - 1)
                out = L.forward_train(None, r0, n)
            if not self.last:
                self._send(out, self.rank + 1)
        # ---- backward: drain ----
        loss = None
        for i, (r0, r1) in enumerate(slices):
            n = r1 - r0
            if self.last:
                loss = L.loss_and_grad(labels[r0:r1], r0, denom=rows, accumulate=i > 0)
            else:
                self._recv(L.grad_buffer(r0, n), self.rank + 1)
            dx = L.backward(r0, n, need_input_grad=not self.first)
            if not self.first:
                self._send(dx, self.rank - 1)
        # ---- update: full-batch dW GEMMs with fused SGD ----
        L.step(lr, rows)
        return loss

    def infer(self, x: Optional[torch.Tensor], rows: int, micro_batches: int) -> Optional[torch.Tensor]:
        """Forward-only relay (what the reference chain does): the last stage returns the
        softmax output for all rows."""
        L = self.layer
        outs = []
        for (r0, r1) in micro_batch_slices(rows, micro_batches):
            if self.first:
                cur = x[r0:r1]
            else:
                cur = torch.empty((r1 - r0, L.expected_input_dim), dtype=L.wire_dtype, device=L.device)
                self._recv(cur, self.rank - 1)
            out = L.process_matrix(cur)
            if not self.last:
                self._send(out.to(L.wire_dtype).contiguous(), self.rank + 1)
            else:
                outs.append(out)
        return torch.cat(outs) if outs else None
