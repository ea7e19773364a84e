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
    def __init__(self, layer, rank: int, world_size: int, group=None):
        self.layer = layer
        self.rank = rank
        self.world = world_size
        self.group = group
        self.first = rank == 0
        self.last = rank == world_size - 1

    def _send(self, t: torch.Tensor, dst: int) -> None:
        dist.send(t, dst=dst, group=self.group)

    def _recv(self, t: torch.Tensor, src: int) -> None:
        dist.recv(t, src=src, group=self.group)

    def train_step(self, x: Optional[torch.Tensor], labels: Optional[torch.Tensor], rows: int,
                   micro_batches: int, lr: float):
        """One GPipe step over ``rows`` examples.  ``x`` is needed on the first stage,
        ``labels`` (int32) on the last.  Returns the loss tensor on the last stage, else None."""
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        # ---- forward: fill ----
        for (r0, r1) in slices:
            n = r1 - r0
            if self.first:
                out = L.forward_train(x[r0:r1], r0)
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
