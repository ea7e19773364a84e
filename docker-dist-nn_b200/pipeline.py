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

import os

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


def pipeline_schedule(kind: str, layers_per_stage: Sequence[int], micro_batches: int,
                      dw_chunk: Optional[int] = None) -> List[List[Tuple[str, int]]]:
    """The op list every stage executes in ONE step, in order, for ``micro_batches`` micro-batches:
    ("F", m) forward of micro-batch m, ("B", m) its backward (loss on the last stage, dX chain, gradient to the
    previous stage), ("W", c) weight-gradient accumulation of chunk c = micro-batches [c * dw_chunk,
    (c + 1) * dw_chunk), ("U", 0) the SGD update.  ``dw_chunk`` None: one full-batch dW GEMM per layer with the
    update fused (a single ("W", -1) op, no "U").

    ``gpipe``: all F, all B, then the W ops, U -- the (S - 1) fill and drain slots are idle.
    ``zb``   : zero-bubble order.  B and W are separate ops, so a stage can do weight-gradient work while it waits
               for its neighbours.  The lists come from a unit-cost simulation of all stages in which an idle
               stage picks, among the ops whose inputs have arrived: the oldest B (it feeds the stage upstream),
               else the oldest F, else the oldest W.  F and B interleave (1F1B-like) and the W chunks fill what
               would be fill / drain bubbles; the bound drops from 3M + 2(S - 1) to about 3M + (S - 1) slots.
               Every op still waits for its real input at run time (flags), so the simulated costs only shape the
               order; any order produced here is deadlock-free because it is one consistent global execution."""
    S, M = len(layers_per_stage), int(micro_batches)
    if kind not in ("gpipe", "zb"):
        raise ValueError(f"schedule must be 'gpipe' or 'zb', got {kind!r}")
    chunks = None
    if dw_chunk:
        chunks = [(c0, min(M, c0 + dw_chunk)) for c0 in range(0, M, dw_chunk)]
    if kind == "gpipe":
        tail = [("W", -1)] if chunks is None else [("W", c) for c in range(len(chunks))] + [("U", 0)]
        return [[("F", m) for m in range(M)] + [("B", m) for m in range(M)] + list(tail) for _ in range(S)]
    if chunks is None:
        raise ValueError("the zero-bubble order needs dw_chunk (weight gradients per chunk of micro-batches)")
    INF = float("inf")
    f_done = [[INF] * M for _ in range(S)]
    b_done = [[INF] * M for _ in range(S)]
    w_done = [[INF] * len(chunks) for _ in range(S)]
    started = [set() for _ in range(S)]
    free_at = [0.0] * S
    order: List[List[Tuple[str, int]]] = [[] for _ in range(S)]
    cost_f = [float(n) for n in layers_per_stage]
    cost_b = [float(n - 1 if i == 0 else n) for i, n in enumerate(layers_per_stage)]  # layer 1 has no dX
    total_ops = S * (2 * M + len(chunks))
    t, done_ops = 0.0, 0
    while done_ops < total_ops:
        progressed = False
        for i in range(S):
            if free_at[i] > t:
                continue

            def ready_b(m):
                src = f_done[i][m] if i == S - 1 else b_done[i + 1][m]
                return ("B", m) not in started[i] and src <= t and f_done[i][m] <= t
            def ready_f(m):
                src = 0.0 if i == 0 else f_done[i - 1][m]
                return ("F", m) not in started[i] and src <= t and (m == 0 or ("F", m - 1) in started[i])
            def ready_w(c):
                return ("W", c) not in started[i] and all(b_done[i][m] <= t for m in range(*chunks[c]))
            op = next((("B", m) for m in range(M) if ready_b(m)), None) \
                or next((("F", m) for m in range(M) if ready_f(m)), None) \
                or next((("W", c) for c in range(len(chunks)) if ready_w(c)), None)
            if op is None:
                continue
            kind_, j = op
            dur = cost_f[i] if kind_ == "F" else (cost_b[i] if kind_ == "B" else
                                                  cost_f[i] * (chunks[j][1] - chunks[j][0]))
            end = t + dur
            (f_done if kind_ == "F" else b_done if kind_ == "B" else w_done)[i][j] = end
            started[i].add(op)
            order[i].append(op)
            free_at[i] = end
            done_ops += 1
            progressed = True
        if not progressed or all(fa > t for fa in free_at):
            nxt = [v for v in free_at if v > t]
            nxt += [v for tab in (f_done, b_done) for row in tab for v in row if t < v < INF]
            if not nxt:
                if not progressed:
                    raise RuntimeError("pipeline_schedule: no op can start (internal error)")
                continue
            t = min(nxt)
    for i in range(S):
        order[i].append(("U", 0))
    return order


class PipelineStage:
    """``transport="nccl"``: dist.send/recv between stages (baseline; also what the CPU/gloo test
    drives).  ``transport="peer"``: the neighbours' receive buffers are mapped with CUDA IPC and the
    GEMM epilogues write into them directly over NVLink; arrival is signalled with 4-byte flags and
    stream memory waits -- no NCCL kernel, no copy and no host synchronisation in the data path."""

    def __init__(self, layer, rank: int, world_size: int, group=None, transport: str = "nccl",
                 schedule: str = "gpipe", dw_chunk: Optional[int] = None,
                 layers_per_stage: Optional[Sequence[int]] = None):
        """``schedule``: "gpipe" (all forwards, all backwards, then the weight update) or "zb" (zero-bubble order:
        forwards and backwards interleave and weight-gradient chunks fill the waits, ``pipeline_schedule``).
        ``dw_chunk``: micro-batches per weight-gradient chunk (``Layer.accumulate_grads``); None = one full-batch
        dW GEMM per layer with the update fused (gpipe only).  The chunk boundaries, not the stage count or the
        schedule, decide the bits of the result.  ``layers_per_stage``: layer counts of ALL stages (only the
        simulated op costs of "zb" use it; default: every stage like this one)."""
        if schedule not in ("gpipe", "zb"):
            raise ValueError(f"schedule must be 'gpipe' or 'zb', got {schedule!r}")
        if schedule == "zb" and not dw_chunk:
            raise ValueError("the zero-bubble order needs dw_chunk (weight gradients per chunk of micro-batches)")
        if schedule == "zb" and transport == "nccl" and world_size > 1 and dist.is_initialized() \
                and dist.get_backend(group) == "nccl":
            raise ValueError("schedule='zb' interleaves sends in both directions: use transport='peer' "
                             "(NCCL send/recv pairs would have to be issued in matching order on both ranks)")
        self.schedule = schedule
        self.dw_chunk = int(dw_chunk) if dw_chunk else None
        self.layers_per_stage = layers_per_stage
        self._op_cache = {}
        self._pending = []
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

        # tile counters of the GEMMs that publish a flag themselves (one per direction: both run on this stream)
        peer = {"key": key, "flags": flags, "M": n_micro, "counters": torch.zeros((2,), dtype=torch.int32, device=dev),
                "fused": bool(getattr(L, "supports_signals", False)) and os.environ.get("DDNN_FUSED_FLAGS", "1") != "0"}
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

    def _ops(self, M: int) -> List[Tuple[str, int]]:
        """This stage's op list for one step (``pipeline_schedule``), cached per micro-batch count."""
        key = (M, self.schedule, self.dw_chunk)
        if self._op_cache.get("key") != key:
            if self.layers_per_stage is not None:
                lps = list(self.layers_per_stage)
            else:  # every stage holds as many layers as this one (the even split the launchers use)
                lps = [len(getattr(self.layer, "layer_params", [None]))] * self.world
            self._op_cache = {"key": key, "ops": pipeline_schedule(self.schedule, lps, M, self.dw_chunk)[self.rank]}
        return self._op_cache["ops"]

    def _weight_ops(self, op, slices, rows: int, lr: float) -> None:
        """("W", -1): full-batch dW GEMMs with the update fused; ("W", c): gradient accumulation of chunk c;
        ("U", 0): the update from the accumulators."""
        L = self.layer
        kind, j = op
        if kind == "U":
            L.apply_update(lr)
        elif j < 0:
            L.step(lr, rows)
        else:
            m0, m1 = j * self.dw_chunk, min(len(slices), (j + 1) * self.dw_chunk)
            L.accumulate_grads(slices[m0][0], slices[m1 - 1][1], first=(j == 0))

    def _train_step_peer(self, x, labels, rows: int, micro_batches: int, lr: float, ready=None):
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        M = len(slices)
        self._peer_setup(rows, M)
        P = self._peer
        self._t += 1
        t = self._t
        fused = P["fused"]
        my_flags = P["flags"].data_ptr()
        cnt = P["counters"].data_ptr()
        has_dx = (not self.first) or len(L.layer_params) > 1   # is there a dX kernel to carry the gradient wait?
        loss = None
        first_send_f = first_send_b = True
        # With fused flags the arrival wait rides in the first GEMM of an op and the departure flag in its last one
        # (ddnn_signals): consecutive GEMMs stay adjacent in the stream and keep their programmatic dependent
        # launch; otherwise a stream memory wait before and a one-thread flag kernel after.
        with torch.cuda.device(L.device):
            in_place = self.first and self._adopt_input(x)
            for op in self._ops(M):
                kind, m = op
                if kind == "F":
                    # the last local layer's epilogue stores into the next GPU's input buffer
                    r0, r1 = slices[m]
                    if ready is not None:
                        torch.cuda.current_stream(L.device).wait_event(ready[m])  # host copy of these rows has landed
                    wait = signal = None
                    if not self.first:
                        if fused:
                            wait = (my_flags + 4 * m, t)
                        else:
                            self._flag_wait(m, t)                # micro-batch m of step t has landed here
                    if not self.last and first_send_f and t > 1:
                        self._flag_wait(2 * M, t - 1)            # rank+1 is done with my step t-1 activations
                    first_send_f = False
                    tgt = None if self.last else P["next_a0"][r0:r1]
                    if fused and not self.last:
                        signal = (P["next_flags"].data_ptr() + 4 * m, t, cnt)
                    kw = {"wait": wait, "signal": signal} if fused else {}
                    L.forward_train(x[r0:r1] if (self.first and not in_place) else None, r0, r1 - r0, out_target=tgt,
                                    **kw)
                    if not self.last and not fused:
                        self._flag_write(P["next_flags"], m, t)
                elif kind == "B":
                    # the dX epilogue stores dZ into the previous GPU's gradient buffer
                    r0, r1 = slices[m]
                    wait = signal = None
                    if self.last:
                        loss = L.loss_and_grad(labels[r0:r1], r0, denom=rows, accumulate=m > 0)
                    elif fused and has_dx:
                        wait = (my_flags + 4 * (M + m), t)
                    else:
                        self._flag_wait(M + m, t)
                    if not self.first and first_send_b and t > 1:
                        self._flag_wait(2 * M + 1, t - 1)        # rank-1 is done with my step t-1 gradients
                    first_send_b = False
                    tgt = None if self.first else P["prev_dz"][r0:r1]
                    if fused and not self.first:
                        signal = (P["prev_flags"].data_ptr() + 4 * (M + m), t, cnt + 4)
                    kw = {"wait": wait, "signal": signal} if fused else {}
                    L.backward(r0, r1 - r0, need_input_grad=not self.first, dx_target=tgt, **kw)
                    if not self.first and not fused:
                        self._flag_write(P["prev_flags"], M + m, t)
                else:
                    self._weight_ops(op, slices, rows, lr)
            # hand the receive buffers back to the neighbours (every reader of them, the dW GEMMs included, is done)
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
        if self.schedule == "zb":
            # F and B interleave, so neighbours may send to each other at the same time: a blocking send on both
            # sides would wait for the other's receive forever; post it and collect the handle at the end of the step
            self._pending.append(dist.isend(t, dst=dst, group=self.group))
        else:
            dist.send(t, dst=dst, group=self.group)

    def _recv(self, t: torch.Tensor, src: int) -> None:
        dist.recv(t, src=src, group=self.group)

    def train_step(self, x: Optional[torch.Tensor], labels: Optional[torch.Tensor], rows: int,
                   micro_batches: int, lr: float, ready=None):
        """One step over ``rows`` examples in ``micro_batches`` micro-batches, in the order of
        ``pipeline_schedule``.  ``x`` is needed on the first stage, ``labels`` (int32) on the last.  Returns
        the loss tensor on the last stage, else None.  ``ready`` (first stage, optional): one CUDA event per
        micro-batch that the forward of that micro-batch waits for -- the host-to-device copy of its rows
        (``fit_stream``)."""
        if self.transport == "peer":
            return self._train_step_peer(x, labels, rows, micro_batches, lr, ready)
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        in_place = self.first and self._adopt_input(x)
        loss = None
        self._pending = []
        for op in self._ops(len(slices)):
            kind, m = op
            if kind == "F":
                r0, r1 = slices[m]
                n = r1 - r0
                if ready is not None:
                    torch.cuda.current_stream(L.device).wait_event(ready[m])
                if self.first:
                    out = L.forward_train(None, r0, n) if in_place else L.forward_train(x[r0:r1], r0)
                else:
                    self._recv(L.input_buffer(r0, n), self.rank - 1)
                    out = L.forward_train(None, r0, n)
                if not self.last:
                    self._send(out, self.rank + 1)
            elif kind == "B":
                r0, r1 = slices[m]
                n = r1 - r0
                if self.last:
                    loss = L.loss_and_grad(labels[r0:r1], r0, denom=rows, accumulate=m > 0)
                else:
                    self._recv(L.grad_buffer(r0, n), self.rank + 1)
                dx = L.backward(r0, n, need_input_grad=not self.first)
                if not self.first:
                    self._send(dx, self.rank - 1)
            else:
                self._weight_ops(op, slices, rows, lr)
        for w in self._pending:
            w.wait()
        self._pending = []
        return loss

    # ------------------------------------------------------------------ host-facing loop
    def fit_stream(self, batches, rows: int, micro_batches: int, lr: float) -> List[float]:
        """Trains on a sequence of HOST batches, one pipeline step each, in order (the batch order of
        run_grpc_inference.py:198-205).  ``batches[i] = (x_host, y_host)``: ``x_host`` (rows, in) pinned, in
        the compute dtype, is read on the FIRST stage only (other stages may pass None); ``y_host`` (rows,)
        pinned int32 on the LAST stage only.  Every stage calls this with the same number of batches.

        First stage: two device slots; a copy stream moves batch i+1 host-to-device micro-batch by
        micro-batch while step i computes, the forward of a micro-batch waits only for its own rows, and
        layer 1 reads the slot in place.  Last stage: labels H2D and the loss D2H of every step are
        stream-ordered; one synchronisation at the end.  Returns the per-step losses on the last stage
        ([] elsewhere)."""
        batches = list(batches)
        if not batches:
            return []
        L = self.layer
        dev = L.device
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        key = (rows, len(slices), L._gen)
        st = getattr(self, "_host", None)
        if st is None or st["key"] != key:
            st = {"key": key, "copy": torch.cuda.Stream(device=dev)}
            if self.first:
                st["slots"] = [torch.empty((rows, L.expected_input_dim), dtype=L.wire_dtype, device=dev)
                               for _ in range(2)]
                st["ready"] = [[torch.cuda.Event() for _ in slices] for _ in range(2)]
                st["done"] = [torch.cuda.Event() for _ in range(2)]
            if self.last:
                st["labels"] = [torch.empty((rows,), dtype=torch.int32, device=dev) for _ in range(2)]
                st["ldone"] = [torch.cuda.Event() for _ in range(2)]
                st["lready"] = [torch.cuda.Event() for _ in range(2)]
            self._host = st
        cur = torch.cuda.current_stream(dev)
        for k in ("done", "ldone"):
            for ev in st.get(k, []):
                ev.record(cur)
        losses_host = torch.empty((len(batches),), dtype=torch.float32).pin_memory() if self.last else None
        for i, (hx, hy) in enumerate(batches):
            s_ = i & 1
            x = y = ready = None
            with torch.cuda.stream(st["copy"]):
                if self.first:
                    if hx.dtype != L.wire_dtype or tuple(hx.shape) != (rows, L.expected_input_dim):
                        raise ValueError(f"fit_stream: host batch must be ({rows}, {L.expected_input_dim}) "
                                         f"{L.wire_dtype}, got {tuple(hx.shape)} {hx.dtype}")
                    x, ready = st["slots"][s_], st["ready"][s_]
                    st["copy"].wait_event(st["done"][s_])      # slot free: step i-2 (its last reader dW_1) is done
                    for m, (r0, r1) in enumerate(slices):
                        x[r0:r1].copy_(hx[r0:r1], non_blocking=True)
                        ready[m].record(st["copy"])
                if self.last:
                    y = st["labels"][s_]
                    st["copy"].wait_event(st["ldone"][s_])
                    y.copy_(hy, non_blocking=True)
                    st["lready"][s_].record(st["copy"])
            if self.last:
                cur.wait_event(st["lready"][s_])
            loss = self.train_step(x, y, rows, micro_batches, lr, ready=ready)
            if self.first:
                st["done"][s_].record(cur)
            if self.last:
                st["ldone"][s_].record(cur)
                losses_host[i:i + 1].copy_(loss, non_blocking=True)
        torch.cuda.synchronize(dev)
        return losses_host.tolist() if self.last else []

    def infer(self, x: Optional[torch.Tensor], rows: int, micro_batches: int) -> Optional[torch.Tensor]:
        """Forward-only relay of ``rows`` examples in ``micro_batches`` chunks -- what the reference chain does
        with one request (grpc_node.py:103-147), pipelined: the last stage returns the output node's activations
        (fp32, (rows, N_out)) for all rows, the other stages None.  Uses the stage's resident buffers (no
        allocation per micro-batch) and the configured transport: with ``peer`` the last local GEMM of a stage
        stores straight into the next GPU's input buffer and the arrival flags ride in the GEMMs."""
        L = self.layer
        L.begin_batch(rows)
        slices = micro_batch_slices(rows, micro_batches)
        M = len(slices)
        if self.transport != "peer":
            for (r0, r1) in slices:
                n = r1 - r0
                if self.first:
                    out = L.forward_train(x[r0:r1], r0, infer=True)
                else:
                    self._recv(L.input_buffer(r0, n), self.rank - 1)
                    out = L.forward_train(None, r0, n, infer=True)
                if not self.last:
                    self._send(out, self.rank + 1)
            return L.output_buffer(0, rows) if self.last else None
        self._peer_setup(rows, M)
        P = self._peer
        self._t += 1
        t = self._t
        fused = P["fused"]
        with torch.cuda.device(L.device):
            in_place = self.first and self._adopt_input(x)
            for m, (r0, r1) in enumerate(slices):
                n = r1 - r0
                wait = signal = None
                if not self.first:
                    if fused:
                        wait = (P["flags"].data_ptr() + 4 * m, t)
                    else:
                        self._flag_wait(m, t)
                if not self.last and m == 0 and t > 1:
                    self._flag_wait(2 * M, t - 1)            # rank+1 is done with what I sent it last time
                tgt = None if self.last else P["next_a0"][r0:r1]
                if fused and not self.last:
                    signal = (P["next_flags"].data_ptr() + 4 * m, t, P["counters"].data_ptr())
                kw = {"wait": wait, "signal": signal} if fused else {}
                L.forward_train(x[r0:r1] if (self.first and not in_place) else None, r0, n, out_target=tgt,
                                infer=True, **kw)
                if not self.last and not fused:
                    self._flag_write(P["next_flags"], m, t)
            # hand the receive buffers back (a training step after this one waits for both flags of step t)
            if not self.first:
                self._flag_write(P["prev_flags"], 2 * M, t)
            if not self.last:
                self._flag_write(P["next_flags"], 2 * M + 1, t)
        return L.output_buffer(0, rows) if self.last else None
