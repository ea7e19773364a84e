#!/usr/bin/env python
"""bench.py -- train samples/s (forward + softmax-CE + backward + SGD) of the docker-dist-nn
layer hot path on N B200s.

    python bench.py --gpus 1 --steps K --warmup W              # our arm
    python bench.py --impl reference --gpus 1 --steps K ...     # CPU arm (oracle port + the reference's own forward)
    torchrun --nproc-per-node N ... bench.py --gpus N ...       # one rank per GPU

A "step" is one pass of the hot path over one batch of synthetic, seeded data.  Default
workload = BASELINE.json configs[3], the config the metric's "% of bf16 tensor peak" and the
north_star's 8-GPU scaling target are quoted on: the 8-layer 16384-wide bf16 FCNN, batch 65536
in 32 micro-batches.  It fits one B200 (one stage holding all 8 layers), and at N = 2 / 4 / 8 the
SAME batch runs through the layer-per-GPU pipeline ("strong" scaling): the per-N values of the
driver's scaling run are the pipeline's scaling curve, and the first losses and a weight checksum
in the line are the same bits at every N.  `roofline` is the tensor-core fraction of the
dominant GEMM.  The small nets (configs[1] MNIST batch 4096 fp32, configs[4] narrow batch 65536
bf16: HBM roofline) and configs[2] (4x8192) are reported in `extras`, each with value, e2e and
roofline.  `--workload mnist4096|narrow65536|wide4x8192` makes one of them the headline; the MNIST-sized
nets are too small to pipeline and run as independent replicas at N > 1.

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for every key.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
if REPO not in sys.path:
    sys.path.insert(0, REPO)



def _host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


if "--impl" in sys.argv and "reference" in sys.argv and os.environ.get("DDNN_REF_REEXEC") != "1":
    # The CPU arm uses every host thread.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers,
    # and BLAS reads the thread count when numpy is first imported, so fix the environment and start over.
    if int(os.environ.get("RANK", "0")) != 0:
        sys.exit(0)  # rank 0 alone runs the CPU arm; the other ranks exit without work
    _n = str(_host_threads())
    os.environ.update({"OMP_NUM_THREADS": _n, "OPENBLAS_NUM_THREADS": _n, "MKL_NUM_THREADS": _n,
                       "DDNN_REF_REEXEC": "1"})
    os.execv(sys.executable, [sys.executable] + sys.argv)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (dims, activations, batch rows, dtype, micro_batches)
    "mnist4096": ([784, 128, 64, 10], ["relu", "relu", "softmax"], 4096, "f32", 1),
    "narrow65536": ([784, 256, 256, 10], ["relu", "relu", "softmax"], 65536, "bf16", 1),
    # micro-batch rows are chosen so that a micro-batch GEMM is a whole number of 148-CTA waves
    # (4096 x 8192 -> 1024 tiles of 128x256 = 6.92 waves; 2048 x 16384 -> 1024 tiles)
    "wide4x8192": ([8192] * 5, ["relu"] * 3 + ["softmax"], 65536, "bf16", 16),
    "wide8x16384": ([16384] * 9, ["relu"] * 7 + ["softmax"], 65536, "bf16", 32),
}
L2_BYTES = 126 * 1024 * 1024


def flops_per_sample(dims):
    """sum_l 6*K*N - 2*K_1*N_1 (layer 1 skips dX)."""
    return sum(6 * k * n for k, n in zip(dims[:-1], dims[1:])) - 2 * dims[0] * dims[1]


def bytes_per_sample(dims, elt):
    """SURVEY.md section 8d activation traffic per sample: sum_l (2K + 3N + K*[l>1]) elements + label."""
    e = 0
    for i, (k, n) in enumerate(zip(dims[:-1], dims[1:])):
        e += 2 * k + 3 * n + (k if i > 0 else 0)
    return e * elt + 4


def load_peaks():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "bf16_tflops": float(p["bf16_tflops"]),
                "bf16_tflops_sustained": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), "source": "measured"}
    except Exception:  # noqa: BLE001
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


# --------------------------------------------------------------------------------------------
# clocks sampler (NVML in-process; nvidia-smi fallback)
# --------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.index = index
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001
            self._h = None

    def _once(self):
        if self._h is None:
            return
        try:
            nv = self._nv
            self.samples.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
            try:
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
            except Exception:  # noqa: BLE001
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
            for bit, name in self.REASONS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:  # noqa: BLE001
            pass

    def start(self):
        def loop():
            while not self._stop.is_set():
                self._once()
                self._stop.wait(0.02)
        self._thr = threading.Thread(target=loop, daemon=True)
        self._thr.start()

    def stop(self):
        self._once()
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=1.0)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------
# CPU arm: the float64 oracle (a port -- the reference has no training code)
# --------------------------------------------------------------------------------------------
def cpu_threads() -> int:
    try:
        from threadpoolctl import threadpool_info
        n = [i.get("num_threads", 0) for i in threadpool_info() if i.get("user_api") == "blas"]
        if n:
            return int(max(n))
    except Exception:  # noqa: BLE001
        pass
    return os.cpu_count() or 1


DEFAULT_WORKLOAD = "wide8x16384"
LR = 0.01
# wide nets: order of a pipeline step and rows per weight-gradient chunk (the same at every GPU count: the chunk
# boundaries are part of the arithmetic, so losses and weights stay bit-identical from 1 to 8 GPUs)
PIPE_SCHEDULE = "zb"
PIPE_DW_CHUNK_ROWS = 8192
# rows of one CPU step (a bounded sample of the workload's batch): the oracle's float64 GEMMs are compute-bound
# from ~100 rows on, and 20 + 5 steps must end within minutes on the host cores
CPU_ROWS = {"mnist4096": 4096, "narrow65536": 8192, "wide4x8192": 256, "wide8x16384": 256}
# ... and layers: every layer of a wide net has the same shape, and a float64 step over all eight 16384 x 16384
# layers spends 12 s of which most is weight-sized memory passes, so the CPU arm times the LAST two (dX, dW + SGD of
# a hidden layer, the softmax-CE output layer) and scales samples/s by their share of the step's FLOPs
CPU_LAYERS = {"wide8x16384": 2}


def line_config(workload: str, world: int):
    """The `config` object of the JSON line -- the same for both arms (`--impl reference` reports on our config)."""
    dims, acts, rows, dtype, micro = WORKLOADS[workload]
    pipe = workload.startswith("wide")
    n_layers = len(acts)
    if pipe:
        par = "single (all layers on one GPU)" if world == 1 else f"pipeline{world} ({n_layers // world} layers per GPU)"
        l2 = f"one batch = {rows * dims[0] * 2 / 2**20:.0f} MiB of inputs, {n_layers * dims[1] * dims[1] * 2 / 2**20:.0f}" \
             f" MiB of bf16 weights > 126 MiB L2"
    else:
        par = "single" if world == 1 else f"replicas{world} (net too small to pipeline)"
        l2 = "inputs rotate over distinct resident batches totalling >= 2 x 126 MiB L2"
    return {"workload": workload, "dims": dims, "activations": acts, "rows_per_step": rows, "micro_batches": micro,
            "lr": LR, "parallelism": par, "l2": l2}


def cpu_train_samples_per_s(workload: str, budget_s: float, steps=None, warmup=1, rows_cap=None):
    """Times oracle.train_step (numpy float64, multi-threaded BLAS) on the host cores on a
    bounded sample of the workload.  Returns (samples/s, description, ms_per_step, steps)."""
    import oracle
    from ddnn_b200.config import make_params, synthetic_batch
    dims, acts, rows, _, _ = WORKLOADS[workload]
    rows = min(rows, rows_cap or CPU_ROWS[workload])
    note, scale = "", 1.0
    keep = CPU_LAYERS.get(workload)
    if keep and keep < len(acts):
        # the sub-net's first layer also computes dX here (in the full net only layer 1 skips it)
        sub = sum(6 * k * n for k, n in zip(dims[-(keep + 1):-1], dims[-keep:]))
        scale = sub / flops_per_sample(dims)
        note = f"; timed on the last {keep} of the {len(acts)} equal layers (incl. their dX), samples/s scaled by " \
               f"their share {scale:.4f} of the step's FLOPs"
        dims, acts = dims[-(keep + 1):], acts[-keep:]
    params = make_params(dims, acts, seed=1)
    x, y = synthetic_batch(rows, dims[0], dims[-1], seed=0)
    def one_step():  # oracle.backward also computes dX of the first layer it is given
        oracle.train_step(params, x, y, LR)
    for _ in range(max(0, warmup)):
        one_step()
    t0 = time.perf_counter()
    n = 0
    while True:
        one_step()
        n += 1
        el = time.perf_counter() - t0
        if (steps is not None and n >= steps) or (steps is None and el >= budget_s):
            break
    desc = f"{n} oracle train steps (numpy float64, {cpu_threads()} BLAS threads) of {workload} at {rows} rows/step{note}"
    return scale * rows * n / el, desc, 1e3 * el / n, n


def reference_forward_baselines(budget_s: float):
    """the reference's OWN code (unmodified copy under baseline/_ref), forward-only, on
    the MNIST FCNN it ships scripts for (BASELINE.json configs[0]).  None when the copy is absent."""
    sys.path.insert(0, os.path.join(REPO, "tools"))
    try:
        import reference_chain as rc
    except Exception:  # noqa: BLE001
        return None
    if not rc.available():
        return {"unavailable": "baseline/_ref is empty (run tools/install_reference.py where /root/reference exists)"}
    from ddnn_b200.config import make_params, synthetic_batch
    dims, acts = WORKLOADS["mnist4096"][0], WORKLOADS["mnist4096"][1]
    params = make_params(dims, acts, seed=1)
    x, _ = synthetic_batch(2560, dims[0], dims[-1], seed=0)
    out = {"model": "784-128-64-10 (configs[0])", "kind": "reference", "cores": cpu_threads(),
           "what": "forward only -- the reference has no backward pass"}
    try:
        v2, calls = rc.time_in_process(params, x, 640, budget_s)
        out["B2_in_process_layer"] = {"value": v2, "unit": "inference samples/s",
                                      "sample": f"{calls} Layer.process_matrix calls of 640 rows, float64"}
    except Exception as e:  # noqa: BLE001
        out["B2_in_process_layer"] = {"error": repr(e)[:200]}
    try:
        v1, rpcs, rpr, _ = rc.time_grpc_chain(params, x, 640, budget_s)
        out["B1_grpc_chain"] = {"value": v1, "unit": "inference samples/s",
                                "sample": f"{rpcs} RPCs of {rpr} rows (under the 4 MiB server cap) through 3 unmodified "
                                          f"grpc_node.py processes on localhost, reference client loop"}
    except Exception as e:  # noqa: BLE001
        out["B1_grpc_chain"] = {"error": repr(e)[:200]}
    return out


def run_reference_arm(args):
    """CPU arm on the box's host cores (rank 0 only; the re-exec at the top of this file gave BLAS every
    thread).  The line's value is the float64 training port of the oracle on a bounded sample of OUR arm's
    config (the reference has no training code); `reference_forward` adds the reference's own
    forward path: the gRPC chain (B1) and the in-process Layer (B2)."""
    workload = args.workload or DEFAULT_WORKLOAD
    v, desc, ms, n = cpu_train_samples_per_s(workload, 0.0, steps=max(1, args.steps), warmup=max(0, args.warmup))
    line = {
        "impl": "reference", "metric": "train samples/sec (fwd+bwd+SGD)", "value": v, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": n, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong" if workload.startswith("wide") else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": line_config(workload, args.gpus),
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cpu_threads(), "kind": "port", "sample": desc,
                         "note": "the reference has no backward/SGD (SURVEY.md 0.1): this is the float64 numpy oracle, "
                                 "pinned bit-exact to the reference's forward"},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if not args.no_extra:
        line["reference_forward"] = reference_forward_baselines(4.0)
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def device_params(dims, acts, device, seed=1):
    """Random-init weights of the architecture, generated ON the device (the wide nets are
    GiBs): W ~ N(0, 2/K) for relu layers, N(0, 1/K) otherwise, b = 0.01."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    out = []
    for k, n, a in zip(dims[:-1], dims[1:], acts):
        std = ((2.0 if a == "relu" else 1.0) / k) ** 0.5
        w = torch.randn((k, n), device=device, generator=g, dtype=torch.float32) * std
        out.append({"weights": w, "biases": torch.full((n,), 0.01, device=device), "activation": a})
    return out


def timed(fn, steps, warmup, device, barrier=None):
    """W untimed calls, then K calls bracketed by (barrier +) synchronize and CUDA events on the
    current stream.  Returns elapsed milliseconds for the K calls."""
    import torch
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize(device)
    if barrier:
        barrier()
        torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(warmup + i)
    e1.record()
    torch.cuda.synchronize(device)
    if barrier:
        barrier()
        torch.cuda.synchronize(device)
    return e0.elapsed_time(e1)


def bench_single(workload, dtype, steps, warmup, device, barrier=None):
    """All layers on one GPU.  Returns a dict with value/e2e/roofline pieces."""
    import torch
    from ddnn_b200 import Trainer, _lib, ops
    dims, acts, rows, wdtype, _ = WORKLOADS[workload]
    dtype = dtype or wdtype
    elt = 4 if dtype == "f32" else 2
    tr = Trainer(device_params(dims, acts, device), dtype=dtype, device=device, input_dim=dims[0])
    tr.lr = LR
    xdt = torch.float32 if dtype == "f32" else torch.bfloat16
    # inputs larger than L2: a pool of distinct resident batches, rotated every step
    batch_bytes = rows * dims[0] * elt
    pool_n = max(2, min(64, -(-2 * L2_BYTES // batch_bytes)))
    g = torch.Generator(device=device).manual_seed(0)
    xs = [torch.rand((rows, dims[0]), device=device, generator=g, dtype=torch.float32).to(xdt) for _ in range(pool_n)]
    ys = [torch.randint(0, dims[-1], (rows,), device=device, generator=g, dtype=torch.int32) for _ in range(pool_n)]

    l0 = _lib.launch_count()
    tr.train_step(xs[0], ys[0])
    torch.cuda.synchronize(device)
    launches_per_step = _lib.launch_count() - l0

    use_graph = workload in ("mnist4096", "narrow65536")
    graphs = []
    if use_graph:
        tr.capture(rows, x_dtype=xdt)  # warms up + validates capture
        for i in range(pool_n):
            gph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gph):
                tr.train_step(xs[i], ys[i])
            graphs.append(gph)

        def step(i):
            graphs[i % pool_n].replay()
    else:
        def step(i):
            tr.train_step(xs[i % pool_n], ys[i % pool_n])

    ms = timed(step, steps, warmup, device, barrier)
    res = {"rows": rows, "dims": dims, "dtype": dtype, "ms_total": ms, "launches_per_step": launches_per_step,
           "pool_batches": pool_n, "pool_bytes": pool_n * batch_bytes, "cuda_graph": use_graph}
    # ---- e2e: HOST buffers, H2D of the batch + D2H of the loss inside the timed region, through
    # the public call Trainer.train_step_host ----
    # host rows in the compute dtype (like the wide nets' e2e): a bf16 model is fed bf16 rows
    hx = [torch.rand((rows, dims[0]), dtype=torch.float32).to(xdt).pin_memory() for _ in range(4)]
    hy = [torch.randint(0, dims[-1], (rows,), dtype=torch.int32).pin_memory() for _ in range(4)]
    # the host loop is PCIe-bound at ~1 ms per step or less: time enough steps for a stable figure
    e2e_steps = max(200 if use_graph else 3, min(steps, 400))
    if use_graph:
        # Trainer.fit_stream: H2D of batch i+1 overlaps the step on batch i; every step's loss comes back
        tr.fit_stream([(hx[i & 3], hy[i & 3]) for i in range(4)])  # untimed: builds the two slot graphs

        def e2e_run(_i):
            losses = tr.fit_stream([(hx[i & 3], hy[i & 3]) for i in range(e2e_steps)])
            assert len(losses) == e2e_steps and all(np.isfinite(losses))
        ms_e2e = timed(e2e_run, 1, 1, device, barrier)
        api = "Trainer.fit_stream (pinned host batches in the compute dtype -> H2D on a copy stream overlapped with the step -> loss D2H)"
    else:
        def e2e_step(i):
            tr.train_step_host(hx[i & 3], hy[i & 3])
        ms_e2e = timed(e2e_step, e2e_steps, min(warmup, 5), device, barrier)
        api = "Trainer.train_step_host (pinned host batch -> H2D -> step -> loss D2H)"
    res.update({"e2e_ms_total": ms_e2e, "e2e_steps": e2e_steps, "h2d": rows * dims[0] * elt + rows * 4, "d2h": 4,
                "e2e_api": api})
    return res, tr, xs, ys


def _kernel_work(L, rows, dtype):
    """Algorithmic (FLOPs, bytes) per launch of every entry point of one train step."""
    elt = 4 if dtype == "f32" else 2
    work = {}
    n_layers = len(L.layer_params)
    for i, p in enumerate(L.layer_params):
        k, n = p["weights"].shape
        last = i == n_layers - 1
        fl = 2.0 * rows * k * n
        work[f"forward L{i + 1} {k}->{n}"] = (fl, rows * k * elt + k * n * elt + rows * n * (4 if last else elt))
        work[f"dX L{i + 1} {n}->{k}"] = (fl, rows * n * elt + k * n * elt + 2 * rows * k * elt)
        work[f"dW+SGD L{i + 1} {k}x{n}"] = (fl, rows * k * elt + rows * n * elt + 2 * k * n * 4 +
                                            (k * n * 2 if dtype == "bf16" else 0))
    k_out, n_out = L.layer_params[-1]["weights"].shape
    # fused skinny output layer: one read of its input, one write of the gradient below it, labels, row losses, dZ
    work[f"output layer fused {k_out}->{n_out}"] = (6.0 * rows * k_out * n_out,
                                                    2 * rows * k_out * elt + rows * 8 + rows * n_out * elt)
    if n_layers >= 2:  # fused two-layer tail: A1 read once, dZ1 written once (+ labels, row losses)
        k1, n2 = L.layer_params[-2]["weights"].shape
        work[f"tail fused {k1}->{n2}->{n_out}"] = (6.0 * rows * (k1 * n2 + n2 * n_out),
                                                   2 * rows * k1 * elt + rows * 8)
    work[f"softmax-CE N={n_out}"] = (5.0 * rows * n_out, rows * n_out * (4 + elt) + rows * 8)
    return work


def kernel_roofline(tr, xs, ys, dtype, steps, device, use_graph):
    """Rank-0 only, no barriers: per-entry-point timing with CUDA events on the launch stream.

    Wide nets (kernels of hundreds of microseconds): in situ -- eager train steps on the rotating
    resident batches with an event pair around every entry-point call.
    Graph-captured narrow nets (kernels of a few microseconds, where an eager launch is slower than
    the kernel): every entry point is captured alone into a CUDA graph per resident batch and the
    graphs are replayed back to back, rotating the batches; intermediates stay where the real step
    leaves them (L2), the input batch comes from HBM.
    Returns one row per entry point with its algorithmic FLOPs and bytes per launch."""
    import torch
    from ddnn_b200 import ops
    L = tr.layer
    rows = xs[0].shape[0]
    work = _kernel_work(L, rows, dtype)
    reps = max(5, min(steps, 50))
    for i in range(3):
        tr.train_step(xs[i % len(xs)], ys[i % len(xs)])
    torch.cuda.synchronize(device)
    acc = {}
    if not use_graph:
        L._probe = []
        for i in range(reps):
            tr.train_step(xs[i % len(xs)], ys[i % len(xs)])
        torch.cuda.synchronize(device)
        spans, L._probe = L._probe, None
        for name, e0, e1 in spans:
            acc.setdefault(name, []).append(e0.elapsed_time(e1))
        how = "CUDA events around the entry point inside eager train steps (in situ)"
    else:
        b = L._bufs
        calls = []
        n_layers = len(L.layer_params)
        xin = xs if xs[0].dtype == L.wire_dtype else None  # layer-1 operands straight from the batch pool
        fused_tail = L.can_fuse_tail2()
        fused_out = L.can_fuse_output_layer() and not fused_tail
        for i, p in enumerate(L.layer_params):
            k, n = p["weights"].shape
            if fused_tail and i >= n_layers - 2:  # what train_step launches for the last two layers
                if i == n_layers - 2:
                    n3 = L.layer_params[-1]["weights"].shape[1]
                    calls.append((f"tail fused {k}->{n}->{n3}", False,
                                  lambda j: L.tail2_kernel(ys[j % len(ys)], rows, ws=L._ws_side)))
                    calls.append(("tail reducers + SGD (side stream in the step)", False,
                                  lambda j: L.tail2_finish(0.0, rows, ws=L._ws_side)))
                continue
            if fused_out and i == n_layers - 1:  # what train_step launches for this layer: one fused entry point
                calls.append((f"output layer fused {k}->{n}", False,
                              lambda j: L.output_layer_step(ys[j % len(ys)], 0.0, rows)))
                continue
            act = "linear" if i == n_layers - 1 else p["activation"]
            src = (lambda j, i=i: xin[j % len(xin)]) if (i == 0 and xin) else (lambda j, i=i: b["a"][i][:rows])
            calls.append((f"forward L{i + 1} {k}->{n}", i == 0 and bool(xin),
                          lambda j, i=i, p=p, act=act, src=src: ops.dense_forward(
                              src(j), L._w(i), p["biases"], act, out=b["a"][i + 1][:rows])))
            if i > 0:
                calls.append((f"dX L{i + 1} {n}->{k}", False,
                              lambda j, i=i: ops.dense_backward_dx(
                                  b["dz"][i][:rows], L._w(i), b["a"][i][:rows],
                                  L.layer_params[i - 1]["activation"], out=b["dz"][i - 1][:rows])))
            calls.append((f"dW+SGD L{i + 1} {k}x{n}", i == 0 and bool(xin),
                          lambda j, i=i, p=p, src=src: ops.dense_backward_dw_sgd(
                              src(j), b["dz"][i][:rows], p["weights"], p["biases"], 0.0, w_shadow=L._shadow[i],
                              workspace=L._ws)))
        if not fused_out and not fused_tail:
            calls.append((f"softmax-CE N={L.layer_params[-1]['weights'].shape[1]}", False,
                          lambda j: ops.softmax_ce(b["a"][-1][:rows], ys[j % len(ys)], b["dz"][-1][:rows],
                                                   b["row_loss"][:rows], b["loss"], 1.0 / rows)))
        for name, rotates, fn in calls:
            graphs = []
            for j in range(len(xs) if rotates else 1):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    fn(j)
                graphs.append(g)
            ms = timed(lambda j: graphs[j % len(graphs)].replay(), reps * 4, 5, device)
            acc[name] = [ms / (reps * 4)]
        how = "each entry point captured alone in a CUDA graph per resident batch, replayed back to back"
    rows_out = []
    for name, ts in acc.items():
        t = max(1e-6, float(np.mean(ts)))
        fl, by = work.get(name, (0.0, 0.0))
        rows_out.append({"kernel": name, "ms": t, "tflops": fl / t / 1e9, "gbs": by / t / 1e6, "flops": fl,
                         "bytes": by, "how": how})
    return rows_out


def roofline_of(ks, dtype, peaks, workload, in_step: bool):
    """`roofline` object for the entry point with the largest share of the step.  Tensor bound (bf16 GEMMs above
    the ridge of 255 FLOP/B) against the measured cuBLAS bf16 rate -- the sustained figure for a kernel timed
    inside a long step, the burst one for a kernel timed alone; HBM bound against the measured copy bandwidth."""
    dom = max(ks, key=lambda r: r["ms"] * r.get("launches", 1))
    tensor = dtype == "bf16" and dom["bytes"] > 0 and dom["flops"] / dom["bytes"] > 255
    if tensor:
        peak = peaks["bf16_tflops_sustained"] if in_step else peaks["bf16_tflops"]
        roof = {"bound": "tensor", "achieved": dom["tflops"], "peak": peak, "unit": "TFLOP/s", "frac": dom["tflops"] / peak,
                "peak_kind": "sustained (kernel timed inside a long step; a pipeline stage that idles between ops runs "
                             "above the sustained clock, so frac can exceed 1 at N > 1: see frac_of_burst)"
                             if in_step else "burst (kernel timed alone)",
                "frac_of_burst": dom["tflops"] / peaks["bf16_tflops"]}
    else:
        roof = {"bound": "hbm", "achieved": dom["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": dom["gbs"] / peaks["hbm_gbs"]}
    traffic = None
    try:
        with open(os.path.join(REPO, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get(workload, {}).get(dom["kernel"])
    except Exception:  # noqa: BLE001
        pass
    total = sum(r["ms"] * r.get("launches", 1) for r in ks)
    roof.update({"traffic": traffic, "algorithmic_bytes": dom["bytes"], "algorithmic_flops": dom["flops"],
                 "kernel": dom["kernel"], "ms_per_launch": dom["ms"], "launches_per_step": dom.get("launches", 1),
                 "peak_source": peaks["source"], "kernel_tflops": dom["tflops"],
                 "step_share": dom["ms"] * dom.get("launches", 1) / total, "timing": dom["how"]})
    return roof


def _compact(ks):
    return [{k: (round(v, 6) if isinstance(v, float) else v) for k, v in r.items()
             if k in ("kernel", "ms", "launches", "tflops", "gbs")} for r in ks]


def small_net_entry(workload, dtype, steps, warmup, device, rank, world, barrier, max_over_ranks, peaks):
    """A net that lives on one GPU (MNIST / narrow): device-resident value, e2e through Trainer.fit_stream,
    per-entry-point roofline.  At world > 1 every rank trains its own replica (no data-path collective)."""
    res, tr, xs, ys = bench_single(workload, dtype, steps, warmup, device, barrier)
    ms_total = max_over_ranks(res["ms_total"])
    e2e_ms = max_over_ranks(res["e2e_ms_total"])
    dims, rows = res["dims"], res["rows"]
    elt = 4 if res["dtype"] == "f32" else 2
    value = world * rows * steps / (ms_total / 1e3)
    fps = flops_per_sample(dims)
    ent = {"workload": workload, "value": value, "unit": "samples/s", "ms_per_step": ms_total / steps, "steps": steps,
           "scaling": "weak", "dtype": res["dtype"], "rows_per_step_per_gpu": rows, "cuda_graph": res["cuda_graph"],
           "launches_per_step": int(res["launches_per_step"]),
           "l2": f"inputs rotate over {res['pool_batches']} distinct resident batches "
                 f"({res['pool_bytes'] / 2**20:.0f} MiB > 126 MiB L2)",
           "tflops": value * fps / 1e12,
           # activation bytes per sample x samples/s over the measured HBM peak
           "step_hbm_bytes_per_sample": bytes_per_sample(dims, elt),
           "step_hbm_frac_of_measured": value * bytes_per_sample(dims, elt) / (world * peaks["hbm_gbs"] * 1e9),
           "e2e": {"value": world * rows * res["e2e_steps"] / (e2e_ms / 1e3), "unit": "samples/s",
                   "h2d_bytes_per_step": res["h2d"], "d2h_bytes_per_step": res["d2h"], "steps": res["e2e_steps"],
                   "api": res["e2e_api"]}}
    if rank == 0:
        ks = kernel_roofline(tr, xs, ys, res["dtype"], steps, device, res["cuda_graph"])
        ent["roofline"] = roofline_of(ks, res["dtype"], peaks, workload, in_step=not res["cuda_graph"])
        ent["kernels"] = _compact(ks)
    del tr, xs, ys
    return ent


def wide_entry(workload, steps, warmup, device, rank, world, barrier, max_over_ranks, peaks, e2e_steps=8):
    """A wide bf16 net through the layer-per-GPU pipeline on `world` GPUs (world = 1: one stage holds every
    layer): the SAME global batch at every world size (strong scaling).  Besides the throughput, the first
    losses and an exact checksum of the post-run weights are reported: the pipeline is built to give the same
    bits at every stage count."""
    import torch
    import torch.distributed as dist
    from ddnn_b200 import Layer, _lib
    from ddnn_b200.pipeline import PipelineStage, stage_assignment
    dims, acts, rows, dtype, micro = WORKLOADS[workload]
    n_layers = len(acts)
    if world > n_layers or n_layers % world:
        return {"workload": workload, "skipped": f"{n_layers} layers do not split over {world} GPUs"}
    lo, hi = stage_assignment(n_layers, world)[rank]
    # every rank draws the full parameter stream so that stage weights do not depend on `world`
    g = torch.Generator(device=device).manual_seed(1)
    params = []
    for i, (k, n, a) in enumerate(zip(dims[:-1], dims[1:], acts)):
        std = ((2.0 if a == "relu" else 1.0) / k) ** 0.5
        w = torch.randn((k, n), device=device, generator=g, dtype=torch.float32) * std
        if lo <= i < hi:
            params.append({"weights": w, "biases": torch.full((n,), 0.01, device=device), "activation": a})
        del w
    layer = Layer.from_params(params, dims[lo], container_name=f"stage{rank}", dtype=dtype, device=device,
                              input_activation=(acts[lo - 1] if lo > 0 else "linear"), is_output=(hi == n_layers))
    del params
    transport = os.environ.get("DDNN_TRANSPORT", "peer") if world > 1 else "local"
    schedule = os.environ.get("DDNN_SCHEDULE", PIPE_SCHEDULE)
    dw_chunk = int(os.environ.get("DDNN_DW_CHUNK", str(max(1, PIPE_DW_CHUNK_ROWS // (rows // micro))))) or None
    stage = PipelineStage(layer, rank, world, transport=transport if world > 1 else "nccl", schedule=schedule,
                          dw_chunk=dw_chunk, layers_per_stage=[b - a for a, b in stage_assignment(n_layers, world)])
    first, last = rank == 0, rank == world - 1
    # batch and labels from their own generators: the same data whatever the stage count
    x = torch.rand((rows, dims[0]), device=device, generator=torch.Generator(device=device).manual_seed(0)
                   ).to(torch.bfloat16) if first else None
    y = torch.randint(0, dims[-1], (rows,), device=device, generator=torch.Generator(device=device).manual_seed(2),
                      dtype=torch.int32) if last else None
    l0 = _lib.launch_count()
    first_losses = []

    def step(i):
        loss = stage.train_step(x, y, rows, micro, LR)
        if i < 3 and last:
            first_losses.append(float(loss.item()))
    ms = max_over_ranks(timed(step, steps, warmup, device, barrier))
    launches = (_lib.launch_count() - l0) // max(1, steps + warmup)
    # exact, order-independent checksum of the fp32 master weights after warmup + steps updates
    sums = [int(p["weights"].view(torch.int32).to(torch.int64).sum().item()) for p in layer.layer_params]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, (sums, first_losses))
        sums = [v for part in gathered for v in part[0]]
        first_losses = gathered[-1][1]
    value = rows * steps / (ms / 1e3)
    fps = flops_per_sample(dims)
    ent = {"workload": workload, "stages": world, "layers_per_stage": hi - lo, "micro_batches": micro,
           "transport": transport, "schedule": schedule, "dw_chunk_micro_batches": dw_chunk, "rows_per_step": rows, "steps": steps, "value": value,
           "unit": "samples/s", "ms_per_step": ms / steps, "scaling": "strong", "dtype": dtype,
           "launches_per_step": int(launches), "tflops_total": value * fps / 1e12,
           "tensor_frac_of_measured_burst": value * fps / 1e12 / (world * peaks["bf16_tflops"]),
           "tensor_frac_of_measured_sustained": value * fps / 1e12 / (world * peaks["bf16_tflops_sustained"]),
           # slots of one GEMM per micro-batch and layer: 3M of work per stage (2M + M(L-1)/L on the first, whose
           # layer 1 has no dX) plus the fill / drain slots the order leaves idle: 2(S-1) GPipe, (S-1) zero-bubble
           "bubble_bound": (3 * micro) / (3 * micro + (2 if schedule == "gpipe" else 1) * (world - 1)),
           "loss_first_steps": first_losses, "loss_first_steps_hex": [float(v).hex() for v in first_losses],
           "weights_checksum": f"{sum(sums) & 0xFFFFFFFFFFFFFFFF:016x}",
           "weights_checksum_per_layer": [f"{v & 0xFFFFFFFFFFFFFFFF:016x}" for v in sums],
           "checksum_of": f"int64 sum of the fp32 master weights' bit patterns after {warmup + steps} steps"}
    # ---- per-entry-point timing, in situ: CUDA events around every call inside 2 more steps (rank 0's stage) ----
    layer._probe = []
    for _ in range(2):
        stage.train_step(x, y, rows, micro, LR)
    torch.cuda.synchronize(device)
    spans, layer._probe = layer._probe, None
    if rank == 0:
        work = {}
        mb = rows // micro
        for i, p in enumerate(layer.layer_params):
            k, n = p["weights"].shape
            is_out = (i == len(layer.layer_params) - 1) and layer.is_output
            work[f"forward L{i + 1} {k}->{n}"] = (2.0 * mb * k * n, mb * k * 2 + k * n * 2 + mb * n * (4 if is_out else 2))
            work[f"dX L{i + 1} {n}->{k}"] = (2.0 * mb * k * n, mb * n * 2 + k * n * 2 + 2 * mb * k * 2)
            work[f"dW+SGD L{i + 1} {k}x{n}"] = (2.0 * rows * k * n, rows * k * 2 + rows * n * 2 + 2 * k * n * 4 + k * n * 2)
            crows = (dw_chunk or micro) * mb
            work[f"dW chunk L{i + 1} {k}x{n}"] = (2.0 * crows * k * n, crows * (k + n) * 2 + 2 * k * n * 4)
            work[f"SGD apply L{i + 1} {k}x{n}"] = (2.0 * k * n, k * n * (4 + 4 + 4 + 2))
        work[f"softmax-CE N={dims[-1]}"] = (5.0 * mb * dims[-1], mb * dims[-1] * 6 + mb * 8)
        acc = {}
        for name, e0, e1 in spans:
            acc.setdefault(name, []).append(e0.elapsed_time(e1))
        ks = []
        for name, ts in acc.items():
            t = max(1e-6, float(np.mean(ts)))
            fl, by = work.get(name, (0.0, 0.0))
            ks.append({"kernel": name, "ms": t, "launches": len(ts) // 2, "tflops": fl / t / 1e9, "gbs": by / t / 1e6,
                       "flops": fl, "bytes": by,
                       "how": "CUDA events around the entry point on its launch stream inside pipeline steps (in situ)"})
        # the GEMM that carries the wait for the next stage's gradient polls inside the kernel: its span is not GEMM time
        waits = f"dX L{len(layer.layer_params)} " if world > 1 else "\0"
        ent["roofline"] = roofline_of([r for r in ks if not r["kernel"].startswith(waits)], dtype, peaks, workload,
                                      in_step=True)
        ent["kernels"] = _compact(ks)
    # ---- e2e: HOST batches through PipelineStage.fit_stream (H2D of the batch on the first stage, micro-batch by
    # micro-batch under the previous step; labels H2D and loss D2H on the last stage) ----
    try:
        hx, hy = [None, None], [None, None]
        if first:  # two pinned host batches (filled from the device batch: host-side RNG of 1 Gi values is slow)
            hx = [torch.empty((rows, dims[0]), dtype=torch.bfloat16, pin_memory=True) for _ in range(2)]
            hx[0].copy_(x)
            hx[1].copy_(x.flip(0))
        if last:
            hy = [torch.empty((rows,), dtype=torch.int32, pin_memory=True) for _ in range(2)]
            hy[0].copy_(y)
            hy[1].copy_(y.flip(0))
        torch.cuda.synchronize(device)
        stage.fit_stream([(hx[i & 1], hy[i & 1]) for i in range(2)], rows, micro, LR)  # untimed: slots, pinned pages

        def e2e_run(_i):
            losses = stage.fit_stream([(hx[i & 1], hy[i & 1]) for i in range(e2e_steps)], rows, micro, LR)
            assert not last or (len(losses) == e2e_steps and all(np.isfinite(losses)))
        e2e_ms = max_over_ranks(timed(e2e_run, 1, 0, device, barrier))
        ent["e2e"] = {"value": rows * e2e_steps / (e2e_ms / 1e3), "unit": "samples/s",
                      "h2d_bytes_per_step": rows * dims[0] * 2 + rows * 4, "d2h_bytes_per_step": 4, "steps": e2e_steps,
                      "api": "PipelineStage.fit_stream (pinned host batch in the compute dtype -> H2D per micro-batch on a "
                             "copy stream, overlapped with the previous step -> pipeline step -> loss D2H)"}
        del hx, hy
    except Exception as e:  # noqa: BLE001
        ent["e2e"] = {"error": repr(e)[:300]}
    del stage, layer, x, y
    torch.cuda.empty_cache()
    return ent


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--dtype", default=None, choices=["f32", "bf16"])
    ap.add_argument("--no-extra", action="store_true", help="headline workload only: no extras, no cpu_baseline")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for cpu_baseline")
    args = ap.parse_args()

    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime
        dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=600))
    barrier = (lambda: dist.barrier(device_ids=[local])) if world > 1 else None

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    peaks = load_peaks()
    steps, warmup = max(1, args.steps), max(3, args.warmup)
    workload = args.workload or DEFAULT_WORKLOAD
    is_wide = workload.startswith("wide")

    sampler = ClockSampler(local)
    sampler.start()
    if is_wide:
        head = wide_entry(workload, steps, warmup, device, rank, world, barrier, max_over_ranks, peaks)
        if "skipped" in head:
            raise SystemExit(f"bench.py: {head['skipped']}")
    else:
        head = small_net_entry(workload, args.dtype, steps, warmup, device, rank, world, barrier, max_over_ranks, peaks)
    clocks = sampler.stop()
    torch.cuda.empty_cache()

    dims = WORKLOADS[workload][0]
    fps = flops_per_sample(dims)
    line = {
        "metric": "train samples/sec (fwd+bwd+SGD)", "value": head["value"], "unit": "samples/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": head["scaling"], "vs_baseline": None, "dtype": head["dtype"], "data": "synthetic",
        "config": line_config(workload, world),
        "clocks": clocks,
        "gpu_launches": int(head["launches_per_step"]) * steps,
        "model_flops_per_sample": fps,
        "tflops": head["value"] * fps / 1e12,
        "tensor_frac_of_measured_burst": head["value"] * fps / 1e12 / (world * peaks["bf16_tflops"]),
        "tensor_frac_of_measured_sustained": head["value"] * fps / 1e12 / (world * peaks["bf16_tflops_sustained"]),
        "peaks": peaks,
    }
    for k in ("e2e", "roofline", "kernels", "stages", "transport", "schedule", "dw_chunk_micro_batches", "bubble_bound", "loss_first_steps",
              "loss_first_steps_hex", "weights_checksum", "weights_checksum_per_layer", "checksum_of", "cuda_graph", "l2",
              "step_hbm_bytes_per_sample", "step_hbm_frac_of_measured"):
        if k in head:
            line[k] = head[k]

    # ---- extras: the other configs, each a full report (value, e2e, roofline); the CPU baseline ----
    if not args.no_extra:
        extras = []
        es, ew = max(steps, 20), max(warmup, 5)
        for wl in ("mnist4096", "narrow65536", "wide4x8192", "wide8x16384"):
            if wl == workload:
                continue
            try:
                if wl.startswith("wide"):
                    if wl == "wide8x16384" or world > 4:
                        continue  # the 16384-wide net is the default headline; 4 layers do not split over 8 GPUs
                    extras.append(wide_entry(wl, min(es, 12), 3, device, rank, world, barrier, max_over_ranks, peaks))
                elif world == 1:  # the small nets do not pipeline: measured on one GPU only
                    extras.append(small_net_entry(wl, None, max(es, 50), ew, device, rank, world, barrier,
                                                  max_over_ranks, peaks))
            except Exception as e:  # noqa: BLE001
                extras.append({"workload": wl, "error": repr(e)[:300]})
            torch.cuda.empty_cache()
        line["extras"] = extras
        if rank == 0 and world == 1:
            v, desc, ms, n = cpu_train_samples_per_s(workload, args.cpu_budget, warmup=1)
            line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cpu_threads(), "kind": "port",
                                    "sample": desc}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
