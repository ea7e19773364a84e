#!/usr/bin/env python
"""bench.py -- train samples/s (forward + softmax-CE + backward + SGD) of the docker-dist-nn
layer hot path on N B200s.

    python bench.py --gpus 1 --steps K --warmup W              # our arm
    python bench.py --impl reference --gpus 1 --steps K ...     # CPU arm (the oracle port)
    torchrun --nproc-per-node N ... bench.py --gpus N ...       # one rank per GPU

A "step" is one pass of the hot path over one batch of synthetic, seeded data.  Default
workload = BASELINE.json configs[1]: the MNIST FCNN 784-128-64-10, batch 4096, fp32 path.
At N > 1 that net is too small to pipeline: every rank trains an
independent replica on its own batches ("weak" scaling, no data-path collective).  The
layer-per-GPU pipeline over NCCL send/recv is exercised on the wide bf16 nets and reported
in the ``pipeline`` object of the same JSON line (config 3: 4x8192; config 4: 8x16384, 32
micro-batches) -- at N = 1 the same nets run as one stage, which is the denominator of the
pipeline's scaling efficiency.

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


def cpu_train_samples_per_s(workload: str, budget_s: float, steps=None, warmup=1, rows_cap=None):
    """Times oracle.train_step (numpy float64, multi-threaded BLAS) on the host cores on a
    bounded sample of the workload.  Returns (samples/s, description, ms_per_step, steps)."""
    import oracle
    from ddnn_b200.config import make_params, synthetic_batch
    dims, acts, rows, _, _ = WORKLOADS[workload]
    if rows_cap:
        rows = min(rows, rows_cap)
    params = make_params(dims, acts, seed=1)
    x, y = synthetic_batch(rows, dims[0], dims[-1], seed=0)
    for _ in range(max(0, warmup)):
        oracle.train_step(params, x, y, 0.01)
    t0 = time.perf_counter()
    n = 0
    while True:
        oracle.train_step(params, x, y, 0.01)
        n += 1
        el = time.perf_counter() - t0
        if (steps is not None and n >= steps) or (steps is None and el >= budget_s):
            break
    desc = f"{n} oracle train steps (numpy float64) of {workload} at {rows} rows/step"
    return rows * n / el, desc, 1e3 * el / n, n


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # other ranks exit 0 without work
    workload = args.workload or "mnist4096"
    # each step a bounded sample: cap rows so that K steps finish within minutes on host cores
    cap = {"mnist4096": 4096, "narrow65536": 8192, "wide4x8192": 64, "wide8x16384": 16}[workload]
    v, desc, ms, n = cpu_train_samples_per_s(workload, 0.0, steps=max(1, args.steps), warmup=max(0, args.warmup),
                                             rows_cap=cap)
    dims = WORKLOADS[workload][0]
    line = {
        "impl": "reference", "metric": "train samples/sec (fwd+bwd+SGD)", "value": v, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": n, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload, "dims": dims, "rows_per_step": min(WORKLOADS[workload][2], cap)},
        "cpu_baseline": {"value": v, "unit": "samples/s", "cores": cpu_threads(), "kind": "port", "sample": desc,
                         "note": "the reference has no backward/SGD (SURVEY.md 0.1) and is Python that cannot "
                                 "travel to the GPU box; this is the float64 numpy oracle pinned to it"},
        "e2e": {"value": v, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
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
    tr.lr = 0.01
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
    hx = [torch.rand((rows, dims[0]), dtype=torch.float32).pin_memory() for _ in range(4)]
    hy = [torch.randint(0, dims[-1], (rows,), dtype=torch.int32).pin_memory() for _ in range(4)]
    e2e_steps = max(3, min(steps, 400))
    if use_graph:
        # Trainer.fit_stream: H2D of batch i+1 overlaps the step on batch i; every step's loss comes back
        tr.fit_stream([(hx[i & 3], hy[i & 3]) for i in range(4)])  # untimed: builds the two slot graphs

        def e2e_run(_i):
            losses = tr.fit_stream([(hx[i & 3], hy[i & 3]) for i in range(e2e_steps)])
            assert len(losses) == e2e_steps and all(np.isfinite(losses))
        ms_e2e = timed(e2e_run, 1, 1, device, barrier)
        api = "Trainer.fit_stream (pinned host batches -> H2D on a copy stream overlapped with the step -> loss D2H)"
    else:
        def e2e_step(i):
            tr.train_step_host(hx[i & 3], hy[i & 3])
        ms_e2e = timed(e2e_step, e2e_steps, min(warmup, 5), device, barrier)
        api = "Trainer.train_step_host (pinned host batch -> H2D -> step -> loss D2H)"
    res.update({"e2e_ms_total": ms_e2e, "e2e_steps": e2e_steps, "h2d": rows * dims[0] * 4 + rows * 4, "d2h": 4,
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


def bench_pipeline(workload, steps, warmup, device, rank, world, barrier):
    """Layer-per-GPU pipeline of a wide net over `world` GPUs (world = 1: one stage)."""
    import torch
    from ddnn_b200 import Layer, _lib
    from ddnn_b200.pipeline import PipelineStage, stage_assignment
    dims, acts, rows, dtype, micro = WORKLOADS[workload]
    n_layers = len(acts)
    if world > n_layers:
        return None
    lo, hi = stage_assignment(n_layers, world)[rank]
    # every rank draws the full parameter stream so that stage weights do not depend on `world`
    g = torch.Generator(device=device).manual_seed(1)
    params = []
    for i, (k, n, a) in enumerate(zip(dims[:-1], dims[1:], acts)):
        std = ((2.0 if a == "relu" else 1.0) / k) ** 0.5
        if lo <= i < hi:
            w = torch.randn((k, n), device=device, generator=g, dtype=torch.float32) * std
            params.append({"weights": w, "biases": torch.full((n,), 0.01, device=device), "activation": a})
    layer = Layer.from_params(params, dims[lo], container_name=f"stage{rank}", dtype=dtype, device=device,
                              input_activation=(acts[lo - 1] if lo > 0 else "linear"), is_output=(hi == n_layers))
    transport = os.environ.get("DDNN_TRANSPORT", "peer") if world > 1 else "nccl"
    stage = PipelineStage(layer, rank, world, transport=transport)
    reserve = int(os.environ.get("DDNN_SM_RESERVE", "0")) if world > 1 else 0
    _lib.set_sm_reserve(reserve)
    gx = torch.Generator(device=device).manual_seed(0)
    x = torch.rand((rows, dims[0]), device=device, generator=gx).to(torch.bfloat16) if rank == 0 else None
    y = torch.randint(0, dims[-1], (rows,), device=device, generator=gx, dtype=torch.int32) if rank == world - 1 else None
    l0 = _lib.launch_count()

    def step(_i):
        stage.train_step(x, y, rows, micro, 0.01)
    ms = timed(step, steps, warmup, device, barrier)
    launches = (_lib.launch_count() - l0) // max(1, steps + warmup)
    _lib.set_sm_reserve(0)
    return {"sm_reserve": reserve, "transport": transport, "workload": workload, "rows": rows, "micro_batches": micro, "ms_total": ms, "dims": dims,
            "launches_per_step": launches, "stages": world, "layers_here": hi - lo}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--dtype", default=None, choices=["f32", "bf16"])
    ap.add_argument("--no-extra", action="store_true", help="skip the pipeline/wide sub-benchmarks and cpu_baseline")
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
        # lazy communicator init (no device_id): each send/recv pair gets its own communicator and
        # stream, so stage-boundary transfers do not serialise behind each other
        import datetime
        dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=240))
    barrier = (lambda: dist.barrier(device_ids=[local])) if world > 1 else None

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    peaks = load_peaks()
    steps, warmup = max(1, args.steps), max(3, args.warmup)
    workload = args.workload or "mnist4096"
    is_pipe_workload = workload.startswith("wide") and world > 1

    sampler = ClockSampler(local)
    sampler.start()
    if is_pipe_workload:
        pr = bench_pipeline(workload, steps, warmup, device, rank, world, barrier)
        res = {"rows": pr["rows"], "dims": pr["dims"], "dtype": "bf16", "ms_total": pr["ms_total"],
               "launches_per_step": pr["launches_per_step"], "cuda_graph": False}
        replicas = 1
    else:
        res, _tr, _xs, _ys = bench_single(workload, args.dtype, steps, warmup, device, barrier)
        res["e2e_ms_total"] = max_over_ranks(res["e2e_ms_total"])
        if rank == 0:
            res["kernels"] = kernel_roofline(_tr, _xs, _ys, res["dtype"], steps, device, res["cuda_graph"])
        del _tr, _xs, _ys
        replicas = world
    ms_total = max_over_ranks(res["ms_total"])
    clocks = sampler.stop()
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier(device_ids=[local])

    dims, rows = res["dims"], res["rows"]
    value = replicas * rows * steps / (ms_total / 1e3)
    fps = flops_per_sample(dims)
    line = {
        "metric": "train samples/sec (fwd+bwd+SGD)", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms_total / steps, "higher_is_better": True,
        "scaling": "strong" if is_pipe_workload else "weak", "vs_baseline": None, "dtype": res["dtype"],
        "data": "synthetic",
        "config": {"workload": workload, "dims": dims, "rows_per_step_per_gpu": rows,
                   "parallelism": (f"pipeline{world}" if is_pipe_workload else
                                   ("single" if world == 1 else f"replicas{world} (net too small to pipeline)")),
                   "l2": f"inputs rotate over {res.get('pool_batches', 1)} distinct resident batches "
                         f"({res.get('pool_bytes', 0) / 2**20:.0f} MiB > 126 MiB L2)",
                   "cuda_graph": res["cuda_graph"], "lr": 0.01},
        "clocks": clocks,
        "gpu_launches": int(res["launches_per_step"]) * steps,
        "model_flops_per_sample": fps,
        "tflops": value * fps / 1e12,
        "tensor_frac_of_measured_burst": value * fps / 1e12 / (world * peaks["bf16_tflops"]),
        # activation bytes per sample x samples/s over the measured HBM peak
        "step_hbm_bytes_per_sample": bytes_per_sample(dims, 4 if res["dtype"] == "f32" else 2),
        "step_hbm_frac_of_measured": value * bytes_per_sample(dims, 4 if res["dtype"] == "f32" else 2) /
                                     (world * peaks["hbm_gbs"] * 1e9),
        "peaks": peaks,
    }
    if "e2e_ms_total" in res:
        e2e_v = replicas * rows * res["e2e_steps"] / (res["e2e_ms_total"] / 1e3)
        line["e2e"] = {"value": e2e_v, "unit": "samples/s", "h2d_bytes_per_step": res["h2d"],
                       "d2h_bytes_per_step": res["d2h"], "steps": res["e2e_steps"],
                       "api": res["e2e_api"]}
    if "kernels" in res:
        ks = res["kernels"]
        dom = max(ks, key=lambda r: r["ms"])
        elt_bound = "tensor" if (res["dtype"] == "bf16" and dom["flops"] / dom["bytes"] > 255) else "hbm"
        if elt_bound == "tensor":
            roof = {"bound": "tensor", "achieved": dom["tflops"], "peak": peaks["bf16_tflops"], "unit": "TFLOP/s",
                    "frac": dom["tflops"] / peaks["bf16_tflops"]}
        else:
            roof = {"bound": "hbm", "achieved": dom["gbs"], "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": dom["gbs"] / peaks["hbm_gbs"]}
        traffic = None
        try:
            with open(os.path.join(REPO, "profiles", "traffic.json")) as f:
                traffic = json.load(f).get(workload, {}).get(dom["kernel"])
        except Exception:  # noqa: BLE001
            pass
        roof.update({"traffic": traffic, "algorithmic_bytes": dom["bytes"], "kernel": dom["kernel"], "ms_per_launch": dom["ms"],
                     "peak_source": peaks["source"], "kernel_tflops": dom["tflops"],
                     "step_share": dom["ms"] / sum(r["ms"] for r in ks),
                     "timing": dom["how"]})
        line["roofline"] = roof
        line["kernels"] = [{k: (round(v, 6) if isinstance(v, float) else v) for k, v in r.items()
                            if k in ("kernel", "ms", "tflops", "gbs")} for r in ks]

    # ---- extras: the wide nets (tensor-core roofline + pipeline), and the CPU baseline ----
    if not args.no_extra and not is_pipe_workload:
        pipes = []
        for wl in ("wide4x8192", "wide8x16384"):
            n_layers = len(WORKLOADS[wl][1])
            if world > n_layers:
                continue
            try:
                psteps = max(2, min(steps, 6 if wl == "wide8x16384" else 12))
                pr = bench_pipeline(wl, psteps, 3, device, rank, world, barrier)
                pms = max_over_ranks(pr["ms_total"])
                v = pr["rows"] * psteps / (pms / 1e3)
                f = flops_per_sample(pr["dims"])
                pipes.append({"workload": wl, "stages": world, "micro_batches": pr["micro_batches"],
                              "sm_reserve": pr["sm_reserve"], "transport": pr["transport"],
                              "rows_per_step": pr["rows"], "steps": psteps, "ms_per_step": pms / psteps,
                              "value": v, "unit": "samples/s", "scaling": "strong", "tflops_total": v * f / 1e12,
                              "tensor_frac_of_measured_burst": v * f / 1e12 / (world * peaks["bf16_tflops"]),
                              "tensor_frac_of_measured_sustained":
                                  v * f / 1e12 / (world * peaks["bf16_tflops_sustained"]),
                              "bubble_bound": (3 * pr["micro_batches"]) /
                                              (3 * pr["micro_batches"] + 2 * (world - 1))})
            except Exception as e:  # noqa: BLE001
                pipes.append({"workload": wl, "error": repr(e)[:300]})
            torch.cuda.empty_cache()
        line["pipeline"] = pipes
    if rank == 0 and not args.no_extra and world == 1:
        v, desc, ms, n = cpu_train_samples_per_s(workload, args.cpu_budget, warmup=1,
                                                 rows_cap={"narrow65536": 8192, "wide4x8192": 64,
                                                           "wide8x16384": 16}.get(workload))
        line["cpu_baseline"] = {"value": v, "unit": "samples/s", "cores": cpu_threads(), "kind": "port", "sample": desc}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
