"""Per-stage timeline of one pipeline step (run under torch.distributed.run, one rank per GPU):
CUDA events around every GEMM / softmax-CE entry point of every micro-batch on every stage, as offsets from a
common step start (all ranks leave a barrier together; clocks of different GPUs are not compared, only offsets on
one GPU).  Gaps between consecutive spans of a stage are what that stage spent waiting for a neighbour's flag.
Writes profiles-style CSV  <out>_rank<r>.csv  (name, start_ms, dur_ms) and prints a summary per rank:
un-probed ms/step, sum of kernel spans, idle before the first backward span (fill), idle inside the backward
phase, idle before the update.

  python -m torch.distributed.run --nproc-per-node N tools/pipeline_timeline.py [workload] [out_prefix]
"""
import datetime
import os
import sys

import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from bench import LR, WORKLOADS  # noqa: E402
from ddnn_b200 import Layer  # noqa: E402
from ddnn_b200.pipeline import PipelineStage, stage_assignment  # noqa: E402


def nvlink_kib(index: int):
    """(tx, rx) KiB summed over this GPU's NVLink links, from `nvidia-smi nvlink -gt d`; None when unavailable."""
    import re
    import subprocess
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(index)], stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, timeout=20).stdout
    except Exception:  # noqa: BLE001
        return None
    tx = sum(int(v) for v in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(v) for v in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return (tx, rx) if (tx or rx) else None


def main():
    workload = sys.argv[1] if len(sys.argv) > 1 else "wide8x16384"
    out = sys.argv[2] if len(sys.argv) > 2 else os.path.join(REPO, "gpurun_out", "timeline")
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=300))
    dims, acts, rows, dtype, micro = WORKLOADS[workload]
    n_layers = len(acts)
    lo, hi = stage_assignment(n_layers, world)[rank]
    g = torch.Generator(device=dev).manual_seed(1)
    params = []
    for i, (k, n, a) in enumerate(zip(dims[:-1], dims[1:], acts)):
        w = torch.randn((k, n), device=dev, generator=g) * ((2.0 if a == "relu" else 1.0) / k) ** 0.5
        if lo <= i < hi:
            params.append({"weights": w, "biases": torch.full((n,), 0.01, device=dev), "activation": a})
        del w
    layer = Layer.from_params(params, dims[lo], dtype=dtype, device=dev, container_name=f"stage{rank}",
                              input_activation=(acts[lo - 1] if lo > 0 else "linear"), is_output=(hi == n_layers))
    stage = PipelineStage(layer, rank, world, transport=os.environ.get("DDNN_TRANSPORT", "peer") if world > 1 else "nccl",
                          schedule=os.environ.get("DDNN_SCHEDULE", "gpipe"),
                          dw_chunk=int(os.environ.get("DDNN_DW_CHUNK", "0")) or None,
                          layers_per_stage=[b - a for a, b in stage_assignment(n_layers, world)])
    x = torch.rand((rows, dims[0]), device=dev, generator=torch.Generator(device=dev).manual_seed(0)).to(torch.bfloat16) \
        if rank == 0 else None
    y = torch.randint(0, dims[-1], (rows,), device=dev, generator=torch.Generator(device=dev).manual_seed(2),
                      dtype=torch.int32) if rank == world - 1 else None

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize(dev)

    for _ in range(2):
        stage.train_step(x, y, rows, micro, LR)
    barrier()
    nv0 = nvlink_kib(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        stage.train_step(x, y, rows, micro, LR)
    e1.record()
    barrier()
    nv1 = nvlink_kib(local)
    plain_ms = e0.elapsed_time(e1) / 3
    nv = ""
    if nv0 and nv1:  # what the fused epilogue stores put on the wire per step (activations out / gradients in, ...)
        nv = f" | NVLink per step: tx {(nv1[0] - nv0[0]) / 3 / 2**20:.3f} GiB, rx {(nv1[1] - nv0[1]) / 3 / 2**20:.3f} GiB"
    layer._probe = []
    barrier()
    start = torch.cuda.Event(enable_timing=True)
    start.record()
    stage.train_step(x, y, rows, micro, LR)
    end = torch.cuda.Event(enable_timing=True)
    end.record()
    barrier()
    spans, layer._probe = layer._probe, None
    rows_out = [(name, start.elapsed_time(a), a.elapsed_time(b)) for name, a, b in spans]
    tag = os.environ.get("DDNN_SCHEDULE", "gpipe") + (f"_chunk{os.environ['DDNN_DW_CHUNK']}" if os.environ.get("DDNN_DW_CHUNK") else "")
    with open(f"{out}_{workload}_{tag}_n{world}_rank{rank}.csv", "w") as f:
        f.write("name,start_ms,dur_ms\n")
        for name, s, d in rows_out:
            f.write(f"{name},{s:.4f},{d:.4f}\n")
    busy = sum(d for _, _, d in rows_out)
    fwd = [r for r in rows_out if r[0].startswith("forward")]
    bwd = [r for r in rows_out if r[0].startswith("dX") or r[0].startswith("softmax")]
    upd = [r for r in rows_out if r[0].startswith("dW") or r[0].startswith("SGD")]

    def phase(rs):
        if not rs:
            return (0.0, 0.0, 0.0)
        t0, t1 = rs[0][1], rs[-1][1] + rs[-1][2]
        return (t0, t1, (t1 - t0) - sum(d for _, _, d in rs))
    f0, f1, fidle = phase(fwd)
    b0, b1, bidle = phase(bwd)
    u0, u1, uidle = phase(upd)
    line = (f"rank {rank}/{world} {workload}: step {plain_ms:.2f} ms un-probed, {start.elapsed_time(end):.2f} ms probed; "
            f"kernels {busy:.2f} ms | forward {f0:.2f}-{f1:.2f} (idle inside {fidle:.2f}) | backward {b0:.2f}-{b1:.2f} "
            f"(idle inside {bidle:.2f}) | update {u0:.2f}-{u1:.2f} (idle inside {uidle:.2f}) | mean GEMM: fwd "
            f"{sum(d for _, _, d in fwd) / max(1, len(fwd)):.4f} ms, dX "
            f"{sum(r[2] for r in bwd if r[0].startswith('dX')) / max(1, len([r for r in bwd if r[0].startswith('dX')])):.4f} ms, "
            f"dW {sum(d for _, _, d in upd) / max(1, len(upd)):.3f} ms{nv}")
    if world > 1:
        lines = [None] * world
        dist.all_gather_object(lines, line)
    else:
        lines = [line]
    if rank == 0:
        print("\n".join(lines), flush=True)
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
