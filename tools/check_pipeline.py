"""Multi-GPU check (run under torchrun, one rank per GPU): the peer-memory transport must give
bit-identical weights and losses to the NCCL send/recv transport, and both must match a
single-GPU run of the same model.   torchrun --nproc-per-node 2 tools/check_pipeline.py"""
import datetime
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import Layer, Trainer  # noqa: E402
from ddnn_b200.config import make_params, synthetic_batch  # noqa: E402
from ddnn_b200.pipeline import PipelineStage, stage_assignment  # noqa: E402


def run(transport, dims, acts, dtype, rows, micro, steps, rank, world, device, schedule="gpipe", dw_chunk=None):
    params = make_params(dims, acts, seed=3)
    lo, hi = stage_assignment(len(acts), world)[rank]
    layer = Layer.from_params(params[lo:hi], dims[lo], dtype=dtype, device=device,
                              input_activation=(acts[lo - 1] if lo > 0 else "linear"), is_output=(hi == len(acts)))
    stage = PipelineStage(layer, rank, world, transport=transport, schedule=schedule, dw_chunk=dw_chunk,
                          layers_per_stage=[b - a for a, b in stage_assignment(len(acts), world)])
    losses = []
    for s in range(steps):
        x, y = synthetic_batch(rows, dims[0], dims[-1], seed=s)
        xd = torch.as_tensor(x, dtype=torch.float32).to(device) if rank == 0 else None
        yd = torch.as_tensor(y, dtype=torch.int32).to(device) if rank == world - 1 else None
        l = stage.train_step(xd, yd, rows, micro, 0.05)
        if l is not None:
            losses.append(float(l.item()))
    torch.cuda.synchronize(device)
    return [p["weights"].clone() for p in layer.layer_params], losses, (lo, hi)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=120))
    ok = True
    for dims, acts, dtype, rows, micro in [([512] * 5, ["relu", "sigmoid", "relu", "softmax"], "bf16", 2048, 4),
                                           ([64, 48, 40, 32, 10], ["relu", "relu", "sigmoid", "softmax"], "f32", 300, 3),
                                           ([256] * 9, ["relu", "sigmoid"] * 3 + ["relu", "softmax"], "bf16", 1024, 5)]:
        if world > len(acts):
            continue
        w_n, l_n, (lo, hi) = run("nccl", dims, acts, dtype, rows, micro, 4, rank, world, device)
        w_p, l_p, _ = run("peer", dims, acts, dtype, rows, micro, 4, rank, world, device)
        same = all(torch.equal(a, b) for a, b in zip(w_n, w_p)) and l_n == l_p
        # chunked weight gradients: the GPipe and the zero-bubble order must agree bit for bit (same chunk boundaries),
        # over NCCL as over peer memory, and stay within rounding of the one-piece gradient
        w_c, l_c, _ = run("nccl", dims, acts, dtype, rows, micro, 4, rank, world, device, "gpipe", 2)
        w_z, l_z, _ = run("peer", dims, acts, dtype, rows, micro, 4, rank, world, device, "zb", 2)
        same = same and all(torch.equal(a, b) for a, b in zip(w_c, w_z)) and l_c == l_z
        chunk_err = max(float((a - b).abs().max() / (b.abs().max() + 1e-30)) for a, b in zip(w_z, w_p))
        same = same and chunk_err < 1e-5
        # single-GPU reference of the same steps (every rank computes it; cheap)
        tr = Trainer(make_params(dims, acts, seed=3), dtype=dtype, device=device)
        ref_l = []
        for s in range(4):
            x, y = synthetic_batch(rows, dims[0], dims[-1], seed=s)
            ref_l.append(float(tr.train_step(torch.as_tensor(x, dtype=torch.float32).to(device),
                                             torch.as_tensor(y, dtype=torch.int32).to(device), 0.05).item()))
        ref_w = [p["weights"] for p in tr.layer.layer_params][lo:hi]
        tol = 1e-5 if dtype == "f32" else 2e-2
        err = max(float((a - b).abs().max() / (b.abs().max() + 1e-30)) for a, b in zip(w_p, ref_w))
        l_ok = (not l_p) or max(abs(a - b) / abs(b) for a, b in zip(l_p, ref_l)) < tol
        flag = torch.tensor([int(same and err < tol and l_ok)], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(f"{dtype} dims={dims} stages={world} micro={micro}: peer==nccl {bool(flag.item())} "
                  f"(rank0: bit-identical={same}, vs single-GPU weights rel err {err:.2e})", flush=True)
        ok = ok and bool(flag.item())
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()
    if rank == 0:
        print("PIPELINE CHECK", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
