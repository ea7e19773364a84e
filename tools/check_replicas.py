"""Data-parallel replicas under torchrun (one rank per GPU): Trainer.train_step_replicas on N shards of a batch must
give every rank the same weights, equal to the single-GPU full-batch step within the fp32 bound.
    torchrun --nproc-per-node 2 tools/check_replicas.py"""
import datetime
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import Trainer  # noqa: E402
from ddnn_b200.config import make_params, synthetic_batch  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", timeout=datetime.timedelta(seconds=120))
    ok = True
    for dtype, dims, rows, tol in (("f32", [784, 128, 64, 10], 1003, 1e-5), ("bf16", [512, 512, 256, 64], 4096, 2e-2)):
        acts = ["relu", "relu", "softmax"]
        dp = Trainer(make_params(dims, acts, seed=1), dtype=dtype, device=dev)
        ref = Trainer(make_params(dims, acts, seed=1), dtype=dtype, device=dev)
        per = -(-rows // world)
        for step in range(3):
            x, y = synthetic_batch(rows, dims[0], dims[-1], seed=step)
            xd = torch.as_tensor(x, dtype=torch.float32).to(dev)
            yd = torch.as_tensor(y, dtype=torch.int32).to(dev)
            sl = slice(rank * per, min(rows, (rank + 1) * per))  # ragged last shard
            l_dp = float(dp.train_step_replicas(xd[sl].contiguous(), yd[sl].contiguous(), 0.05).item())
            l_ref = float(ref.train_step(xd, yd, 0.05).item())
            ok = ok and abs(l_dp - l_ref) < tol * abs(l_ref)
        for a, b in zip(dp.layer.layer_params, ref.layer.layer_params):
            err = float((a["weights"] - b["weights"]).abs().max() / b["weights"].abs().max())
            ok = ok and err < tol
            gathered = [torch.empty_like(a["weights"]) for _ in range(world)]
            dist.all_gather(gathered, a["weights"])
            ok = ok and all(torch.equal(g, gathered[0]) for g in gathered)  # replicas stay bit-identical to each other
        flag = torch.tensor([int(ok)], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
        if rank == 0:
            print(f"{dtype} {dims} on {world} replicas: {'ok' if ok else 'MISMATCH'}", flush=True)
    dist.barrier(device_ids=[local])
    dist.destroy_process_group()
    if rank == 0:
        print("REPLICA CHECK", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
