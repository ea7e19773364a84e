"""Runs a few eager (un-graphed) train steps of one workload on cuda:0 -- the short command
that ncu wraps.   python tools/profile_step.py <workload> [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from ddnn_b200 import Trainer, _lib  # noqa: E402


def main():
    wl = sys.argv[1] if len(sys.argv) > 1 else "mnist4096"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    dims, acts, rows, dtype, _ = bench.WORKLOADS[wl]
    if len(sys.argv) > 3:
        rows = int(sys.argv[3])
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    tr = Trainer(bench.device_params(dims, acts, dev), dtype=dtype, device=dev, input_dim=dims[0])
    g = torch.Generator(device=dev).manual_seed(0)
    x = torch.rand((rows, dims[0]), device=dev, generator=g).to(torch.float32 if dtype == "f32" else torch.bfloat16)
    y = torch.randint(0, dims[-1], (rows,), device=dev, generator=g, dtype=torch.int32)
    for _ in range(steps):
        loss = tr.train_step(x, y, 0.01)
    torch.cuda.synchronize()
    print(f"{wl}: {steps} steps, loss {loss.item():.5f}, {_lib.launch_count()} libddnn launches")


if __name__ == "__main__":
    main()
