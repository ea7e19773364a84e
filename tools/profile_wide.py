"""The three GEMM shapes of the 16384-wide (or WIDTH-wide) net, one launch each after a warm-up, for ncu:
forward of one micro-batch (MB x W x W), dX of one micro-batch, dW + SGD over the whole batch (W x W, reduction over
ROWS rows).     ncu --set full -k regex:umma -s 3 -c 3 python tools/profile_wide.py [width] [rows] [micro_rows]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import ops  # noqa: E402


def main():
    width = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    mb = int(sys.argv[3]) if len(sys.argv) > 3 else 2048
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    wm = torch.randn((width, width), device=dev, generator=g) * (2.0 / width) ** 0.5
    w = wm.to(torch.bfloat16)
    bias = torch.full((width,), 0.01, device=dev)
    x = torch.rand((rows, width), device=dev, generator=g).to(torch.bfloat16)
    dz = (torch.randn((rows, width), device=dev, generator=g) / rows).to(torch.bfloat16)
    out = torch.empty((mb, width), dtype=torch.bfloat16, device=dev)
    ws = torch.empty((max(256, ops.dw_workspace_bytes(rows, width, width, torch.bfloat16)),), dtype=torch.uint8, device=dev)
    reps = int(os.environ.get("REPS", "2"))
    chunk = int(os.environ.get("DW_CHUNK", "0"))  # > 0: only the dW GEMM, over the first DW_CHUNK rows (ncu experiments)
    if chunk:
        for _ in range(reps):
            ops.dense_backward_dw_sgd(x[:chunk], dz[:chunk], wm, bias, 0.0, w_shadow=w, workspace=ws)
        torch.cuda.synchronize()
        return
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    for r in range(reps):  # the last repetition is the one ncu keeps (-s skips the first)
        ev[0].record()
        ops.dense_forward(x[:mb], w, bias, "relu", out=out)
        ev[1].record()
        ops.dense_backward_dx(dz[:mb], w, x[:mb], "relu", out=out)
        ev[2].record()
        ops.dense_backward_dw_sgd(x, dz, wm, bias, 0.0, w_shadow=w, workspace=ws)
        ev[3].record()
    torch.cuda.synchronize()
    fl = 2.0 * width * width
    print(f"forward {ev[0].elapsed_time(ev[1]):.3f} ms ({fl * mb / ev[0].elapsed_time(ev[1]) / 1e9:.0f} TFLOP/s), "
          f"dX {ev[1].elapsed_time(ev[2]):.3f} ms ({fl * mb / ev[1].elapsed_time(ev[2]) / 1e9:.0f}), "
          f"dW+SGD {ev[2].elapsed_time(ev[3]):.3f} ms ({fl * rows / ev[2].elapsed_time(ev[3]) / 1e9:.0f})", flush=True)


if __name__ == "__main__":
    main()
