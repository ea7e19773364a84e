"""Experiment behind the chunked weight gradient: the full-batch dW GEMM of a wide layer (W x W output,
reduction over ROWS rows) against the same reduction cut into row chunks, each chunk its own GEMM with a
read-modify-write epilogue on the fp32 master (the SGD epilogue with lr = 0 moves the same bytes an accumulating
epilogue would).  Each variant runs back to back for about SECS seconds, variants alternate ROUNDS times, so every
figure is taken at the clocks the power cap leaves after warm-up.
    python tools/dw_chunking.py [width] [rows] [secs] [rounds]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import ops  # noqa: E402


def main():
    width = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    rows = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    secs = float(sys.argv[3]) if len(sys.argv) > 3 else 1.5
    rounds = int(sys.argv[4]) if len(sys.argv) > 4 else 2
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    wm = torch.randn((width, width), device=dev, generator=g) * (2.0 / width) ** 0.5
    w = wm.to(torch.bfloat16)
    bias = torch.full((width,), 0.01, device=dev)
    x = torch.rand((rows, width), device=dev, generator=g).to(torch.bfloat16)
    dz = (torch.randn((rows, width), device=dev, generator=g) / rows).to(torch.bfloat16)
    out = torch.empty((2048, width), dtype=torch.bfloat16, device=dev)
    ws = torch.empty((max(256, ops.dw_workspace_bytes(rows, width, width, torch.bfloat16)),), dtype=torch.uint8, device=dev)

    def dw(chunk):
        for r0 in range(0, rows, chunk):
            ops.dense_backward_dw_sgd(x[r0:r0 + chunk], dz[r0:r0 + chunk], wm, bias, 0.0, w_shadow=w, workspace=ws)

    def fwd(_):
        for r0 in range(0, rows, 2048):
            ops.dense_forward(x[r0:r0 + 2048], w, bias, "relu", out=out)

    variants = [("forward, 2048-row micro-batches", fwd, 0)] + \
               [(f"dW+SGD in chunks of {c} rows", dw, c) for c in (rows, rows // 2, rows // 4, rows // 8, rows // 16, rows // 32)]
    fl = 2.0 * rows * width * width
    acc = {n: [] for n, _, _ in variants}
    for _ in range(rounds):
        for name, fn, arg in variants:
            fn(arg)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            n = 0
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            while time.perf_counter() - t0 < secs:
                fn(arg)
                n += 1
                torch.cuda.synchronize()
            e1.record()
            torch.cuda.synchronize()
            acc[name].append(e0.elapsed_time(e1) / n)
    for name, ts in acc.items():
        t = sum(ts) / len(ts)
        print(f"{width}^2 x {rows}: {name}: {t:.2f} ms per pass over the batch ({fl / t / 1e9:.0f} TFLOP/s)  rounds: "
              + ", ".join(f"{v:.2f}" for v in ts), flush=True)


if __name__ == "__main__":
    main()
