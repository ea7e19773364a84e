"""Diagnostic for the tcgen05 GEMMs on a GPU box: per-shape error summary + where errors sit.
Run under a timeout:  timeout 300 python tools/umma_check.py > gpurun_out/umma_check.log 2>&1"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import _lib, ops  # noqa: E402


def report(tag, got, ref):
    got = got.double().cpu().numpy()
    err = np.abs(got - ref)
    scale = np.abs(ref).max() + 1e-300
    bad = err > 1e-2 * scale
    print(f"{tag}: rel_err={err.max() / scale:.3e} bad={bad.sum()}/{bad.size}", flush=True)
    if bad.any():
        rows, cols = np.nonzero(bad)
        print("   bad rows (first 20):", sorted(set(rows.tolist()))[:20])
        print("   bad cols (first 20):", sorted(set(cols.tolist()))[:20])
        print("   sample got/ref:", got[rows[0], cols[0]], ref[rows[0], cols[0]])


def main():
    _lib.set_engine(_lib.ENGINE_TCGEN05)
    g = torch.Generator(device="cuda").manual_seed(0)
    for (b, k, n) in [(128, 64, 64), (128, 128, 64), (128, 64, 128), (128, 64, 256), (256, 256, 256), (300, 200, 72),
                      (1024, 784, 256)]:
        x = (torch.rand((b, k), device="cuda", generator=g) - 0.25).to(torch.bfloat16)
        w = (torch.randn((k, n), device="cuda", generator=g) / k ** 0.5).to(torch.bfloat16)
        dz = torch.randn((b, n), device="cuda", generator=g).to(torch.bfloat16)
        bias = torch.zeros(n, device="cuda")
        xr, wr, dzr = (t.double().cpu().numpy() for t in (x, w, dz))
        print(f"--- B={b} K={k} N={n}", flush=True)
        out = ops.dense_forward(x, w, bias, "linear", out_dtype=torch.float32)
        torch.cuda.synchronize()
        report("  fwd (A K-major, B MN-major)", out, xr @ wr)
        dx = ops.dense_backward_dx(dz, w, None, "linear")
        torch.cuda.synchronize()
        report("  dx  (A K-major, B K-major) ", dx, dzr @ wr.T)
        dw, db = ops.dense_backward_dw(x, dz)
        torch.cuda.synchronize()
        report("  dw  (A MN-major, B MN-major)", dw, xr.T @ dzr)
        report("  db", db, dzr.sum(axis=0))
    print("launches", _lib.launch_count())


if __name__ == "__main__":
    main()
