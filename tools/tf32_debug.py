"""Pattern tests of the tcgen05 tf32 GEMMs (forward / dX / dW) that localise layout mistakes."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import _lib, ops  # noqa: E402


def dev(a):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).cuda()


def show(name, got, ref):
    got = got.double().cpu().numpy()
    err = np.abs(got - ref).max() / max(1e-30, np.abs(ref).max())
    print(f"{name}: rel err {err:.3e}")
    if err > 1e-5:
        print("   got ", np.array2string(got[:3, :6], precision=4), "\n   ref ", np.array2string(ref[:3, :6], precision=4))
        bad = np.argwhere(np.abs(got - ref) > 1e-4 * np.abs(ref).max())
        print("   first bad", bad[:4].tolist(), "n bad", len(bad), "of", got.size)


_lib.set_f32_tensor_cores(2)
rng = np.random.default_rng(0)
for (b, k, n) in [(128, 32, 32), (128, 64, 32), (256, 32, 64), (4096, 784, 128), (300, 100, 36)]:
    x = rng.standard_normal((b, k))
    w = rng.standard_normal((k, n))
    bias = np.zeros(n)
    show(f"fwd ones {b}x{k}x{n}", ops.dense_forward(dev(np.ones((b, k))), dev(np.ones((k, n))), dev(bias), "linear"),
         np.full((b, n), float(k)))
    show(f"fwd rand {b}x{k}x{n}", ops.dense_forward(dev(x), dev(w), dev(bias), "linear"), x @ w)
    dz = rng.standard_normal((b, n))
    show(f"dx  rand {b}x{k}x{n}", ops.dense_backward_dx(dev(dz), dev(w), dev(np.ones((b, k))), "linear"), dz @ w.T)
    dw, db = ops.dense_backward_dw(dev(x), dev(dz))
    show(f"dw  rand {b}x{k}x{n}", dw, x.T @ dz)
    torch.cuda.synchronize()
