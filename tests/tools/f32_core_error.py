"""Prints the normalised max error of the fp32 path (both inner-product cores) against the float64
oracle: single GEMMs and the 3-layer gradient chain of tests/test_gpu_layer.py."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import oracle  # noqa: E402
from ddnn_b200 import Trainer, _lib, ops  # noqa: E402
from ddnn_b200.config import make_params, synthetic_batch  # noqa: E402


def rel(a, b):
    a = a.double().cpu().numpy() if hasattr(a, "cpu") else np.asarray(a)
    return float(np.abs(a - b).max() / np.abs(b).max())


def dev(a, dt=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dt).cuda()


for tc in (2, 1, 0):
    _lib.set_f32_tensor_cores(tc)
    rng = np.random.default_rng(0)
    x = rng.random((4096, 784)) - 0.25
    w = rng.standard_normal((784, 128)) / 28
    dz = rng.standard_normal((4096, 128)) / 4096
    print({2: "tcgen05", 1: "mma.sync", 0: "ffma    "}[tc], "fwd 4096x784x128 %.2e" % rel(ops.dense_forward(dev(x), dev(w), dev(np.zeros(128)), "linear"), x @ w),
          "dW %.2e" % rel(ops.dense_backward_dw(dev(x), dev(dz))[0], x.T @ dz), end="  ")
    dims, acts = [784, 128, 64, 10], ["relu", "relu", "softmax"]
    params = make_params(dims, acts, seed=1)
    xb, yb = synthetic_batch(96, 784, 10, seed=0)
    a, zs = oracle.forward_all(params, xb)
    loss, dws, dbs, _ = oracle.backward(params, a, zs, yb)
    t = Trainer(params)
    L = t.layer
    L.begin_batch(96)
    L.forward_train(dev(xb), 0)
    L.loss_and_grad(dev(yb, torch.int32), 0)
    L.backward(0, 96, need_input_grad=False)
    print("chain dW errs", ["%.2e" % rel(g[0], r) for g, r in zip(L.gradients(), dws)])
# error growth with the reduction length (no split-K on forward): one tile column, K up to 16384
for tc in (2, 0):
    _lib.set_f32_tensor_cores(tc)
    out = []
    for k in (256, 1024, 4096, 16384):
        rng = np.random.default_rng(k)
        x = rng.random((512, k))
        w = rng.standard_normal((k, 128)) / np.sqrt(k)
        out.append("K=%d %.2e" % (k, rel(ops.dense_forward(dev(x), dev(w), dev(np.zeros(128)), "linear"), x @ w)))
    print({2: "tcgen05", 0: "ffma    "}[tc], "forward 512xKx128:", "  ".join(out))
_lib.set_f32_tensor_cores(None)
