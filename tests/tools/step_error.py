"""Per-layer normalised errors of weights, biases and their UPDATES after n fp32 MNIST steps at batch 4096, for the
three fp32 inner products (0 FFMA, 1 mma.sync, 2 tcgen05 + mma.sync tail)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import oracle  # noqa: E402
from ddnn_b200 import Trainer, _lib  # noqa: E402
from ddnn_b200.config import make_params, synthetic_batch  # noqa: E402

dims, acts = [784, 128, 64, 10], ["relu", "relu", "softmax"]
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
x, y = synthetic_batch(4096, 784, 10, seed=0)
ref = [dict(p) for p in make_params(dims, acts, seed=1)]
init = [dict(p) for p in make_params(dims, acts, seed=1)]
for _ in range(steps):
    oracle.train_step(ref, x, y, 0.01)


def rel(a, b):
    return float(np.abs(np.asarray(a) - np.asarray(b)).max() / max(1e-300, np.abs(np.asarray(b)).max()))


for mode in (0, 1, 2):
    _lib.set_f32_tensor_cores(mode)
    t = Trainer(make_params(dims, acts, seed=1))
    xd = torch.as_tensor(x, dtype=torch.float32).cuda()
    yd = torch.as_tensor(y, dtype=torch.int32).cuda()
    for _ in range(steps):
        t.train_step(xd, yd, 0.01)
    out = []
    for g, r, i0 in zip(t.export_params(), ref, init):
        out.append("W %.1e (upd %.1e)  b %.1e (upd %.1e)" % (
            rel(g["weights"], r["weights"]), rel(g["weights"] - i0["weights"], r["weights"] - i0["weights"]),
            rel(g["biases"], r["biases"]), rel(g["biases"] - i0["biases"], r["biases"] - i0["biases"])))
    print("mode", mode, " | ".join(out))
_lib.set_f32_tensor_cores(None)
