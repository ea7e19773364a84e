"""Times the tcgen05 tf32 entry points alone (CUDA-graph replays) for a few shapes."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import _lib, ops  # noqa: E402

_lib.set_f32_tensor_cores(int(os.environ.get("MODE", "2")))
_lib.check(_lib.lib.ddnn_set_cluster_split_k(int(os.environ.get("CLUSTER", "1"))))
dev = torch.device("cuda", 0)
shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(4096, 784, 128), (4096, 128, 64), (16384, 784, 128), (4096, 2048, 512)]
for (b, k, n) in shapes:
    x = torch.rand((b, k), device=dev)
    w = torch.randn((k, n), device=dev) / k ** 0.5
    bias = torch.zeros(n, device=dev)
    out = torch.empty((b, n), device=dev)
    dz = torch.randn((b, n), device=dev)
    dx = torch.empty((b, k), device=dev)
    res = []
    for name, fn in (("fwd", lambda: ops.dense_forward(x, w, bias, "relu", out=out)),
                     ("dx", lambda: ops.dense_backward_dx(dz, w, x, "relu", out=dx)),
                     ("dw", lambda: ops.dense_backward_dw(x, dz))):
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(10):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1000 / 50
        res.append(f"{name} {us:7.1f} us ({2.0 * b * k * n / us / 1e6:6.1f} TF)")
    print(f"{b}x{k}x{n}: " + "  ".join(res))
