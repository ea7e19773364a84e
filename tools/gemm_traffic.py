"""One 8192^3 forward GEMM (CTA-pair kernel) a few times -- wrapped by ncu to read DRAM bytes per launch."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.rand((n, n), device="cuda", generator=g).to(torch.bfloat16)
w = (torch.randn((n, n), device="cuda", generator=g) / 90).to(torch.bfloat16)
b = torch.zeros(n, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(3):
    ops.dense_forward(x, w, b, "relu")
e0.record()
for _ in range(10):
    ops.dense_forward(x, w, b, "relu")
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"hints={os.environ.get('DDNN_L2_HINTS', '1')} n={n}: {ms * 1e3:.1f} us, {2 * n ** 3 / ms / 1e9:.1f} TFLOP/s")
