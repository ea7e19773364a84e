"""One-GPU experiment behind the pipeline's per-boundary overhead: a stage's micro-batch GEMMs
timed (a) as the plain stream-ordered chain the single-stage run enjoys (programmatic dependent launch on), (b) with
programmatic dependent launch off, (c) with what a stage boundary puts between two GEMMs today -- a one-thread flag
kernel after, a stream memory wait before -- the flags being already satisfied, so only the launch-chain effect is
measured.    python tools/boundary_cost.py [width] [micro_rows] [micro_batches]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddnn_b200 import _lib, ops  # noqa: E402


def main():
    width = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
    mb = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    m = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    w = (torch.randn((width, width), device=dev, generator=g) * (2.0 / width) ** 0.5).to(torch.bfloat16)
    bias = torch.full((width,), 0.01, device=dev)
    x = torch.rand((m * mb, width), device=dev, generator=g).to(torch.bfloat16)
    out = torch.empty_like(x)
    flags = torch.full((4,), 1 << 20, dtype=torch.int32, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream

    def chain(boundary):
        for i in range(m):
            if boundary:
                _lib.check(_lib.lib.ddnn_flag_wait(flags.data_ptr(), 1, st))
            ops.dense_forward(x[i * mb:(i + 1) * mb], w, bias, "relu", out=out[i * mb:(i + 1) * mb])
            if boundary:
                _lib.check(_lib.lib.ddnn_flag_write(flags.data_ptr() + 4, 1 << 20, st))

    def timed(boundary, reps=3):
        chain(boundary)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            chain(boundary)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (reps * m)

    fl = 2.0 * mb * width * width
    variants = (("stream-ordered chain, PDL on", 1, False), ("PDL off", 0, False),
                ("flag wait + flag kernel between GEMMs, PDL on", 1, True),
                ("flag wait + flag kernel between GEMMs, PDL off", 0, True))
    for _ in range(int(os.environ.get("WARM", "40"))):  # reach the clocks the power cap leaves before measuring
        chain(False)
    acc = {v[0]: [] for v in variants}
    for _ in range(int(os.environ.get("ROUNDS", "4"))):  # interleaved: drift hits every variant alike
        for name, pdl, boundary in variants:
            _lib.lib.ddnn_set_pdl(pdl)
            acc[name].append(timed(boundary))
    for name, ts in acc.items():
        t = sum(ts) / len(ts)
        print(f"{width}x{width}, {mb}-row micro-batches: {name}: {t * 1e3:.1f} us per GEMM, {fl / t / 1e9:.0f} TFLOP/s  "
              f"(rounds: {', '.join(f'{v * 1e3:.1f}' for v in ts)})", flush=True)
    _lib.lib.ddnn_set_pdl(1)


if __name__ == "__main__":
    main()
