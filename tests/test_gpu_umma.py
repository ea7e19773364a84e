"""The tcgen05/TMA/TMEM GEMM family, forced on, against the float64 oracle.

Two comparisons per case: against the oracle on the SAME bf16-rounded operands (isolates the
kernel: only fp32 accumulation order differs, bound 1e-4 normalised for K up to 8192, plus one
bf16 rounding of the output where the output is bf16) and against the oracle on the original
float64 operands (the end-to-end bf16 bound)."""
import numpy as np
import pytest
import torch

import oracle
from gpu_util import BF16_ACT_TOL, BF16_GRAD_TOL, rel_err, to_dev, to_np

pytestmark = pytest.mark.gpu

ACC_TOL = 1e-4        # fp32 accumulation vs float64 on identical operands
BF16_OUT_TOL = 6e-3   # + one round-to-bf16 of the result (2^-8 relative, normalised)

# (B, K, N): one tile, many tiles, ragged M/N/K tails, narrow, K not a multiple of 64
SHAPES = [(128, 64, 64), (128, 256, 256), (256, 512, 128), (300, 200, 72), (1024, 784, 256), (96, 1024, 520), (3, 128, 64),
          (2048, 2048, 2048)]


@pytest.fixture(autouse=True)
def _force_tcgen05():
    from ddnn_b200 import _lib
    _lib.set_engine(_lib.ENGINE_TCGEN05)
    yield
    _lib.set_engine(_lib.ENGINE_AUTO)


def _ops(seed, b, k, n):
    rng = np.random.default_rng(seed)
    x = rng.random((b, k)) - 0.25
    w = rng.standard_normal((k, n)) / np.sqrt(k)
    bias = rng.standard_normal(n) * 0.1
    dz = rng.standard_normal((b, n)) / b
    return x, w, bias, dz


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("act", ["relu", "sigmoid", "linear"])
def test_forward(cuda_device, shape, act):
    from ddnn_b200 import ops
    b, k, n = shape
    x, w, bias, _ = _ops(10, b, k, n)
    xb, wb = to_dev(x, torch.bfloat16), to_dev(w, torch.bfloat16)
    ref_same = oracle.activate(np.dot(to_np(xb), to_np(wb)) + bias, act)
    got32 = ops.dense_forward(xb, wb, to_dev(bias), act, out_dtype=torch.float32)
    assert rel_err(to_np(got32), ref_same) < ACC_TOL
    got16 = ops.dense_forward(xb, wb, to_dev(bias), act)
    assert got16.dtype == torch.bfloat16
    assert rel_err(to_np(got16), ref_same) < BF16_OUT_TOL
    assert rel_err(to_np(got16), oracle.activate(np.dot(x, w) + bias, act)) < BF16_ACT_TOL


def test_forward_softmax_output(cuda_device):
    from ddnn_b200 import ops
    x, w, bias, _ = _ops(11, 64, 256, 512)
    xb, wb = to_dev(x, torch.bfloat16), to_dev(w, torch.bfloat16)
    got = ops.dense_forward(xb, wb, to_dev(bias), "softmax")
    ref = oracle.activate(np.dot(to_np(xb), to_np(wb)) + bias, "softmax")
    assert got.dtype == torch.float32 and rel_err(to_np(got), ref) < ACC_TOL


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("act_prev", ["relu", "sigmoid", "linear"])
def test_backward_dx(cuda_device, shape, act_prev):
    from ddnn_b200 import ops
    b, k, n = shape
    x, w, _, dz = _ops(12, b, k, n)
    rng = np.random.default_rng(13)
    zprev = rng.standard_normal((b, k))
    ab = to_dev(oracle.activate(zprev, act_prev), torch.bfloat16)
    dzb, wb = to_dev(dz, torch.bfloat16), to_dev(w, torch.bfloat16)
    a_r = to_np(ab)
    deriv = {"relu": (a_r > 0).astype(np.float64), "sigmoid": a_r * (1 - a_r), "linear": np.ones_like(a_r)}[act_prev]
    ref_same = np.dot(to_np(dzb), to_np(wb).T) * deriv
    got = ops.dense_backward_dx(dzb, wb, ab, act_prev)
    assert rel_err(to_np(got), ref_same) < BF16_OUT_TOL


@pytest.mark.parametrize("shape", SHAPES + [(4096, 784, 256), (8192, 128, 64)])
def test_backward_dw_and_sgd(cuda_device, shape):
    from ddnn_b200 import ops
    b, k, n = shape
    x, w, bias, dz = _ops(14, b, k, n)
    xb, dzb = to_dev(x, torch.bfloat16), to_dev(dz, torch.bfloat16)
    ref_dw = np.dot(to_np(xb).T, to_np(dzb))
    ref_db = to_np(dzb).sum(axis=0)
    dw, db = ops.dense_backward_dw(xb, dzb)
    assert rel_err(to_np(dw), ref_dw) < ACC_TOL
    assert rel_err(to_np(db), ref_db) < ACC_TOL
    assert rel_err(to_np(dw), np.dot(x.T, dz)) < BF16_GRAD_TOL
    # fused SGD epilogue: fp32 master updated in place, bf16 shadow refreshed from it
    wm, bm = to_dev(w), to_dev(bias)
    sh = torch.zeros((k, n), dtype=torch.bfloat16, device="cuda")
    ops.dense_backward_dw_sgd(xb, dzb, wm, bm, 0.5, w_shadow=sh)
    assert rel_err(to_np(wm), to_np(to_dev(w)) - 0.5 * ref_dw) < 1e-5
    assert rel_err(to_np(bm), to_np(to_dev(bias)) - 0.5 * ref_db) < 1e-5
    assert torch.equal(sh, wm.to(torch.bfloat16))


def test_dw_deterministic_and_linear(cuda_device):
    """Size-independent properties at a size the oracle would not finish quickly:
    run-to-run bit equality, and dW(x, 2*dz) == 2*dW(x, dz) exactly (power-of-two scaling)."""
    from ddnn_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.rand((16384, 1024), device="cuda", generator=g).to(torch.bfloat16)
    dz = torch.randn((16384, 512), device="cuda", generator=g).to(torch.bfloat16)
    a, da = ops.dense_backward_dw(x, dz)
    b, _ = ops.dense_backward_dw(x, dz)
    assert torch.equal(a, b)
    c, dc = ops.dense_backward_dw(x, dz * 2)
    assert torch.equal(c, a * 2) and torch.equal(dc, da * 2)


def test_forward_full_size_checksum(cuda_device):
    """8192-wide layer (config 3 shape): rows of X that are one-hot pick rows of W exactly."""
    from ddnn_b200 import ops
    n = 8192
    g = torch.Generator(device="cuda").manual_seed(1)
    w = (torch.randn((n, n), device="cuda", generator=g) / 90).to(torch.bfloat16)
    idx = torch.randint(0, n, (512,), device="cuda", generator=g)
    x = torch.zeros((512, n), device="cuda", dtype=torch.bfloat16)
    x[torch.arange(512, device="cuda"), idx] = 1
    out = ops.dense_forward(x, w, torch.zeros(n, device="cuda"), "linear")
    assert torch.equal(out, w[idx])
    # and a dense probe against a float64 dot on a few columns
    xd = torch.rand((256, n), device="cuda", generator=g).to(torch.bfloat16)
    out = ops.dense_forward(xd, w, torch.zeros(n, device="cuda"), "linear", out_dtype=torch.float32)
    cols = [0, 1, 255, 256, 4097, 8191]
    ref = xd.double().cpu().numpy() @ w[:, cols].double().cpu().numpy()
    assert rel_err(out[:, cols].double().cpu().numpy(), ref) < ACC_TOL


@pytest.mark.parametrize("bkn", [(4096, 512, 4096), (4000, 320, 4104)])
def test_cta_pair_kernels(cuda_device, bkn):
    """Problems with >= 2 waves of 256x256 tiles run on the CTA-pair (tcgen05 cta_group::2) kernel:
    forward, dX and dW/SGD against the oracle on the bf16-rounded operands, ragged M/N tails included,
    and bit-identical to the single-CTA kernel (same k order)."""
    from ddnn_b200 import _lib, ops
    b, k, n = bkn
    rng = np.random.default_rng(20)
    # forward: D[b, n]
    x = to_dev(rng.random((b, k)) - 0.25, torch.bfloat16)
    w = to_dev(rng.standard_normal((k, n)) / np.sqrt(k), torch.bfloat16)
    bias = to_dev(rng.standard_normal(n) * 0.1)
    ref = oracle.activate(np.dot(to_np(x), to_np(w)) + to_np(bias), "relu")
    got = ops.dense_forward(x, w, bias, "relu", out_dtype=torch.float32)
    assert rel_err(to_np(got), ref) < ACC_TOL
    # dX: D[b, k2] with k2 = n (wide input side), reduction over n2 = k
    dz = to_dev(rng.standard_normal((b, k)) / b, torch.bfloat16)
    w2 = to_dev(rng.standard_normal((n, k)) / np.sqrt(k), torch.bfloat16)     # layer n -> k
    a_prev = to_dev(oracle.activate(rng.standard_normal((b, n)), "relu"), torch.bfloat16)
    ref_dx = np.dot(to_np(dz), to_np(w2).T) * (to_np(a_prev) > 0)
    got_dx = ops.dense_backward_dx(dz, w2, a_prev, "relu")
    assert rel_err(to_np(got_dx), ref_dx) < BF16_OUT_TOL
    # dW + fused SGD: D[n, n2] = X2^T dZ2 with a short reduction (rows = k)
    x2 = to_dev(rng.random((k, n)), torch.bfloat16)
    dz2 = to_dev(rng.standard_normal((k, n)) / k, torch.bfloat16)
    ref_dw = np.dot(to_np(x2).T, to_np(dz2))
    dw, db = ops.dense_backward_dw(x2, dz2)
    assert rel_err(to_np(dw), ref_dw) < ACC_TOL and rel_err(to_np(db), to_np(dz2).sum(axis=0)) < ACC_TOL
    wm = to_dev(rng.standard_normal((n, n)) * 0.01)
    w0 = to_np(wm).copy()
    bm = torch.zeros(n, device="cuda")
    sh = torch.zeros((n, n), dtype=torch.bfloat16, device="cuda")
    ops.dense_backward_dw_sgd(x2, dz2, wm, bm, 0.5, w_shadow=sh)
    assert rel_err(to_np(wm), w0 - 0.5 * ref_dw) < 1e-5 and torch.equal(sh, wm.to(torch.bfloat16))
    # single-CTA kernel on the same inputs: identical bits
    _lib.set_cta_pairs(False)
    try:
        assert torch.equal(ops.dense_forward(x, w, bias, "relu", out_dtype=torch.float32), got)
        assert torch.equal(ops.dense_backward_dx(dz, w2, a_prev, "relu"), got_dx)
        assert torch.equal(ops.dense_backward_dw(x2, dz2)[0], dw)
    finally:
        _lib.set_cta_pairs(True)


@pytest.mark.timeout(180)
@pytest.mark.parametrize("bkn", [(512, 256, 256), (4096, 512, 4096)])
def test_signals_folded_into_the_gemm(cuda_device, bkn):
    """ddnn_dense_forward_signalled / ddnn_dense_backward_dx_signalled (single-CTA and CTA-pair kernels): same bits as
    the plain entry points; the last CTA publishes done_flag and leaves the tile counter at zero; a kernel whose
    wait_flag is not yet set starts, polls, and finishes once another stream sets the flag."""
    from ddnn_b200 import _lib, ops
    b, k, n = bkn
    rng = np.random.default_rng(30)
    x = to_dev(rng.random((b, k)) - 0.25, torch.bfloat16)
    w = to_dev(rng.standard_normal((k, n)) / np.sqrt(k), torch.bfloat16)
    bias = to_dev(rng.standard_normal(n) * 0.1)
    flags = torch.zeros((4,), dtype=torch.int32, device="cuda")
    counter = torch.zeros((1,), dtype=torch.int32, device="cuda")
    plain = ops.dense_forward(x, w, bias, "relu")
    flags[0] = 7
    sig = ops.make_signals(flags.data_ptr(), 7, flags.data_ptr() + 4, 42, counter.data_ptr())
    for _ in range(2):  # twice: the counter must have been left at zero
        got = ops.dense_forward(x, w, bias, "relu", signals=sig)
        torch.cuda.synchronize()
        assert torch.equal(got, plain) and int(flags[1]) == 42 and int(counter[0]) == 0
        flags[1] = 0
    # dX: wait only (no publish)
    dz = to_dev(rng.standard_normal((b, n)) / b, torch.bfloat16)
    a_prev = to_dev(oracle.activate(rng.standard_normal((b, k)), "relu"), torch.bfloat16)
    plain_dx = ops.dense_backward_dx(dz, w, a_prev, "relu")
    got_dx = ops.dense_backward_dx(dz, w, a_prev, "relu", signals=ops.make_signals(flags.data_ptr(), 7))
    assert torch.equal(got_dx, plain_dx)
    # a wait that is satisfied only later, from another stream
    if b * n <= 512 * 256:  # few CTAs: the flag kernel of the other stream always finds a free SM
        # the flag kernel must already be loaded: the first launch of a kernel loads its module lazily, and that can
        # wait for the device to go idle -- which it never does while the GEMM below polls for this very flag
        scratch_flag = torch.zeros((1,), dtype=torch.int32, device="cuda")
        _lib.check(_lib.lib.ddnn_flag_write(scratch_flag.data_ptr(), 1, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        side = torch.cuda.Stream()
        sig = ops.make_signals(flags.data_ptr() + 8, 5, flags.data_ptr() + 12, 9, counter.data_ptr())
        with torch.cuda.stream(side):
            late = ops.dense_forward(x, w, bias, "relu", signals=sig)
        import time
        time.sleep(0.05)
        assert int(flags[3].item()) == 0  # still polling: nothing published
        _lib.check(_lib.lib.ddnn_flag_write(flags.data_ptr() + 8, 5, torch.cuda.current_stream().cuda_stream))
        side.synchronize()
        assert torch.equal(late, plain) and int(flags[3]) == 9
