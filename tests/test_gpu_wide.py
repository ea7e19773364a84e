"""The wide configurations at their OWN widths (BASELINE.json configs[2]: 8192, configs[3]: 16384), the shapes the
north_star's tensor-peak and 8-GPU targets are quoted on: every tcgen05 GEMM layout (forward, dX, dW + fused SGD, all
on the CTA-pair kernel at these sizes), the wide softmax-CE, and one full train step, against the float64 oracle.

Batches are a few hundred rows (not multiples of the 128 / 256-row tiles) so that the float64 products finish in
seconds on the host: K and N, the dimensions that set tile counts, wave quantisation, split plans and accumulation
length, are the real ones.  Kernel checks use the oracle on the SAME bf16-rounded operands (only the fp32
accumulation order differs: ACC_TOL); the train step is compared end to end within the bf16 bounds."""
import numpy as np
import pytest
import torch

import oracle
from gpu_util import BF16_ACT_TOL, BF16_GRAD_TOL, bf16_update_tol, elementwise_ok, rel_err, to_np

pytestmark = pytest.mark.gpu

ACC_TOL = 1e-4        # fp32 accumulation (chains up to 16384 long) vs float64 on identical operands
BF16_OUT_TOL = 6e-3   # + one round-to-bf16 of the result
WIDTHS = [8192, 16384]


def _dev_ops(width, rows, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = (torch.rand((rows, width), device="cuda", generator=g) - 0.25).to(torch.bfloat16)
    w = (torch.randn((width, width), device="cuda", generator=g) * (2.0 / width) ** 0.5).to(torch.bfloat16)
    bias = torch.randn((width,), device="cuda", generator=g) * 0.1
    return x, w, bias


@pytest.mark.parametrize("width", WIDTHS)
def test_forward_at_full_width(cuda_device, width):
    from ddnn_b200 import ops
    rows = 300
    x, w, bias = _dev_ops(width, rows, 1)
    ref = oracle.activate(np.dot(to_np(x), to_np(w)) + to_np(bias), "relu")
    got32 = ops.dense_forward(x, w, bias, "relu", out_dtype=torch.float32)
    assert rel_err(to_np(got32), ref) < ACC_TOL
    assert elementwise_ok(to_np(got32), ref, rtol=1e-4, atol=1e-4 * float(np.abs(ref).max()))
    got16 = ops.dense_forward(x, w, bias, "relu")
    assert got16.dtype == torch.bfloat16 and rel_err(to_np(got16), ref) < BF16_OUT_TOL


@pytest.mark.parametrize("width", WIDTHS)
def test_backward_dx_at_full_width(cuda_device, width):
    from ddnn_b200 import ops
    rows = 300
    _, w, _ = _dev_ops(width, rows, 2)
    g = torch.Generator(device="cuda").manual_seed(3)
    dz = (torch.randn((rows, width), device="cuda", generator=g) / rows).to(torch.bfloat16)
    a_prev = torch.relu(torch.randn((rows, width), device="cuda", generator=g)).to(torch.bfloat16)
    ref = np.dot(to_np(dz), to_np(w).T) * (to_np(a_prev) > 0)
    got = ops.dense_backward_dx(dz, w, a_prev, "relu")
    assert rel_err(to_np(got), ref) < BF16_OUT_TOL


@pytest.mark.parametrize("width", WIDTHS)
def test_backward_dw_sgd_at_full_width(cuda_device, width):
    """dW = X^T dZ with the reduction over 1000 rows (not a multiple of the 64-row k-block), the full width x
    width output (4096 / 16384 tiles of 128x256), SGD fused into the epilogue, bf16 shadow refreshed."""
    from ddnn_b200 import ops
    rows = 1000
    x, w, bias = _dev_ops(width, rows, 4)
    g = torch.Generator(device="cuda").manual_seed(5)
    dz = (torch.randn((rows, width), device="cuda", generator=g) / rows).to(torch.bfloat16)
    wm = w.float().contiguous()
    w0 = to_np(wm)
    bm = bias.clone()
    sh = torch.zeros((width, width), dtype=torch.bfloat16, device="cuda")
    ops.dense_backward_dw_sgd(x, dz, wm, bm, 0.5, w_shadow=sh)
    ref_dw = np.dot(to_np(x).T, to_np(dz))
    got_upd = to_np(wm) - w0                       # what the epilogue applied: -lr * dW (+ fp32 rounding of W - lr dW)
    assert rel_err(got_upd, -0.5 * ref_dw) < 2e-3  # fp32 ulp of |W| ~ 0.05 against |lr dW| ~ 1e-4: rounding of the sum
    assert rel_err(to_np(wm), w0 - 0.5 * ref_dw) < 1e-6
    assert rel_err(to_np(bm), to_np(bias) - 0.5 * to_np(dz).sum(axis=0)) < 1e-5
    assert torch.equal(sh, wm.to(torch.bfloat16))
    del w0, ref_dw, got_upd
    # gradients-only form (fp32 store epilogue) on the same operands
    dw, db = ops.dense_backward_dw(x, dz)
    assert rel_err(to_np(dw), np.dot(to_np(x).T, to_np(dz))) < ACC_TOL
    assert rel_err(to_np(db), to_np(dz).sum(axis=0)) < ACC_TOL


@pytest.mark.parametrize("width", WIDTHS)
def test_softmax_ce_at_full_width(cuda_device, width):
    """Row softmax + CE + dZ over 8192 / 16384 classes (one CTA per row), fp32 and bf16 gradient output."""
    from ddnn_b200 import ops
    rows = 300
    g = torch.Generator(device="cuda").manual_seed(6)
    z = torch.randn((rows, width), device="cuda", generator=g) * 3.0
    y = torch.randint(0, width, (rows,), device="cuda", generator=g, dtype=torch.int32)
    loss_ref, p_ref, dz_ref = oracle.softmax_ce(to_np(z), y.cpu().numpy().astype(np.int64), denom=4 * rows)
    for dt, tol in ((torch.float32, 1e-5), (torch.bfloat16, BF16_OUT_TOL)):
        dz = torch.empty((rows, width), dtype=dt, device="cuda")
        p = torch.empty((rows, width), dtype=torch.float32, device="cuda")
        row_loss = torch.empty((rows,), device="cuda")
        loss = torch.zeros((1,), device="cuda")
        ops.softmax_ce(z, y, dz, row_loss, loss, 1.0 / (4 * rows), p=p)
        assert abs(loss.item() - loss_ref) < 1e-5 * abs(loss_ref)
        assert rel_err(to_np(p), p_ref) < 1e-5
        assert rel_err(to_np(dz), dz_ref) < tol
    out = ops.softmax_rows(z)
    assert rel_err(to_np(out), p_ref) < 1e-5 and elementwise_ok(to_np(out), p_ref, rtol=1e-4, atol=1e-9)


def test_train_step_of_an_8192_wide_net(cuda_device):
    """One full Trainer step (forward, softmax-CE over 8192 classes, dX, dW + SGD) of a 2-layer 8192-wide bf16 net,
    300 rows, against oracle.train_step in float64 from the same fp32 weights: loss, post-step weights, and the
    UPDATE itself within the bf16 update bound (bf16_update_tol in gpu_util.py: rounding + ReLU gate flips)."""
    from ddnn_b200 import Trainer
    width, rows, lr = 8192, 300, 0.5
    g = torch.Generator(device="cuda").manual_seed(7)
    params = []
    for act in ("relu", "softmax"):
        w = torch.randn((width, width), device="cuda", generator=g) * ((2.0 if act == "relu" else 1.0) / width) ** 0.5
        params.append({"weights": w, "biases": torch.full((width,), 0.01, device="cuda"), "activation": act})
    ref = [{"weights": to_np(p["weights"]), "biases": to_np(p["biases"]), "activation": p["activation"]} for p in params]
    before = [r["weights"].copy() for r in ref]
    x = torch.rand((rows, width), device="cuda", generator=g)
    y = torch.randint(0, width, (rows,), device="cuda", generator=g, dtype=torch.int32)
    t = Trainer(params, dtype="bf16")
    got = t.train_step(x, y, lr).item()
    want = oracle.train_step(ref, to_np(x), y.cpu().numpy().astype(np.int64), lr)
    assert abs(got - want) < 2e-2 * want
    for i, (p, r, b0) in enumerate(zip(t.layer.layer_params, ref, before)):
        gw = to_np(p["weights"])
        assert rel_err(gw, r["weights"]) < BF16_ACT_TOL
        # layer 1 sits below a ReLU (gate flips, see bf16_update_tol); the output layer's dZ has no gate above it
        assert rel_err(gw - b0, r["weights"] - b0) < bf16_update_tol(rows, below_relu=(i == 0)), i
