"""Parity of every C-ABI entry point with the float64 oracle on the same seeded inputs."""
import numpy as np
import pytest
import torch

import oracle
from gpu_util import BF16_ACT_TOL, BF16_GRAD_TOL, F32_TOL, cond_ok, rel_err, to_dev, to_np

pytestmark = pytest.mark.gpu

ACTS = ["relu", "sigmoid", "linear", "softmax"]
# (B, K, N): MNIST layers, ragged / non-multiple-of-tile, N = 10, B = 1, wide-ish
SHAPES = [(64, 784, 128), (33, 128, 64), (50, 64, 10), (1, 7, 3), (257, 100, 130), (128, 256, 256), (5, 3, 1)]


def _data(b, k, n, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.random((b, k)) - 0.25
    w = rng.standard_normal((k, n)) / np.sqrt(k)
    bias = rng.standard_normal(n) * 0.1
    return x, w, bias


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("act", ACTS)
def test_forward_f32(cuda_device, shape, act):
    from ddnn_b200 import ops
    b, k, n = shape
    x, w, bias = _data(b, k, n)
    ref = oracle.activate(np.dot(x, w) + bias, act)
    got = ops.dense_forward(to_dev(x), to_dev(w), to_dev(bias), act)
    assert rel_err(to_np(got), ref) < F32_TOL
    if act != "softmax":  # every element within 1e-5 of its own condition sum (relu / sigmoid / identity are 1-Lipschitz)
        assert cond_ok(to_np(got), ref, np.dot(np.abs(x), np.abs(w)) + np.abs(bias), F32_TOL)


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("act_prev", ["relu", "sigmoid", "linear"])
def test_backward_dx_f32(cuda_device, shape, act_prev):
    from ddnn_b200 import ops
    b, k, n = shape
    rng = np.random.default_rng(1)
    dz = rng.standard_normal((b, n))
    w = rng.standard_normal((k, n)) / np.sqrt(n)
    zprev = rng.standard_normal((b, k))
    aprev = oracle.activate(zprev, act_prev)
    ref = oracle.ddnn_oracle._act_backward(np.dot(dz, w.T), aprev, zprev, act_prev)
    got = ops.dense_backward_dx(to_dev(dz), to_dev(w), to_dev(aprev), act_prev)
    assert rel_err(to_np(got), ref) < F32_TOL


def test_backward_dx_rejects_hidden_softmax(cuda_device):
    from ddnn_b200 import _lib, ops
    t = torch.zeros((4, 4), device="cuda")
    with pytest.raises(_lib.DdnnError) as e:
        ops.dense_backward_dx(t, t, t, "softmax")
    assert e.value.status == _lib.ERR_ACTIVATION


@pytest.mark.parametrize("shape", SHAPES + [(4096, 784, 128), (1000, 10, 10)])
def test_backward_dw_f32(cuda_device, shape):
    from ddnn_b200 import ops
    b, k, n = shape
    rng = np.random.default_rng(2)
    x = rng.random((b, k))
    dz = rng.standard_normal((b, n)) / b
    dw, db = ops.dense_backward_dw(to_dev(x), to_dev(dz))
    assert rel_err(to_np(dw), np.dot(x.T, dz)) < F32_TOL
    assert rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL
    assert cond_ok(to_np(dw), np.dot(x.T, dz), np.dot(np.abs(x).T, np.abs(dz)), F32_TOL)
    assert cond_ok(to_np(db), dz.sum(axis=0), np.abs(dz).sum(axis=0), F32_TOL)


@pytest.mark.parametrize("shape", [(64, 784, 128), (50, 64, 10), (257, 100, 130)])
def test_dw_sgd_fused_update_f32(cuda_device, shape):
    from ddnn_b200 import ops
    b, k, n = shape
    rng = np.random.default_rng(3)
    x, dz = rng.random((b, k)), rng.standard_normal((b, n)) / b
    w, bias = rng.standard_normal((k, n)), rng.standard_normal(n)
    wd, bd = to_dev(w), to_dev(bias)
    ops.dense_backward_dw_sgd(to_dev(x), to_dev(dz), wd, bd, 0.1)
    assert rel_err(to_np(wd), w - 0.1 * np.dot(x.T, dz)) < F32_TOL
    assert rel_err(to_np(bd), bias - 0.1 * dz.sum(axis=0)) < F32_TOL


def test_dw_is_run_to_run_deterministic(cuda_device):
    from ddnn_b200 import ops
    rng = np.random.default_rng(4)
    x, dz = to_dev(rng.random((4096, 784))), to_dev(rng.standard_normal((4096, 128)))
    a, _ = ops.dense_backward_dw(x, dz)
    b, _ = ops.dense_backward_dw(x, dz)
    assert torch.equal(a, b)


@pytest.mark.parametrize("bn", [(64, 10), (7, 2), (33, 100), (16, 1000), (9, 1024), (5, 1030), (4, 8192), (3, 4099)])
def test_softmax_ce(cuda_device, bn):
    from ddnn_b200 import ops
    b, n = bn
    rng = np.random.default_rng(5)
    z = rng.standard_normal((b, n)) * 3
    y = rng.integers(0, n, size=b)
    loss, p, dz = oracle.softmax_ce(z, y, denom=b * 2)
    zd = to_dev(z)
    dzd = torch.empty_like(zd)
    pd = torch.empty_like(zd)
    rl = torch.empty((b,), device="cuda")
    ls = torch.zeros((1,), device="cuda")
    ops.softmax_ce(zd, to_dev(y, torch.int32), dzd, rl, ls, 1.0 / (b * 2), p=pd)
    assert rel_err(to_np(pd), p) < F32_TOL
    assert rel_err(to_np(dzd), dz) < F32_TOL
    assert abs(ls.item() - loss) < F32_TOL * abs(loss) + 1e-7
    # accumulate adds
    ops.softmax_ce(zd, to_dev(y, torch.int32), dzd, rl, ls, 1.0 / (b * 2), accumulate=True)
    assert abs(ls.item() - 2 * loss) < 2 * F32_TOL * abs(loss) + 1e-6
    # softmax rows (grpc_node.py:68-71), in place
    ops.softmax_rows(zd, out=zd)
    assert rel_err(to_np(zd), oracle.activate(z, "softmax")) < F32_TOL


def test_activate_matches_reference_grid(cuda_device, golden_small):
    from ddnn_b200 import ops
    a = golden_small["activations"]
    grid = np.array(a["grid"])
    for func in ("relu", "sigmoid", "softmax", "unknown"):
        got = to_np(ops.activate(to_dev(grid), func))
        np.testing.assert_allclose(got, np.array(a["outputs"][func]), rtol=1e-5, atol=1e-30)


def test_empty_batch_is_a_no_op(cuda_device):
    from ddnn_b200 import ops
    x = torch.empty((0, 8), device="cuda")
    w = torch.zeros((8, 4), device="cuda")
    out = ops.dense_forward(x, w, torch.zeros(4, device="cuda"), "relu")
    assert out.shape == (0, 4)


def test_cast_roundtrip(cuda_device):
    from ddnn_b200 import ops
    t = torch.randn(1003, device="cuda")
    b = ops.cast(t, torch.bfloat16)
    assert torch.equal(b, t.to(torch.bfloat16))
    assert torch.equal(ops.cast(b, torch.float32), b.float())


# ---- bf16 operands on the SIMT engine (narrow / misaligned layers of a bf16 model) -------------

@pytest.mark.parametrize("shape", [(64, 784, 128), (50, 64, 10), (257, 100, 130)])
def test_bf16_simt_engine(cuda_device, shape):
    from ddnn_b200 import _lib, ops
    b, k, n = shape
    x, w, bias = _data(b, k, n, seed=6)
    rng = np.random.default_rng(7)
    dz = rng.standard_normal((b, n)) / b
    _lib.set_engine(_lib.ENGINE_SIMT)
    try:
        xb, wb, dzb = (to_dev(t, torch.bfloat16) for t in (x, w, dz))
        # compare against the oracle evaluated on the bf16-rounded operands: isolates kernel error
        xr, wr, dzr = (to_np(t) for t in (xb, wb, dzb))
        got = ops.dense_forward(xb, wb, to_dev(bias), "relu", out_dtype=torch.float32)
        assert rel_err(to_np(got), oracle.activate(np.dot(xr, wr) + bias, "relu")) < F32_TOL
        got = ops.dense_forward(xb, wb, to_dev(bias), "relu")
        assert rel_err(to_np(got), oracle.activate(np.dot(x, w) + bias, "relu")) < BF16_ACT_TOL
        dw, db = ops.dense_backward_dw(xb, dzb)
        assert rel_err(to_np(dw), np.dot(xr.T, dzr)) < F32_TOL
        assert rel_err(to_np(dw), np.dot(x.T, dz)) < BF16_GRAD_TOL
        assert rel_err(to_np(db), dzr.sum(axis=0)) < F32_TOL
        dx = ops.dense_backward_dx(dzb, wb, None, "linear")
        assert rel_err(to_np(dx), np.dot(dz, w.T)) < BF16_ACT_TOL
    finally:
        _lib.set_engine(_lib.ENGINE_AUTO)


@pytest.mark.parametrize("shape", [(50, 64, 10), (257, 100, 130), (33, 20, 5)])
def test_simt_engine_f32_forced(cuda_device, shape):
    """The generic register-tiled kernels on shapes that AUTO routes to the skinny kernels."""
    from ddnn_b200 import _lib, ops
    b, k, n = shape
    x, w, bias = _data(b, k, n, seed=8)
    rng = np.random.default_rng(9)
    dz = rng.standard_normal((b, n)) / b
    _lib.set_engine(_lib.ENGINE_SIMT)
    try:
        got = ops.dense_forward(to_dev(x), to_dev(w), to_dev(bias), "sigmoid")
        assert rel_err(to_np(got), oracle.activate(np.dot(x, w) + bias, "sigmoid")) < F32_TOL
        dw, db = ops.dense_backward_dw(to_dev(x), to_dev(dz))
        assert rel_err(to_np(dw), np.dot(x.T, dz)) < F32_TOL and rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL
        dx = ops.dense_backward_dx(to_dev(dz), to_dev(w), None, "linear")
        assert rel_err(to_np(dx), np.dot(dz, w.T)) < F32_TOL
    finally:
        _lib.set_engine(_lib.ENGINE_AUTO)


@pytest.mark.parametrize("bkn", [(65536, 256, 10), (4096, 64, 10), (777, 300, 16), (100, 1000, 1), (64, 5, 13)])
@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_skinny_output_layer_kernels(cuda_device, bkn, dt):
    """N <= 16 layers (the 10-class softmax layer): forward, dX, dW/db at narrow-config sizes."""
    from ddnn_b200 import ops
    b, k, n = bkn
    tdt = torch.float32 if dt == "f32" else torch.bfloat16
    rng = np.random.default_rng(10)
    x = rng.random((b, k)) - 0.25
    w = rng.standard_normal((k, n)) / np.sqrt(k)
    bias = rng.standard_normal(n) * 0.1
    dz = rng.standard_normal((b, n)) / b
    a_prev = oracle.activate(rng.standard_normal((b, k)), "relu")
    xd, wd, dzd, ad = (to_dev(t, tdt) for t in (x, w, dz, a_prev))
    xr, wr, dzr, ar = (to_np(t) for t in (xd, wd, dzd, ad))  # operands as the kernel sees them
    got = ops.dense_forward(xd, wd, to_dev(bias), "linear", out_dtype=torch.float32)
    assert rel_err(to_np(got), np.dot(xr, wr) + bias) < F32_TOL
    dw, db = ops.dense_backward_dw(xd, dzd)
    assert rel_err(to_np(dw), np.dot(xr.T, dzr)) < F32_TOL
    assert rel_err(to_np(db), dzr.sum(axis=0)) < F32_TOL
    dx = ops.dense_backward_dx(dzd, wd, ad, "relu")
    ref = np.dot(dzr, wr.T) * (ar > 0)
    assert rel_err(to_np(dx), ref) < (F32_TOL if dt == "f32" else 6e-3)


@pytest.mark.parametrize("shape", [(4096, 784, 128), (333, 128, 64), (4096, 64, 128), (1000, 100, 132),
                                   (2048, 2048, 256), (777, 36, 32), (5000, 260, 68), (20000, 64, 4096),
                                   (256, 4096, 64), (1, 784, 128), (5, 64, 32)])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_f32_inner_product_cores(cuda_device, shape, mode):
    """fp32 forward, dX and dW on 16-byte-aligned shapes with each inner product: FFMA (0), error-compensated
    3xTF32 on mma.sync (1) and on tcgen05.mma.kind::tf32 (2, the default: hi/lo operand split in shared memory,
    correction terms in their own TMEM accumulator).  All must stay within the fp32 bound of the float64 oracle;
    shapes cover ragged M / N / K tiles (zero-filled TMA boxes), every BN and the split-K dW plan."""
    from ddnn_b200 import _lib, ops
    b, k, n = shape
    if mode == 1 and max(shape) >= 4096 and shape != (4096, 784, 128) and shape != (4096, 64, 128):
        pytest.skip("the mma.sync variant has no bound on its accumulation chain (comparison only)")
    x, w, bias = _data(b, k, n, seed=12)
    rng = np.random.default_rng(13)
    dz = rng.standard_normal((b, n)) / b
    a_prev = np.maximum(rng.standard_normal((b, k)), 0.0)
    _lib.set_f32_tensor_cores(mode)
    try:
        got = ops.dense_forward(to_dev(x), to_dev(w), to_dev(bias), "relu")
        assert rel_err(to_np(got), oracle.activate(np.dot(x, w) + bias, "relu")) < F32_TOL
        dw, db = ops.dense_backward_dw(to_dev(x), to_dev(dz))
        assert rel_err(to_np(dw), np.dot(x.T, dz)) < F32_TOL
        assert rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL
        dx = ops.dense_backward_dx(to_dev(dz), to_dev(w), to_dev(a_prev), "relu")
        assert rel_err(to_np(dx), np.dot(dz, w.T) * (a_prev > 0)) < F32_TOL
    finally:
        _lib.set_f32_tensor_cores(None)


@pytest.mark.parametrize("case", [(4096, 64, 10, "f32", "relu"), (1000, 16, 3, "f32", "sigmoid"),
                                  (300, 256, 16, "f32", "linear"), (777, 128, 1, "f32", "relu"),
                                  (65536, 256, 10, "bf16", "relu"), (513, 64, 13, "bf16", "sigmoid"),
                                  (1000, 48, 16, "bf16", "linear"), (130, 256, 2, "bf16", "relu"),
                                  (20000, 128, 9, "bf16", "sigmoid"), (513, 24, 5, "bf16", "relu")])
def test_fused_output_layer_step(cuda_device, case):
    """ddnn_output_layer_train: logits + softmax-CE + dX + dW/db (+ SGD) of a skinny output layer in one
    pass, against the oracle evaluated on the operands as the kernel sees them.  bf16 with K % 16 == 0 runs the
    mma.sync variant (128-row tiles, ragged last tile, K < 256 leaves warps without features), K = 24 the FFMA one."""
    from ddnn_b200 import ops
    b, k, n, dt, act_prev = case
    tdt = torch.float32 if dt == "f32" else torch.bfloat16
    rng = np.random.default_rng(30)
    a = oracle.activate(rng.standard_normal((b, k)), act_prev)
    w = rng.standard_normal((k, n)) / np.sqrt(k)
    bias = rng.standard_normal(n) * 0.1
    y = rng.integers(0, n, size=b)
    ad, wd = to_dev(a, tdt), to_dev(w, tdt)
    ar, wr = to_np(ad), to_np(wd)
    assert ops.output_layer_supported(ad, k, n)
    loss, _, dz = oracle.softmax_ce(ar @ wr + bias, y)
    zprev_sign = {"relu": (ar > 0).astype(np.float64), "sigmoid": ar * (1 - ar), "linear": np.ones_like(ar)}[act_prev]
    ref_dx = (dz @ wr.T) * zprev_sign
    ws = torch.empty(max(256, ops.output_layer_workspace_bytes(b, k, n, tdt)), dtype=torch.uint8, device="cuda")
    row_loss = torch.empty(b, device="cuda")
    ls = torch.zeros(1, device="cuda")
    dzd = torch.empty((b, n), dtype=tdt, device="cuda")
    dxd = torch.empty((b, k), dtype=tdt, device="cuda")
    dw = torch.empty((k, n), device="cuda")
    db = torch.empty(n, device="cuda")
    bd = to_dev(bias)
    ops.output_layer_train(ad, wd, to_dev(y, torch.int32), row_loss, ls, 1.0 / b, act_prev, ws, bias=bd, dz=dzd,
                           dz_prev=dxd, dw_out=dw, db_out=db)
    out_tol = F32_TOL if dt == "f32" else 6e-3   # bf16 outputs carry one rounding
    assert abs(ls.item() - loss) < F32_TOL * abs(loss) + 1e-7
    assert rel_err(to_np(dzd), dz) < out_tol
    assert rel_err(to_np(dxd), ref_dx) < out_tol
    assert rel_err(to_np(dw), ar.T @ dz) < F32_TOL
    assert rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL
    # fused SGD update, twice (the second call finds the first call's partials in the workspace)
    wm = to_dev(wr)
    sh = to_dev(wr, torch.bfloat16) if dt == "bf16" else None
    for step in range(2):
        w_before, b_before = to_np(wm).copy(), to_np(bd).copy()
        w_used = to_np(sh) if dt == "bf16" else w_before
        _, _, dz2 = oracle.softmax_ce(ar @ w_used + b_before, y)
        ops.output_layer_train(ad, sh if dt == "bf16" else wm, to_dev(y, torch.int32), row_loss, ls, 1.0 / b, act_prev,
                               ws, w_master=wm, bias=bd, lr=0.3, w_shadow=sh)
        assert rel_err(to_np(wm), w_before - 0.3 * (ar.T @ dz2)) < F32_TOL
        assert rel_err(to_np(bd), b_before - 0.3 * dz2.sum(axis=0)) < F32_TOL
        if sh is not None:
            assert torch.equal(sh, wm.to(torch.bfloat16))


@pytest.mark.parametrize("case", [(4096, 64, 10, "f32"), (20000, 256, 10, "bf16")])
def test_fused_output_layer_two_call_form(cuda_device, case):
    """apply = DDNN_OUTLAYER_DEFER + ddnn_output_layer_finish on another stream gives bit-identical weights, bias
    and loss to the single call."""
    from ddnn_b200 import ops
    b, k, n, dt = case
    tdt = torch.float32 if dt == "f32" else torch.bfloat16
    rng = np.random.default_rng(31)
    a = to_dev(np.maximum(rng.standard_normal((b, k)), 0), tdt)
    w0 = rng.standard_normal((k, n)) / np.sqrt(k)
    y = to_dev(rng.integers(0, n, size=b), torch.int32)
    ws = torch.empty(max(256, ops.output_layer_workspace_bytes(b, k, n, tdt)), dtype=torch.uint8, device="cuda")
    res = []
    for deferred in (False, True):
        wm, bias = to_dev(w0), to_dev(np.full(n, 0.01))
        sh = wm.to(torch.bfloat16) if dt == "bf16" else None
        row_loss, loss = torch.empty(b, device="cuda"), torch.zeros(1, device="cuda")
        dxd = torch.empty((b, k), dtype=tdt, device="cuda")
        wop = sh if dt == "bf16" else wm
        if not deferred:
            ops.output_layer_train(a, wop, y, row_loss, loss, 1.0 / b, "relu", ws, w_master=wm, bias=bias, lr=0.1,
                                   w_shadow=sh, dz_prev=dxd)
        else:
            ops.output_layer_train(a, wop, y, row_loss, None, 1.0 / b, "relu", ws, bias=bias, dz_prev=dxd, defer=True)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                ops.output_layer_finish(b, k, n, tdt, row_loss, loss, 1.0 / b, ws, w_master=wm, bias=bias, lr=0.1,
                                        w_shadow=sh)
            torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        res.append((wm.clone(), bias.clone(), loss.clone(), dxd.clone()))
    for x0, x1 in zip(*res):
        assert torch.equal(x0, x1)
    assert not torch.equal(res[0][0], to_dev(w0))  # the update happened


@pytest.mark.parametrize("case", [(4096, 128, 64, 10, "relu", "relu"), (1000, 32, 16, 10, "relu", "sigmoid"),
                                  (77, 100, 36, 3, "sigmoid", "linear"), (5000, 128, 64, 16, "linear", "relu"),
                                  (33, 4, 4, 1, "relu", "relu"), (20000, 64, 48, 7, "relu", "sigmoid")])
@pytest.mark.parametrize("mode", [0, 2])
def test_fused_two_layer_tail(cuda_device, case, mode):
    """ddnn_tail2_train / ddnn_tail2_finish (hidden layer + output layer, forward + softmax-CE + backward + SGD in
    one kernel and its reducers) against the float64 oracle run layer by layer: loss, dZ1, the four gradients,
    and the in-place update; FFMA variant (mode 0) and 3xTF32 mma.sync variant (default).  Shapes cover ragged
    last tiles, more tiles than SMs (persistent loop), padded K1 / N2 / N3 and every activation pair."""
    from ddnn_b200 import _lib, ops
    _lib.set_f32_tensor_cores(mode)
    try:
        _tail2_case(ops, case)
    finally:
        _lib.set_f32_tensor_cores(None)


def _tail2_case(ops, case):
    b, k1, n2, n3, act1, act2 = case
    rng = np.random.default_rng(40)
    a1 = oracle.activate(rng.standard_normal((b, k1)), act1)
    params = [{"weights": rng.standard_normal((k1, n2)) / np.sqrt(k1), "biases": rng.standard_normal(n2) * 0.1,
               "activation": act2},
              {"weights": rng.standard_normal((n2, n3)) / np.sqrt(n2), "biases": rng.standard_normal(n3) * 0.1,
               "activation": "softmax"}]
    y = rng.integers(0, n3, size=b)
    acts, zs = oracle.forward_all(params, a1)
    loss, dws, dbs, da1 = oracle.backward(params, acts, zs, y)
    f1 = {"relu": (a1 > 0).astype(np.float64), "sigmoid": a1 * (1 - a1), "linear": np.ones_like(a1)}[act1]
    ws = torch.empty(max(256, ops.tail2_workspace_bytes(b, k1, n2, n3, torch.float32)), dtype=torch.uint8, device="cuda")
    assert ws.numel() > 256
    a1d, yd = to_dev(a1), to_dev(y, torch.int32)
    w2, b2, w3, b3 = (to_dev(params[0]["weights"]), to_dev(params[0]["biases"]), to_dev(params[1]["weights"]),
                      to_dev(params[1]["biases"]))
    row_loss, ls = torch.empty(b, device="cuda"), torch.zeros(1, device="cuda")
    dz1 = torch.empty((b, k1), device="cuda")
    ops.tail2_train(a1d, w2, b2, w3, b3, yd, row_loss, 1.0 / b, act1, act2, ws, dz1=dz1)
    grads = (torch.empty((k1, n2), device="cuda"), torch.empty(n2, device="cuda"), torch.empty((n2, n3), device="cuda"),
             torch.empty(n3, device="cuda"))
    ops.tail2_finish(b, row_loss, ls, 1.0 / b, ws, w2, b2, w3, b3, apply=False, grads=grads)
    assert abs(ls.item() - loss) < F32_TOL * abs(loss) + 1e-7
    assert rel_err(to_np(dz1), da1 * f1) < F32_TOL
    for got, ref in zip(grads, (dws[0], dbs[0], dws[1], dbs[1])):
        assert rel_err(to_np(got), ref) < F32_TOL
    # in-place SGD through the same partials
    ops.tail2_finish(b, row_loss, None, 1.0 / b, ws, w2, b2, w3, b3, lr=0.2)
    assert rel_err(to_np(w2), params[0]["weights"] - 0.2 * dws[0]) < F32_TOL
    assert rel_err(to_np(b2), params[0]["biases"] - 0.2 * dbs[0]) < F32_TOL
    assert rel_err(to_np(w3), params[1]["weights"] - 0.2 * dws[1]) < F32_TOL
    assert rel_err(to_np(b3), params[1]["biases"] - 0.2 * dbs[1]) < F32_TOL
    # shapes the kernel does not take: no workspace, DDNN_ERR_UNSUPPORTED
    assert ops.tail2_workspace_bytes(b, 256, 64, 10, torch.float32) == 0
    assert ops.tail2_workspace_bytes(b, 128, 64, 10, torch.bfloat16) == 0


@pytest.mark.timeout(120)
@pytest.mark.parametrize("shape", [(4096, 784, 128), (1000, 300, 72), (512, 128, 64), (3000, 256, 256), (96, 64, 48)])
def test_cluster_split_k_dw_f32(cuda_device, shape):
    """fp32 dW with the split-K reduction inside the kernel (thread-block cluster + distributed shared memory, SGD
    applied by the same kernel): gradients and fused update against the float64 oracle and against the workspace +
    reducer path (same products, different summation order of the partial tiles), run-to-run bit-identical."""
    from ddnn_b200 import _lib, ops
    b, k, n = shape
    rng = np.random.default_rng(40)
    x, dz = rng.random((b, k)), rng.standard_normal((b, n)) / b
    w, bias = rng.standard_normal((k, n)), rng.standard_normal(n)
    xd, dzd = to_dev(x), to_dev(dz)
    dw, db = ops.dense_backward_dw(xd, dzd)
    dw2, db2 = ops.dense_backward_dw(xd, dzd)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)
    assert rel_err(to_np(dw), np.dot(x.T, dz)) < F32_TOL and rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL
    assert cond_ok(to_np(dw), np.dot(x.T, dz), np.dot(np.abs(x).T, np.abs(dz)), F32_TOL)
    wd, bd = to_dev(w), to_dev(bias)
    ops.dense_backward_dw_sgd(xd, dzd, wd, bd, 0.1)
    assert rel_err(to_np(wd), w - 0.1 * np.dot(x.T, dz)) < F32_TOL
    assert rel_err(to_np(bd), bias - 0.1 * dz.sum(axis=0)) < F32_TOL
    _lib.check(_lib.lib.ddnn_set_cluster_split_k(0))
    try:
        dw_ws, db_ws = ops.dense_backward_dw(xd, dzd)
    finally:
        _lib.check(_lib.lib.ddnn_set_cluster_split_k(1))
    assert rel_err(to_np(dw), to_np(dw_ws)) < 2e-6 and rel_err(to_np(db), to_np(db_ws)) < 2e-6


@pytest.mark.timeout(120)
@pytest.mark.parametrize("shape", [(4096, 784, 128), (4096, 784, 256), (1000, 1024, 72), (2048, 2000, 512)])
@pytest.mark.parametrize("act", ["relu", "sigmoid"])
def test_cluster_split_k_forward_f32(cuda_device, shape, act):
    """fp32 forward with few output tiles and a long reduction: the reduction is split over a thread-block cluster,
    the partial tiles are summed through distributed shared memory and bias + activation run on the sum.  Against the
    float64 oracle (normalised and per-element condition-sum bounds) and against the plain kernel (summation order
    only); run-to-run bit-identical.  Shapes without a cluster plan simply take the plain kernel in both runs."""
    from ddnn_b200 import _lib, ops
    b, k, n = shape
    x, w, bias = _data(b, k, n, seed=41)
    xd, wd, bd = to_dev(x), to_dev(w), to_dev(bias)
    ref = oracle.activate(np.dot(x, w) + bias, act)
    got = ops.dense_forward(xd, wd, bd, act)
    assert torch.equal(got, ops.dense_forward(xd, wd, bd, act))
    assert rel_err(to_np(got), ref) < F32_TOL
    assert cond_ok(to_np(got), ref, np.dot(np.abs(x), np.abs(w)) + np.abs(bias), F32_TOL)
    _lib.check(_lib.lib.ddnn_set_cluster_split_k(0))
    try:
        plain = ops.dense_forward(xd, wd, bd, act)
    finally:
        _lib.check(_lib.lib.ddnn_set_cluster_split_k(1))
    assert rel_err(to_np(got), to_np(plain)) < 2e-6
