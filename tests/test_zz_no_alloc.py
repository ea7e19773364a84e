"""ddnn_dense_backward_dw_sgd_static: the weight-gradient / SGD step without a caller workspace.

Some hosts of libddnn.so have no way to give it device memory -- the GPU's memory already
belongs to their framework's private pool so a plain cudaMalloc fails, or they are a thin
FFI stub that can only pass on pointers it received, or allocation is banned on their request
path because it stalls the device.  For them the split-K scratch is an 8 MiB array inside the
library.  These tests drive that entry point through the C ABI exactly as such a host would
(inputs and outputs only, no workspace argument) and hold its results against the regular
entry point, which gets a workspace from torch.

The file name sorts last on purpose: under `pytest -x` every other test has already run when
this one starts, and it must pass both on its own in a fresh process and after the rest of
the suite has used the library (descriptor caches warm, other kernels launched before it).
"""
import numpy as np
import pytest
import torch

from gpu_util import F32_TOL, rel_err, to_dev, to_np

pytestmark = pytest.mark.gpu


def _static_call(x, dz, w=None, bias=None, dw=None, db=None, lr=0.0, shadow=None):
    """One call of the no-workspace entry point through ctypes; raises on a non-zero status."""
    from ddnn_b200 import _lib
    b, k = x.shape
    n = dz.shape[1]
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    apply = 1 if w is not None else 0
    rc = _lib.lib.ddnn_dense_backward_dw_sgd_static(
        ptr(x), ptr(dz), ptr(w), ptr(shadow), ptr(bias), ptr(dw), ptr(db), float(lr), apply, b, k, n,
        _lib.F32 if x.dtype == torch.float32 else _lib.BF16, torch.cuda.current_stream().cuda_stream)
    _lib.check(rc)


def test_two_calls_in_a_row_match_the_workspace_entry_point(cuda_device):
    """Back-to-back calls with different layers: the second call finds the first call's partial
    sums still in the shared scratch and must not be affected by them.  Gradients are compared
    with the regular entry point (same kernels, caller-owned workspace) and with float64."""
    from ddnn_b200 import ops
    rng = np.random.default_rng(21)
    for (b, k, n) in [(4096, 784, 128), (512, 64, 10)]:          # MNIST layer 1, then the ragged N = 10 layer
        x = rng.random((b, k))
        dz = rng.standard_normal((b, n)) / b
        xd, dzd = to_dev(x), to_dev(dz)
        dw = torch.empty((k, n), device="cuda")
        db = torch.empty((n,), device="cuda")
        _static_call(xd, dzd, dw=dw, db=db)
        ref_dw, ref_db = ops.dense_backward_dw(xd, dzd)
        assert rel_err(to_np(dw), to_np(ref_dw)) < F32_TOL
        assert rel_err(to_np(db), to_np(ref_db)) < F32_TOL
        assert rel_err(to_np(dw), x.T @ dz) < F32_TOL
        assert rel_err(to_np(db), dz.sum(axis=0)) < F32_TOL


def test_fused_sgd_twice_matches_the_workspace_entry_point(cuda_device):
    """Two SGD updates in a row on the same weights: the fused update through the built-in
    scratch tracks the regular entry point step for step."""
    from ddnn_b200 import ops
    rng = np.random.default_rng(22)
    b, k, n = 1000, 300, 70
    w0, b0 = rng.standard_normal((k, n)), rng.standard_normal(n)
    w_s, b_s, w_r, b_r = to_dev(w0), to_dev(b0), to_dev(w0), to_dev(b0)
    for step in range(2):
        xd = to_dev(rng.random((b, k)))
        dzd = to_dev(rng.standard_normal((b, n)) / b)
        _static_call(xd, dzd, w=w_s, bias=b_s, lr=0.25)
        ops.dense_backward_dw_sgd(xd, dzd, w_r, b_r, 0.25)
        assert rel_err(to_np(w_s), to_np(w_r)) < F32_TOL, step
        assert rel_err(to_np(b_s), to_np(b_r)) < F32_TOL, step


def test_bf16_layer_through_the_static_scratch(cuda_device):
    """A bf16 layer (tcgen05 split-K partials + column-sum partials share the scratch)."""
    from ddnn_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.rand((2048, 512), device="cuda", generator=g).to(torch.bfloat16)
    dz = (torch.randn((2048, 256), device="cuda", generator=g) / 2048).to(torch.bfloat16)
    for _ in range(2):
        dw = torch.empty((512, 256), device="cuda")
        db = torch.empty((256,), device="cuda")
        _static_call(x, dz, dw=dw, db=db)
        ref_dw, ref_db = ops.dense_backward_dw(x, dz)
        assert rel_err(to_np(dw), to_np(ref_dw)) < F32_TOL
        assert rel_err(to_np(db), to_np(ref_db)) < F32_TOL


def test_layers_beyond_the_scratch_limit_are_refused(cuda_device):
    """A layer whose split-K plan does not fit in DDNN_STATIC_SCRATCH_BYTES is rejected with a
    readable message and nothing is written."""
    from ddnn_b200 import _lib
    b, k, n = 256, 4096, 4096
    assert _lib.lib.ddnn_dw_workspace_bytes(b, k, n, _lib.F32) > _lib.STATIC_SCRATCH_BYTES
    x = torch.zeros((b, k), device="cuda")
    dz = torch.zeros((b, n), device="cuda")
    dw = torch.full((k, n), 7.0, device="cuda")
    db = torch.full((n,), 7.0, device="cuda")
    with pytest.raises(_lib.DdnnError) as e:
        _static_call(x, dz, dw=dw, db=db)
    assert e.value.status == _lib.ERR_WORKSPACE
    assert "built-in scratch holds" in e.value.text and "caller-owned workspace" in e.value.text
    torch.cuda.synchronize()
    assert float(dw.min()) == 7.0 and float(db.max()) == 7.0


def test_no_device_memory_is_taken_by_the_call(cuda_device):
    """What the entry point is for: after its first call on a device (which looks up the built-in array's
    address), further calls leave the device's free memory and torch's allocator untouched -- the driver hands
    out nothing and nothing is returned, however many layers go through."""
    rng = np.random.default_rng(23)
    xd, dzd = to_dev(rng.random((2048, 256))), to_dev(rng.standard_normal((2048, 64)) / 2048)
    dw = torch.empty((256, 64), device="cuda")
    db = torch.empty((64,), device="cuda")
    _static_call(xd, dzd, dw=dw, db=db)      # one-time set-up on this device happens here at the latest
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    st0 = torch.cuda.memory_stats()
    for _ in range(5):
        _static_call(xd, dzd, dw=dw, db=db)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    st1 = torch.cuda.memory_stats()
    assert free1 == free0                                                     # the driver handed out nothing
    for counter in ("allocation.all.allocated", "segment.all.allocated"):     # and torch was asked for nothing
        assert st1[counter] == st0[counter], counter


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs a second GPU")
def test_every_device_gets_its_own_scratch_address(cuda_device):
    """The array lives once per GPU and its address differs from device to device: a call with cuda:1 current must
    use cuda:1's copy (the look-up is per device ordinal), and give the regular entry point's result there."""
    from ddnn_b200 import ops
    rng = np.random.default_rng(24)
    x, dz = rng.random((1024, 128)), rng.standard_normal((1024, 32)) / 1024
    for dev in ("cuda:1", "cuda:0"):
        with torch.cuda.device(dev):
            xd, dzd = to_dev(x, device=dev), to_dev(dz, device=dev)
            dw = torch.empty((128, 32), device=dev)
            db = torch.empty((32,), device=dev)
            _static_call(xd, dzd, dw=dw, db=db)
            ref_dw, ref_db = ops.dense_backward_dw(xd, dzd)
            assert rel_err(to_np(dw), to_np(ref_dw)) < F32_TOL and rel_err(to_np(db), to_np(ref_db)) < F32_TOL
            assert rel_err(to_np(dw), x.T @ dz) < F32_TOL
