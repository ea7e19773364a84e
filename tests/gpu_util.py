import numpy as np
import torch


def rel_err(got, ref) -> float:
    """max |got - ref| / max |ref|  (normalised max error; scale-free, robust near zero)."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    if ref.size == 0:
        return 0.0
    return float(np.max(np.abs(got - ref)) / (np.max(np.abs(ref)) + 1e-300))


def to_dev(a, dtype=torch.float32, device="cuda"):
    return torch.as_tensor(np.ascontiguousarray(a), dtype=dtype).to(device).contiguous()


def to_np(t):
    return t.detach().double().cpu().numpy()


# north_star: fp32 path within 1e-5 relative of the float64 reference
F32_TOL = 1e-5
# bf16 operands (8-bit mantissa, eps = 2^-8) with fp32 accumulation: activations within 2e-2 of
# the float64 oracle (normalised max error), weight gradients within 3e-2
BF16_ACT_TOL = 2e-2
BF16_GRAD_TOL = 3e-2


def bf16_update_tol(rows: int, below_relu: bool) -> float:
    """Bound (normalised max error) on a weight UPDATE of the bf16 path against the float64 oracle.  Rounding
    alone (operands to bf16, eps = 2^-9 each, fp32 accumulation) stays within BF16_GRAD_TOL.  A layer BELOW a ReLU
    has a second, larger term: the bf16 forward error (~2e-3 of a pre-activation at K = 8192) flips the gates of
    pre-activations that close to zero, and a flipped gate adds or removes ONE whole term x * dz of the ``rows``-term
    sum behind each element of dW.  The largest such term is ~5 sigma(dz) * max(x) while max |dW| over the matrix
    is ~5.5 * sqrt(rows) * rms(x) * sigma(dz), i.e. the error is <= ~1.6 / sqrt(rows) of the tensor's maximum
    (9% at 300 rows, 0.6% at 65536): 2 / sqrt(rows) is allowed for it."""
    return BF16_GRAD_TOL + (2.0 / float(rows) ** 0.5 if below_relu else 0.0)


def elementwise_ok(got, ref, rtol: float, atol: float) -> bool:
    """Elementwise check beside the normalised-max one: |got - ref| <= rtol * |ref| + atol for EVERY element, so
    an element far below the tensor's maximum cannot hide a large relative error behind it.  ``atol`` is stated
    by the caller in the units of the tensor (it covers elements that cancel to ~0)."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    return bool(np.all(np.abs(got - ref) <= rtol * np.abs(ref) + atol))


def cond_ok(got, ref, cond, tol: float) -> bool:
    """Elementwise bound for a sum of products: |got - ref| <= tol * cond for EVERY element, where ``cond`` is the
    same expression evaluated on absolute values (sum_k |a_k| |b_k|, the condition sum of that element).  Unlike a
    plain relative check it stays meaningful for elements that cancel to ~0, and unlike the normalised-max check it
    does not let a small element hide behind the tensor's largest one."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape == np.asarray(cond).shape
    return bool(np.all(np.abs(got - ref) <= tol * np.asarray(cond) + 1e-300))
