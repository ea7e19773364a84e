"""Thin torch-tensor wrappers over the C ABI (include/ddnn.h).

torch is used for device memory and the current stream only; every computation below is
one call into libddnn.so.  Tensors must be CUDA, contiguous and of the dtype stated.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import ACT_LINEAR, ACT_SOFTMAX, BF16, F32, check, lib

_DT = {torch.float32: F32, torch.bfloat16: BF16}


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _req(t: torch.Tensor, dtype=None, name="tensor") -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must live on a CUDA device: libddnn has no CPU path")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")


def make_signals(wait_flag: Optional[int] = None, wait_value: int = 0, done_flag: Optional[int] = None,
                 done_value: int = 0, counter: Optional[int] = None):
    """A ``ddnn_signals`` for ``dense_forward`` / ``dense_backward_dx`` (raw device addresses of int32 flags): the
    kernel waits for ``*wait_flag >= wait_value`` before loading its operand and/or its last CTA stores
    ``done_value`` to ``*done_flag`` (``counter``: 4 zeroed bytes on the device, one per stream)."""
    if wait_flag is None and done_flag is None:
        return None
    return _lib.Signals(wait_flag, int(wait_value), done_flag, int(done_value), counter)


def dense_forward(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, act, out: Optional[torch.Tensor] = None,
                  out_dtype: Optional[torch.dtype] = None, signals=None) -> torch.Tensor:
    """act(x @ w + bias).  x (B,K), w (K,N) share a dtype (fp32/bf16); bias fp32."""
    _req(x, name="x"); _req(w, x.dtype, "w"); _req(bias, torch.float32, "bias")
    b, k = x.shape
    k2, n = w.shape
    if k != k2 or bias.numel() != n:
        raise ValueError(f"expected input dim {k2}, got {k}")
    act_id = act if isinstance(act, int) else _lib.activation_id(act)
    if out is None:
        od = out_dtype or (torch.float32 if act_id == ACT_SOFTMAX else x.dtype)
        out = torch.empty((b, n), dtype=od, device=x.device)
    _req(out, name="out")
    if tuple(out.shape) != (b, n):  # out may be a neighbour GPU's buffer: never hand a mismatched slice to the kernel
        raise ValueError(f"dense_forward: out has shape {tuple(out.shape)}, expected {(b, n)}")
    with torch.cuda.device(x.device):
        if signals is None:
            check(lib.ddnn_dense_forward(_p(x), _p(w), _p(bias), _p(out), b, k, n, act_id, _DT[x.dtype],
                                         _DT[out.dtype], _stream(x)))
        else:
            check(lib.ddnn_dense_forward_signalled(_p(x), _p(w), _p(bias), _p(out), b, k, n, act_id, _DT[x.dtype],
                                                   _DT[out.dtype], signals, _stream(x)))
    return out


def softmax_rows(z: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _req(z, torch.float32, "z")
    out = torch.empty_like(z) if out is None else out
    _req(out, torch.float32, "out")
    with torch.cuda.device(z.device):
        check(lib.ddnn_softmax_rows(_p(z), _p(out), z.shape[0], z.shape[1], _stream(z)))
    return out


def activate(v: torch.Tensor, act, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _req(v, torch.float32, "values")
    if v.dim() != 2:
        raise ValueError("values must be 2-D (rows, features)")
    out = torch.empty_like(v) if out is None else out
    act_id = act if isinstance(act, int) else _lib.activation_id(act)
    with torch.cuda.device(v.device):
        check(lib.ddnn_activate(_p(v), _p(out), v.shape[0], v.shape[1], act_id, _stream(v)))
    return out


def softmax_ce(z: torch.Tensor, labels: torch.Tensor, dz: torch.Tensor, row_loss: torch.Tensor, loss: torch.Tensor,
               inv_denom: float, p: Optional[torch.Tensor] = None, accumulate: bool = False) -> None:
    """dz = (softmax(z) - onehot(labels)) * inv_denom ; loss (+)= inv_denom * sum(-log p[y])."""
    _req(z, torch.float32, "z"); _req(labels, torch.int32, "labels"); _req(dz, name="dz")
    _req(row_loss, torch.float32, "row_loss"); _req(loss, torch.float32, "loss")
    if p is not None:
        _req(p, torch.float32, "p")
    b, n = z.shape
    if labels.numel() != b or dz.shape != z.shape or row_loss.numel() < b:
        raise ValueError("softmax_ce: shape mismatch")
    with torch.cuda.device(z.device):
        check(lib.ddnn_softmax_ce(_p(z), _p(labels), _p(p), _p(dz), _p(row_loss), _p(loss), b, n, float(inv_denom),
                                  _DT[dz.dtype], 1 if accumulate else 0, _stream(z)))


def dense_backward_dx(dz: torch.Tensor, w: torch.Tensor, a_prev: Optional[torch.Tensor], act_prev,
                      out: Optional[torch.Tensor] = None, signals=None) -> torch.Tensor:
    """(dz @ w.T) * f'(a_prev) with f' of the previous layer's activation."""
    _req(dz, name="dz"); _req(w, dz.dtype, "w")
    b, n = dz.shape
    k, n2 = w.shape
    if n != n2:
        raise ValueError(f"expected gradient width {n2}, got {n}")
    act_id = act_prev if isinstance(act_prev, int) else _lib.activation_id(act_prev)
    if a_prev is not None:
        _req(a_prev, dz.dtype, "a_prev")
    if out is None:
        out = torch.empty((b, k), dtype=dz.dtype, device=dz.device)
    _req(out, dz.dtype, "out")
    if tuple(out.shape) != (b, k):
        raise ValueError(f"dense_backward_dx: out has shape {tuple(out.shape)}, expected {(b, k)}")
    if a_prev is not None and act_id != ACT_LINEAR and tuple(a_prev.shape) != (b, k):
        raise ValueError(f"dense_backward_dx: a_prev has shape {tuple(a_prev.shape)}, expected {(b, k)}")
    ap = _p(a_prev if act_id != ACT_LINEAR else None)
    with torch.cuda.device(dz.device):
        if signals is None:
            check(lib.ddnn_dense_backward_dx(_p(dz), _p(w), ap, _p(out), b, k, n, act_id, _DT[dz.dtype], _stream(dz)))
        else:
            check(lib.ddnn_dense_backward_dx_signalled(_p(dz), _p(w), ap, _p(out), b, k, n, act_id, _DT[dz.dtype],
                                                       signals, _stream(dz)))
    return out


def dw_workspace_bytes(b: int, k: int, n: int, dtype: torch.dtype) -> int:
    return int(lib.ddnn_dw_workspace_bytes(b, k, n, _DT[dtype]))


def dense_backward_dw(x: torch.Tensor, dz: torch.Tensor, workspace: Optional[torch.Tensor] = None
                      ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Gradients only: (x.T @ dz, column sums of dz), both fp32."""
    _req(x, name="x"); _req(dz, x.dtype, "dz")
    b, k = x.shape
    n = dz.shape[1]
    dw = torch.empty((k, n), dtype=torch.float32, device=x.device)
    db = torch.empty((n,), dtype=torch.float32, device=x.device)
    ws = _workspace(workspace, b, k, n, x)
    with torch.cuda.device(x.device):
        check(lib.ddnn_dense_backward_dw_sgd(_p(x), _p(dz), None, None, None, _p(dw), _p(db), 0.0, 0, b, k, n,
                                             _DT[x.dtype], _p(ws), ws.numel(), _stream(x)))
    return dw, db


def dense_backward_dw_sgd(x: torch.Tensor, dz: torch.Tensor, w_master: torch.Tensor, bias: torch.Tensor, lr: float,
                          w_shadow: Optional[torch.Tensor] = None, workspace: Optional[torch.Tensor] = None) -> None:
    """Fused update: w_master -= lr * x.T @ dz ; bias -= lr * colsum(dz) ; shadow = bf16(w_master)."""
    _req(x, name="x"); _req(dz, x.dtype, "dz"); _req(w_master, torch.float32, "w_master")
    _req(bias, torch.float32, "bias")
    if w_shadow is not None:
        _req(w_shadow, torch.bfloat16, "w_shadow")
    b, k = x.shape
    n = dz.shape[1]
    if tuple(w_master.shape) != (k, n) or dz.shape[0] != b:
        raise ValueError("dense_backward_dw_sgd: shape mismatch")
    ws = _workspace(workspace, b, k, n, x)
    with torch.cuda.device(x.device):
        check(lib.ddnn_dense_backward_dw_sgd(_p(x), _p(dz), _p(w_master), _p(w_shadow), _p(bias), None, None, float(lr),
                                             1, b, k, n, _DT[x.dtype], _p(ws), ws.numel(), _stream(x)))


def dense_backward_dw_accumulate(x: torch.Tensor, dz: torch.Tensor, dw_acc: torch.Tensor, db_acc: torch.Tensor,
                                 first: bool, workspace: Optional[torch.Tensor] = None) -> None:
    """dw_acc (+)= x.T @ dz ; db_acc (+)= colsum(dz) for one row chunk of the batch (``first`` overwrites)."""
    _req(x, name="x"); _req(dz, x.dtype, "dz"); _req(dw_acc, torch.float32, "dw_acc"); _req(db_acc, torch.float32, "db_acc")
    b, k = x.shape
    n = dz.shape[1]
    if tuple(dw_acc.shape) != (k, n) or db_acc.numel() != n or dz.shape[0] != b:
        raise ValueError("dense_backward_dw_accumulate: shape mismatch")
    ws = _workspace(workspace, b, k, n, x)
    with torch.cuda.device(x.device):
        check(lib.ddnn_dense_backward_dw_accumulate(_p(x), _p(dz), _p(dw_acc), _p(db_acc), 1 if first else 0, b, k, n,
                                                    _DT[x.dtype], _p(ws), ws.numel(), _stream(x)))


def sgd_apply(w_master: torch.Tensor, bias: torch.Tensor, dw: torch.Tensor, db: torch.Tensor, lr: float,
              w_shadow: Optional[torch.Tensor] = None) -> None:
    """w_master -= lr * dw ; bias -= lr * db ; shadow = bf16(w_master): the update after the last gradient chunk."""
    _req(w_master, torch.float32, "w_master"); _req(bias, torch.float32, "bias")
    _req(dw, torch.float32, "dw"); _req(db, torch.float32, "db")
    if w_shadow is not None:
        _req(w_shadow, torch.bfloat16, "w_shadow")
    k, n = w_master.shape
    if tuple(dw.shape) != (k, n) or db.numel() != n or bias.numel() != n:
        raise ValueError("sgd_apply: shape mismatch")
    with torch.cuda.device(w_master.device):
        check(lib.ddnn_sgd_apply(_p(w_master), _p(w_shadow), _p(bias), _p(dw), _p(db), float(lr), k, n,
                                 _stream(w_master)))


def output_layer_supported(a: torch.Tensor, k: int, n: int) -> bool:
    """Shapes the fused output-layer step handles (ddnn_output_layer_train): N <= 16, K <= 256, rows of
    the input 16-byte aligned and at most 64 16-byte chunks long."""
    e = 4 if a.dtype == torch.float32 else 8
    return 1 <= n <= 16 and 1 <= k <= 256 and k % e == 0 and a.data_ptr() % 16 == 0


def output_layer_workspace_bytes(b: int, k: int, n: int, dtype: torch.dtype) -> int:
    return int(lib.ddnn_output_layer_workspace_bytes(b, k, n, _DT[dtype]))


def output_layer_train(a: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, row_loss: torch.Tensor,
                       loss: Optional[torch.Tensor], inv_denom: float, act_prev, workspace: torch.Tensor,
                       w_master: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None, lr: float = 0.0,
                       w_shadow: Optional[torch.Tensor] = None, dz: Optional[torch.Tensor] = None,
                       dz_prev: Optional[torch.Tensor] = None, dw_out: Optional[torch.Tensor] = None,
                       db_out: Optional[torch.Tensor] = None, accumulate_loss: bool = False,
                       defer: bool = False) -> None:
    """Fused skinny output layer: logits + softmax-CE + dX + dW/db (+ SGD when ``w_master`` is given).
    ``defer=True`` runs the fused kernel only; ``output_layer_finish`` (any stream ordered after this call, same
    workspace) then reduces the loss and the split partials and applies the update."""
    _req(a, name="a"); _req(w, a.dtype, "w"); _req(labels, torch.int32, "labels")
    _req(row_loss, torch.float32, "row_loss"); _req(bias, torch.float32, "bias")
    b, k = a.shape
    n = w.shape[1]
    apply = 2 if defer else (1 if w_master is not None else 0)
    act_id = act_prev if isinstance(act_prev, int) else _lib.activation_id(act_prev)
    with torch.cuda.device(a.device):
        check(lib.ddnn_output_layer_train(_p(a), _p(w), _p(labels), _p(dz), _p(row_loss), _p(loss), _p(dz_prev),
                                          _p(w_master), _p(w_shadow), _p(bias), _p(dw_out), _p(db_out), float(lr),
                                          apply, b, k, n, act_id, float(inv_denom), 1 if accumulate_loss else 0,
                                          _DT[a.dtype], _p(workspace), workspace.numel(), _stream(a)))


def output_layer_finish(rows: int, k: int, n: int, dtype: torch.dtype, row_loss: torch.Tensor,
                        loss: Optional[torch.Tensor], inv_denom: float, workspace: torch.Tensor,
                        w_master: Optional[torch.Tensor] = None, bias: Optional[torch.Tensor] = None, lr: float = 0.0,
                        w_shadow: Optional[torch.Tensor] = None, dw_out: Optional[torch.Tensor] = None,
                        db_out: Optional[torch.Tensor] = None, accumulate_loss: bool = False) -> None:
    """Second half of ``output_layer_train(..., defer=True)`` on the CURRENT stream of the workspace's device."""
    _req(row_loss, torch.float32, "row_loss"); _req(bias, torch.float32, "bias")
    with torch.cuda.device(workspace.device):
        check(lib.ddnn_output_layer_finish(_p(row_loss), _p(loss), _p(w_master), _p(w_shadow), _p(bias), _p(dw_out),
                                           _p(db_out), float(lr), 1 if w_master is not None else 0, rows, k, n,
                                           float(inv_denom), 1 if accumulate_loss else 0, _DT[dtype], _p(workspace),
                                           workspace.numel(), _stream(workspace)))


def tail2_workspace_bytes(b: int, k1: int, n2: int, n3: int, dtype: torch.dtype) -> int:
    """0 when the fused two-layer tail does not take this shape / dtype."""
    return int(lib.ddnn_tail2_workspace_bytes(b, k1, n2, n3, _DT[dtype]))


def tail2_train(a1: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor, w3: torch.Tensor, b3: torch.Tensor,
                labels: torch.Tensor, row_loss: torch.Tensor, inv_denom: float, act1, act2,
                workspace: torch.Tensor, dz1: Optional[torch.Tensor] = None) -> None:
    """Fused last two layers (``ddnn_tail2_train``): row losses, dZ1 and the split partials in ``workspace``."""
    _req(a1, torch.float32, "a1"); _req(w2, torch.float32, "w2"); _req(w3, torch.float32, "w3")
    _req(b2, torch.float32, "b2"); _req(b3, torch.float32, "b3")
    _req(labels, torch.int32, "labels"); _req(row_loss, torch.float32, "row_loss")
    b, k1 = a1.shape
    n2, n3 = w2.shape[1], w3.shape[1]
    ids = [a if isinstance(a, int) else _lib.activation_id(a) for a in (act1, act2)]
    with torch.cuda.device(a1.device):
        check(lib.ddnn_tail2_train(_p(a1), _p(w2), _p(b2), _p(w3), _p(b3), _p(labels), _p(dz1), _p(row_loss), ids[0],
                                   ids[1], b, k1, n2, n3, float(inv_denom), _DT[a1.dtype], _p(workspace),
                                   workspace.numel(), _stream(a1)))


def tail2_finish(rows: int, row_loss: torch.Tensor, loss: Optional[torch.Tensor], inv_denom: float,
                 workspace: torch.Tensor, w2: torch.Tensor, b2: torch.Tensor, w3: torch.Tensor, b3: torch.Tensor,
                 lr: float = 0.0, apply: bool = True, grads=None, accumulate_loss: bool = False) -> None:
    """Second half on the CURRENT stream: loss + both reductions; SGD in place (``apply``) or the gradients into
    ``grads = (dW2, db2, dW3, db3)``."""
    k1, n2 = w2.shape
    n3 = w3.shape[1]
    g = grads if grads is not None else (None, None, None, None)
    with torch.cuda.device(workspace.device):
        check(lib.ddnn_tail2_finish(_p(row_loss), _p(loss), _p(w2), _p(b2), _p(w3), _p(b3), _p(g[0]), _p(g[1]),
                                    _p(g[2]), _p(g[3]), float(lr), 1 if apply else 0, rows, k1, n2, n3,
                                    float(inv_denom), 1 if accumulate_loss else 0, _DT[w2.dtype], _p(workspace),
                                    workspace.numel(), _stream(workspace)))


def _workspace(ws: Optional[torch.Tensor], b: int, k: int, n: int, like: torch.Tensor) -> torch.Tensor:
    need = dw_workspace_bytes(b, k, n, like.dtype)
    if ws is None or ws.numel() < need:
        ws = torch.empty((max(need, 256),), dtype=torch.uint8, device=like.device)
    return ws


def cast(t: torch.Tensor, dtype: torch.dtype, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """fp32 <-> bf16 through the library's cast kernels."""
    _req(t, name="t")
    if out is None:
        out = torch.empty(t.shape, dtype=dtype, device=t.device)
    _req(out, dtype, "out")
    with torch.cuda.device(t.device):
        if t.dtype == torch.float32 and dtype == torch.bfloat16:
            check(lib.ddnn_cast_f32_to_bf16(_p(t), _p(out), t.numel(), _stream(t)))
        elif t.dtype == torch.bfloat16 and dtype == torch.float32:
            check(lib.ddnn_cast_bf16_to_f32(_p(t), _p(out), t.numel(), _stream(t)))
        else:
            raise TypeError(f"cast {t.dtype} -> {dtype} not supported")
    return out
