"""ctypes binding of libddnn.so -- the only way Python reaches the kernels.

There is no CPU fallback: if the shared object is missing (and cannot be built because
nvcc is absent) importing this module raises, and every wrapper raises on a non-zero
status.  ``SYMBOLS`` lists every entry point ``include/ddnn.h`` declares.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libddnn.so"

# status codes / enums (include/ddnn.h)
OK, ERR_DIM, ERR_ACTIVATION, ERR_CUDA, ERR_ARG, ERR_WORKSPACE, ERR_UNSUPPORTED = range(7)
ACT_LINEAR, ACT_RELU, ACT_SIGMOID, ACT_SOFTMAX = range(4)
F32, BF16 = 0, 1
ENGINE_AUTO, ENGINE_SIMT, ENGINE_TCGEN05 = 0, 1, 2
STATIC_SCRATCH_BYTES = 8 * 1024 * 1024  # DDNN_STATIC_SCRATCH_BYTES

_vp, _i64, _i32, _f32, _sz = C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_size_t



class Signals(C.Structure):
    """``ddnn_signals`` (include/ddnn.h): flag wait / flag publish folded into a GEMM kernel."""
    _fields_ = [("wait_flag", C.c_void_p), ("wait_value", C.c_int32), ("done_flag", C.c_void_p),
                ("done_value", C.c_int32), ("counter", C.c_void_p)]


SYMBOLS = {
    "ddnn_version": (C.c_int, []),
    "ddnn_last_error": (C.c_char_p, []),
    "ddnn_activation_from_name": (C.c_int, [C.c_char_p]),
    "ddnn_set_engine": (C.c_int, [_i32]),
    "ddnn_set_sm_reserve": (C.c_int, [_i32]),
    "ddnn_set_f32_tensor_cores": (C.c_int, [_i32]),
    "ddnn_set_cta_pairs": (C.c_int, [_i32]),
    "ddnn_set_pdl": (C.c_int, [_i32]),
    "ddnn_set_cluster_split_k": (C.c_int, [_i32]),
    "ddnn_dense_forward": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _i32, _vp]),
    "ddnn_dense_forward_signalled": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _i32,
                                               C.POINTER(Signals), _vp]),
    "ddnn_dense_backward_dx_signalled": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32,
                                                   C.POINTER(Signals), _vp]),
    "ddnn_softmax_rows": (C.c_int, [_vp, _vp, _i64, _i64, _vp]),
    "ddnn_activate": (C.c_int, [_vp, _vp, _i64, _i64, _i32, _vp]),
    "ddnn_softmax_ce": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _f32, _i32, _i32, _vp]),
    "ddnn_dense_backward_dx": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _i64, _i64, _i32, _i32, _vp]),
    "ddnn_dense_backward_dw_sgd": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _f32, _i32, _i64, _i64, _i64,
                                             _i32, _vp, _sz, _vp]),
    "ddnn_dw_workspace_bytes": (_sz, [_i64, _i64, _i64, _i32]),
    "ddnn_dense_backward_dw_accumulate": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i64, _i64, _i64, _i32, _vp, _sz, _vp]),
    "ddnn_sgd_apply": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _f32, _i64, _i64, _vp]),
    "ddnn_dense_backward_dw_sgd_static": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _f32, _i32, _i64, _i64, _i64,
                                                    _i32, _vp]),
    "ddnn_output_layer_train": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f32, _i32,
                                          _i64, _i64, _i64, _i32, _f32, _i32, _i32, _vp, _sz, _vp]),
    "ddnn_output_layer_finish": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _f32, _i32, _i64, _i64, _i64, _f32, _i32,
                                           _i32, _vp, _sz, _vp]),
    "ddnn_output_layer_workspace_bytes": (_sz, [_i64, _i64, _i64, _i32]),
    "ddnn_tail2_workspace_bytes": (_sz, [_i64, _i64, _i64, _i64, _i32]),
    "ddnn_tail2_train": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i64, _i64, _i64, _i64, _f32,
                                   _i32, _vp, _sz, _vp]),
    "ddnn_tail2_finish": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _f32, _i32, _i64, _i64, _i64,
                                    _i64, _f32, _i32, _i32, _vp, _sz, _vp]),
    "ddnn_cast_f32_to_bf16": (C.c_int, [_vp, _vp, _i64, _vp]),
    "ddnn_cast_bf16_to_f32": (C.c_int, [_vp, _vp, _i64, _vp]),
    "ddnn_ipc_export": (C.c_int, [_vp, _vp, C.POINTER(C.c_int64)]),
    "ddnn_ipc_import": (C.c_int, [_vp, _i64, C.POINTER(C.c_void_p)]),
    "ddnn_flag_write": (C.c_int, [_vp, _i32, _vp]),
    "ddnn_flag_wait": (C.c_int, [_vp, _i32, _vp]),
    "ddnn_launch_count": (_i64, []),
}


class DdnnError(RuntimeError):
    def __init__(self, status: int, text: str):
        super().__init__(f"libddnn status {status}: {text}")
        self.status = status
        self.text = text


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        try:
            from .build import build
            build()
        except Exception as e:  # noqa: BLE001
            raise RuntimeError(
                f"libddnn.so is missing at {LIB_PATH} and could not be built ({e}). "
                "There is no CPU fallback; run `python docker-dist-nn_b200/build.py`.") from e
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def last_error() -> str:
    return (lib.ddnn_last_error() or b"").decode("utf-8", "replace")


def check(status: int) -> None:
    if status != OK:
        raise DdnnError(status, last_error())


def activation_id(name) -> int:
    """grpc_node.py:62-73 string -> enum; unknown strings (and None) are linear."""
    if name is None:
        return ACT_LINEAR
    return lib.ddnn_activation_from_name(str(name).encode())


def launch_count() -> int:
    return int(lib.ddnn_launch_count())


def set_engine(engine: int) -> None:
    check(lib.ddnn_set_engine(engine))


def set_pdl(on: bool) -> None:
    """Programmatic dependent launch between the kernels of a step (default on)."""
    check(lib.ddnn_set_pdl(1 if on else 0))


def set_cta_pairs(on: bool) -> None:
    check(lib.ddnn_set_cta_pairs(1 if on else 0))


F32_TC_FFMA, F32_TC_MMA_SYNC, F32_TC_TCGEN05 = 0, 1, 2
F32_TC_DEFAULT = F32_TC_TCGEN05  # DDNN_F32_TC_DEFAULT in include/ddnn.h


def set_f32_tensor_cores(mode=None) -> None:
    """Inner product of the aligned fp32 GEMMs: 0 FFMA, 1 3xTF32 on mma.sync, 2 3xTF32 on tcgen05
    (``None`` restores the library default)."""
    check(lib.ddnn_set_f32_tensor_cores(F32_TC_DEFAULT if mode is None else int(mode)))


def set_sm_reserve(sms: int) -> None:
    check(lib.ddnn_set_sm_reserve(sms))


if os.environ.get("DDNN_PDL", "1") == "0":  # A/B switch for measurements
    set_pdl(False)
