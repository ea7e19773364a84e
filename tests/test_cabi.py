"""The C-ABI library loads on a machine without a GPU and exports every symbol that
include/ddnn.h declares (no compute calls here)."""
import ctypes
import os
import re

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(REPO, "include", "ddnn.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ddnn_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported():
    from ddnn_b200 import _lib
    names = _declared()
    assert len(names) >= 14
    raw = ctypes.CDLL(str(_lib.LIB_PATH))
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/ddnn.h but not exported"
    # and the Python binding covers exactly the declared set
    assert sorted(_lib.SYMBOLS) == names


def test_host_only_entry_points():
    from ddnn_b200 import _lib
    assert _lib.lib.ddnn_version() >= 100
    assert _lib.activation_id("relu") == _lib.ACT_RELU
    assert _lib.activation_id("sigmoid") == _lib.ACT_SIGMOID
    assert _lib.activation_id("softmax") == _lib.ACT_SOFTMAX
    for other in ("linear", "unknown", "tanh", "", None):  # grpc_node.py:72-73
        assert _lib.activation_id(other) == _lib.ACT_LINEAR
    assert _lib.lib.ddnn_dw_workspace_bytes(4096, 784, 128, _lib.F32) > 0
    assert _lib.lib.ddnn_dw_workspace_bytes(0, 784, 128, _lib.F32) == 0
    # the fused output layer only needs scratch for shapes it accepts (N <= 16, K <= 256); a wide layer must
    # not size the caller's workspace (regression: 512 GiB asked for a 16384 x 16384 layer)
    assert 0 < _lib.lib.ddnn_output_layer_workspace_bytes(65536, 256, 10, _lib.BF16) < 16 << 20
    assert _lib.lib.ddnn_output_layer_workspace_bytes(65536, 16384, 16384, _lib.BF16) == 0
    assert _lib.lib.ddnn_output_layer_workspace_bytes(2048, 8192, 8192, _lib.BF16) == 0
    # argument errors are reported through the status code + ddnn_last_error, never thrown
    rc = _lib.lib.ddnn_dense_forward(None, None, None, None, 4, 0, 3, 0, 0, 0, None)
    assert rc == _lib.ERR_DIM and "bad dimensions" in _lib.last_error()
    rc = _lib.lib.ddnn_set_engine(9)
    assert rc == _lib.ERR_ARG
    assert _lib.lib.ddnn_set_f32_tensor_cores(3) == _lib.ERR_ARG and "mode 3" in _lib.last_error()
    assert _lib.lib.ddnn_set_f32_tensor_cores(_lib.F32_TC_DEFAULT) == 0 and _lib.lib.ddnn_set_pdl(1) == 0
    # the fused two-layer tail only takes small fp32 shapes: capability query and status code, no GPU needed
    assert _lib.lib.ddnn_tail2_workspace_bytes(4096, 128, 64, 10, _lib.F32) > 0
    for bad in ((4096, 256, 64, 10, _lib.F32), (4096, 128, 128, 10, _lib.F32), (4096, 128, 64, 17, _lib.F32),
                (4096, 126, 64, 10, _lib.F32), (4096, 128, 64, 10, _lib.BF16)):
        assert _lib.lib.ddnn_tail2_workspace_bytes(*bad) == 0
    rc = _lib.lib.ddnn_tail2_train(None, None, None, None, None, None, None, None, 1, 1, 4096, 256, 64, 10, 1.0, _lib.F32,
                                   None, 0, None)
    assert rc == _lib.ERR_UNSUPPORTED and "K1 <= 128" in _lib.last_error()


def test_product_does_not_import_oracle():
    pkg = os.path.join(REPO, "docker-dist-nn_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(root, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
