"""gRPC compatibility shim: wire format, servicer behaviour and relay chain
on CPU with an oracle-backed processor; the GPU ``Layer`` behind the same service under ``-m gpu``."""
import os
import struct
import sys

import grpc
import numpy as np
import pytest

import oracle
from ddnn_b200 import grpc_shim as G
from ddnn_b200.config import make_params

REF_SRC = "/root/reference/src"


class OracleProcessor:
    """Duck-typed stand-in for Layer (tests only): the float64 oracle behind ``process_matrix``."""

    def __init__(self, params, in_dim, name="node", next_nodes=None):
        self.layer_params, self.expected_input_dim = params, in_dim
        self.container_name, self.next_nodes = name, list(next_nodes or [])

    def process_matrix(self, x):
        return oracle.process_matrix(self.layer_params, x, self.expected_input_dim, self.container_name)


def test_wire_format_known_answer():
    """proto3: Matrix.rows = field 1 (length-delimited), Row.values = field 1 packed doubles
    (/root/reference/src/proto/dist_nn.proto:5-11)."""
    m = G.numpy_to_matrix(np.array([[1.0, 2.0], [3.0, 4.5]]))
    row = lambda vals: b"\x0a" + bytes([8 * len(vals)]) + b"".join(struct.pack("<d", v) for v in vals)  # noqa: E731
    r1, r2 = row([1.0, 2.0]), row([3.0, 4.5])
    expect = b"\x0a" + bytes([len(r1)]) + r1 + b"\x0a" + bytes([len(r2)]) + r2
    assert m.SerializeToString() == expect
    back = G.matrix_to_numpy(G.Matrix.FromString(expect))
    assert back.dtype == np.float64 and back.tolist() == [[1.0, 2.0], [3.0, 4.5]]
    assert G.Matrix().SerializeToString() == b"" and G.METHOD == "/grpc_dist_nn.LayerService/Process"


def test_server_relay_chain_and_errors():
    """Two shim nodes chained through NEXT_NODES (grpc_node.py:120-135 relay), driven by the client call
    of run_grpc_inference.py:135-147; dim mismatch -> INVALID_ARGUMENT; empty request -> empty Matrix."""
    params = make_params([6, 5, 4, 3], ["relu", "sigmoid", "softmax"], seed=2)
    last = OracleProcessor(params[2:], 4, "layer_container_1")
    s2, p2 = G.start_server(last, 0, host="127.0.0.1")
    first = OracleProcessor(params[:2], 6, "layer_container_0", [{"host": "127.0.0.1", "port": str(p2)}])
    s1, p1 = G.start_server(first, 0, host="127.0.0.1")
    try:
        x = np.random.default_rng(0).random((7, 6))
        with grpc.insecure_channel(f"127.0.0.1:{p1}") as ch:
            stub = G.make_stub(ch)
            resp = stub(G.numpy_to_matrix(x), timeout=10)
            out = np.vstack([r.values for r in resp.rows])
            assert np.array_equal(out, oracle.process_matrix(params, x, 6))
            assert len(stub(G.Matrix(), timeout=10).rows) == 0          # grpc_node.py:108-111
            with pytest.raises(grpc.RpcError) as e:                      # grpc_node.py:149-153
                stub(G.numpy_to_matrix(np.zeros((2, 5))), timeout=10)
            assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
            assert e.value.details() == "(layer_container_0) Layer 1: expected input dim 6, got 5"
        # downstream node gone: the first node relays the failure code (grpc_node.py:136-140)
        s2.stop(0)
        with grpc.insecure_channel(f"127.0.0.1:{p1}") as ch:
            with pytest.raises(grpc.RpcError) as e:
                G.make_stub(ch)(G.numpy_to_matrix(x), timeout=10)
            assert e.value.code() == grpc.StatusCode.UNAVAILABLE
    finally:
        s1.stop(0)
        s2.stop(0)


@pytest.mark.skipif(not os.path.isdir(REF_SRC), reason="reference tree only exists in the build container")
def test_reference_generated_client_talks_to_the_shim():
    """The unmodified reference's generated stubs (dist_nn_pb2 / dist_nn_pb2_grpc) call the shim: proves the
    dynamically built messages and the method path are wire-compatible."""
    sys.path.insert(0, os.path.join(REF_SRC, "proto"))
    try:
        import dist_nn_pb2
        import dist_nn_pb2_grpc
    finally:
        sys.path.pop(0)
    params = make_params([5, 4, 3], ["relu", "softmax"], seed=4)
    srv, port = G.start_server(OracleProcessor(params, 5), 0, host="127.0.0.1")
    try:
        x = np.random.default_rng(1).random((4, 5))
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = dist_nn_pb2_grpc.LayerServiceStub(ch)
            req = dist_nn_pb2.Matrix(rows=[dist_nn_pb2.Row(values=r) for r in x.tolist()])
            resp = stub.Process(req, timeout=10)
        out = np.vstack([r.values for r in resp.rows])
        assert np.array_equal(out, oracle.process_matrix(params, x, 5))
        assert req.SerializeToString() == G.numpy_to_matrix(x).SerializeToString()
    finally:
        srv.stop(0)


@pytest.mark.gpu
def test_gpu_layer_behind_the_service(cuda_device):
    from ddnn_b200 import Layer
    params = make_params([784, 32, 16, 10], ["relu", "relu", "softmax"], seed=1)
    layer = Layer.from_params(params, 784, container_name="gpu_node")
    srv, port = G.start_server(layer, 0, host="127.0.0.1")
    try:
        x = np.random.default_rng(2).random((64, 784))
        with grpc.insecure_channel(f"127.0.0.1:{port}") as ch:
            stub = G.make_stub(ch)
            out = np.vstack([r.values for r in stub(G.numpy_to_matrix(x), timeout=30).rows])
            ref = oracle.process_matrix(params, x, 784)
            assert np.max(np.abs(out - ref)) / np.max(np.abs(ref)) < 1e-5
            assert np.array_equal(out.argmax(axis=1), ref.argmax(axis=1))
            with pytest.raises(grpc.RpcError) as e:
                stub(G.numpy_to_matrix(np.zeros((2, 3))), timeout=30)
            assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
            assert e.value.details() == "(gpu_node) Layer 1: expected input dim 784, got 3"
    finally:
        srv.stop(0)
