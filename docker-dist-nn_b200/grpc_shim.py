"""gRPC compatibility shim: ``grpc_dist_nn.LayerService/Process`` served by the GPU ``Layer``.

Lets an existing docker-dist-nn deployment point its client (run_grpc_inference.py) or an
upstream node's NEXT_NODES at a B200-backed node without changing a line on their side: same
service name, same messages (``Row{repeated double values = 1}``, ``Matrix{repeated Row rows = 1}``,
/root/reference/src/proto/dist_nn.proto:1-15), same behaviour as the reference servicer
(/root/reference/src/grpc_node.py:99-158): empty request -> empty Matrix; ValueError ->
INVALID_ARGUMENT; anything else -> INTERNAL; forwarding to ``next_nodes[0]`` with a 10 s timeout
and relaying its reply or its status code.

The message classes are built at import time from a descriptor written here (protoc and the
reference's generated stubs are not needed); the wire format is identical, which
``tests/test_grpc_shim.py`` checks byte for byte and, where the reference tree is present, against
the reference's own generated client.
"""
from __future__ import annotations

from concurrent import futures
from typing import Callable, Optional

import grpc
import numpy as np
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

SERVICE = "grpc_dist_nn.LayerService"
METHOD = f"/{SERVICE}/Process"


def _build_messages():
    fd = descriptor_pb2.FileDescriptorProto()
    fd.name = "ddnn_b200/dist_nn_compat.proto"
    fd.package = "grpc_dist_nn"
    fd.syntax = "proto3"
    row = fd.message_type.add()
    row.name = "Row"
    f = row.field.add()
    f.name, f.number = "values", 1
    f.label = descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED
    f.type = descriptor_pb2.FieldDescriptorProto.TYPE_DOUBLE
    mat = fd.message_type.add()
    mat.name = "Matrix"
    f = mat.field.add()
    f.name, f.number = "rows", 1
    f.label = descriptor_pb2.FieldDescriptorProto.LABEL_REPEATED
    f.type = descriptor_pb2.FieldDescriptorProto.TYPE_MESSAGE
    f.type_name = ".grpc_dist_nn.Row"
    pool = descriptor_pool.DescriptorPool()  # private pool: no clash if the reference's stubs are also imported
    pool.Add(fd)
    get = getattr(message_factory, "GetMessageClass", None)
    if get is None:  # protobuf < 4.21
        factory = message_factory.MessageFactory(pool)
        return (factory.GetPrototype(pool.FindMessageTypeByName("grpc_dist_nn.Row")),
                factory.GetPrototype(pool.FindMessageTypeByName("grpc_dist_nn.Matrix")))
    return (get(pool.FindMessageTypeByName("grpc_dist_nn.Row")),
            get(pool.FindMessageTypeByName("grpc_dist_nn.Matrix")))


Row, Matrix = _build_messages()


def matrix_to_numpy(msg) -> np.ndarray:
    """grpc_node.py:107,113: rows -> list of lists -> float64 array."""
    return np.array([list(r.values) for r in msg.rows], dtype=np.float64)


def numpy_to_matrix(a: np.ndarray):
    """grpc_node.py:126-127 / :146-147."""
    return Matrix(rows=[Row(values=row) for row in np.asarray(a, dtype=np.float64).tolist()])


def make_stub(channel: grpc.Channel) -> Callable:
    """``stub(matrix, timeout=...) -> Matrix`` for LayerService/Process on ``channel``."""
    return channel.unary_unary(METHOD, request_serializer=Matrix.SerializeToString,
                               response_deserializer=Matrix.FromString)


class LayerServiceServicer:
    """Same control flow as grpc_node.py:99-158 around any object with ``container_name``,
    ``next_nodes`` and ``process_matrix(ndarray) -> ndarray`` (the GPU ``Layer``)."""

    def __init__(self, layer_processor):
        self.layer_processor = layer_processor

    def Process(self, request, context):
        lp = self.layer_processor
        try:
            input_list = [list(row.values) for row in request.rows]
            if not input_list:
                print(f"({lp.container_name}) Received empty input matrix.")
                return Matrix()
            if len({len(r) for r in input_list}) != 1:
                # the reference's np.array/np.dot raise ValueError on ragged rows -> INVALID_ARGUMENT (grpc_node.py:149-153)
                raise ValueError(f"({lp.container_name}) input rows have different lengths")
            output = lp.process_matrix(np.array(input_list))
            if lp.next_nodes:
                nxt = lp.next_nodes[0]
                addr = f"{nxt['host']}:{nxt['port']}"
                try:
                    with grpc.insecure_channel(addr) as channel:
                        return make_stub(channel)(numpy_to_matrix(output), timeout=10)
                except grpc.RpcError as e:
                    print(f"({lp.container_name}) ERROR: gRPC call to {addr} failed: {e.code()} - {e.details()}")
                    context.set_code(e.code())
                    context.set_details(f"Failed to forward request to {addr}: {e.details()}")
                    return Matrix()
            return numpy_to_matrix(output)
        except ValueError as ve:
            print(f"({lp.container_name}) ERROR processing matrix: {ve}")
            context.set_code(grpc.StatusCode.INVALID_ARGUMENT)
            context.set_details(str(ve))
            return Matrix()
        except Exception as e:  # noqa: BLE001
            print(f"({lp.container_name}) UNEXPECTED ERROR: {e}")
            context.set_code(grpc.StatusCode.INTERNAL)
            context.set_details(f"An internal error occurred: {e}")
            return Matrix()


def add_servicer_to_server(servicer: LayerServiceServicer, server: grpc.Server) -> None:
    handler = grpc.unary_unary_rpc_method_handler(servicer.Process, request_deserializer=Matrix.FromString,
                                                  response_serializer=Matrix.SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(SERVICE, {"Process": handler}),))


def start_server(layer_processor, port: Optional[int] = None, host: str = "0.0.0.0", max_workers: int = 10):
    """Starts the service (grpc_node.py:161-178 without the blocking wait).  ``port`` None/0 picks a free
    port.  Returns (server, bound_port).  Like the reference no message-size options are set, so the
    default 4 MiB receive cap applies."""
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers))
    add_servicer_to_server(LayerServiceServicer(layer_processor), server)
    bound = server.add_insecure_port(f"{host}:{port or 0}")
    server.start()
    return server, bound


def serve() -> None:
    """Drop-in for ``python grpc_node.py``: env-configured GPU Layer behind LayerService."""
    from .layer import Layer
    layer = Layer()
    server, port = start_server(layer, layer.listen_port)
    print(f"({layer.container_name}) gRPC LayerService (B200) started on 0.0.0.0:{port}")
    server.wait_for_termination()


if __name__ == "__main__":
    serve()
