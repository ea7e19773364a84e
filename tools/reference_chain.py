"""Times the reference's OWN CPU implementation of the forward path from the unmodified
copy under baseline/_ref/ (tools/install_reference.py).  Forward-only: that is all the reference does
(SURVEY.md section 0.1).  Used by `bench.py --impl reference` only.

B2  in-process  `Layer.process_matrix`  (/root/reference/src/grpc_node.py:75-97) on the batches;
B1  the chain: one unmodified `grpc_node.py` process per layer on localhost with the environment contract of
    /root/reference/src/run_grpc_fcnn.py:101-106 (ports 5101 + 100*i like :130), driven by the reference
    client's own `run_batch_inference` (/root/reference/src/run_grpc_inference.py:112-158) over sequential
    chunks (:198-205) that stay under the server's default 4 MiB receive cap (grpc_node.py:169).
"""
from __future__ import annotations

import contextlib
import io
import json
import os
import socket
import subprocess
import sys
import tempfile
import time

import numpy as np

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SRC = os.path.join(REPO, "baseline", "_ref", "src")


def available() -> bool:
    return os.path.exists(os.path.join(REF_SRC, "grpc_node.py"))


def _neurons(params_layer):
    w = np.asarray(params_layer["weights"])
    return [{"weights": w[:, j].tolist(), "bias": float(params_layer["biases"][j]),
             "activation": params_layer["activation"]} for j in range(w.shape[1])]


def _import_ref():
    cwd = os.getcwd()
    os.chdir(REF_SRC)  # grpc_node.py:10 puts cwd/proto on sys.path
    try:
        if REF_SRC not in sys.path:
            sys.path.insert(0, REF_SRC)
        import grpc_node  # noqa: PLC0415
        import run_grpc_inference  # noqa: PLC0415
    finally:
        os.chdir(cwd)
    return grpc_node, run_grpc_inference


def time_in_process(params, x: np.ndarray, rows_per_call: int, budget_s: float):
    """B2: the reference Layer holding ALL layers, process_matrix on sequential chunks of x, repeated for
    about budget_s.  Returns (samples/s, calls)."""
    grpc_node, _ = _import_ref()
    env = {"CONTAINER_NAME": "ref", "EXPECTED_INPUT_DIM": str(x.shape[1]),
           "NEURONS_CONFIG": json.dumps({f"layer_{i + 1}": _neurons(p) for i, p in enumerate(params)}),
           "NEXT_NODES": "[]"}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    os.environ.pop("NEURONS_FILE_CONFIG", None)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            layer = grpc_node.Layer()
    finally:
        for k, v in old.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)
    chunks = [x[a:a + rows_per_call] for a in range(0, x.shape[0], rows_per_call)]
    n = calls = 0
    with contextlib.redirect_stdout(io.StringIO()):  # process_matrix prints per call (grpc_node.py:96)
        layer.process_matrix(chunks[0])
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < budget_s:
            for c in chunks:
                layer.process_matrix(c)
                n += c.shape[0]
                calls += 1
        el = time.perf_counter() - t0
    return n / el, calls


def _free_port(p):
    with socket.socket() as s:
        try:
            s.bind(("127.0.0.1", p))
            return p
        except OSError:
            s.bind(("127.0.0.1", 0))
            return s.getsockname()[1]


def time_grpc_chain(params, x: np.ndarray, rows_per_rpc: int, budget_s: float, timeout_s: float = 30.0):
    """B1: one unmodified grpc_node.py process per layer, relay chain, reference client loop.
    Returns (samples/s, rpcs, rows_per_rpc) or raises RuntimeError when the chain does not come up."""
    _, client = _import_ref()
    tmp = tempfile.mkdtemp(prefix="ddnn_ref_chain_")
    ports = [_free_port(5101 + 100 * i) for i in range(len(params))]
    procs = []
    try:
        dim = x.shape[1]
        for i, p in enumerate(params):
            cfg = os.path.join(tmp, f"layer{i}.json")
            with open(cfg, "w") as f:
                json.dump({"layer_1": _neurons(p)}, f)
            nxt = [{"host": "127.0.0.1", "port": str(ports[i + 1])}] if i + 1 < len(params) else []
            env = dict(os.environ, CONTAINER_NAME=f"fcnn_layer_container_{i}", LISTEN_PORT=str(ports[i]),
                       EXPECTED_INPUT_DIM=str(dim), NEURONS_FILE_CONFIG=cfg, NEXT_NODES=json.dumps(nxt))
            env.pop("NEURONS_CONFIG", None)
            procs.append(subprocess.Popen([sys.executable, "grpc_node.py"], cwd=REF_SRC, env=env,
                                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL))
            dim = np.asarray(p["weights"]).shape[1]
        deadline = time.time() + timeout_s
        for port in ports:  # readiness poll like run_grpc_fcnn.py:157-172
            while True:
                try:
                    socket.create_connection(("127.0.0.1", port), timeout=0.5).close()
                    break
                except OSError:
                    if time.time() > deadline or any(pr.poll() is not None for pr in procs):
                        raise RuntimeError("reference chain did not start")
                    time.sleep(0.1)
        addr = f"127.0.0.1:{ports[0]}"
        chunks = [x[a:a + rows_per_rpc].tolist() for a in range(0, x.shape[0], rows_per_rpc)]
        out, _ = client.run_batch_inference(chunks[0], addr, timeout_s)
        if out is None:
            raise RuntimeError("reference chain returned no result")
        n = rpcs = 0
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < budget_s:
            for c in chunks:
                out, _ = client.run_batch_inference(c, addr, timeout_s)
                if out is None:
                    raise RuntimeError("reference chain failed mid-run")
                n += len(c)
                rpcs += 1
        el = time.perf_counter() - t0
        return n / el, rpcs, rows_per_rpc, out
    finally:
        for pr in procs:
            pr.terminate()
        for pr in procs:
            try:
                pr.wait(timeout=5)
            except subprocess.TimeoutExpired:
                pr.kill()
        import shutil
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    sys.path.insert(0, REPO)
    from ddnn_b200.config import make_params, synthetic_batch
    import oracle
    params = make_params([784, 128, 64, 10], ["relu", "relu", "softmax"], seed=1)
    x, _ = synthetic_batch(2560, 784, 10, seed=0)
    v2, calls = time_in_process(params, x, 640, 3.0)
    print(f"B2 in-process Layer.process_matrix: {v2:,.0f} samples/s ({calls} calls of 640 rows)")
    v1, rpcs, rpr, out = time_grpc_chain(params, x, 640, 5.0)
    ref = oracle.process_matrix(params, x[-640:], 784)
    print(f"B1 3-node gRPC chain: {v1:,.0f} samples/s ({rpcs} RPCs of {rpr} rows); "
          f"max |chain - oracle| = {np.max(np.abs(out - ref)):.2e}")
