"""The reference's own forward path, run unmodified from baseline/_ref (tools/install_reference.py), against the
oracle: the in-process `Layer` (grpc_node.py:75-97) and the 3-process gRPC relay chain (grpc_node.py:99-178 driven
by run_grpc_inference.py:112-158).  Pins the oracle to outputs of the reference itself on the machine the tests
run on (besides the committed golden vectors), and keeps the CPU baseline of `bench.py --impl reference` honest.
Skipped where the copy is absent."""
import os
import sys

import numpy as np
import pytest

import oracle
from ddnn_b200.config import make_params, synthetic_batch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
import reference_chain as rc  # noqa: E402

pytestmark = pytest.mark.skipif(not rc.available(), reason="baseline/_ref is not installed")


def test_in_process_reference_layer_and_grpc_chain_equal_the_oracle():
    dims, acts = [784, 32, 16, 10], ["relu", "relu", "softmax"]
    params = make_params(dims, acts, seed=1)
    x, _ = synthetic_batch(96, 784, 10, seed=4)
    v2, calls = rc.time_in_process(params, x, 48, 0.05)
    assert v2 > 0 and calls >= 2
    v1, rpcs, rows_per_rpc, last_out = rc.time_grpc_chain(params, x, 48, 0.05)
    assert v1 > 0 and rpcs >= 2 and rows_per_rpc == 48
    want = oracle.process_matrix(params, x[-48:], 784)
    np.testing.assert_allclose(last_out, want, rtol=0, atol=1e-15)  # float64 both sides; BLAS blocking may differ by 1 ulp
