"""The oracle against the reference's golden vectors (tests/golden, minted by running the
unmodified reference) and against torch autograd for the parts the reference lacks."""
import json
import os

import numpy as np
import pytest
import torch

import oracle
from ddnn_b200.config import make_params, node_config_to_params
from conftest import GOLDEN


def test_kat_bit_exact(golden_small):
    k = golden_small["kat"]
    params = oracle.build_layer_params(k["neurons_config"])
    out = oracle.process_matrix(params, np.array(k["input"]), k["expected_input_dim"])
    assert out.tolist() == k["output"]
    # the survey's hand-checked value
    assert out[0, 0] == pytest.approx(0.3529721910540915, abs=1e-16)


def test_activations_bit_exact(golden_small):
    a = golden_small["activations"]
    grid = np.array(a["grid"])
    with np.errstate(over="ignore"):
        for func, expect in a["outputs"].items():
            got = oracle.activate(grid, func)
            assert np.array_equal(got, np.array(expect)), func
    # anything that is not relu/sigmoid/softmax is the identity (grpc_node.py:72-73)
    assert a["outputs"]["unknown"] == a["grid"] and a["outputs"]["tanh"] == a["grid"]


def test_small_net_prefixes_bit_exact(golden_small):
    s = golden_small["small_net"]
    x = np.array(s["input"])
    for upto, expect in s["prefix_outputs"].items():
        nc = {f"layer_{i + 1}": s["config"]["layers"][i]["neurons"] for i in range(int(upto))}
        out = oracle.process_matrix(oracle.build_layer_params(nc), x, 5)
        assert np.array_equal(out, np.array(expect)), f"prefix {upto}"
    # two nodes [2, 2] chained == one node with four layers
    nc_a = {f"layer_{i + 1}": s["config"]["layers"][i]["neurons"] for i in range(2)}
    nc_b = {f"layer_{i + 1}": s["config"]["layers"][2 + i]["neurons"] for i in range(2)}
    mid = oracle.process_matrix(oracle.build_layer_params(nc_a), x, 5)
    out = oracle.process_matrix(oracle.build_layer_params(nc_b), mid, 6)
    assert np.array_equal(out, np.array(s["chained_2_2"]))


def test_manual_nn_second_statement(golden_small):
    s = golden_small["small_net"]
    cfg = s["manual_nn_config"]
    nc = {f"layer_{i + 1}": cfg["layers"][i]["neurons"] for i in range(3)}
    out = oracle.process_matrix(oracle.build_layer_params(nc), np.array(s["input"]), 5)
    np.testing.assert_allclose(out, np.array(s["manual_nn_outputs"]), rtol=1e-12, atol=1e-15)


def test_layer_key_order_and_empty_layers():
    n = lambda w, a: [{"weights": w, "bias": 0.0, "activation": a}]
    cfg = {"layer_10": n([2.0], "linear"), "layer_2": n([3.0], "relu"), "layer_3": []}
    p = oracle.build_layer_params(cfg)
    assert [q["activation"] for q in p] == ["relu", "linear"]  # int sort, not string sort; empty skipped
    assert node_config_to_params(cfg)[0]["weights"].tolist() == p[0]["weights"].tolist()
    # activation of the FIRST neuron wins; missing key -> linear
    mixed = {"layer_1": [{"weights": [1.0], "bias": 0.0}, {"weights": [1.0], "bias": 0.0, "activation": "relu"}]}
    assert oracle.build_layer_params(mixed)[0]["activation"] == "linear"


def test_dim_mismatch_message(golden_small):
    k = golden_small["kat"]
    params = oracle.build_layer_params(k["neurons_config"])
    with pytest.raises(ValueError) as e:
        oracle.process_matrix(params, np.zeros((2, 3)), 2, "golden")
    assert str(e.value) == golden_small["edges"]["dim_mismatch_message"]


def test_mnist_shapes_bit_exact(golden_mnist):
    for tag, dims in (("mnist_128_64", [784, 128, 64, 10]), ("mnist_32_16", [784, 32, 16, 10])):
        params = make_params(dims, ["relu", "relu", "softmax"], seed=1)
        np.testing.assert_array_equal(np.array([float(np.sum(p["weights"])) for p in params]),
                                      golden_mnist[f"{tag}_w_checksum"])
        x = golden_mnist[f"{tag}_x"]
        assert np.array_equal(oracle.process_matrix(params, x, 784), golden_mnist[f"{tag}_out"])
        assert np.array_equal(oracle.process_matrix(params[:1], x, 784), golden_mnist[f"{tag}_h1"])


def test_stage_partition_matches_reference(golden_small):
    m = golden_small["mappings"]
    from ddnn_b200.config import params_to_layers
    layers = params_to_layers(make_params(m["dims"], ["relu"] * 5 + ["softmax"], seed=m["seed"]))
    for dist_s, expect in m["by_distribution"].items():
        got = oracle.stage_partition(layers, json.loads(dist_s), len(m["examples"][0]["input"]))
        assert sorted(map(str, got)) == sorted(expect)
        for c, e in expect.items():
            g = got[int(c)]
            assert g["expected_input"] == e["expected_input"]
            assert {k: len(v) for k, v in g["neurons_config"].items()} == e["layer_keys"]
            nxt = e["next_nodes"][0]["host"].rsplit("_", 1)[1] if e["next_nodes"] else None
            assert (None if g["next"] is None else str(g["next"])) == nxt
    with pytest.raises(ValueError) as e:
        oracle.stage_partition(layers, [1], 4)
    assert str(e.value) == m["bad_sum_message"]


def test_batch_slices_reference_order():
    assert oracle.batch_slices(10, 4) == [(0, 4), (4, 8), (8, 10)]
    assert oracle.batch_slices(10, None) == [(0, 10)]
    assert oracle.batch_slices(10, 10) == [(0, 10)] and oracle.batch_slices(10, 99) == [(0, 10)]
    assert oracle.batch_slices(0, 4) == []
    assert oracle.batch_slices(8, 4) == [(0, 4), (4, 8)]


# ---- the parts with no reference implementation: cross-check with torch autograd (float64) ----

def _torch_model(params):
    ws = [torch.tensor(p["weights"], dtype=torch.float64, requires_grad=True) for p in params]
    bs = [torch.tensor(p["biases"], dtype=torch.float64, requires_grad=True) for p in params]
    return ws, bs


def _torch_forward(ws, bs, acts, x):
    h = x
    for i, (w, b, a) in enumerate(zip(ws, bs, acts)):
        z = h @ w + b
        if i == len(ws) - 1:
            return z  # logits; CE applies log-softmax
        h = torch.relu(z) if a == "relu" else torch.sigmoid(z) if a == "sigmoid" else z
    return h


@pytest.mark.parametrize("acts", [["relu", "sigmoid", "unknown", "softmax"], ["sigmoid", "relu", "linear"],
                                  ["softmax"]])
def test_backward_matches_autograd(acts):
    dims = [6, 9, 5, 7, 4][: len(acts) + 1]
    params = make_params(dims, acts, seed=5)
    rng = np.random.default_rng(2)
    x = rng.random((13, dims[0])) - 0.3
    y = rng.integers(0, dims[-1], size=13)
    a, zs = oracle.forward_all(params, x)
    loss, dws, dbs, dx = oracle.backward(params, a, zs, y)
    ws, bs = _torch_model(params)
    xt = torch.tensor(x, requires_grad=True)
    tl = torch.nn.functional.cross_entropy(_torch_forward(ws, bs, acts, xt), torch.tensor(y))
    tl.backward()
    assert loss == pytest.approx(tl.item(), rel=1e-12)
    for i in range(len(params)):
        np.testing.assert_allclose(dws[i], ws[i].grad.numpy(), rtol=1e-10, atol=1e-14)
        np.testing.assert_allclose(dbs[i], bs[i].grad.numpy(), rtol=1e-10, atol=1e-14)
    np.testing.assert_allclose(dx, xt.grad.numpy(), rtol=1e-10, atol=1e-14)


def test_train_steps_match_torch_sgd():
    acts = ["relu", "relu", "softmax"]
    dims = [12, 10, 8, 5]
    params = make_params(dims, acts, seed=9)
    ws, bs = _torch_model(params)
    opt = torch.optim.SGD(ws + bs, lr=0.05)
    rng = np.random.default_rng(4)
    for step in range(5):
        x = rng.random((16, 12))
        y = rng.integers(0, 5, size=16)
        l_or = oracle.train_step(params, x, y, 0.05)
        opt.zero_grad()
        l_t = torch.nn.functional.cross_entropy(_torch_forward(ws, bs, acts, torch.tensor(x)), torch.tensor(y))
        l_t.backward()
        opt.step()
        assert l_or == pytest.approx(l_t.item(), rel=1e-12)
    for p, w, b in zip(params, ws, bs):
        np.testing.assert_allclose(p["weights"], w.detach().numpy(), rtol=1e-11, atol=1e-14)
        np.testing.assert_allclose(p["biases"], b.detach().numpy(), rtol=1e-11, atol=1e-14)


def test_micro_batch_sum_equals_full_batch():
    """softmax_ce with denom = global batch: per-micro-batch gradients add up to the full one."""
    params = make_params([7, 6, 4], ["relu", "softmax"], seed=1)
    rng = np.random.default_rng(0)
    x, y = rng.random((12, 7)), rng.integers(0, 4, size=12)
    a, zs = oracle.forward_all(params, x)
    loss, dws, dbs, _ = oracle.backward(params, a, zs, y)
    acc_w = [np.zeros_like(d) for d in dws]
    acc_l = 0.0
    for s in (slice(0, 5), slice(5, 12)):
        am, zm = oracle.forward_all(params, x[s])
        l, dw, db, _ = oracle.backward(params, am, zm, y[s], denom=12)
        acc_l += l
        for i in range(2):
            acc_w[i] += dw[i]
    assert acc_l == pytest.approx(loss, rel=1e-12)
    for i in range(2):
        np.testing.assert_allclose(acc_w[i], dws[i], rtol=1e-11, atol=1e-15)


def test_hidden_softmax_and_bad_output_activation():
    params = make_params([5, 4, 3], ["softmax", "softmax"], seed=2)
    rng = np.random.default_rng(1)
    x, y = rng.random((6, 5)), rng.integers(0, 3, size=6)
    a, zs = oracle.forward_all(params, x)
    loss, dws, _, _ = oracle.backward(params, a, zs, y)
    ws, bs = _torch_model(params)
    h = torch.softmax(torch.tensor(x) @ ws[0] + bs[0], dim=-1)
    tl = torch.nn.functional.cross_entropy(h @ ws[1] + bs[1], torch.tensor(y))
    tl.backward()
    np.testing.assert_allclose(dws[0], ws[0].grad.numpy(), rtol=1e-10, atol=1e-14)
    bad = make_params([5, 3], ["relu"], seed=2)
    a, zs = oracle.forward_all(bad, x)
    with pytest.raises(ValueError):
        oracle.backward(bad, a, zs, y)


def test_training_restatement_tracks_the_reference_models_own_trajectory():
    """tests/golden/golden_sgd_trajectory.npz was minted (make_golden.py) from the reference's OWN model class and
    loss (scripts/generate_mnist_pytorch.py:22-37: FCNN, nn.CrossEntropyLoss) under torch autograd + plain SGD in
    float64.  The oracle's softmax-CE / backward / SGD restatement must reproduce its 16 losses and final weights."""
    from ddnn_b200.config import batch_slices, synthetic_batch
    g = dict(np.load(os.path.join(GOLDEN, "golden_sgd_trajectory.npz")))
    params = [{"weights": g[f"w{i}_init"].astype(np.float64), "biases": g[f"b{i}_init"].astype(np.float64),
               "activation": a} for i, a in enumerate(["relu", "relu", "softmax"])]
    n, bs = int(g["rows"]), int(g["batch_size"])
    x, y = synthetic_batch(n, 784, 10, seed=int(g["data_seed"]))
    losses = []
    for _ in range(int(g["epochs"])):
        for (a, b) in batch_slices(n, bs):
            losses.append(oracle.train_step(params, x[a:b], y[a:b], float(g["lr"])))
    np.testing.assert_allclose(losses, g["losses"], rtol=1e-12)
    for i, p in enumerate(params):
        np.testing.assert_allclose(p["weights"], g[f"w{i}_final"], rtol=1e-10, atol=1e-14)
        np.testing.assert_allclose(p["biases"], g[f"b{i}_final"], rtol=1e-10, atol=1e-14)
