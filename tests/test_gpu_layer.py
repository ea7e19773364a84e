"""The reference-facing API (`Layer`, `Trainer`) on the GPU: forward against the reference's
golden vectors, training against the float64 oracle's N-step weight trajectory."""
import json
import os

import numpy as np
import pytest
import torch

import oracle
from ddnn_b200.config import make_params, synthetic_batch
from gpu_util import BF16_ACT_TOL, F32_TOL, bf16_update_tol, rel_err, to_dev, to_np

pytestmark = pytest.mark.gpu


def _env_layer(neurons_config, dim, **extra):
    from ddnn_b200 import Layer
    env = {"CONTAINER_NAME": "golden", "EXPECTED_INPUT_DIM": str(dim), "NEURONS_CONFIG": json.dumps(neurons_config),
           "NEXT_NODES": "[]", **extra}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    os.environ.pop("NEURONS_FILE_CONFIG", None)
    try:
        return Layer()
    finally:
        for k, v in old.items():
            os.environ.pop(k, None) if v is None else os.environ.__setitem__(k, v)


def test_kat_through_env_contract(cuda_device, golden_small):
    k = golden_small["kat"]
    L = _env_layer(k["neurons_config"], k["expected_input_dim"])
    assert L.container_name == "golden" and L.expected_input_dim == 2 and L.next_nodes == []
    assert tuple(L.layer_params[0]["weights"].shape) == (2, 3)  # (in, out), grpc_node.py:51
    out = L.process_matrix(np.array(k["input"]))
    assert out.dtype == np.float64
    np.testing.assert_allclose(out, np.array(k["output"]), rtol=F32_TOL)


def test_file_config_and_missing_config(cuda_device, golden_small, tmp_path):
    from ddnn_b200 import Layer
    k = golden_small["kat"]
    f = tmp_path / "n.json"
    f.write_text(json.dumps(k["neurons_config"]))
    old = {v: os.environ.pop(v, None) for v in ("NEURONS_CONFIG", "NEURONS_FILE_CONFIG")}
    try:
        os.environ["NEURONS_FILE_CONFIG"] = str(f)
        os.environ["EXPECTED_INPUT_DIM"] = "2"
        L = Layer()
        np.testing.assert_allclose(L.process_matrix(np.array(k["input"])), np.array(k["output"]), rtol=F32_TOL)
        os.environ["NEURONS_FILE_CONFIG"] = str(tmp_path / "absent.json")
        with pytest.raises(SystemExit) as e:   # grpc_node.py:32-34
            Layer()
        assert e.value.code == 1
        os.environ.pop("NEURONS_FILE_CONFIG")
        with pytest.raises(SystemExit):        # grpc_node.py:38-41
            Layer()
    finally:
        os.environ.pop("EXPECTED_INPUT_DIM", None)
        for k_, v in old.items():
            if v is not None:
                os.environ[k_] = v


def test_small_net_prefixes_and_errors(cuda_device, golden_small):
    s = golden_small["small_net"]
    x = np.array(s["input"])
    for upto, expect in s["prefix_outputs"].items():
        nc = {f"layer_{i + 1}": s["config"]["layers"][i]["neurons"] for i in range(int(upto))}
        out = _env_layer(nc, 5).process_matrix(x)
        assert rel_err(out, np.array(expect)) < F32_TOL, f"prefix {upto}"
    nc = {f"layer_{i + 1}": s["config"]["layers"][i]["neurons"] for i in range(4)}
    L = _env_layer(nc, 5)
    with pytest.raises(ValueError) as e:
        L.process_matrix(np.zeros((2, 3)))
    assert str(e.value) == "(golden) Layer 1: expected input dim 5, got 3"  # grpc_node.py:84 text
    assert L.process_matrix(np.zeros((0, 5))).shape == (0, 3)           # empty in -> empty out
    # CUDA tensor in -> CUDA tensor out
    out = L.process_matrix(to_dev(x))
    assert out.is_cuda and rel_err(to_np(out), np.array(s["prefix_outputs"]["4"])) < F32_TOL
    # activate(): reference semantics incl. unknown -> identity
    a = golden_small["activations"]
    for func in ("relu", "sigmoid", "softmax", "tanh"):
        np.testing.assert_allclose(L.activate(np.array(a["grid"]), func), np.array(a["outputs"][func]), rtol=1e-5,
                                   atol=1e-30)


def test_mnist_forward_golden(cuda_device, golden_mnist):
    from ddnn_b200 import Layer
    for tag, dims in (("mnist_128_64", [784, 128, 64, 10]), ("mnist_32_16", [784, 32, 16, 10])):
        params = make_params(dims, ["relu", "relu", "softmax"], seed=1)
        L = Layer.from_params(params, 784)
        out = L.process_matrix(golden_mnist[f"{tag}_x"])
        assert rel_err(out, golden_mnist[f"{tag}_out"]) < F32_TOL
        assert np.array_equal(out.argmax(axis=1), golden_mnist[f"{tag}_out"].argmax(axis=1))
        Lb = Layer.from_params(params, 784, dtype="bf16")
        assert rel_err(Lb.process_matrix(golden_mnist[f"{tag}_x"]), golden_mnist[f"{tag}_out"]) < BF16_ACT_TOL


@pytest.mark.parametrize("dims,acts", [([784, 128, 64, 10], ["relu", "relu", "softmax"]),
                                        ([784, 32, 16, 10], ["relu", "relu", "softmax"]),
                                        ([20, 17, 9, 5], ["sigmoid", "relu", "linear"])])
def test_gradients_match_oracle(cuda_device, dims, acts):
    from ddnn_b200 import Trainer
    params = make_params(dims, acts, seed=1)
    x, y = synthetic_batch(96, dims[0], dims[-1], seed=0)
    a, zs = oracle.forward_all(params, x)
    loss, dws, dbs, _ = oracle.backward(params, a, zs, y)
    t = Trainer(params)
    L = t.layer
    L.begin_batch(96)
    L.forward_train(to_dev(x), 0)
    gl = L.loss_and_grad(to_dev(y, torch.int32), 0)
    L.backward(0, 96, need_input_grad=False)
    assert abs(gl.item() - loss) < F32_TOL * loss
    for (dw, db), rw, rb in zip(L.gradients(), dws, dbs):
        assert rel_err(to_np(dw), rw) < F32_TOL
        assert rel_err(to_np(db), rb) < F32_TOL


def test_n_step_weight_trajectory_f32(cuda_device):
    """Sequential unshuffled batches with a ragged tail (run_grpc_inference.py:198-205 order),
    20 SGD steps: losses and post-step weights track the float64 oracle within 1e-5."""
    from ddnn_b200 import Trainer
    from ddnn_b200.config import batch_slices
    dims, acts = [784, 128, 64, 10], ["relu", "relu", "softmax"]
    params = make_params(dims, acts, seed=1)
    ref = [dict(p) for p in make_params(dims, acts, seed=1)]
    x, y = synthetic_batch(1000, 784, 10, seed=0)
    t = Trainer(params)
    got_losses = t.fit((x, y), epochs=2, batch_size=96, lr=0.05)
    ref_losses = []
    for _ in range(2):
        for (a, b) in batch_slices(1000, 96):
            ref_losses.append(oracle.train_step(ref, x[a:b], y[a:b], 0.05))
    assert len(got_losses) == len(ref_losses) == 22
    np.testing.assert_allclose(got_losses, ref_losses, rtol=F32_TOL)
    for g, r in zip(t.export_params(), ref):
        assert rel_err(g["weights"], r["weights"]) < F32_TOL
        assert rel_err(g["biases"], r["biases"]) < F32_TOL


def test_trajectory_of_the_reference_model_class_f32(cuda_device):
    """The 16-step SGD trajectory minted from the reference's own FCNN class + nn.CrossEntropyLoss
    (scripts/generate_mnist_pytorch.py:22-37; tests/golden/make_golden.py, torch autograd in float64): a fixture
    for softmax-CE / backward / SGD that does not come from oracle/.  Same initial weights, same seeded data, the
    reference client's batch order; losses and final weights within the fp32 bound."""
    from conftest import GOLDEN
    from ddnn_b200 import Trainer
    g = dict(np.load(os.path.join(GOLDEN, "golden_sgd_trajectory.npz")))
    params = [{"weights": g[f"w{i}_init"], "biases": g[f"b{i}_init"], "activation": a}
              for i, a in enumerate(["relu", "relu", "softmax"])]
    x, y = synthetic_batch(int(g["rows"]), 784, 10, seed=int(g["data_seed"]))
    t = Trainer(params)
    got = t.fit((x, y), epochs=int(g["epochs"]), batch_size=int(g["batch_size"]), lr=float(g["lr"]))
    np.testing.assert_allclose(got, g["losses"], rtol=F32_TOL)
    for i, p in enumerate(t.export_params()):
        assert rel_err(p["weights"], g[f"w{i}_final"]) < F32_TOL
        assert rel_err(p["biases"], g[f"b{i}_final"]) < F32_TOL


def test_cuda_graph_step_equals_eager(cuda_device):
    from ddnn_b200 import Trainer
    dims, acts = [784, 32, 16, 10], ["relu", "relu", "softmax"]
    x, y = synthetic_batch(256, 784, 10, seed=3)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    eager = Trainer(make_params(dims, acts, seed=1))
    graph = Trainer(make_params(dims, acts, seed=1))
    graph.capture(256, lr=0.05)
    sx, sy = graph.static_inputs
    sx.copy_(xd); sy.copy_(yd)
    for _ in range(3):
        le = eager.train_step(xd, yd, 0.05).item()
        lg = graph.replay().item()
        assert le == lg
    for a, b in zip(eager.export_params(), graph.export_params()):
        assert np.array_equal(a["weights"], b["weights"])
    # host entry point goes through the same graph
    l = graph.train_step_host(torch.as_tensor(x, dtype=torch.float32).pin_memory(),
                              torch.as_tensor(y, dtype=torch.int32).pin_memory())
    assert np.isfinite(l)


@pytest.mark.parametrize("dtype,dims", [("f32", [784, 128, 64, 10]), ("bf16", [784, 256, 256, 10]),
                                        ("f32", [100, 200, 150, 12])])
def test_programmatic_dependent_launch_is_transparent(cuda_device, dtype, dims):
    """ddnn_set_pdl(0 / 1): chaining the kernels of a step with programmatic dependent launch (every kernel waits
    for its predecessor before touching global memory) must not change a single bit of losses or weights, eagerly
    and in a captured graph."""
    from ddnn_b200 import Trainer, _lib
    acts = ["relu", "relu", "softmax"]
    x, y = synthetic_batch(512, dims[0], dims[-1], seed=5)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    results = []
    try:
        for pdl in (False, True):
            _lib.set_pdl(pdl)
            t = Trainer(make_params(dims, acts, seed=2), dtype=dtype)
            losses = [t.train_step(xd, yd, 0.05).item() for _ in range(3)]
            t.capture(512, lr=0.05)
            sx, sy = t.static_inputs
            sx.copy_(xd); sy.copy_(yd)
            losses += [t.replay().item() for _ in range(3)]
            results.append((losses, t.export_params()))
    finally:
        _lib.set_pdl(True)
    assert results[0][0] == results[1][0]
    for a, b in zip(results[0][1], results[1][1]):
        assert np.array_equal(a["weights"], b["weights"]) and np.array_equal(a["biases"], b["biases"])


def test_bf16_training_tracks_loss_curve(cuda_device):
    """bf16 tensor-core path: looser bound, matching loss curve (north_star)."""
    from ddnn_b200 import Trainer
    dims, acts = [256, 256, 128, 64], ["relu", "relu", "softmax"]
    params = make_params(dims, acts, seed=2)
    ref = [dict(p) for p in make_params(dims, acts, seed=2)]
    x, y = synthetic_batch(512, 256, 64, seed=1)
    t = Trainer(params, dtype="bf16")
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    for step in range(10):
        lg = t.train_step(xd, yd, 0.1).item()
        lr_ = oracle.train_step(ref, x, y, 0.1)
        assert abs(lg - lr_) < 2e-2 * lr_, (step, lg, lr_)
    for g, r in zip(t.export_params(), ref):
        assert rel_err(g["weights"], r["weights"]) < BF16_ACT_TOL
    # the bf16 shadow is exactly the rounded master
    for p, sh in zip(t.layer.layer_params, t.layer._shadow):
        assert torch.equal(sh, p["weights"].to(torch.bfloat16))


def test_two_stage_chain_equals_single_node(cuda_device):
    """[2, 1] split on one device: stage 2's dX applies stage 1's activation derivative; the
    deferred full-batch dW step over 3 micro-batches equals the single-node step."""
    from ddnn_b200 import Layer, Trainer
    dims, acts = [64, 48, 40, 10], ["relu", "sigmoid", "softmax"]
    x, y = synthetic_batch(90, 64, 10, seed=5)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    single = Trainer(make_params(dims, acts, seed=4))
    single.train_step(xd, yd, 0.1)
    p = make_params(dims, acts, seed=4)
    s0 = Layer.from_params(p[:2], 64, is_output=False)
    s1 = Layer.from_params(p[2:], 40, is_output=True, input_activation="sigmoid")
    s0.begin_batch(90); s1.begin_batch(90)
    for r0 in (0, 30, 60):
        out = s0.forward_train(xd[r0:r0 + 30], r0)
        s1.forward_train(out, r0)
    for i, r0 in enumerate((0, 30, 60)):
        s1.loss_and_grad(yd[r0:r0 + 30], r0, denom=90, accumulate=i > 0)
        dx = s1.backward(r0, 30)
        s0.grad_buffer(r0, 30).copy_(dx)
        s0.backward(r0, 30, need_input_grad=False)
    s0.step(0.1); s1.step(0.1)
    chained = s0.export_params() + s1.export_params()
    for a, b in zip(single.export_params(), chained):
        assert rel_err(b["weights"], a["weights"]) < 1e-6
        assert rel_err(b["biases"], a["biases"]) < 1e-6


def test_direct_output_targets_and_flags(cuda_device):
    """What the peer transport does, on one GPU and ONE stream (so every flag is already set when its
    wait executes -- two streams waiting on each other on one GPU may share a hardware queue and
    deadlock): stage 0's forward epilogue writes straight into stage 1's input buffer (out_target),
    stage 1's dX epilogue writes into stage 0's gradient buffer (dx_target), ddnn_flag_write /
    ddnn_flag_wait bracket every hand-over.  Result must equal the single-node step to fp32 rounding."""
    from ddnn_b200 import Layer, Trainer, _lib
    dims, acts = [64, 48, 40, 10], ["relu", "sigmoid", "softmax"]
    x, y = synthetic_batch(96, 64, 10, seed=6)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    single = Trainer(make_params(dims, acts, seed=4))
    single.train_step(xd, yd, 0.1)
    p = make_params(dims, acts, seed=4)
    s0 = Layer.from_params(p[:2], 64, is_output=False)
    s1 = Layer.from_params(p[2:], 40, is_output=True, input_activation="sigmoid")
    s0.begin_batch(96); s1.begin_batch(96)
    flags = torch.zeros(8, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    fw = lambda i, v: _lib.check(_lib.lib.ddnn_flag_write(flags.data_ptr() + 4 * i, v, st))  # noqa: E731
    fwait = lambda i, v: _lib.check(_lib.lib.ddnn_flag_wait(flags.data_ptr() + 4 * i, v, st))  # noqa: E731
    chunks = [(0, 32), (32, 64), (64, 96)]
    for m, (r0, r1) in enumerate(chunks):
        s0.forward_train(xd[r0:r1], r0, out_target=s1.input_buffer(r0, r1 - r0))
        fw(m, 1)
        fwait(m, 1)
        s1.forward_train(None, r0, r1 - r0)
    for m, (r0, r1) in enumerate(chunks):
        s1.loss_and_grad(yd[r0:r1], r0, denom=96, accumulate=m > 0)
        s1.backward(r0, r1 - r0, dx_target=s0.grad_buffer(r0, r1 - r0))
        fw(4 + m, 2)
        fwait(4 + m, 1)   # GEQ: 2 >= 1
        s0.backward(r0, r1 - r0, need_input_grad=False)
    s0.step(0.1); s1.step(0.1)
    torch.cuda.synchronize()
    assert flags.tolist() == [1, 1, 1, 0, 2, 2, 2, 0]
    # (the single-node Trainer runs its output layer through the fused kernel: same math, different
    # summation order, so the comparison is to fp32 rounding rather than bit for bit)
    for a, b in zip(single.export_params(), s0.export_params() + s1.export_params()):
        assert rel_err(b["weights"], a["weights"]) < 1e-6 and rel_err(b["biases"], a["biases"]) < 1e-6


def layers_from_cfg(cfg):
    from ddnn_b200.config import layers_to_params
    return layers_to_params(cfg["layers"])


def test_json_export_roundtrip_and_eval_driver(cuda_device, tmp_path):
    """train on the GPU, write the model back in the reference's JSON
    schema, reload it (bit-identical predictions), serve it through the oracle (= the reference forward)
    and run the run_grpc_inference.py-compatible evaluation driver on an examples file."""
    from ddnn_b200 import Trainer
    from ddnn_b200 import infer
    from ddnn_b200.config import load_model_config, make_config
    dims, acts = [20, 16, 12, 4], ["relu", "sigmoid", "softmax"]
    cfg = make_config(dims, acts, seed=6, layer_distribution=[2, 1], learning_rate=0.2, epochs=3, batch_size=32)
    x, y = synthetic_batch(100, 20, 4, seed=2)
    t = Trainer(cfg)
    losses = t.fit((x, y))                       # epochs / batch size / lr from the config
    ref = [dict(p) for p in layers_from_cfg(cfg)]
    ref_losses = [oracle.train_step(ref, x[a:b], y[a:b], 0.2) for _ in range(3) for (a, b) in
                  [(0, 32), (32, 64), (64, 96), (96, 100)]]
    np.testing.assert_allclose(losses, ref_losses, rtol=F32_TOL)
    path = tmp_path / "trained.json"
    t.save(str(path))
    saved = json.loads(path.read_text())
    assert [l["nodes"] for l in saved["layers"]] == [16, 12, 4] and saved["layer_distribution"] == [2, 1]
    assert saved["layers"][0]["type"] == "hidden" and saved["layers"][-1]["type"] == "output"
    assert saved["layers"][1]["neurons"][0]["activation"] == "sigmoid" and len(saved["layers"][0]["neurons"][0]["weights"]) == 20
    # reload on the GPU: identical predictions
    t2 = Trainer(str(path))
    assert np.array_equal(t.predict(x), t2.predict(x))
    # the reference's forward (oracle, pinned bit-exact to grpc_node.py) on the exported file
    ref_params = oracle.build_layer_params({f"layer_{i + 1}": l["neurons"] for i, l in enumerate(saved["layers"])})
    ref_out = oracle.process_matrix(ref_params, x, 20)
    assert rel_err(t.predict(x), ref_out) < F32_TOL
    # evaluation driver: reference CLI semantics (sequential batches, ragged tail, argmax == label)
    examples = [{"input": x[i].tolist(), "label": int(y[i])} for i in range(100)]
    inputs = tmp_path / "examples.json"
    inputs.write_text(json.dumps({"examples": examples}))
    correct, n, total, per_batch = infer.evaluate(str(path), examples, batch_size=32)
    assert n == 100 and len(per_batch) == 4
    assert correct == int((ref_out.argmax(axis=1) == y).sum())
    c1, n1, _, pb1 = infer.evaluate(str(path), examples, input_index=7)
    assert n1 == 100 and len(pb1) == 1 and c1 == int(ref_out[7].argmax() == y[7])
    assert infer.main(["--config", str(path), "--inputs", str(inputs), "--batch-size", "40"]) == 0
    assert infer.main(["200", "--config", str(path), "--inputs", str(inputs)]) == 1  # index out of range


def test_default_bench_configuration_matches_oracle(cuda_device):
    """BASELINE.json configs[1] at its own size: 784-128-64-10, batch 4096, fp32, the captured 7-launch step of
    bench.py (tcgen05 3xTF32 forward / dW, fused two-layer tail, side-stream reducers): 3 steps against the float64
    oracle, losses and every weight within the fp32 bound."""
    from ddnn_b200 import Trainer
    dims, acts = [784, 128, 64, 10], ["relu", "relu", "softmax"]
    ref = [dict(p) for p in make_params(dims, acts, seed=1)]
    x, y = synthetic_batch(4096, 784, 10, seed=0)
    t = Trainer(make_params(dims, acts, seed=1))
    assert t.layer.begin_batch(4096) is None and t.layer.can_fuse_tail2()
    t.capture(4096, lr=0.01)
    sx, sy = t.static_inputs
    sx.copy_(to_dev(x)); sy.copy_(to_dev(y, torch.int32))
    for _ in range(3):
        got = t.replay().item()
        want = oracle.train_step(ref, x, y, 0.01)
        assert abs(got - want) < F32_TOL * want
    for g, r in zip(t.export_params(), ref):
        assert rel_err(g["weights"], r["weights"]) < F32_TOL
        # Biases are 0.01 +- 1e-4, so 1e-5 of max|b| is 1e-7 absolute = a fraction of ONE dZ element times lr: a single
        # ReLU gate that flips because a pre-activation is closer to zero than the fp32 forward error (786 432 gates
        # per step here) moves a bias by that much -- in any fp32 implementation (measured: the same 1.0e-5 / 2.9e-5
        # with mma.sync and tcgen05 products and with either tail kernel, whatever their forward accuracy).
        assert rel_err(g["biases"], r["biases"]) < 5 * F32_TOL


def test_narrow_configuration_tracks_oracle_at_full_size(cuda_device):
    """BASELINE.json configs[4] at its own size: 784-256-256-10, batch 65536, bf16 (pair-kernel GEMMs with 8
    epilogue warps, split-K dW with the ones-MMA bias gradient, output layer on mma.sync): two steps against the
    float64 oracle within the bf16 bounds; the bias gradients (the fused paths) within the gradient bound."""
    from ddnn_b200 import Trainer
    dims, acts = [784, 256, 256, 10], ["relu", "relu", "softmax"]
    ref = [dict(p) for p in make_params(dims, acts, seed=1)]
    before = [dict(p) for p in make_params(dims, acts, seed=1)]
    x, y = synthetic_batch(65536, 784, 10, seed=0)
    t = Trainer(make_params(dims, acts, seed=1), dtype="bf16")
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    lr = 0.5  # large enough for the update to stand out of the bf16 noise
    for _ in range(2):
        got = t.train_step(xd, yd, lr).item()
        want = oracle.train_step(ref, x, y, lr)
        assert abs(got - want) < 2e-2 * want
    for i, (g, r, b0) in enumerate(zip(t.export_params(), ref, before)):
        assert rel_err(g["weights"], r["weights"]) < BF16_ACT_TOL
        # the UPDATE itself (difference to the initial weights), not just the weights it barely moves: bound from
        # bf16 rounding plus, below a ReLU, gate flips (gpu_util.bf16_update_tol; two steps are compared, so twice)
        tol = 2 * bf16_update_tol(65536, below_relu=(i < 2))
        assert rel_err(g["weights"] - b0["weights"], r["weights"] - b0["weights"]) < tol, i
        assert rel_err(g["biases"] - b0["biases"], r["biases"] - b0["biases"]) < tol, i


@pytest.mark.parametrize("dtype,dims,cap,tail", [("f32", [784, 128, 64, 10], 4096, 3540),
                                                 ("f32", [784, 128, 64, 10], 4096, 1249),
                                                 ("bf16", [784, 256, 256, 10], 65536, 36000)])
def test_ragged_last_batch_needs_more_split_partials(cuda_device, dtype, dims, cap, tail):
    """Split plans are not monotone in the batch: the fused two-layer tail runs 147 CTAs at 4096 rows and 148 at
    3540, the fused bf16 output layer 512 blocks at 65536 rows and up to 592 around 36000, the tf32 split-K dW more
    splits at 1249 rows than at 4096.  A full batch followed by such a ragged last one (the order of
    run_grpc_inference.py:198-205) must run through the buffers sized for the full batch -- and agree with the
    oracle -- instead of failing with DDNN_ERR_WORKSPACE."""
    from ddnn_b200 import Trainer
    acts = ["relu", "relu", "softmax"]
    ref = [dict(p) for p in make_params(dims, acts, seed=1)]
    t = Trainer(make_params(dims, acts, seed=1), dtype=dtype)
    x, y = synthetic_batch(cap + tail, dims[0], dims[-1], seed=0)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    tol = F32_TOL if dtype == "f32" else 2e-2
    for (a, b) in ((0, cap), (cap, cap + tail)):
        got = t.train_step(xd[a:b], yd[a:b], 0.05).item()
        want = oracle.train_step(ref, x[a:b], y[a:b], 0.05)
        assert abs(got - want) < tol * want
    for g, r in zip(t.export_params(), ref):
        assert rel_err(g["weights"], r["weights"]) < (F32_TOL if dtype == "f32" else BF16_ACT_TOL)


@pytest.mark.parametrize("dtype,dims", [("f32", [784, 128, 64, 10]), ("bf16", [784, 256, 256, 10])])
def test_input_batch_is_read_in_place(cuda_device, dtype, dims):
    """Layer 1 (forward and dW_1) reads the caller's batch where it lies when it has the compute dtype (no
    device-to-device staging copy); a batch that cannot be adopted (here: a strided view) still goes through the
    node's own buffer.  Both routes give the same bits."""
    from ddnn_b200 import Trainer
    acts = ["relu", "relu", "softmax"]
    cdt = torch.float32 if dtype == "f32" else torch.bfloat16
    x, y = synthetic_batch(512, dims[0], dims[-1], seed=5)
    xd, yd = to_dev(x).to(cdt).contiguous(), to_dev(y, torch.int32)
    wide = torch.zeros((512, dims[0] + 16), dtype=cdt, device=xd.device)
    wide[:, :dims[0]] = xd
    strided = wide[:, :dims[0]]
    assert not strided.is_contiguous()
    t_in, t_cp = (Trainer(make_params(dims, acts, seed=1), dtype=dtype) for _ in range(2))
    for _ in range(3):
        l_in = t_in.train_step(xd, yd, 0.05).item()
        assert t_in.layer._a0_ext is xd
        l_cp = t_cp.train_step(strided, yd, 0.05).item()
        assert t_cp.layer._a0_ext is None
        assert l_in == l_cp
    for a, b in zip(t_in.export_params(), t_cp.export_params()):
        assert np.array_equal(a["weights"], b["weights"]) and np.array_equal(a["biases"], b["biases"])


@pytest.mark.parametrize("dtype,dims,rows,chunk", [("f32", [784, 128, 64, 10], 1000, 384),
                                                   ("bf16", [512, 512, 256, 64], 4096, 1024)])
def test_chunked_weight_gradient_and_update(cuda_device, dtype, dims, rows, chunk):
    """ddnn_dense_backward_dw_accumulate + ddnn_sgd_apply (Layer.step(chunk_rows=...)): the gradient summed over row
    chunks (ragged last chunk) and applied once equals the one-piece fused update up to fp32 rounding of the
    partial sums, is bit-reproducible, and tracks the float64 oracle."""
    from ddnn_b200 import Trainer
    acts = ["relu", "sigmoid", "softmax"]
    x, y = synthetic_batch(rows, dims[0], dims[-1], seed=11)
    xd, yd = to_dev(x), to_dev(y, torch.int32)
    ref = [dict(p) for p in make_params(dims, acts, seed=4)]
    want = oracle.train_step(ref, x, y, 0.1)
    outs = []
    for chunk_rows in (None, chunk, chunk):
        t = Trainer(make_params(dims, acts, seed=4), dtype=dtype)
        L = t.layer
        L.begin_batch(rows)
        L.forward_train(xd, 0)
        got = L.loss_and_grad(yd, 0).item()
        L.backward(0, rows, need_input_grad=False)
        L.step(0.1, rows, chunk_rows=chunk_rows)
        assert abs(got - want) < (F32_TOL if dtype == "f32" else 2e-2) * want
        outs.append(t.export_params())
    one_piece, chunked, again = outs
    for a, b, c, r in zip(one_piece, chunked, again, ref):
        assert np.array_equal(b["weights"], c["weights"]) and np.array_equal(b["biases"], c["biases"])
        assert rel_err(b["weights"], a["weights"]) < 1e-6 and rel_err(b["biases"], a["biases"]) < 1e-6
        assert rel_err(b["weights"], r["weights"]) < (F32_TOL if dtype == "f32" else BF16_ACT_TOL)


def test_single_stage_pipeline_host_loop_and_relay(cuda_device):
    """PipelineStage on ONE stage (what bench.py times at N = 1): `fit_stream` (pinned host batches, H2D per
    micro-batch on a copy stream, the slot read in place) gives the same losses and weights as stepping on device
    tensors, in both orders with equal gradient chunks; `infer` (forward-only relay on the resident buffers) equals
    `process_matrix`."""
    from ddnn_b200 import Layer
    from ddnn_b200.pipeline import PipelineStage
    dims, acts = [256, 256, 128, 64], ["relu", "sigmoid", "softmax"]
    rows, micro = 1024, 8
    batches = [synthetic_batch(rows, dims[0], dims[-1], seed=s) for s in range(3)]

    def stage(schedule, dw_chunk):
        L = Layer.from_params(make_params(dims, acts, seed=2), dims[0], dtype="bf16", is_output=True)
        return PipelineStage(L, 0, 1, schedule=schedule, dw_chunk=dw_chunk)

    dev_stage, host_stage, zb_stage = stage("gpipe", 2), stage("gpipe", 2), stage("zb", 2)
    dev_losses = []
    for x, y in batches:
        xd = to_dev(x, torch.bfloat16)
        dev_losses.append(dev_stage.train_step(xd, to_dev(y, torch.int32), rows, micro, 0.05).item())
    host = [(torch.as_tensor(x, dtype=torch.float32).to(torch.bfloat16).pin_memory(),
             torch.as_tensor(y, dtype=torch.int32).pin_memory()) for x, y in batches]
    assert host_stage.fit_stream(host, rows, micro, 0.05) == dev_losses
    assert zb_stage.fit_stream(host, rows, micro, 0.05) == dev_losses
    for a, b, c in zip(dev_stage.layer.layer_params, host_stage.layer.layer_params, zb_stage.layer.layer_params):
        assert torch.equal(a["weights"], b["weights"]) and torch.equal(a["weights"], c["weights"])
        assert torch.equal(a["biases"], b["biases"]) and torch.equal(a["biases"], c["biases"])
    xq = to_dev(batches[0][0], torch.bfloat16)
    relay = dev_stage.infer(xq, rows, 3)
    direct = dev_stage.layer.process_matrix(xq)
    assert torch.equal(relay, direct)
    ref = oracle.process_matrix(dev_stage.layer.export_params(), to_np(xq), dims[0])
    assert rel_err(to_np(relay), ref) < BF16_ACT_TOL


def test_fit_stream_takes_host_rows_in_the_compute_dtype(cuda_device):
    """A bf16 model fed bf16 host rows (half the PCIe bytes, slot read in place) and fed the same values as fp32 rows
    (cast on the device) trains to the same bits."""
    from ddnn_b200 import Trainer
    dims, acts = [784, 256, 256, 10], ["relu", "relu", "softmax"]
    data = [synthetic_batch(512, 784, 10, seed=s) for s in range(4)]
    hb = [(torch.as_tensor(x, dtype=torch.float32).to(torch.bfloat16).pin_memory(),
           torch.as_tensor(y, dtype=torch.int32).pin_memory()) for x, y in data]
    hf = [(xb.float().pin_memory(), yb) for xb, yb in hb]   # the same (bf16-representable) values as fp32
    ta, tb = (Trainer(make_params(dims, acts, seed=1), dtype="bf16") for _ in range(2))
    assert ta.fit_stream(hb, lr=0.05) == tb.fit_stream(hf, lr=0.05)
    for a, b in zip(ta.export_params(), tb.export_params()):
        assert np.array_equal(a["weights"], b["weights"])
