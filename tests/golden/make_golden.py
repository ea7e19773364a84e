"""Generates the golden vectors under tests/golden/ by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (it reads /root/reference, which does not exist on the GPU
box):   python tests/golden/make_golden.py

The reference ships no tests or fixtures, so these are minted from its
own code: ``Layer`` / ``Layer.activate`` / ``Layer.process_matrix`` of
/root/reference/src/grpc_node.py, ``LayerServiceServicer.Process`` for the empty-request
behaviour, ``calculate_layer_mappings`` of /root/reference/src/run_grpc_fcnn.py (imported
with a stub ``docker`` module: the SDK is not installed) and ``forward_pass`` of
/root/reference/scripts/manual_nn.py as a second statement of the forward.
"""
import contextlib
import io
import json
import os
import sys
import types

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)


def import_reference():
    os.chdir(os.path.join(REF, "src"))  # grpc_node.py:10 adds cwd/proto to sys.path
    sys.path.insert(0, os.path.join(REF, "src"))
    sys.modules.setdefault("docker", types.ModuleType("docker"))
    # `docker.DockerClient` / `docker.models...` appear in annotations evaluated at import
    d = sys.modules["docker"]
    d.DockerClient = object
    d.models = types.SimpleNamespace(containers=types.SimpleNamespace(Container=object),
                                     networks=types.SimpleNamespace(Network=object))
    d.errors = types.SimpleNamespace(APIError=Exception, NotFound=Exception)
    import grpc_node
    import run_grpc_fcnn
    return grpc_node, run_grpc_fcnn


def ref_layer(grpc_node, neurons_config, expected_dim, name="golden"):
    env = {"CONTAINER_NAME": name, "EXPECTED_INPUT_DIM": str(expected_dim),
           "NEURONS_CONFIG": json.dumps(neurons_config), "NEXT_NODES": "[]"}
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    os.environ.pop("NEURONS_FILE_CONFIG", None)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            return grpc_node.Layer()
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def quiet(fn, *a):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a)


def load_manual_forward():
    """scripts/manual_nn.py executes file reads at import (:73-99); take only the function
    definitions above them."""
    src = open(os.path.join(REF, "scripts", "manual_nn.py")).read()
    head = src.split("CONFIG_FILE =")[0]
    ns = {}
    exec(compile(head, "manual_nn_head", "exec"), ns)
    return ns["forward_pass"]


def load_reference_model_class():
    """scripts/generate_mnist_pytorch.py downloads MNIST and trains at import; take only its model definition
    (`class FCNN`, :22-34) -- the reference's own statement of the 784-128-64-10 network whose loss is
    `nn.CrossEntropyLoss()` on the logits (:37)."""
    src = open(os.path.join(REF, "scripts", "generate_mnist_pytorch.py")).read()
    body = src[src.index("class FCNN"):src.index("# Instantiate the model")]
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    ns = {"torch": torch, "nn": nn, "F": F}
    exec(compile(body, "generate_mnist_pytorch_FCNN", "exec"), ns)
    return ns["FCNN"]


def sgd_trajectory():
    """a6-a8 (softmax-CE, backward, SGD) have no implementation in the reference's src/; the one place the
    reference states training semantics is its torch script.  This mints an N-step trajectory from THAT model
    class and loss (float64, optimizer swapped from Adam to plain SGD as north_star mandates) on seeded synthetic
    data in the reference client's batch order (sequential, ragged tail: run_grpc_inference.py:198-205).  The
    fixture therefore comes from reference-authored code + torch autograd, not from oracle/."""
    import torch
    from ddnn_b200.config import batch_slices, synthetic_batch
    FCNN = load_reference_model_class()
    crit = torch.nn.CrossEntropyLoss()
    lr, bs, epochs, n = 0.05, 128, 2, 1000

    def run(data_seed):
        torch.manual_seed(0)
        model = FCNN()
        # start from float32-representable weights (what a GPU fp32 run can be given exactly), then go to float64
        for prm in model.parameters():
            prm.data = prm.data.float().double()
        opt = torch.optim.SGD(model.parameters(), lr=lr)
        x, y = synthetic_batch(n, 784, 10, seed=data_seed)
        fcs = [model.fc1, model.fc2, model.fc3]
        out = {"lr": np.float64(lr), "batch_size": np.int64(bs), "epochs": np.int64(epochs), "rows": np.int64(n),
               "data_seed": np.int64(data_seed)}
        for i, fc in enumerate(fcs):  # exporter convention (pytorch.py:68-103): one row of fc.weight per neuron -> W = weight.T
            out[f"w{i}_init"] = fc.weight.detach().numpy().T.astype(np.float32)
            out[f"b{i}_init"] = fc.bias.detach().numpy().astype(np.float32)
        losses, min_z = [], np.inf
        xt, yt = torch.as_tensor(x, dtype=torch.float64), torch.as_tensor(y, dtype=torch.long)
        for _ in range(epochs):
            for (a, b) in batch_slices(n, bs):
                with torch.no_grad():  # smallest |pre-activation| of a ReLU gate in this step
                    z1 = model.fc1(xt[a:b])
                    z2 = model.fc2(torch.relu(z1))
                    min_z = min(min_z, float(z1.abs().min()), float(z2.abs().min()))
                opt.zero_grad()
                loss = crit(model(xt[a:b]), yt[a:b])
                loss.backward()
                opt.step()
                losses.append(float(loss.item()))
        out["losses"] = np.array(losses)
        out["min_abs_preactivation"] = np.float64(min_z)
        for i, fc in enumerate(fcs):
            out[f"w{i}_final"] = fc.weight.detach().numpy().T.copy()
            out[f"b{i}_final"] = fc.bias.detach().numpy().copy()
        return out

    # A ReLU gate whose pre-activation is closer to zero than the fp32 forward error (~1e-6 here) flips in ANY fp32
    # implementation and moves a weight by lr * x * dz / B ~ 1e-5: with 384 000 gates along the trajectory a few
    # are always that marginal (seed 21: one at 5e-8).  The data seed is the first one whose trajectory has none
    # below 2e-6, so that the fixture tests the arithmetic and not the luck of a gate.
    for data_seed in range(21, 400):
        out = run(data_seed)
        if float(out["min_abs_preactivation"]) >= 2e-6:
            break
    else:
        raise RuntimeError("no data seed without a marginal ReLU gate")
    losses = out["losses"]
    np.savez_compressed(os.path.join(HERE, "golden_sgd_trajectory.npz"), **out)
    print("wrote", os.path.join(HERE, "golden_sgd_trajectory.npz"), f"({len(losses)} steps)")


def main():
    sgd_trajectory()
    grpc_node, fcnn = import_reference()
    from ddnn_b200.config import make_config, make_params, params_to_layers

    out = {}

    # ---- 1. the SURVEY section 4 known-answer test: 2 -> 3 relu -> 2 softmax ----------------
    sample = json.load(open(os.path.join(REF, "config", "config_sample.json")))
    inputs = json.load(open(os.path.join(REF, "config", "example_inputs", "example_inputs_sample.json")))
    kat_cfg = {"layer_1": sample["layers"][0]["neurons"], "layer_2": sample["layers"][2]["neurons"]}
    x = np.array(inputs["examples"][0]["input"])
    L = ref_layer(grpc_node, kat_cfg, 2)
    out["kat"] = {"neurons_config": kat_cfg, "expected_input_dim": 2, "input": x.tolist(),
                  "output": quiet(L.process_matrix, x).tolist()}

    # ---- 2. activations on a grid incl. extremes (grpc_node.py:62-73) ----------------------
    grid = np.array([[-1e3, -30.0, -1.5, -0.0, 0.0, 1e-8, 0.5, 2.0, 30.0, 700.0],
                     [3.0, 3.0, 3.0, 3.0, 3.0, 3.0, 3.0, 3.0, 3.0, 3.0],
                     [-5.0, -4.0, -3.0, -2.0, -1.0, 0.0, 1.0, 2.0, 3.0, 4.0]])
    acts = {}
    with np.errstate(over="ignore"):
        for f in ("relu", "sigmoid", "softmax", "linear", "unknown", "tanh"):
            acts[f] = L.activate(grid, f).tolist()
    out["activations"] = {"grid": grid.tolist(), "outputs": acts}

    # ---- 3. small multi-activation net, every prefix as one node ----------------------------
    dims = [5, 7, 6, 4, 3]
    act_names = ["relu", "sigmoid", "unknown", "softmax"]
    cfg = make_config(dims, act_names, seed=7)
    rng = np.random.default_rng(11)
    xs = rng.random((9, 5)) * 2 - 0.5
    prefixes = {}
    for upto in range(1, len(dims)):
        nc = {f"layer_{i + 1}": cfg["layers"][i]["neurons"] for i in range(upto)}
        prefixes[str(upto)] = quiet(ref_layer(grpc_node, nc, 5).process_matrix, xs).tolist()
    # per-stage split [2, 2]: second node fed by the first node's output
    nc_a = {f"layer_{i + 1}": cfg["layers"][i]["neurons"] for i in range(2)}
    nc_b = {f"layer_{i + 1}": cfg["layers"][2 + i]["neurons"] for i in range(2)}
    mid = quiet(ref_layer(grpc_node, nc_a, 5).process_matrix, xs)
    chained = quiet(ref_layer(grpc_node, nc_b, 6).process_matrix, mid)
    manual = load_manual_forward()
    # manual_nn has no sigmoid (:17-21): check it on a relu/linear/softmax net instead
    cfg_m = make_config([5, 7, 4, 3], ["relu", "linear", "softmax"], seed=8)
    man_out = [manual(cfg_m, row).tolist() for row in xs]
    nc_m = {f"layer_{i + 1}": cfg_m["layers"][i]["neurons"] for i in range(3)}
    ref_m = quiet(ref_layer(grpc_node, nc_m, 5).process_matrix, xs)
    assert np.allclose(np.array(man_out), ref_m, rtol=1e-12, atol=1e-15), "manual_nn disagrees with Layer"
    out["small_net"] = {"config": cfg, "dims": dims, "activations": act_names, "input": xs.tolist(),
                        "prefix_outputs": prefixes, "chained_2_2": chained.tolist(),
                        "manual_nn_config": cfg_m, "manual_nn_outputs": man_out}

    # ---- 4. error / edge behaviour -------------------------------------------------------------
    try:
        quiet(L.process_matrix, np.zeros((2, 3)))
        msg = None
    except ValueError as e:
        msg = str(e)

    class Ctx:
        def set_code(self, c): self.code = c
        def set_details(self, d): self.details = d
    sv = grpc_node.LayerServiceServicer(L)
    empty = quiet(sv.Process, grpc_node.dist_nn_pb2.Matrix(), Ctx())
    out["edges"] = {"dim_mismatch_message": msg, "empty_request_rows": len(empty.rows)}

    # ---- 5. calculate_layer_mappings (run_grpc_fcnn.py:176-252) -----------------------------------
    layers6 = params_to_layers(make_params([4, 5, 6, 7, 3, 2, 2], ["relu"] * 5 + ["softmax"], seed=3))
    examples = [{"input": [0.0, 0.1, 0.2, 0.3], "label": 1}]
    maps = {}
    for dist in ([6], [1, 1, 1, 1, 1, 1], [2, 2, 2], [3, 0, 3], [1, 2, 0, 3], [0, 6], [4, 2]):
        m = fcnn.calculate_layer_mappings(layers6, dist, examples)
        # neurons_config is large: keep the layer keys and the neuron counts only
        slim = {str(k): {"container_name": v["container_name"], "listen_port": v["listen_port"],
                         "expected_input": v["expected_input"], "next_nodes": v["next_nodes"],
                         "layer_keys": {lk: len(lv) for lk, lv in v["neurons_config"].items()}}
                for k, v in m.items()}
        maps[json.dumps(dist)] = slim
    try:
        fcnn.calculate_layer_mappings(layers6, [1], examples)
        bad = None
    except ValueError as e:
        bad = str(e)
    out["mappings"] = {"dims": [4, 5, 6, 7, 3, 2, 2], "seed": 3, "examples": examples, "by_distribution": maps,
                       "bad_sum_message": bad}

    with open(os.path.join(HERE, "golden_small.json"), "w") as f:
        json.dump(out, f)

    # ---- 6. MNIST-shaped nets (the scripts' 784-128-64-10 and the notebook's 784-32-16-10) --------
    big = {}
    for tag, d in (("mnist_128_64", [784, 128, 64, 10]), ("mnist_32_16", [784, 32, 16, 10])):
        params = make_params(d, ["relu", "relu", "softmax"], seed=1)
        xb = np.random.default_rng(0).random((32, 784))
        Lb = grpc_node.Layer.__new__(grpc_node.Layer)  # multi-MB JSON through an env var is pointless
        Lb.container_name, Lb.expected_input_dim = "golden", 784
        Lb.layer_params = [dict(p) for p in params]
        big[f"{tag}_x"] = xb
        big[f"{tag}_out"] = quiet(Lb.process_matrix, xb)
        Lb.layer_params = [dict(p) for p in params[:1]]
        big[f"{tag}_h1"] = quiet(Lb.process_matrix, xb)
        big[f"{tag}_w_checksum"] = np.array([float(np.sum(p["weights"])) for p in params])
    np.savez_compressed(os.path.join(HERE, "golden_mnist.npz"), **big)
    print("wrote", os.path.join(HERE, "golden_small.json"), os.path.join(HERE, "golden_mnist.npz"))


if __name__ == "__main__":
    main()
