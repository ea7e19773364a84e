"""Model / input JSON schemas, stage partitioning and batch ordering of docker-dist-nn.

Pure host code (numpy only).  Keeps the reference's file formats so any file the
reference accepts loads here unchanged:

* model    ``{"layers": [{"type", "nodes", "neurons": [{"weights", "bias", "activation"}]}],
              "layer_distribution": [...]}``  -- /root/reference/config/config_sample.json:1-33,
              read at /root/reference/src/run_grpc_fcnn.py:265-266
* per node ``{"layer_<n>": [neurons...]}``    -- run_grpc_fcnn.py:213-214, grpc_node.py:43-55
* inputs   ``{"examples": [{"input": [...], "label": int}]}`` -- run_grpc_inference.py:35-52

Optional top-level training keys (ignored by the reference, which only reads ``layers``
and ``layer_distribution``): ``learning_rate`` (0.01), ``epochs`` (1), ``batch_size``
(None = one batch), ``seed``, ``micro_batches`` (1), ``dtype`` ("f32" | "bf16").
"""
from __future__ import annotations

import json
import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np

# run_grpc_fcnn.py:20-23
CONTAINER_NAME_PREFIX = "layer_container_"
BASE_PORT_FIRST_LAYER = 5100
BASE_PORT_INCREMENT = 100

TRAIN_DEFAULTS = {"learning_rate": 0.01, "epochs": 1, "batch_size": None, "seed": 0, "micro_batches": 1,
                  "dtype": "f32"}


# ---------------------------------------------------------------------------
# model config
# ---------------------------------------------------------------------------

def load_model_config(src) -> Dict[str, Any]:
    """Reads a model file (path, JSON string or dict).  ``layers`` must be top level
    (run_grpc_fcnn.py:265); the notebook exporter's ``{"model": {"layers": ...}}``
    nesting (Centralized_MNIST_Experimentation.ipynb:494-506) is un-nested for
    convenience.  Training keys get their defaults."""
    if isinstance(src, dict):
        cfg = dict(src)
    else:
        text = str(src)
        if text.lstrip().startswith("{"):
            cfg = json.loads(text)
        else:
            with open(text, "r") as f:
                cfg = json.load(f)
    if "layers" not in cfg and isinstance(cfg.get("model"), dict) and "layers" in cfg["model"]:
        inner = cfg["model"]
        cfg = {**{k: v for k, v in cfg.items() if k != "model"}, **inner}
    if "layers" not in cfg:
        raise KeyError("layers")
    for k, v in TRAIN_DEFAULTS.items():
        cfg.setdefault(k, v)
    if "layer_distribution" not in cfg:
        # The reference defaults to [1] (run_grpc_fcnn.py:23), which fails its own sum check
        # for any multi-layer model; one layer per stage is the usable default.
        cfg["layer_distribution"] = [1] * len(cfg["layers"])
    return cfg


def neurons_to_params(neurons: Sequence[dict]) -> Dict[str, Any]:
    """grpc_node.py:51-53 for one layer: weights (K, N) = per-neuron rows transposed,
    biases (N,), activation of the FIRST neuron (default ``linear``)."""
    w = np.array([n["weights"] for n in neurons], dtype=np.float64).T
    b = np.array([n["bias"] for n in neurons], dtype=np.float64)
    if w.ndim != 2:
        raise ValueError("neurons of one layer must all have the same number of weights")
    return {"weights": w, "biases": b, "activation": neurons[0].get("activation", "linear")}


def node_config_to_params(neurons_config_data: Dict[str, list]) -> List[Dict[str, Any]]:
    """grpc_node.py:43-55: ``layer_<n>`` keys sorted by integer suffix, empty layers skipped."""
    out = []
    for key in sorted(neurons_config_data, key=lambda x: int(x.split("_")[1])):
        neurons = neurons_config_data[key]
        if not neurons:
            continue
        out.append(neurons_to_params(neurons))
    return out


def layers_to_params(layers: Sequence[dict]) -> List[Dict[str, Any]]:
    return [neurons_to_params(layer.get("neurons", [])) for layer in layers if layer.get("neurons")]


def params_to_layers(params: Sequence[Dict[str, Any]]) -> List[dict]:
    """Inverse of ``layers_to_params``: the exporter schema of
    Centralized_MNIST_Experimentation.ipynb:462-506 (one neuron per output column,
    ``type`` hidden/output, ``nodes`` = N)."""
    layers = []
    for i, p in enumerate(params):
        w = np.asarray(p["weights"], dtype=np.float64)
        b = np.asarray(p["biases"], dtype=np.float64)
        act = p.get("activation", "linear")
        neurons = [{"weights": w[:, j].tolist(), "bias": float(b[j]), "activation": act} for j in range(w.shape[1])]
        layers.append({"type": "output" if i == len(params) - 1 else "hidden", "nodes": int(w.shape[1]),
                       "neurons": neurons})
    return layers


def make_params(dims: Sequence[int], activations: Sequence[str], seed: int = 1) -> List[Dict[str, Any]]:
    """Seeded, dimensionally consistent weights: W ~ N(0, 2/K) for relu
    layers (He), N(0, 1/K) otherwise, b = 0.01, drawn as float64 from
    ``np.random.default_rng(seed)``.  The shipped sample config is inconsistent
    (SURVEY.md section 0.5), so tests and benchmarks use these."""
    if len(activations) != len(dims) - 1:
        raise ValueError("need one activation per layer")
    rng = np.random.default_rng(seed)
    params = []
    for k, n, act in zip(dims[:-1], dims[1:], activations):
        std = math.sqrt((2.0 if act == "relu" else 1.0) / k)
        params.append({"weights": rng.standard_normal((k, n)) * std, "biases": np.full((n,), 0.01),
                       "activation": act})
    return params


def make_config(dims: Sequence[int], activations: Sequence[str], seed: int = 1,
                layer_distribution: Optional[Sequence[int]] = None, **train_keys) -> Dict[str, Any]:
    """A full model dict in the reference schema (small nets only: ~20 bytes/param as JSON)."""
    cfg = {"layers": params_to_layers(make_params(dims, activations, seed))}
    if layer_distribution is not None:
        cfg["layer_distribution"] = list(layer_distribution)
    cfg.update(train_keys)
    return cfg


# ---------------------------------------------------------------------------
# stage partitioning
# ---------------------------------------------------------------------------

def calculate_layer_mappings(layers_config: List[Dict[str, Any]], distribution: List[int],
                             input_matrix_example: List[Dict[str, Any]]) -> Dict[int, Dict[str, Any]]:
    """Same contract as run_grpc_fcnn.py:176-252: container index -> {container_name,
    listen_port, expected_input, neurons_config, next_nodes}.  The initial input width is
    ``len(examples[0]["input"])``; each later container expects the ``nodes`` of the last
    layer before it; zero-layer containers are skipped and bridged over."""
    if sum(distribution) != len(layers_config):
        raise ValueError("Sum of layer_distribution does not match the number of layers in config.")
    initial = 0
    if input_matrix_example and isinstance(input_matrix_example[0].get("input"), list):
        initial = len(input_matrix_example[0]["input"])
    mappings: Dict[int, Dict[str, Any]] = {}
    g = 0
    cur = initial
    n = len(distribution)
    for c in range(n):
        cnt = distribution[c]
        if cnt == 0:
            continue
        neurons_cfg = {}
        last_out = cur
        for i in range(cnt):
            if g >= len(layers_config):
                raise IndexError("Layer distribution requests more layers than available in config.")
            layer = layers_config[g]
            neurons_cfg[f"layer_{i + 1}"] = layer.get("neurons", [])
            last_out = layer.get("nodes", 0)
            g += 1
        nxt = c + 1
        while nxt < n and distribution[nxt] == 0:
            nxt += 1
        next_nodes = []
        if nxt < n:
            next_nodes = [{"host": f"{CONTAINER_NAME_PREFIX}{nxt}",
                           "port": str(BASE_PORT_FIRST_LAYER + nxt * BASE_PORT_INCREMENT + 1)}]
        mappings[c] = {
            "container_name": f"{CONTAINER_NAME_PREFIX}{c}",
            "listen_port": BASE_PORT_FIRST_LAYER + c * BASE_PORT_INCREMENT + 1,
            "expected_input": cur,
            "neurons_config": neurons_cfg,
            "next_nodes": next_nodes,
        }
        cur = last_out
    return mappings


def stage_layer_ranges(num_layers: int, distribution: Sequence[int]) -> List[Tuple[int, int]]:
    """[first, last) global layer index per NON-EMPTY stage, in pipeline order."""
    if sum(distribution) != num_layers:
        raise ValueError("Sum of layer_distribution does not match the number of layers in config.")
    out, g = [], 0
    for cnt in distribution:
        if cnt:
            out.append((g, g + cnt))
        g += cnt
    return out


def even_distribution(num_layers: int, num_stages: int) -> List[int]:
    """Consecutive layers split as evenly as possible (earlier stages get the remainder)."""
    if num_stages < 1 or num_stages > num_layers:
        raise ValueError(f"cannot place {num_layers} layers on {num_stages} stages")
    base, rem = divmod(num_layers, num_stages)
    return [base + (1 if s < rem else 0) for s in range(num_stages)]


# ---------------------------------------------------------------------------
# inputs + batch ordering
# ---------------------------------------------------------------------------

def load_example_inputs(src) -> List[Dict[str, Any]]:
    """run_grpc_inference.py:35-52."""
    if isinstance(src, dict):
        data = src
    else:
        with open(src, "r") as f:
            data = json.load(f)
    return data.get("examples", [])


def examples_to_arrays(examples: Sequence[Dict[str, Any]]) -> Tuple[np.ndarray, np.ndarray]:
    """Stacks ``input`` rows (float64, like ``np.array(list-of-lists)`` at grpc_node.py:113)
    and labels (int64; missing labels become -1)."""
    x = np.array([ex.get("input") for ex in examples], dtype=np.float64)
    y = np.array([(-1 if ex.get("label") is None else ex.get("label")) for ex in examples], dtype=np.int64)
    return x, y


def batch_slices(num_examples: int, batch_size: Optional[int]) -> List[Tuple[int, int]]:
    """The only batch order the reference has (run_grpc_inference.py:183-205): sequential,
    unshuffled ``[b*bs, min((b+1)*bs, N))`` chunks, ``ceil(N/bs)`` of them, ragged last batch
    kept; a single batch when ``batch_size`` is None or >= N."""
    if num_examples <= 0:
        return []
    if batch_size is None or batch_size >= num_examples:
        return [(0, num_examples)]
    nb = math.ceil(num_examples / batch_size)
    return [(b * batch_size, min(b * batch_size + batch_size, num_examples)) for b in range(nb)]


def micro_batch_slices(rows: int, micro_batches: int) -> List[Tuple[int, int]]:
    """Splits one batch into ``micro_batches`` consecutive row ranges (the first ones take
    the remainder); empty ranges are dropped."""
    m = max(1, min(micro_batches, rows)) if rows > 0 else 1
    base, rem = divmod(rows, m)
    out, r = [], 0
    for i in range(m):
        n = base + (1 if i < rem else 0)
        if n:
            out.append((r, r + n))
        r += n
    return out


def synthetic_batch(rows: int, in_dim: int, num_classes: int, seed: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """SURVEY.md section 8d synthetic data: X ~ U[0,1) (the /255 scaling of
    generate_mnist_pytorch.py:10), labels ~ U{0..C-1}, ``default_rng(seed)``."""
    rng = np.random.default_rng(seed)
    x = rng.random((rows, in_dim))
    y = rng.integers(0, num_classes, size=(rows,))
    return x, y
