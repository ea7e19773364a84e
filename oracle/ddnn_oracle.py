"""Float64 numpy restatement of the docker-dist-nn layer hot path.  TEST INFRASTRUCTURE ONLY.

Pinning status
--------------
* forward (``build_layer_params``, ``activate``, ``process_matrix``): PINNED.  The
  reference ships no tests or golden vectors, so the pin is the
  reference itself: ``tests/golden/make_golden.py`` imports the unmodified
  ``/root/reference/src/grpc_node.py`` ``Layer`` in the build container, runs it on
  seeded inputs and commits the outputs under ``tests/golden/``;
  ``tests/test_oracle.py`` checks this file against those vectors bit for bit.
* ``stage_partition``: PINNED the same way against
  ``/root/reference/src/run_grpc_fcnn.py:176-252`` (``calculate_layer_mappings``).
* ``batch_slices``: restates ``/root/reference/src/run_grpc_inference.py:198-205``.
* softmax-CE / backward / SGD: **parity unpinned** by the reference -- the reference
  has no loss, no gradients and no weight update anywhere in ``src/``.
  These follow from the forward definition by calculus, use the loss semantics of
  ``/root/reference/scripts/generate_mnist_pytorch.py:37`` (mean-reduced
  cross-entropy on integer labels) and plain SGD as BASELINE.json's north_star
  demands; ``tests/test_oracle.py`` cross-checks them against torch autograd in
  float64 on CPU.

All arithmetic is float64, like the reference (JSON floats -> ``np.array``).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

LayerParams = Dict[str, object]


# ---------------------------------------------------------------------------
# forward -- follows /root/reference/src/grpc_node.py
# ---------------------------------------------------------------------------

def build_layer_params(neurons_config_data: Dict[str, list]) -> List[LayerParams]:
    """grpc_node.py:43-55.  Keys ``layer_<n>`` sorted by integer suffix, empty layers
    skipped, ``weights`` = per-neuron rows transposed to (in, out), activation taken
    from the FIRST neuron, default ``linear``."""
    params: List[LayerParams] = []
    for key in sorted(neurons_config_data, key=lambda x: int(x.split("_")[1])):
        neurons = neurons_config_data[key]
        if not neurons:
            continue
        w = np.array([n["weights"] for n in neurons], dtype=np.float64).T
        b = np.array([n["bias"] for n in neurons], dtype=np.float64)
        act = neurons[0].get("activation", "linear")
        params.append({"weights": w, "biases": b, "activation": act})
    return params


def activate(values: np.ndarray, func: str) -> np.ndarray:
    """grpc_node.py:62-73.  relu / sigmoid (not overflow-guarded, like the reference) /
    row softmax with max subtraction; every other string is the identity."""
    if func == "relu":
        return np.maximum(0, values)
    if func == "sigmoid":
        return 1 / (1 + np.exp(-values))
    if func == "softmax":
        e = np.exp(values - np.max(values, axis=-1, keepdims=True))
        return e / np.sum(e, axis=-1, keepdims=True)
    return values


def process_matrix(layer_params: Sequence[LayerParams], input_matrix: np.ndarray,
                   expected_input_dim: int, container_name: str = "layer") -> np.ndarray:
    """grpc_node.py:75-97.  Chains ``Z = X.W + b`` -> ``activate`` over the node's
    layers; raises the reference's ValueError text on an input-width mismatch."""
    cur = input_matrix
    expected = expected_input_dim
    for i, p in enumerate(layer_params):
        if cur.shape[1] != expected:
            raise ValueError(
                f"({container_name}) Layer {i + 1}: expected input dim {expected}, got {cur.shape[1]}")
        z = np.dot(cur, p["weights"]) + p["biases"]
        cur = activate(z, p["activation"])
        expected = cur.shape[1]
    return cur


def forward_all(layer_params: Sequence[LayerParams], x: np.ndarray) -> Tuple[List[np.ndarray], List[np.ndarray]]:
    """Same arithmetic as ``process_matrix`` but keeps every layer's input-side
    activation ``A[l]`` (A[0] = x) and pre-activation ``Z[l]`` for the backward pass."""
    acts = [np.asarray(x, dtype=np.float64)]
    zs = []
    for p in layer_params:
        z = np.dot(acts[-1], p["weights"]) + p["biases"]
        zs.append(z)
        acts.append(activate(z, p["activation"]))
    return acts, zs


# ---------------------------------------------------------------------------
# loss / backward / update -- no reference implementation (see module docstring)
# ---------------------------------------------------------------------------

def softmax_ce(z_last: np.ndarray, labels: np.ndarray, denom: Optional[int] = None
               ) -> Tuple[float, np.ndarray, np.ndarray]:
    """Mean softmax cross-entropy on the output node's pre-activation ``z_last``.

    Returns (loss, p, dZ) with ``p`` = the node's softmax output (grpc_node.py:68-71),
    ``loss = -sum_i log p[i, y_i] / denom`` and ``dZ = (p - onehot(y)) / denom``;
    ``denom`` defaults to the number of rows (generate_mnist_pytorch.py:37,
    ``nn.CrossEntropyLoss`` default ``reduction='mean'``).  A pipeline passes the
    GLOBAL batch size as ``denom`` for every micro-batch so the sum over
    micro-batches equals the full-batch mean."""
    z = np.asarray(z_last, dtype=np.float64)
    n = z.shape[0]
    denom = n if denom is None else denom
    m = np.max(z, axis=-1, keepdims=True)
    e = np.exp(z - m)
    s = np.sum(e, axis=-1, keepdims=True)
    p = e / s
    rows = np.arange(n)
    logp = (z[rows, labels] - m[:, 0]) - np.log(s[:, 0])
    loss = float(-np.sum(logp) / denom) if n else 0.0
    dz = p.copy()
    dz[rows, labels] -= 1.0
    dz /= denom
    return loss, p, dz


def _act_backward(d_a: np.ndarray, a: np.ndarray, z: np.ndarray, func: str) -> np.ndarray:
    """dZ from dA through the derivative of grpc_node.py:62-73's activations.
    relu: 1[Z>0] (0 at Z==0, matching np.maximum(0,.) subgradient used by torch);
    sigmoid: A(1-A); softmax (hidden): A*(dA - sum(dA*A)); anything else: identity."""
    if func == "relu":
        return d_a * (z > 0)
    if func == "sigmoid":
        return d_a * a * (1.0 - a)
    if func == "softmax":
        return a * (d_a - np.sum(d_a * a, axis=-1, keepdims=True))
    return d_a


def backward(layer_params: Sequence[LayerParams], acts: Sequence[np.ndarray], zs: Sequence[np.ndarray],
             labels: np.ndarray, denom: Optional[int] = None):
    """Gradients of the mean softmax-CE wrt every layer's W and b.

    The output layer must be ``softmax`` (the loss is -log of the node's output) or a
    linear/unknown activation (its output is then treated as logits, the torch script's
    convention).  Returns (loss, [dW_l (K,N)], [db_l (N,)], dX of layer 1's input)."""
    last = layer_params[-1]["activation"]
    if last in ("relu", "sigmoid"):
        raise ValueError(f"output activation {last!r} has no softmax-CE loss")
    loss, _, dz = softmax_ce(zs[-1], labels, denom)
    d_ws: List[np.ndarray] = [None] * len(layer_params)
    d_bs: List[np.ndarray] = [None] * len(layer_params)
    for l in range(len(layer_params) - 1, -1, -1):
        d_ws[l] = np.dot(acts[l].T, dz)
        d_bs[l] = np.sum(dz, axis=0)
        d_a = np.dot(dz, layer_params[l]["weights"].T)
        if l > 0:
            dz = _act_backward(d_a, acts[l], zs[l - 1], layer_params[l - 1]["activation"])
        else:
            dz = d_a
    return loss, d_ws, d_bs, dz


def sgd_step(layer_params: Sequence[LayerParams], d_ws, d_bs, lr: float) -> None:
    """Plain SGD in place: W <- W - lr*dW, b <- b - lr*db (no momentum / decay)."""
    for p, dw, db in zip(layer_params, d_ws, d_bs):
        p["weights"] = p["weights"] - lr * dw
        p["biases"] = p["biases"] - lr * db


def train_step(layer_params: Sequence[LayerParams], x: np.ndarray, labels: np.ndarray, lr: float) -> float:
    """One full-batch step: forward, softmax-CE, backward, SGD.  Returns the loss
    measured BEFORE the update."""
    acts, zs = forward_all(layer_params, x)
    loss, d_ws, d_bs, _ = backward(layer_params, acts, zs, labels)
    sgd_step(layer_params, d_ws, d_bs, lr)
    return loss


# ---------------------------------------------------------------------------
# host-side conventions
# ---------------------------------------------------------------------------

def batch_slices(num_examples: int, batch_size: Optional[int]) -> List[Tuple[int, int]]:
    """run_grpc_inference.py:183-205: sequential, unshuffled [start, end) chunks,
    ceil(N/bs) of them, ragged last batch kept; one batch when bs is None or >= N."""
    if num_examples <= 0:
        return []
    if batch_size is None or batch_size >= num_examples:
        return [(0, num_examples)]
    nb = math.ceil(num_examples / batch_size)
    return [(b * batch_size, min(b * batch_size + batch_size, num_examples)) for b in range(nb)]


def stage_partition(layers_config: Sequence[dict], distribution: Sequence[int], initial_input_dim: int
                    ) -> Dict[int, dict]:
    """run_grpc_fcnn.py:176-252 without the Docker naming/ports: container index ->
    {expected_input, neurons_config {layer_<i>: neurons}, next (index of the next
    non-empty container or None)}.  Zero-layer containers are skipped; the expected
    input of a container is the ``nodes`` field of the last layer before it."""
    if sum(distribution) != len(layers_config):
        raise ValueError("Sum of layer_distribution does not match the number of layers in config.")
    out: Dict[int, dict] = {}
    g = 0
    cur_dim = initial_input_dim
    n = len(distribution)
    for c in range(n):
        cnt = distribution[c]
        if cnt == 0:
            continue
        cfg = {}
        last_out = cur_dim
        for i in range(cnt):
            layer = layers_config[g]
            cfg[f"layer_{i + 1}"] = layer.get("neurons", [])
            last_out = layer.get("nodes", 0)
            g += 1
        nxt = None
        j = c + 1
        while j < n and distribution[j] == 0:
            j += 1
        if j < n:
            nxt = j
        out[c] = {"expected_input": cur_dim, "neurons_config": cfg, "next": nxt}
        cur_dim = last_out
    return out
