"""Host logic: schemas, stage partitioning (vs the reference's calculate_layer_mappings
outputs in tests/golden), batch ordering."""
import json
import os

import numpy as np
import pytest

from ddnn_b200 import config as C

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# verbatim structure of /root/reference/config/config_sample.json (inconsistent dims on purpose,
# SURVEY.md section 0.5) -- it must PARSE, even though it cannot run
SAMPLE = {"layers": [
    {"type": "hidden", "nodes": 3, "neurons": [{"weights": [0.1, 0.2], "bias": 0.3, "activation": "relu"},
                                                {"weights": [0.4, 0.5], "bias": 0.6, "activation": "relu"},
                                                {"weights": [0.7, 0.8], "bias": 0.9, "activation": "relu"}]},
    {"type": "output", "nodes": 2, "neurons": [{"weights": [1.0, 1.1, 1.2], "bias": 0.5, "activation": "softmax"},
                                                {"weights": [1.5, 1.3, 1.1], "bias": 0.8, "activation": "softmax"}]}]}


def test_load_defaults_and_extra_keys():
    cfg = C.load_model_config(SAMPLE)
    assert cfg["learning_rate"] == 0.01 and cfg["epochs"] == 1 and cfg["micro_batches"] == 1
    assert cfg["layer_distribution"] == [1, 1]
    cfg2 = C.load_model_config(json.dumps({**SAMPLE, "learning_rate": 0.5, "layer_distribution": [2]}))
    assert cfg2["learning_rate"] == 0.5 and cfg2["layer_distribution"] == [2]
    nested = C.load_model_config({"model": SAMPLE, "inference_metrics": {}})
    assert len(nested["layers"]) == 2
    with pytest.raises(KeyError):
        C.load_model_config({"nope": 1})


def test_params_layout_and_roundtrip():
    p = C.layers_to_params(SAMPLE["layers"])
    assert p[0]["weights"].shape == (2, 3) and p[0]["weights"][1, 2] == 0.8  # (in, out) like grpc_node.py:51
    assert p[1]["activation"] == "softmax" and p[1]["biases"].tolist() == [0.5, 0.8]
    back = C.params_to_layers(p)
    assert back[0]["neurons"][2]["weights"] == [0.7, 0.8] and back[1]["type"] == "output" and back[1]["nodes"] == 2
    again = C.layers_to_params(back)
    for a, b in zip(p, again):
        assert np.array_equal(a["weights"], b["weights"]) and np.array_equal(a["biases"], b["biases"])


def test_mappings_match_reference(golden_small):
    m = golden_small["mappings"]
    layers = C.params_to_layers(C.make_params(m["dims"], ["relu"] * 5 + ["softmax"], seed=m["seed"]))
    for dist_s, expect in m["by_distribution"].items():
        got = C.calculate_layer_mappings(layers, json.loads(dist_s), m["examples"])
        slim = {str(k): {"container_name": v["container_name"], "listen_port": v["listen_port"],
                         "expected_input": v["expected_input"], "next_nodes": v["next_nodes"],
                         "layer_keys": {lk: len(lv) for lk, lv in v["neurons_config"].items()}}
                for k, v in got.items()}
        assert slim == expect, dist_s
    with pytest.raises(ValueError) as e:
        C.calculate_layer_mappings(layers, [1], m["examples"])
    assert str(e.value) == m["bad_sum_message"]


def test_stage_ranges_and_even_distribution():
    assert C.stage_layer_ranges(6, [3, 0, 3]) == [(0, 3), (3, 6)]
    assert C.even_distribution(8, 8) == [1] * 8 and C.even_distribution(8, 3) == [3, 3, 2]
    assert C.even_distribution(4, 2) == [2, 2]
    with pytest.raises(ValueError):
        C.even_distribution(2, 3)
    from ddnn_b200.pipeline import stage_assignment
    assert stage_assignment(4, 2, [2, 2]) == [(0, 2), (2, 4)]
    assert stage_assignment(4, 2, [1, 1, 1, 1]) == [(0, 2), (2, 4)]  # wrong stage count -> even split
    assert stage_assignment(4, 2, [3, 0, 1]) == [(0, 3), (3, 4)]


def test_batch_and_micro_batch_slices():
    assert C.batch_slices(10, 4) == [(0, 4), (4, 8), (8, 10)]
    assert C.batch_slices(3, None) == [(0, 3)] and C.batch_slices(0, 2) == []
    assert C.micro_batch_slices(10, 4) == [(0, 3), (3, 6), (6, 8), (8, 10)]
    assert C.micro_batch_slices(2, 8) == [(0, 1), (1, 2)]
    assert C.micro_batch_slices(8, 1) == [(0, 8)]


def test_examples_and_synthetic():
    ex = [{"input": [0.1, 0.2], "label": 1}, {"input": [0.3, 0.4]}]
    x, y = C.examples_to_arrays(ex)
    assert x.dtype == np.float64 and x.shape == (2, 2) and y.tolist() == [1, -1]
    x, y = C.synthetic_batch(5, 3, 4, seed=0)
    x2, _ = C.synthetic_batch(5, 3, 4, seed=0)
    assert np.array_equal(x, x2) and x.min() >= 0 and x.max() < 1 and y.max() < 4
