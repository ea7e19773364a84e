"""CPU oracle for the docker-dist-nn hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker / the timed CPU baseline.
The product package (``docker-dist-nn_b200``) never imports it and fails loudly when
its CUDA library is missing.
"""
from .ddnn_oracle import (  # noqa: F401
    activate,
    build_layer_params,
    process_matrix,
    forward_all,
    softmax_ce,
    backward,
    sgd_step,
    train_step,
    batch_slices,
    stage_partition,
)
