"""docker-dist-nn_b200 -- the docker-dist-nn layer hot path (dense forward / backward /
softmax-CE / SGD) as hand-written sm_100a kernels behind a C ABI, with the reference's
``Layer`` API and JSON schemas on top.

The directory name is not a Python identifier; ``import ddnn_b200`` (repo-root shim) loads
this package under that name.  ``config`` is importable without a GPU; everything that
computes imports ``_lib`` and therefore needs libddnn.so (no CPU fallback).
"""
from . import config  # noqa: F401  (pure host code, always importable)

__all__ = ["config", "Layer", "Trainer", "PipelineStage", "ops"]


def __getattr__(name):  # lazy: keep `import ddnn_b200.config` free of torch/CUDA
    if name == "Layer":
        from .layer import Layer
        return Layer
    if name == "Trainer":
        from .trainer import Trainer
        return Trainer
    if name == "PipelineStage":
        from .pipeline import PipelineStage
        return PipelineStage
    if name in ("ops", "_lib", "layer", "trainer", "pipeline", "build", "infer", "grpc_shim"):
        import importlib
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
