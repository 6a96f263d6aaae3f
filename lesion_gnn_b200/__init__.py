"""lesion_gnn_b200: the message-passing hot path of zacharielegault/lesion-gnn on NVIDIA B200 (sm_100a).

PyG-compatible layers (`lesion_gnn_b200.nn`), the reference's models on top of them (`.models`), on-device graph
construction (`.transforms`, `.graph`) and the training step (`.train`), all calling hand-written CUDA kernels
through the C ABI of include/lgk.h (liblgk.so).  CUDA only -- there is no CPU or PyTorch fallback.
"""
from . import _lib, data, graph, models, nn, ops, train, transforms  # noqa: F401
from .graph import CSRGraph

__version__ = "0.1.0"
__all__ = ["CSRGraph", "nn", "models", "ops", "graph", "transforms", "train", "data"]
