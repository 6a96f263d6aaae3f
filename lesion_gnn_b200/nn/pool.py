"""Global pooling with PyG's signatures: global_{mean,max,add}_pool(x, batch, size=None)
(reference models/gat.py:56, gin.py:33, pointnet.py:38) plus the fused mean||max read-out of BASELINE.json
config 3.  `batch` is the sorted int64 graph-id vector of PyG collation, or a `Segments` object that already
carries the graph pointer (no device->host read at all)."""
from __future__ import annotations

from typing import Optional, Union

import torch
from torch import Tensor

from .. import compile as _compile
from .. import ops
from ..graph import as_ptr
from . import functional as F


class Segments:
    """Graph segmentation of a collated batch: int32 ptr[B+1] on the device and B on the host."""

    __slots__ = ("ptr", "num_graphs", "batch")

    def __init__(self, ptr: Tensor, num_graphs: int, batch: Optional[Tensor] = None):
        self.ptr, self.num_graphs, self.batch = ptr, int(num_graphs), batch

    @staticmethod
    def from_batch(batch: Tensor, num_graphs: int) -> "Segments":
        b = batch if batch.dtype == torch.int64 else batch.long()
        return Segments(ops.batch_ptr(b.contiguous(), num_graphs), num_graphs, batch)


def _ptr(batch: Union[Tensor, Segments, None], n: int, size: Optional[int], device) -> Tensor:
    if isinstance(batch, Segments):
        return batch.ptr
    if batch is None:  # PyG: a single graph
        return torch.tensor([0, n], dtype=torch.int32, device=device)
    return as_ptr(batch, size)[0]


def _pool_eager(x: Tensor, batch, size, mode: str) -> Tensor:
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    squeeze = x.dim() == 1
    x2 = x.unsqueeze(-1) if squeeze else x
    out = F.segment_pool(x2, _ptr(batch, x2.size(0), size, x.device), mode)
    return out.squeeze(-1) if squeeze else out


def _pool_fake(mode: str):
    def fake(x, batch, num_graphs):
        if x.dim() == 1:
            return x.new_empty((num_graphs,))
        return x.new_empty((num_graphs, x.size(1) * (2 if mode == "mean+max" else 1)))
    return fake


# while torch.compile traces a model, a pooling call is ONE opaque operator (lesion_gnn_b200/compile.py); the number of
# graphs is an integer argument -- without `size` it is read from the data first (`batch[-1] + 1`, a graph break exactly as
# PyG's own `int(batch.max()) + 1`)
_POOL_OPS = {mode: _compile.opaque_function(
    (lambda mode: lambda x, batch, num_graphs: _pool_eager(x, batch, int(num_graphs), mode))(mode), _pool_fake(mode))
    for mode in ("sum", "mean", "max", "mean+max")}


def _pool(x: Tensor, batch, size, mode: str) -> Tensor:
    if _compile.is_compiling() and isinstance(batch, Tensor):
        if size is None:
            size = int(batch[-1].item()) + 1
        return _POOL_OPS[mode](x, batch, ints=[size])
    return _pool_eager(x, batch, size, mode)


def global_add_pool(x: Tensor, batch, size: Optional[int] = None) -> Tensor:
    return _pool(x, batch, size, "sum")


def global_mean_pool(x: Tensor, batch, size: Optional[int] = None) -> Tensor:
    return _pool(x, batch, size, "mean")


def global_max_pool(x: Tensor, batch, size: Optional[int] = None) -> Tensor:
    return _pool(x, batch, size, "max")


def global_mean_max_pool(x: Tensor, batch, size: Optional[int] = None) -> Tensor:
    """cat([global_mean_pool, global_max_pool], dim=1) in one pass over x"""
    return _pool(x, batch, size, "mean+max")


class SortAggregation(_compile.OpaqueModule, torch.nn.Module):
    """torch_geometric.nn.aggr.SortAggregation(k) as the reference calls it (models/drgnet.py:37,59):
    `sort_pool(x, batch)` -> [B, k * F], nodes of each graph ordered by their last channel (descending, stable),
    first k kept, zero padded.  One lgk_sort_pool_fwd launch instead of to_dense_batch + sort + gather."""

    def __init__(self, k: int):
        super().__init__()
        self.k = int(k)
        self._register_opaque()

    def forward(self, x: Tensor, index=None, ptr: Optional[Tensor] = None, dim_size: Optional[int] = None, dim: int = -2) -> Tensor:
        if _compile.is_compiling() and isinstance(index, Tensor) and ptr is None and dim in (-2, 0):
            if dim_size is None:
                dim_size = int(index[-1].item()) + 1
            return self._dispatch(x, index, ints=[dim_size])
        return self._forward_eager(x, index, dim_size, ptr, dim)

    def _fake(self, x: Tensor, index, dim_size) -> Tensor:
        return x.new_empty((dim_size, self.k * x.size(1)))

    def _forward_eager(self, x: Tensor, index=None, dim_size: Optional[int] = None, ptr: Optional[Tensor] = None, dim: int = -2) -> Tensor:
        if not x.is_cuda:
            raise RuntimeError("x must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
        if dim not in (-2, 0) or x.dim() != 2:
            raise NotImplementedError("SortAggregation is implemented for [N, F] node features")
        return F.sort_pool(x, _ptr(index, x.size(0), dim_size, x.device), self.k)

    def __repr__(self) -> str:
        return f"SortAggregation(k={self.k})"
