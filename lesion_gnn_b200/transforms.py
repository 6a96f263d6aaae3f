"""On-device graph construction for a whole collated batch: the work the reference does per sample on the CPU
inside Dataset.__getitem__ (torch_geometric.transforms.KNNGraph via reference transforms.py:19-23,
configs/config.py:47, datasets/datamodule.py:43-48) and the reference's own GaussianDistance transform
(transforms.py:32-79)."""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor

from . import ops
from .graph import as_ptr
from .nn.pool import Segments


def knn_graph(pos: Tensor, k: int, batch=None, loop: bool = False, flow: str = "source_to_target",
              num_edges: Optional[int] = None) -> Tensor:
    """torch_cluster.knn_graph semantics: [2, E] int64, row 0 = neighbour (source), row 1 = centre (target),
    edges grouped by centre in ascending distance; ties to the lower index."""
    if flow != "source_to_target":
        raise NotImplementedError("only flow='source_to_target' is on the reference's path")
    if not pos.is_cuda:
        raise RuntimeError("pos must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    if pos.dim() == 1:
        pos = pos.unsqueeze(-1)
    pos = pos.contiguous()
    if isinstance(batch, Segments):
        ptr = batch.ptr
    elif batch is None:
        ptr = torch.tensor([0, pos.size(0)], dtype=torch.int32, device=pos.device)
    else:
        ptr = as_ptr(batch)[0]
    return ops.knn_graph_raw(pos, ptr, k, loop, num_edges)


def radius_graph(pos: Tensor, r: float, batch=None, loop: bool = False, max_num_neighbors: int = 32,
                 flow: str = "source_to_target") -> Tensor:
    """torch_cluster.radius_graph semantics (torch_geometric.transforms.RadiusGraph; reference scripts/sweep.py:113-117):
    [2, E] int64, row 0 = neighbour (source), row 1 = centre (target); neighbours with |p_j - p_i| < r in ascending
    index order, at most `max_num_neighbors` per centre."""
    if flow != "source_to_target":
        raise NotImplementedError("only flow='source_to_target' is on the reference's path")
    if not pos.is_cuda:
        raise RuntimeError("pos must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    if pos.dim() == 1:
        pos = pos.unsqueeze(-1)
    pos = pos.contiguous()
    if isinstance(batch, Segments):
        ptr = batch.ptr
    elif batch is None:
        ptr = torch.tensor([0, pos.size(0)], dtype=torch.int32, device=pos.device)
    else:
        ptr = as_ptr(batch)[0]
    return ops.radius_graph_raw(pos, ptr, r, loop, max_num_neighbors)


def knn_num_edges(counts, k: int, loop: bool) -> int:
    """E of knn_graph from host-side graph sizes (avoids the device->host read)."""
    total = 0
    for n in counts:
        n = int(n)
        total += n * max(min(k, n) if loop else min(k, n - 1), 0)
    return total


class GaussianDistance:
    """reference transforms.py:32-79, EDGE_WEIGHT_REPLACE mode: edge_weight = exp(-d^2/2s^2)/sqrt(2 pi s^2)."""

    def __init__(self, sigma: float, dtype: torch.dtype = torch.float32):
        self.sigma, self.dtype = sigma, dtype

    def __call__(self, pos: Tensor, edge_index: Tensor) -> Tensor:
        if edge_index.numel() == 0:
            return torch.empty(0, dtype=self.dtype, device=pos.device)
        return ops.gaussian_distance(pos.contiguous(), edge_index.contiguous(), self.sigma).to(self.dtype)
