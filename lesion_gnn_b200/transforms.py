"""On-device graph construction for a whole collated batch: the work the reference does per sample on the CPU
inside Dataset.__getitem__ (torch_geometric.transforms.KNNGraph via reference transforms.py:19-23,
configs/config.py:47, datasets/datamodule.py:43-48) and the reference's own GaussianDistance transform
(transforms.py:32-79)."""
from __future__ import annotations

import dataclasses
import warnings
from enum import Enum
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


# ---------------------------------------------------------------------------------------------------------------
# Loader-level transform API: `get_transform(config)(data) -> data` exactly as reference
# transforms.py:19-23 builds it and datasets/datamodule.py:43-48 applies it -- but on a whole collated batch that is
# already on the device, instead of per sample on the CPU inside Dataset.__getitem__.
# ---------------------------------------------------------------------------------------------------------------


@dataclasses.dataclass(kw_only=True)
class TransformConfig:
    """reference transforms.py:13-16"""
    name: str
    kwargs: dict = dataclasses.field(default_factory=dict)


class SaveAs(Enum):
    """reference transforms.py:26-29"""
    EDGE_WEIGHT_REPLACE = "edge_weight_replace"
    EDGE_ATTR_CAT = "edge_attr_cat"
    EDGE_ATTR_REPLACE = "edge_attr_replace"


class Data:
    """The attributes of torch_geometric.data.Data / Batch that the reference's transforms and models touch
    (x, pos, edge_index, edge_attr, edge_weight, y, batch, adj_t; reference datasets/nodes/lesions.py:170-176,
    models/gat.py:86-92).  `batch` / `num_graphs` describe a collated batch; a single graph leaves them None."""

    _FIELDS = ("x", "pos", "edge_index", "edge_attr", "edge_weight", "y", "batch", "adj_t")

    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, edge_weight=None, batch=None,
                 num_graphs: Optional[int] = None, counts=None, adj_t=None):
        self.x, self.edge_index, self.edge_attr, self.y, self.pos = x, edge_index, edge_attr, y, pos
        self.edge_weight, self.batch, self.adj_t = edge_weight, batch, adj_t
        self.num_graphs = num_graphs
        self.counts = None if counts is None else list(counts)  # host-side graph sizes (avoids device->host reads)
        self.edges_grouped_by_dst = False  # set by KNNGraph / RadiusGraph: ToSparseTensor then takes the sorted-input build

    @property
    def num_nodes(self) -> int:
        for t in (self.x, self.pos):
            if t is not None:
                return int(t.size(0))
        raise ValueError("Data holds neither x nor pos")

    def clone(self) -> "Data":
        d = Data(num_graphs=self.num_graphs, counts=self.counts)
        for f in self._FIELDS:
            v = getattr(self, f)
            setattr(d, f, v.clone() if isinstance(v, Tensor) else v)
        d.edges_grouped_by_dst = getattr(self, "edges_grouped_by_dst", False)
        return d

    def to(self, device, non_blocking: bool = False) -> "Data":
        d = Data(num_graphs=self.num_graphs, counts=self.counts)
        for f in self._FIELDS:
            v = getattr(self, f)
            setattr(d, f, v.to(device, non_blocking=non_blocking) if isinstance(v, Tensor) else v)
        d.edges_grouped_by_dst = getattr(self, "edges_grouped_by_dst", False)
        return d

    @staticmethod
    def from_host_batch(hb, device=None) -> "Data":
        """a collated data.HostBatch as a Data object (optionally moved to `device`)"""
        d = Data(x=hb.x, pos=hb.pos, y=hb.y, batch=hb.batch, edge_index=hb.edge_index, edge_weight=hb.edge_weight,
                 num_graphs=hb.num_graphs, counts=hb.counts)
        d.edges_grouped_by_dst = bool(getattr(hb, "dst_sorted", False))
        return d if device is None else d.to(device)

    def segments(self) -> Optional[Segments]:
        if self.batch is None:
            return None
        if self.num_graphs is None:
            return Segments(as_ptr(self.batch)[0], as_ptr(self.batch)[1], self.batch)
        return Segments.from_batch(self.batch, self.num_graphs)


class BaseTransform:
    def __call__(self, data: Data) -> Data:  # pragma: no cover
        raise NotImplementedError

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}()"


class Compose(BaseTransform):
    """torch_geometric.transforms.Compose (reference datasets/datamodule.py:46)"""

    def __init__(self, transforms):
        self.transforms = list(transforms)

    def __call__(self, data: Data) -> Data:
        for t in self.transforms:
            data = t(data)
        return data


class KNNGraph(BaseTransform):
    """torch_geometric.transforms.KNNGraph(k=6, loop=False, force_undirected=False, flow='source_to_target') --
    reference configs/config.py:47 `TransformConfig("KNNGraph", {"k": 6, "loop": True})`: sets data.edge_index from
    data.pos (per graph of data.batch) and clears data.edge_attr; one lgk_knn_graph launch for the whole batch."""

    def __init__(self, k: int = 6, loop: bool = False, force_undirected: bool = False, flow: str = "source_to_target",
                 cosine: bool = False, num_workers: int = 1):
        if force_undirected or cosine:
            raise NotImplementedError("force_undirected / cosine are not on the reference's path")
        self.k, self.loop, self.flow = int(k), bool(loop), flow

    def __call__(self, data: Data) -> Data:
        data.edge_attr = None
        num_edges = knn_num_edges(data.counts, self.k, self.loop) if data.counts is not None else None
        data.edge_index = knn_graph(data.pos, self.k, data.segments(), loop=self.loop, flow=self.flow, num_edges=num_edges)
        data.edges_grouped_by_dst = self.flow == "source_to_target"  # lgk_knn_graph emits the k neighbours of node 0, 1, 2, ...
        return data

    def __repr__(self) -> str:
        return f"KNNGraph(k={self.k})"


class RadiusGraph(BaseTransform):
    """torch_geometric.transforms.RadiusGraph(r, loop=False, max_num_neighbors=32) (reference scripts/sweep.py:113-117)"""

    def __init__(self, r: float, loop: bool = False, max_num_neighbors: int = 32, flow: str = "source_to_target",
                 num_workers: int = 1):
        self.r, self.loop, self.max_num_neighbors, self.flow = float(r), bool(loop), int(max_num_neighbors), flow

    def __call__(self, data: Data) -> Data:
        data.edge_attr = None
        data.edge_index = radius_graph(data.pos, self.r, data.segments(), loop=self.loop,
                                       max_num_neighbors=self.max_num_neighbors, flow=self.flow)
        data.edges_grouped_by_dst = self.flow == "source_to_target"
        return data

    def __repr__(self) -> str:
        return f"RadiusGraph(r={self.r})"


class ToSparseTensor(BaseTransform):
    """torch_geometric.transforms.ToSparseTensor() as reference datasets/datamodule.py:45 appends it when compile=False:
    data.adj_t = the destination-sorted CSR (+ CSC) of the batch with edge_weight as its values; edge_index and
    edge_weight are removed (models read `data.adj_t`, models/gat.py:87-88)."""

    def __init__(self, attr: Optional[str] = "edge_weight", remove_edge_index: bool = True):
        self.attr, self.remove_edge_index = attr, remove_edge_index

    def __call__(self, data: Data) -> Data:
        from .graph import CSRGraph
        value = getattr(data, self.attr) if self.attr else None
        # edges made by KNNGraph / RadiusGraph on the device are grouped by destination: the sorted-input build applies
        data.adj_t = CSRGraph.from_edge_index(data.edge_index, data.num_nodes, value,
                                              dst_sorted=bool(getattr(data, "edges_grouped_by_dst", False)))
        if self.remove_edge_index:
            data.edge_index = None
            if self.attr:
                setattr(data, self.attr, None)
        return data


class GaussianDistance(BaseTransform):
    """reference transforms.py:32-79: edge weight exp(-d^2 / 2 sigma^2) / sqrt(2 pi sigma^2) of every edge, saved as
    data.edge_weight (EDGE_WEIGHT_REPLACE), appended to data.edge_attr (EDGE_ATTR_CAT) or replacing it
    (EDGE_ATTR_REPLACE).  A graph without edges warns and is returned unchanged (transforms.py:51-53).
    Also callable as `GaussianDistance(sigma)(pos, edge_index) -> weights` (tensor form)."""

    def __init__(self, sigma: float, save_as: SaveAs = SaveAs.EDGE_WEIGHT_REPLACE, dtype: torch.dtype = torch.float32):
        self.sigma, self.dtype = sigma, dtype
        self.save_as = SaveAs(save_as) if not isinstance(save_as, SaveAs) else save_as

    def weights(self, pos: Tensor, edge_index: Tensor) -> Tensor:
        if edge_index.numel() == 0:
            return torch.empty(0, dtype=self.dtype, device=pos.device)
        return ops.gaussian_distance(pos.contiguous(), edge_index.contiguous(), self.sigma).to(self.dtype)

    def __call__(self, data, edge_index: Optional[Tensor] = None):
        if isinstance(data, Tensor):  # tensor form: (pos, edge_index)
            return self.weights(data, edge_index)
        if data.edge_index is None or data.edge_index.numel() == 0:
            warnings.warn("The graph has no edges, returning the original data object.")
            return data
        dist = self.weights(data.pos, data.edge_index)
        pseudo = data.edge_attr
        if self.save_as is SaveAs.EDGE_WEIGHT_REPLACE:
            data.edge_weight = dist
        elif self.save_as is SaveAs.EDGE_ATTR_CAT:
            dist = dist.view(-1, 1)
            if pseudo is not None:
                pseudo = pseudo.view(-1, 1) if pseudo.dim() == 1 else pseudo
                data.edge_attr = torch.cat([pseudo, dist.type_as(pseudo)], dim=-1)
            else:
                data.edge_attr = dist
        else:
            data.edge_attr = dist.view(-1, 1)
        return data

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}(sigma={self.sigma})"


_TRANSFORMS = {"KNNGraph": KNNGraph, "RadiusGraph": RadiusGraph, "ToSparseTensor": ToSparseTensor,
               "GaussianDistance": GaussianDistance, "Compose": Compose}


def get_transform(config, **kwargs) -> BaseTransform:
    """reference transforms.py:19-23: `get_transform(TransformConfig(name, kwargs))`; a bare name with keyword arguments
    is accepted too.  Names are those of torch_geometric.transforms the reference's configs and sweep use."""
    if isinstance(config, str):
        name, kw = config, kwargs
    else:
        name, kw = config.name, dict(config.kwargs, **kwargs)
    if name not in _TRANSFORMS:
        raise NotImplementedError(f"transform {name!r} is not on the reference's path (have: {sorted(_TRANSFORMS)})")
    return _TRANSFORMS[name](**kw)
