"""Synthetic lesion-graph batches with the shapes SURVEY.md 8d fixes for BASELINE.json's five configs, and the
PyG-style collation (reference datasets/datamodule.py:64-69: x concatenated, `batch` = sorted graph ids).

Everything here is host-side data plumbing (CPU tensors, optionally pinned); the graph itself (kNN edges, CSR)
is built on the device by lesion_gnn_b200.transforms / lesion_gnn_b200.graph.
"""
from __future__ import annotations

import dataclasses
import math
from typing import Optional

import torch
from torch import Tensor

SEED = 1234  # the reference's seed (configs/config.py:70)


@dataclasses.dataclass
class HostBatch:
    x: Tensor                 # [N, F] float32
    pos: Tensor               # [N, 2] float64 (pixel-scale centroids, reference lesions.py:175)
    batch: Tensor             # [N] int64, sorted
    y: Tensor                 # [B] int64 DR grade 0..4
    counts: list              # nodes per graph (host)
    edge_index: Optional[Tensor] = None   # [2, E] int64
    edge_weight: Optional[Tensor] = None  # [E] float32

    @property
    def dst_sorted(self) -> bool:
        """edge_index[1] non-decreasing (kNN graphs collated sample by sample are): checked once on the host and cached;
        lets the device build its CSR by boundary detection (graph.CSRGraph.from_edge_index(dst_sorted=True))"""
        ei = self.edge_index
        if ei is None:
            return False
        key = (ei.data_ptr(), ei.size(1))
        cached = self.__dict__.get("_dst_sorted")
        if cached is None or cached[0] != key:
            d = ei[1]
            cached = (key, bool(ei.size(1) == 0 or bool((d[1:] >= d[:-1]).all())) if not ei.is_cuda else (key, False))
            self.__dict__["_dst_sorted"] = cached
        return cached[1]

    @property
    def num_graphs(self) -> int:
        return len(self.counts)

    @property
    def num_nodes(self) -> int:
        return self.x.size(0)

    def pin(self) -> "HostBatch":
        for f in dataclasses.fields(self):
            v = getattr(self, f.name)
            if isinstance(v, Tensor) and torch.cuda.is_available():
                setattr(self, f.name, v.pin_memory())
        return self

    def h2d_bytes(self) -> int:
        return sum(v.numel() * v.element_size() for v in (self.x, self.batch, self.y, self.edge_index, self.edge_weight)
                   if v is not None)


@dataclasses.dataclass
class GraphSample:
    """one graph as the reference's datasets yield it (torch_geometric.data.Data: x, pos, edge_index, y and the optional
    edge weight of GaussianDistance; reference datasets/nodes/lesions.py:170-176)"""
    x: Tensor
    pos: Optional[Tensor] = None
    edge_index: Optional[Tensor] = None
    y: Optional[Tensor] = None
    edge_weight: Optional[Tensor] = None


def collate(samples) -> HostBatch:
    """PyG DataLoader collation (Batch.from_data_list; reference datasets/datamodule.py:64-69): node
    tensors concatenated, every graph's edge_index shifted by the number of nodes before it, `batch` = the graph id of
    every node (ascending, so graphs are contiguous slabs), y concatenated.  Host-side index work: exact."""
    samples = list(samples)
    if not samples:
        raise ValueError("collate() needs at least one graph")
    counts = [int(s.x.size(0)) for s in samples]
    offsets = [0]
    for n in counts:
        offsets.append(offsets[-1] + n)
    x = torch.cat([s.x for s in samples], dim=0)
    pos = None
    if all(s.pos is not None for s in samples):
        pos = torch.cat([s.pos for s in samples], dim=0)
    has_edges = [s.edge_index is not None for s in samples]
    if any(has_edges) and not all(has_edges):
        raise ValueError("either every graph or none carries edge_index")
    edge_index = edge_weight = None
    if all(has_edges):
        for s, n in zip(samples, counts):
            if s.edge_index.numel() and (int(s.edge_index.min()) < 0 or int(s.edge_index.max()) >= n):
                raise ValueError("edge_index refers to a node outside its graph")
        edge_index = torch.cat([s.edge_index.long() + o for s, o in zip(samples, offsets)], dim=1)
        has_w = [s.edge_weight is not None for s in samples]
        if any(has_w) and not all(has_w):
            raise ValueError("either every graph or none carries edge_weight")
        if all(has_w):
            edge_weight = torch.cat([s.edge_weight for s in samples])
    batch = torch.repeat_interleave(torch.arange(len(samples)), torch.tensor(counts, dtype=torch.long))
    ys = [s.y for s in samples]
    y = torch.cat([t.reshape(-1) for t in ys]) if all(t is not None for t in ys) else torch.zeros(len(samples), dtype=torch.long)
    if pos is None:
        pos = torch.zeros(x.size(0), 2, dtype=torch.float64)
    return HostBatch(x=x, pos=pos, batch=batch, y=y, counts=counts, edge_index=edge_index, edge_weight=edge_weight)


def ptr_from_counts(counts) -> Tensor:
    """PyG Batch.ptr: int64 [B + 1] cumulative node counts"""
    return torch.cat([torch.zeros(1, dtype=torch.long), torch.tensor(list(counts), dtype=torch.long).cumsum(0)])


def node_counts(num_graphs: int, spec, gen: torch.Generator) -> list:
    """spec = int (fixed size) or 'ddr': ceil(LogNormal(ln 30, 0.9)) clipped to [1, 600] -- long-tailed like the
    reference's DDR lesion graphs (notebooks/viz.ipynb cells 4-6); node 0 of each graph is the background."""
    if isinstance(spec, int):
        return [spec] * num_graphs
    if isinstance(spec, (list, tuple)):  # explicit sizes (a rank's shard of a global batch)
        if len(spec) != num_graphs:
            raise ValueError("one node count per graph")
        return [int(n) for n in spec]
    if spec == "ddr":
        z = torch.randn(num_graphs, generator=gen, dtype=torch.float64)
        n = torch.exp(math.log(30.0) + 0.9 * z).ceil().clamp(1, 600).long()
        return n.tolist()
    raise ValueError(spec)


def synth_batch(num_graphs: int, nodes, feat: int, seed: int = SEED, layout: str = "uniform",
                num_classes: int = 5) -> HostBatch:
    """features ~ N(0,1) fp32; pos ~ U[0,1536)^2 fp64 ('uniform', node order random) or a jittered grid in
    raster order ('raster': how superpixel / connected-component labelling numbers its regions)."""
    gen = torch.Generator().manual_seed(seed)
    counts = node_counts(num_graphs, nodes, gen)
    n_total = int(sum(counts))
    x = torch.randn(n_total, feat, generator=gen, dtype=torch.float32)
    if layout == "uniform":
        pos = torch.rand(n_total, 2, generator=gen, dtype=torch.float64) * 1536.0
    elif layout == "raster":
        chunks = []
        for n in counts:
            side = max(1, math.ceil(math.sqrt(n)))
            idx = torch.arange(n, dtype=torch.float64)
            cell = 1536.0 / side
            gx, gy = idx % side, torch.div(idx, side, rounding_mode="floor")
            jit = torch.rand(n, 2, generator=gen, dtype=torch.float64)
            chunks.append(torch.stack([(gx + 0.15 + 0.7 * jit[:, 0]) * cell, (gy + 0.15 + 0.7 * jit[:, 1]) * cell], dim=1))
        pos = torch.cat(chunks) if chunks else torch.zeros(0, 2, dtype=torch.float64)
    else:
        raise ValueError(layout)
    batch = torch.repeat_interleave(torch.arange(num_graphs), torch.tensor(counts, dtype=torch.long))
    y = torch.randint(0, num_classes, (num_graphs,), generator=gen)
    return HostBatch(x=x, pos=pos, batch=batch, y=y, counts=counts)


def shard_counts(counts: list, world_size: int) -> list:
    """Data-parallel partition of a batch over ranks: contiguous ranges of graphs, balanced by
    node count rather than graph count (long-tailed sizes).  Returns [(g_lo, g_hi)] per rank."""
    total = sum(counts)
    bounds, acc, g = [0], 0, 0
    for r in range(1, world_size):
        target = total * r / world_size
        while g < len(counts) and acc + counts[g] / 2 <= target:
            acc += counts[g]
            g += 1
        bounds.append(g)
    bounds.append(len(counts))
    for r in range(1, len(bounds)):
        bounds[r] = max(bounds[r], bounds[r - 1])
    return [(bounds[r], bounds[r + 1]) for r in range(world_size)]


def take_graphs(hb: HostBatch, g_lo: int, g_hi: int) -> HostBatch:
    """sub-batch of graphs [g_lo, g_hi) with node ids and graph ids re-based (what collation of that subset
    would have produced)"""
    n_lo = int(sum(hb.counts[:g_lo]))
    n_hi = int(sum(hb.counts[:g_hi]))
    out = HostBatch(x=hb.x[n_lo:n_hi].clone(), pos=hb.pos[n_lo:n_hi].clone(), batch=hb.batch[n_lo:n_hi] - g_lo,
                    y=hb.y[g_lo:g_hi].clone(), counts=list(hb.counts[g_lo:g_hi]))
    if hb.edge_index is not None:
        dst = hb.edge_index[1]
        m = (dst >= n_lo) & (dst < n_hi)
        out.edge_index = hb.edge_index[:, m] - n_lo
        if hb.edge_weight is not None:
            out.edge_weight = hb.edge_weight[m].clone()
    return out
