"""Graph structure shared by all layers of a step.

`CSRGraph` is the stand-in for the `adj_t` SparseTensor the reference passes to its models when
`compile=False` (reference datasets/datamodule.py:44-45, models/gat.py:87-90): destination-sorted CSR plus the
CSC transpose index for the backward pass, int32, built once per batch by lgk_csr_build and shared by every
layer and both passes.  Layers accept either a `CSRGraph` or the plain `[2, E]` int64 `edge_index`; the latter
is converted on first use and memoised (keyed on the tensor's storage, shape and version counter) so that
L layers x (fwd + bwd) still cost one build.
"""
from __future__ import annotations

import os

from collections import OrderedDict
from typing import Optional

import torch
from torch import Tensor

from . import ops


_SORTED_MIN_EDGES = int(os.environ.get("LGK_CSR_SORTED_MIN_EDGES", str(1 << 20)))


class CSRGraph:
    """rowptr/col/perm: CSR by destination, entries ordered by (dst, src, original edge id).
    colptr/row/csr2csc: CSC by source, entries ordered by CSR slot.  value: edge weights in CSR order."""

    __slots__ = ("num_nodes", "num_edges", "rowptr", "col", "perm", "colptr", "row", "csr2csc", "value", "_norm", "_ready", "_quads")

    def __init__(self, num_nodes, num_edges, rowptr, col, perm, colptr=None, row=None, csr2csc=None, value=None):
        self.num_nodes, self.num_edges = int(num_nodes), int(num_edges)
        self.rowptr, self.col, self.perm = rowptr, col, perm
        self.colptr, self.row, self.csr2csc = colptr, row, csr2csc
        self.value = value
        self._norm = {}
        self._quads = {}  # row-blocked forms of the CSR ("csr") / CSC ("csc"), built on first use (quads())
        self._ready = None  # event recorded on the stream that built the graph (from_edge_index(..., stream=))

    @staticmethod
    def from_edge_index(edge_index: Tensor, num_nodes: int, edge_weight: Optional[Tensor] = None,
                        with_csc: bool = True, stream: Optional["torch.cuda.Stream"] = None,
                        dst_sorted: bool = False) -> "CSRGraph":
        """`dst_sorted`: edge_index[1] is known to be non-decreasing (the output order of knn_graph / radius_graph and of
        collated per-sample kNN graphs): the faster boundary-detection build runs (ops.csr_build).
        `stream`: build on that stream, forked from the current one (the build overlaps whatever the current
        stream does next, e.g. the in_proj transform, which does not need the graph); the first layer that uses
        the graph joins it (`join()`, called by `as_graph`)."""
        if not edge_index.is_cuda:
            raise RuntimeError("edge_index must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
        if stream is not None:
            stream.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(stream):
                g = CSRGraph.from_edge_index(edge_index, num_nodes, edge_weight, with_csc, dst_sorted=dst_sorted)
                g._ready = torch.cuda.Event()
                g._ready.record(stream)
            return g
        if edge_index.dtype != torch.int64:
            edge_index = edge_index.long()
        edge_index = edge_index.contiguous()
        # measured (B200): sorted-input build 605 vs 798 us at E = 20.5 M, but 24.8 vs 22.3 us at E = 69 k (one memset node
        # more, same kernel count): below ~1 M edges the general build stays
        use_sorted = dst_sorted and edge_index.size(1) >= _SORTED_MIN_EDGES
        rowptr, col, perm, colptr, row, csr2csc = ops.csr_build(edge_index, num_nodes, with_csc, dst_sorted=use_sorted)
        g = CSRGraph(num_nodes, edge_index.size(1), rowptr, col, perm, colptr, row, csr2csc)
        if edge_weight is not None:
            g.value = g.permute_edge_values(edge_weight)
        return g

    def join(self) -> "CSRGraph":
        """make the current stream wait for a graph that was built on another stream (no-op otherwise)"""
        ev = self._ready
        if ev is not None:
            self._ready = None
            cur = torch.cuda.current_stream()
            cur.wait_event(ev)
            for t in (self.rowptr, self.col, self.perm, self.colptr, self.row, self.csr2csc, self.value):
                if t is not None:
                    t.record_stream(cur)  # allocated on the build stream, used on this one
        return self

    def permute_edge_values(self, edge_weight: Tensor) -> Tensor:
        """per-edge values in original edge order -> CSR slot order (SparseTensor `value`)"""
        if edge_weight.dim() != 1 or edge_weight.numel() != self.num_edges:
            raise ValueError("edge_weight must have shape [E]")
        return ops.gather_f32(edge_weight.detach().float().contiguous(), self.perm)

    def with_value(self, edge_weight: Optional[Tensor]) -> "CSRGraph":
        """same structure, different edge weights (shares the index tensors)"""
        g = CSRGraph(self.num_nodes, self.num_edges, self.rowptr, self.col, self.perm, self.colptr, self.row, self.csr2csc)
        g._quads = self._quads  # same structure
        g.value = None if edge_weight is None else self.permute_edge_values(edge_weight)
        return g

    def gcn_norm(self, improved: bool, add_self_loops: bool):
        """(dinv, selfw) of gcn_norm for this graph's weights; memoised per (improved, add_self_loops)."""
        key = (bool(improved), bool(add_self_loops))
        if key not in self._norm:
            self._norm[key] = ops.gcn_norm(self.rowptr, self.col, self.value, 2.0 if improved else 1.0, add_self_loops)
        return self._norm[key]

    def inv_degree(self) -> Tensor:
        """1 / max(in-degree, 1) per node (the mean aggregation's backward scale); computed once per graph"""
        t = self._norm.get("inv_degree")
        if t is None:
            t = (self.rowptr[1:] - self.rowptr[:-1]).clamp(min=1).float().reciprocal()
            self._norm["inv_degree"] = t
        return t

    def quads(self, which: str = "csr") -> "ops.Quads":
        """row-blocked form of the CSR (forward aggregation) or the CSC (backward) for the windowed aggregation kernel
        (ops.spmm(..., quads=)); built once per graph and direction, shared by every layer"""
        q = self._quads.get(which)
        if q is None:
            if which == "csr":
                q = ops.quad_build(self.rowptr, self.col)
            elif which == "csc":
                if self.colptr is None:
                    raise RuntimeError("this CSRGraph was built without the CSC transpose index")
                q = ops.quad_build(self.colptr, self.row)
            else:
                raise ValueError(which)
            self._quads[which] = q
        return q

    def to_edge_index(self) -> Tensor:
        """[2, E] int64 in CSR order (row 0 = src, row 1 = dst)"""
        deg = (self.rowptr[1:] - self.rowptr[:-1]).long()
        dst = torch.repeat_interleave(torch.arange(self.num_nodes, device=self.col.device), deg)
        return torch.stack([self.col.long(), dst])


_CACHE: "OrderedDict[tuple, tuple]" = OrderedDict()
_CACHE_SIZE = 16


def _key(t: Tensor, extra=()) -> tuple:
    return (t.data_ptr(), tuple(t.shape), t._version, t.device.index) + tuple(extra)


def clear_cache() -> None:
    _CACHE.clear()


def _from_adj_t(adj) -> Optional[Tensor]:
    """(edge_index, value) of an `adj_t`-style sparse adjacency (rows = destinations, reference models/gat.py:87-90):
    a torch_sparse.SparseTensor-like object with `.csr()` -> (rowptr, col, value), or a torch sparse CSR / COO tensor."""
    if isinstance(adj, Tensor):
        if adj.layout == torch.sparse_csr:
            rowptr, col, value = adj.crow_indices(), adj.col_indices(), adj.values()
        elif adj.layout == torch.sparse_coo:
            adj = adj.coalesce()
            idx = adj.indices()
            return torch.stack([idx[1], idx[0]]), adj.values()
        else:
            return None
    elif hasattr(adj, "csr") and callable(adj.csr):
        rowptr, col, value = adj.csr()
    else:
        return None
    deg = (rowptr[1:] - rowptr[:-1]).long()
    dst = torch.repeat_interleave(torch.arange(rowptr.numel() - 1, device=col.device), deg)
    return torch.stack([col.long(), dst]), value


def as_graph(edge_index, num_nodes: int, edge_weight: Optional[Tensor] = None) -> CSRGraph:
    """CSRGraph for `edge_index`: the [2,E] tensor, a CSRGraph, or an `adj_t` (torch_sparse.SparseTensor-like object
    or torch sparse CSR/COO tensor whose rows are destinations); memoised per tensor."""
    if not isinstance(edge_index, CSRGraph) and not (isinstance(edge_index, Tensor) and edge_index.layout == torch.strided):
        conv = _from_adj_t(edge_index)
        if conv is None:
            raise TypeError(f"cannot interpret {type(edge_index).__name__} as a graph")
        if edge_weight is not None:
            raise ValueError("an adj_t carries its own values; do not pass edge_weight as well")
        ei, value = conv
        attr = getattr(edge_index, "_lgk_graph", None)  # memoised on the object when it allows attributes
        if attr is not None and attr.num_nodes == num_nodes:
            return attr.join()
        g = CSRGraph.from_edge_index(ei.contiguous(), num_nodes, None if value is None or value.dim() != 1 else value.float())
        try:
            edge_index._lgk_graph = g
        except AttributeError:
            pass
        return g
    if isinstance(edge_index, CSRGraph):
        if edge_weight is not None:
            raise ValueError("pass edge weights through CSRGraph.value when handing a CSRGraph to a layer")
        if edge_index.num_nodes != num_nodes:
            raise ValueError(f"graph has {edge_index.num_nodes} nodes but x has {num_nodes} rows")
        return edge_index.join()
    k = _key(edge_index, (int(num_nodes),))
    hit = _CACHE.get(k)
    if hit is None:
        g = CSRGraph.from_edge_index(edge_index, num_nodes)
        _CACHE[k] = (g, edge_index, {})  # keep the tensor alive so its storage cannot be recycled under the key
        while len(_CACHE) > _CACHE_SIZE:
            _CACHE.popitem(last=False)
        hit = _CACHE[k]
    else:
        _CACHE.move_to_end(k)
    g, _, weighted = hit
    if edge_weight is None:
        return g
    wk = _key(edge_weight)
    if wk not in weighted:
        weighted.clear()
        weighted[wk] = (g.with_value(edge_weight), edge_weight)
    return weighted[wk][0]


def as_ptr(batch: Tensor, size: Optional[int] = None):
    """(int32 graph pointer [B+1], B) for a sorted `batch` vector, memoised per tensor.  With size=None this
    reads batch[-1] on the host once (PyG's global_*_pool does the same through batch.max())."""
    if not batch.is_cuda:
        raise RuntimeError("batch must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    k = _key(batch, ("ptr", size))
    hit = _CACHE.get(k)
    if hit is None:
        if size is None:
            size = int(batch[-1].item()) + 1 if batch.numel() > 0 else 0
        b = batch if batch.dtype == torch.int64 else batch.long()
        ptr = ops.batch_ptr(b.contiguous(), size)
        _CACHE[k] = ((ptr, int(size)), batch, {})
        while len(_CACHE) > _CACHE_SIZE:
            _CACHE.popitem(last=False)
        hit = _CACHE[k]
    else:
        _CACHE.move_to_end(k)
    return hit[0]
