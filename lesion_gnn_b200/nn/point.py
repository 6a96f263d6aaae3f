"""PointNet++ operators and the unsorted segment reduce with torch_geometric's signatures:
`fps`, `radius`, `PointNetConv` (reference models/pointnet.py:6,13-29) and `scatter` on unsorted labels (reference
datasets/nodes/lesions.py:88-93, `extract_features_by_cc`).  Everything runs on lgk kernels: farthest point sampling
and the radius query in knn.cu, the edge MLP on the lgk transforms, the max aggregation on the segment-reduce pooling
kernel (radius() emits edges grouped by destination), the unsorted reduce as a CSR build + row gather + segment reduce
(deterministic: no floating-point atomics)."""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch
from torch import Tensor

from .. import _lib, ops
from ..graph import as_ptr
from . import functional as F
from .pool import Segments


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _segments(batch, n: int, device) -> Tuple[Tensor, int]:
    """(int32 ptr [B+1], B) for a sorted batch vector, a Segments object or None (one graph)"""
    if isinstance(batch, Segments):
        return batch.ptr, batch.num_graphs
    if batch is None:
        return torch.tensor([0, n], dtype=torch.int32, device=device), 1
    return as_ptr(batch)


def fps(x: Tensor, batch=None, ratio: float = 0.5, random_start: bool = True, batch_size: Optional[int] = None) -> Tensor:
    """torch_cluster.fps / torch_geometric.nn.fps: indices of ceil(ratio * n_g) farthest-point samples of every graph
    (reference models/pointnet.py:22).  random_start=False starts at the first node of each graph."""
    if not x.is_cuda:
        raise RuntimeError("x must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    if x.dtype not in (torch.float32, torch.float64):
        raise TypeError("fps takes float32 or float64 points")
    x = x.contiguous()
    n = x.size(0)
    ptr, B = _segments(batch, n, x.device)
    deg = (ptr[1:] - ptr[:-1]).to(x.dtype)
    # torch_cluster: out_ptr = (deg.to(ratio.dtype) * ratio).ceil().long().cumsum(0), ratio a tensor of x.dtype
    m = (deg * torch.tensor(ratio, dtype=x.dtype, device=x.device)).ceil().long()
    out_ptr = torch.cat([m.new_zeros(1), m.cumsum(0)])
    total = int(out_ptr[-1].item())  # the sample count depends on the data: one device->host read, as in torch_cluster
    start = None
    if random_start:
        u = torch.rand(B, device=x.device, dtype=torch.float64)
        start = (u * (ptr[1:] - ptr[:-1]).double()).long().contiguous()
    out = torch.empty(total, dtype=torch.int64, device=x.device)
    dist = torch.empty(max(n, 1), dtype=x.dtype, device=x.device)
    if total > 0:
        _lib.lib().call("lgk_fps", x.data_ptr(), 0 if x.dtype == torch.float32 else 1, x.size(1), ptr.data_ptr(), out_ptr.data_ptr(),
                        None if start is None else start.data_ptr(), B, n, dist.data_ptr(), out.data_ptr(), _stream())
    return out


def radius(x: Tensor, y: Tensor, r: float, batch_x=None, batch_y=None, max_num_neighbors: int = 32,
           num_workers: int = 1, batch_size: Optional[int] = None) -> Tensor:
    """torch_cluster.radius: [2, E] int64, row 0 = index into y (query), row 1 = index into x (point), grouped by query
    in ascending point index, at most max_num_neighbors per query (reference models/pointnet.py:23)"""
    if not (x.is_cuda and y.is_cuda):
        raise RuntimeError("x and y must be CUDA tensors: lesion_gnn_b200 has no CPU path")
    if x.dtype != y.dtype or x.dtype not in (torch.float32, torch.float64):
        raise TypeError("radius takes float32 or float64 points of one dtype")
    x, y = x.contiguous(), y.contiguous()
    ptr_x, B = _segments(batch_x, x.size(0), x.device)
    ptr_y, By = _segments(batch_y, y.size(0), y.device)
    if By > B:
        raise ValueError("batch_y refers to more graphs than batch_x")
    if By < B:  # trailing graphs without queries
        ptr_y = torch.cat([ptr_y, ptr_y[-1:].expand(B - By)]).contiguous()
    L = _lib.lib()
    dt = 0 if x.dtype == torch.float32 else 1
    ny = y.size(0)
    counts = torch.empty(max(ny, 1), dtype=torch.int32, device=x.device)
    L.call("lgk_radius", x.data_ptr(), y.data_ptr(), dt, x.size(1), ptr_x.data_ptr(), ptr_y.data_ptr(), B, x.size(0), ny, float(r),
           int(max_num_neighbors), counts.data_ptr(), None, None, None, _stream())
    incl = torch.cumsum(counts[:ny], 0, dtype=torch.int64)
    E = int(incl[-1].item()) if ny > 0 else 0
    out = torch.empty(2, E, dtype=torch.int64, device=x.device)
    if E > 0:
        offsets = (incl - counts[:ny]).contiguous()
        L.call("lgk_radius", x.data_ptr(), y.data_ptr(), dt, x.size(1), ptr_x.data_ptr(), ptr_y.data_ptr(), B, x.size(0), ny, float(r),
               int(max_num_neighbors), None, offsets.data_ptr(), out[0].data_ptr(), out[1].data_ptr(), _stream())
    return out


def scatter(src: Tensor, index: Tensor, dim: int = 0, dim_size: Optional[int] = None, reduce: str = "sum") -> Tensor:
    """torch_geometric.utils.scatter(src, index, dim=0, dim_size, reduce) for UNSORTED labels -- reference
    datasets/nodes/lesions.py:93: features [H*W, C] reduced per connected component.  sum / mean / max; empty segments
    give 0.  Rows are grouped by label with the CSR builder (stable, ascending row id), gathered into label order and
    reduced by the segment-reduce pooling kernel.  Differentiable in `src`."""
    if dim not in (0, -src.dim()):
        raise NotImplementedError("scatter is implemented along dim 0")
    if not src.is_cuda:
        raise RuntimeError("src must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    reduce = {"add": "sum"}.get(reduce, reduce)
    if reduce not in ("sum", "mean", "max"):
        raise ValueError(f"unsupported reduce {reduce!r}")
    squeeze = src.dim() == 1
    x = (src.unsqueeze(-1) if squeeze else src.reshape(src.size(0), -1)).contiguous()
    n = x.size(0)
    if dim_size is None:
        dim_size = int(index.max().item()) + 1 if n > 0 else 0
    if n == 0 or dim_size == 0:
        return src.new_zeros((dim_size,) + tuple(src.shape[1:]))
    # label l <- rows {i : index[i] = l}: a "graph" with an edge (src = i) -> (dst = index[i])
    nn_ = max(n, dim_size)
    ei = torch.stack([torch.arange(n, device=x.device), index.long()]).contiguous()
    rowptr, col, _, _, _, _ = ops.csr_build(ei, nn_, with_csc=False)
    order = col.long()                                    # rows in label order, ascending row id inside a label
    xs = _GatherRows.apply(x, order)
    out = F.segment_pool(xs, rowptr[: dim_size + 1].contiguous(), reduce)
    out = out.squeeze(-1) if squeeze else out.reshape((dim_size,) + tuple(src.shape[1:]))
    return out


class _GatherRows(torch.autograd.Function):
    """y = x[order] with `order` a permutation: the backward is the gather by the inverse permutation (no atomics)"""

    @staticmethod
    def forward(ctx, x, order):
        ctx.save_for_backward(order)
        ctx.n = x.size(0)
        return ops.gather_rows(x, order)

    @staticmethod
    def backward(ctx, dy):
        (order,) = ctx.saved_tensors
        if order.numel() != ctx.n:
            raise RuntimeError("_GatherRows backward needs a permutation")
        inv = torch.empty_like(order)
        inv[order] = torch.arange(order.numel(), device=order.device)
        return ops.gather_rows(dy.contiguous(), inv), None


class _EdgeConcat(torch.autograd.Function):
    """cat[x_src[src], pos_src[src] - pos_dst[dst]] per edge; gradient w.r.t. x_src = deterministic scatter-sum by src"""

    @staticmethod
    def forward(ctx, x_src, pos_src, pos_dst, src, dst):
        E = src.numel()
        Fdim = 0 if x_src is None else x_src.size(1)
        D = pos_src.size(1)
        out = torch.empty(E, Fdim + D, dtype=torch.float32, device=pos_src.device)
        _lib.lib().call("lgk_edge_concat", None if x_src is None else x_src.data_ptr(), pos_src.data_ptr(), pos_dst.data_ptr(),
                        src.data_ptr(), dst.data_ptr(), out.data_ptr(), E, Fdim, D, _stream())
        ctx.save_for_backward(src)
        ctx.shape = (None if x_src is None else x_src.size(0), Fdim)
        return out

    @staticmethod
    def backward(ctx, dmsg):
        (src,) = ctx.saved_tensors
        nx, Fdim = ctx.shape
        dx = None
        if nx is not None and ctx.needs_input_grad[0]:
            with torch.no_grad():
                dx = scatter(dmsg[:, :Fdim].contiguous(), src, 0, nx, "sum")
        return dx, None, None, None, None


class PointNetConv(torch.nn.Module):
    """torch_geometric.nn.PointNetConv(local_nn, global_nn=None, add_self_loops=True) as the reference uses it
    (models/pointnet.py:18: add_self_loops=False, bipartite inputs): x_i' = global_nn(max_j local_nn(cat[x_j, pos_j - pos_i])).
    `edge_index` must be grouped by destination (row 1 ascending), which `radius()` guarantees."""

    def __init__(self, local_nn: Optional[torch.nn.Module] = None, global_nn: Optional[torch.nn.Module] = None,
                 add_self_loops: bool = True):
        super().__init__()
        if add_self_loops:
            raise NotImplementedError("add_self_loops=True is not on the reference's path (bipartite set abstraction)")
        self.local_nn, self.global_nn, self.add_self_loops = local_nn, global_nn, add_self_loops

    def forward(self, x: Union[Tensor, Tuple[Optional[Tensor], Optional[Tensor]], None], pos, edge_index: Tensor) -> Tensor:
        x_src = x[0] if isinstance(x, tuple) else x
        pos_src, pos_dst = pos if isinstance(pos, tuple) else (pos, pos)
        src, dst = edge_index[0].contiguous(), edge_index[1].contiguous()
        n_dst = pos_dst.size(0)
        msg = _EdgeConcat.apply(None if x_src is None else x_src.float().contiguous(), pos_src.float().contiguous(),
                                pos_dst.float().contiguous(), src, dst)
        if self.local_nn is not None:
            msg = self.local_nn(msg)
        # max over the incoming edges of every destination: contiguous segments (edges are grouped by destination)
        ptr = ops.batch_ptr(dst, n_dst)
        out = F.segment_pool(msg, ptr, "max")
        if self.global_nn is not None:
            out = self.global_nn(out)
        return out


def ceil_ratio_counts(counts, ratio: float, dtype=torch.float32):
    """host-side ceil(ratio * n_g) exactly as torch_cluster computes it (in the dtype of the points)"""
    c = torch.tensor(list(counts), dtype=dtype)
    return (c * torch.tensor(ratio, dtype=dtype)).ceil().long().tolist()


__all__ = ["fps", "radius", "scatter", "PointNetConv"]
