"""Op-for-op CPU restatement of the PyG 2.5.1 operators on lesion-gnn's hot path.

TEST INFRASTRUCTURE (see oracle/__init__.py) -- parity unpinned: PyG is absent
from /root/reference and from this image, so each function restates the
published torch-geometric 2.5.1 / torch-scatter 2.1.2 / torch-sparse 0.6.18 /
torch-cluster 1.6.3 algorithm and cites the reference call site that lands in
it.  Everything is plain PyTorch on CPU, dtype-generic (fp64 for the checker,
fp32 for the timed CPU baseline), written in the gather -> elementwise ->
scatter order PyG itself uses so that fp32 rounding resembles the real path.

Conventions (PyG `flow="source_to_target"`): edge_index[0] = src j,
edge_index[1] = dst i; messages flow j -> i and are reduced per i.
"""
from __future__ import annotations

import math
from itertools import pairwise
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor

# ----------------------------------------------------------------------------
# torch_geometric.utils.scatter / softmax / self-loop helpers
# ----------------------------------------------------------------------------


def _broadcast(index: Tensor, src: Tensor) -> Tensor:
    # torch_geometric.utils.scatter.broadcast for dim=0
    size = [1] * src.dim()
    size[0] = -1
    return index.view(size).expand_as(src)


def scatter(src: Tensor, index: Tensor, dim_size: int, reduce: str = "sum") -> Tensor:
    """torch_geometric.utils.scatter(src, index, dim=0, dim_size, reduce).

    Lands here from: global_mean_pool (reference gat.py:56, gin.py:33),
    global_max_pool (pointnet.py:38), MessagePassing aggregation (gat.py:51,
    gin.py:31, drgnet.py:55) and extract_features_by_cc (lesions.py:93).
    """
    size = (dim_size,) + tuple(src.shape[1:])
    if reduce in ("sum", "add"):
        return src.new_zeros(size).scatter_add_(0, _broadcast(index, src), src)
    if reduce == "mean":
        count = src.new_zeros(dim_size)
        count.scatter_add_(0, index, src.new_ones(src.size(0)))
        count = count.clamp(min=1)
        out = src.new_zeros(size).scatter_add_(0, _broadcast(index, src), src)
        return out / count.view((-1,) + (1,) * (src.dim() - 1))
    if reduce in ("max", "min"):
        # empty segments stay 0 (include_self=False on a zero-initialised out)
        return src.new_zeros(size).scatter_reduce_(
            0, _broadcast(index, src), src, reduce=f"a{reduce}", include_self=False
        )
    raise ValueError(reduce)


def softmax(src: Tensor, index: Tensor, num_nodes: int) -> Tensor:
    """torch_geometric.utils.softmax(src, index, num_nodes=N, dim=0).

    Used by GATConv.edge_update (reference gat.py:31,51).  max is detached,
    denominator gets +1e-16.
    """
    src_max = scatter(src.detach(), index, num_nodes, reduce="max")
    out = src - src_max.index_select(0, index)
    out = out.exp()
    out_sum = scatter(out, index, num_nodes, reduce="sum") + 1e-16
    return out / out_sum.index_select(0, index)


def remove_self_loops(edge_index: Tensor, edge_attr: Optional[Tensor] = None):
    mask = edge_index[0] != edge_index[1]
    edge_index = edge_index[:, mask]
    return edge_index, (None if edge_attr is None else edge_attr[mask])


def add_self_loops(edge_index: Tensor, edge_attr: Optional[Tensor], fill_value: float, num_nodes: int):
    """Appends N loops (i, i) AFTER the existing edges (PyG order)."""
    loop_index = torch.arange(num_nodes, dtype=edge_index.dtype, device=edge_index.device)
    loop_index = loop_index.unsqueeze(0).repeat(2, 1)
    if edge_attr is not None:
        loop_attr = edge_attr.new_full((num_nodes,) + tuple(edge_attr.shape[1:]), fill_value)
        edge_attr = torch.cat([edge_attr, loop_attr], dim=0)
    return torch.cat([edge_index, loop_index], dim=1), edge_attr


def add_remaining_self_loops(edge_index: Tensor, edge_attr: Optional[Tensor], fill_value: float, num_nodes: int):
    """torch_geometric.utils.add_remaining_self_loops: keeps the weight of an
    existing loop (last one wins on duplicates), adds the missing ones with
    `fill_value`; result = non-loop edges, then N loop edges."""
    row, col = edge_index[0], edge_index[1]
    mask = row != col
    loop_index = torch.arange(num_nodes, dtype=edge_index.dtype, device=edge_index.device)
    loop_index = loop_index.unsqueeze(0).repeat(2, 1)
    if edge_attr is not None:
        loop_attr = edge_attr.new_full((num_nodes,) + tuple(edge_attr.shape[1:]), fill_value)
        inv_mask = ~mask
        loop_attr[row[inv_mask]] = edge_attr[inv_mask]
        edge_attr = torch.cat([edge_attr[mask], loop_attr], dim=0)
    edge_index = torch.cat([edge_index[:, mask], loop_index], dim=1)
    return edge_index, edge_attr


def gcn_norm(edge_index: Tensor, edge_weight: Optional[Tensor], num_nodes: int, improved: bool = False,
             add_self_loops_: bool = True, dtype: torch.dtype = torch.float32):
    """torch_geometric.nn.conv.gcn_conv.gcn_norm, Tensor path."""
    fill_value = 2.0 if improved else 1.0
    if add_self_loops_:
        edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight, fill_value, num_nodes)
    if edge_weight is None:
        edge_weight = torch.ones((edge_index.size(1),), dtype=dtype, device=edge_index.device)
    row, col = edge_index[0], edge_index[1]
    deg = scatter(edge_weight, col, num_nodes, reduce="sum")
    deg_inv_sqrt = deg.pow(-0.5)
    deg_inv_sqrt.masked_fill_(deg_inv_sqrt == float("inf"), 0)
    edge_weight = deg_inv_sqrt[row] * edge_weight * deg_inv_sqrt[col]
    return edge_index, edge_weight


def propagate(x: Tensor, edge_index: Tensor, num_nodes: int, aggr: str = "sum",
              edge_weight: Optional[Tensor] = None) -> Tensor:
    """MessagePassing.propagate with message = [w_e *] x_j, aggregation `aggr`."""
    src, dst = edge_index[0], edge_index[1]
    msg = x.index_select(0, src)
    if edge_weight is not None:
        msg = edge_weight.view(-1, *([1] * (msg.dim() - 1))) * msg
    return scatter(msg, dst, num_nodes, reduce=aggr)


# ----------------------------------------------------------------------------
# Layers: same constructor arguments, parameter names and forward signatures
# as PyG 2.5.1 (state_dicts interchange).
# ----------------------------------------------------------------------------


def glorot_(w: Tensor) -> Tensor:
    a = math.sqrt(6.0 / (w.size(-2) + w.size(-1)))
    with torch.no_grad():
        return w.uniform_(-a, a)


class PygLinear(nn.Module):
    """torch_geometric.nn.dense.linear.Linear: weight [out, in], glorot init, zero bias."""

    def __init__(self, in_channels: int, out_channels: int, bias: bool = True):
        super().__init__()
        self.weight = nn.Parameter(glorot_(torch.empty(out_channels, in_channels)))
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None

    def forward(self, x: Tensor) -> Tensor:
        return F.linear(x, self.weight, self.bias)


class GCNConv(nn.Module):
    """PyG GCNConv: D^-1/2 (A + I) D^-1/2 (X W^T) + b.  SURVEY 8a row a-6."""

    def __init__(self, in_channels, out_channels, improved=False, cached=False, add_self_loops=True,
                 normalize=True, bias=True):
        super().__init__()
        self.improved, self.add_self_loops, self.normalize = improved, add_self_loops, normalize
        self.lin = PygLinear(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None

    def forward(self, x, edge_index, edge_weight=None):
        n = x.size(0)
        if self.normalize:
            edge_index, edge_weight = gcn_norm(edge_index, edge_weight, n, self.improved, self.add_self_loops, x.dtype)
        x = self.lin(x)
        out = propagate(x, edge_index, n, "sum", edge_weight)
        if self.bias is not None:
            out = out + self.bias
        return out


class GATConv(nn.Module):
    """PyG GATConv (int in_channels => shared `lin`).  SURVEY 8a row a-2;
    reference call site gat.py:31 / gat.py:51."""

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2, dropout=0.0,
                 add_self_loops=True, edge_dim=None, fill_value="mean", bias=True):
        super().__init__()
        assert edge_dim is None, "edge features are not on the reference's path"
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.negative_slope, self.dropout = concat, negative_slope, dropout
        self.add_self_loops = add_self_loops
        self.lin = PygLinear(in_channels, heads * out_channels, bias=False)
        self.att_src = nn.Parameter(glorot_(torch.empty(1, heads, out_channels)))
        self.att_dst = nn.Parameter(glorot_(torch.empty(1, heads, out_channels)))
        if bias:
            self.bias = nn.Parameter(torch.zeros(heads * out_channels if concat else out_channels))
        else:
            self.bias = None

    def forward(self, x, edge_index, edge_attr=None, size=None, return_attention_weights=None):
        H, C = self.heads, self.out_channels
        n = x.size(0)
        h = self.lin(x).view(-1, H, C)
        alpha_src = (h * self.att_src).sum(dim=-1)
        alpha_dst = (h * self.att_dst).sum(dim=-1)
        if self.add_self_loops:
            edge_index, _ = remove_self_loops(edge_index)
            edge_index, _ = add_self_loops(edge_index, None, 0.0, n)
        src, dst = edge_index[0], edge_index[1]
        alpha = alpha_src.index_select(0, src) + alpha_dst.index_select(0, dst)
        alpha = F.leaky_relu(alpha, self.negative_slope)
        alpha = softmax(alpha, dst, n)
        alpha = F.dropout(alpha, p=self.dropout, training=self.training)
        msg = alpha.unsqueeze(-1) * h.index_select(0, src)
        out = scatter(msg, dst, n, reduce="sum")
        out = out.view(-1, H * C) if self.concat else out.mean(dim=1)
        if self.bias is not None:
            out = out + self.bias
        if return_attention_weights:
            return out, (edge_index, alpha)
        return out


class PygBatchNorm(nn.Module):
    """torch_geometric.nn.norm.BatchNorm: wraps BatchNorm1d in `.module`."""

    def __init__(self, in_channels: int):
        super().__init__()
        self.module = nn.BatchNorm1d(in_channels)

    def forward(self, x):
        return self.module(x)


class MLP(nn.Module):
    """torch_geometric.nn.models.MLP(channel_list, act, dropout, norm, plain_last=True).

    Reference: gin.py:23 `MLP([d1, d2, d2], act="ELU", dropout=p)` and
    drgnet.py:48 `MLP([dense, 128, C], dropout=0.5, norm=None, act=F.elu)`.
    """

    def __init__(self, channel_list, act="ELU", dropout=0.0, norm="batch_norm", plain_last=True, bias=True):
        super().__init__()
        self.channel_list = list(channel_list)
        self.plain_last = plain_last
        n_lin = len(channel_list) - 1
        self.dropout = [dropout] * n_lin
        if plain_last:
            self.dropout[-1] = 0.0
        self.act = (lambda t: F.elu(t)) if (act == "ELU" or act is F.elu or act == "elu") else (lambda t: F.relu(t))
        self.lins = nn.ModuleList()
        for d1, d2 in pairwise(channel_list):
            lin = PygLinear(d1, d2, bias=bias)
            # PyG MLP uses Linear's default 'kaiming_uniform' initialiser
            nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5))
            if lin.bias is not None:
                bound = 1 / math.sqrt(d1)
                nn.init.uniform_(lin.bias, -bound, bound)
            self.lins.append(lin)
        self.norms = nn.ModuleList()
        for d in (channel_list[1:-1] if plain_last else channel_list[1:]):
            self.norms.append(PygBatchNorm(d) if norm is not None else nn.Identity())

    def forward(self, x):
        for i, (lin, norm) in enumerate(zip(self.lins, self.norms)):
            x = lin(x)
            x = norm(x)
            x = self.act(x)
            x = F.dropout(x, p=self.dropout[i], training=self.training)
        if self.plain_last:
            x = self.lins[-1](x)
            x = F.dropout(x, p=self.dropout[-1], training=self.training)
        return x


class GINConv(nn.Module):
    """PyG GINConv: nn((1 + eps) x_i + sum_j x_j).  SURVEY 8a row a-4; reference gin.py:23,31."""

    def __init__(self, nn_module: nn.Module, eps: float = 0.0, train_eps: bool = False):
        super().__init__()
        self.nn = nn_module
        self.initial_eps = eps
        if train_eps:
            self.eps = nn.Parameter(torch.tensor([eps], dtype=torch.float32))
        else:
            self.register_buffer("eps", torch.tensor([eps], dtype=torch.float32))

    def forward(self, x, edge_index, size=None):
        out = propagate(x, edge_index, x.size(0), "sum")
        out = out + (1 + self.eps.to(x.dtype)) * x
        return self.nn(out)


class SAGEConv(nn.Module):
    """PyG SAGEConv: W_l aggr_j x_j + b_l + W_r x_i.  SURVEY 8a row a-7."""

    def __init__(self, in_channels, out_channels, aggr="mean", normalize=False, root_weight=True, project=False,
                 bias=True):
        super().__init__()
        assert not project
        self.aggr, self.normalize, self.root_weight = aggr, normalize, root_weight
        self.lin_l = PygLinear(in_channels, out_channels, bias=bias)
        self.lin_r = PygLinear(in_channels, out_channels, bias=False) if root_weight else None
        for lin in (self.lin_l, self.lin_r):
            if lin is not None:  # SAGEConv's Linear uses the kaiming_uniform default
                nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5))

    def forward(self, x, edge_index, size=None):
        out = propagate(x, edge_index, x.size(0), {"add": "sum"}.get(self.aggr, self.aggr))
        out = self.lin_l(out)
        if self.lin_r is not None:
            out = out + self.lin_r(x)
        if self.normalize:
            out = F.normalize(out, p=2.0, dim=-1)
        return out


class GraphConv(nn.Module):
    """PyG GraphConv: W_rel sum_j e_ji x_j + b + W_root x_i.  SURVEY 8a row a-5; reference drgnet.py:31,33,55."""

    def __init__(self, in_channels, out_channels, aggr="add", bias=True):
        super().__init__()
        self.aggr = aggr
        self.lin_rel = PygLinear(in_channels, out_channels, bias=bias)
        self.lin_root = PygLinear(in_channels, out_channels, bias=False)
        for lin in (self.lin_rel, self.lin_root):
            nn.init.kaiming_uniform_(lin.weight, a=math.sqrt(5))

    def forward(self, x, edge_index, edge_weight=None, size=None):
        out = propagate(x, edge_index, x.size(0), {"add": "sum"}.get(self.aggr, self.aggr), edge_weight)
        out = self.lin_rel(out)
        return out + self.lin_root(x)


# ----------------------------------------------------------------------------
# Global pooling (reference gat.py:56, gin.py:33, pointnet.py:38, drgnet.py:37,59)
# ----------------------------------------------------------------------------


def _num_graphs(batch: Tensor, size: Optional[int]) -> int:
    if size is not None:
        return int(size)
    return int(batch.max()) + 1 if batch.numel() > 0 else 0


def global_add_pool(x, batch, size=None):
    return scatter(x, batch, _num_graphs(batch, size), "sum")


def global_mean_pool(x, batch, size=None):
    return scatter(x, batch, _num_graphs(batch, size), "mean")


def global_max_pool(x, batch, size=None):
    return scatter(x, batch, _num_graphs(batch, size), "max")


def sort_aggregation(x: Tensor, batch: Tensor, k: int, size: Optional[int] = None) -> Tensor:
    """torch_geometric.nn.aggr.SortAggregation(k): per graph, sort nodes by the
    LAST channel descending, keep the first k (zero-padded), flatten to [B, k*F].
    Reference drgnet.py:37,59."""
    B = _num_graphs(batch, size)
    Fdim = x.size(1)
    counts = torch.bincount(batch, minlength=B)
    nmax = int(counts.max()) if B > 0 else 0
    ptr = torch.zeros(B + 1, dtype=torch.long)
    ptr[1:] = counts.cumsum(0)
    fill = float(x.detach().min()) - 1.0
    dense = x.new_full((B, nmax, Fdim), fill)
    pos = torch.arange(x.size(0)) - ptr[batch]
    dense[batch, pos] = x
    _, perm = dense[:, :, -1].sort(dim=-1, descending=True, stable=True)
    dense = dense.gather(1, perm.unsqueeze(-1).expand(-1, -1, Fdim))
    if nmax >= k:
        dense = dense[:, :k].contiguous()
    else:
        dense = torch.cat([dense, dense.new_full((B, k - nmax, Fdim), fill)], dim=1)
    dense = torch.where(dense == fill, torch.zeros_like(dense), dense)
    return dense.view(B, k * Fdim)


# ----------------------------------------------------------------------------
# Graph construction (reference transforms.py:19-23 -> KNNGraph; datamodule.py:45 -> ToSparseTensor)
# ----------------------------------------------------------------------------


def batch_to_ptr(batch: Tensor, num_graphs: int) -> Tensor:
    counts = torch.bincount(batch, minlength=num_graphs)
    ptr = torch.zeros(num_graphs + 1, dtype=torch.long)
    ptr[1:] = counts.cumsum(0)
    return ptr


def knn_graph(pos: Tensor, k: int, batch: Optional[Tensor] = None, loop: bool = False) -> Tensor:
    """torch_cluster.knn_graph(pos, k, batch, loop, flow='source_to_target').

    For every centre i: its k nearest points of the same graph by squared L2
    (including itself iff loop; k+1 searched and the self match dropped
    otherwise), emitted grouped by centre, ascending distance (ties: lower
    index first).  Returns [2, E] int64 with row 0 = neighbour (src), row 1 =
    centre (dst).  Graphs with fewer than k candidates yield fewer edges.
    """
    n = pos.size(0)
    if batch is None:
        batch = torch.zeros(n, dtype=torch.long)
    B = int(batch.max()) + 1 if n > 0 else 0
    ptr = batch_to_ptr(batch, B)
    kk = k if loop else k + 1
    srcs, dsts = [], []
    for g in range(B):
        lo, hi = int(ptr[g]), int(ptr[g + 1])
        p = pos[lo:hi]
        ng = hi - lo
        if ng == 0:
            continue
        d = (p.unsqueeze(1) - p.unsqueeze(0)).pow(2).sum(-1)  # [centre, cand]
        kq = min(kk, ng)
        # stable ascending sort == ties broken by lower index
        order = torch.sort(d, dim=1, stable=True).indices[:, :kq]
        centre = torch.arange(ng).unsqueeze(1).expand(-1, kq)
        s, c = order.reshape(-1), centre.reshape(-1)
        if not loop:
            keep = s != c
            s, c = s[keep], c[keep]
        srcs.append(s + lo)
        dsts.append(c + lo)
    if not srcs:
        return torch.zeros(2, 0, dtype=torch.long)
    return torch.stack([torch.cat(srcs), torch.cat(dsts)])


def radius_graph(pos: Tensor, r: float, batch: Optional[Tensor] = None, loop: bool = False,
                 max_num_neighbors: int = 32) -> Tensor:
    """torch_cluster.radius_graph: all j with |p_j - p_i| < r ... (<= max_num_neighbors per centre, index order)."""
    n = pos.size(0)
    if batch is None:
        batch = torch.zeros(n, dtype=torch.long)
    B = int(batch.max()) + 1 if n > 0 else 0
    ptr = batch_to_ptr(batch, B)
    srcs, dsts = [], []
    for g in range(B):
        lo, hi = int(ptr[g]), int(ptr[g + 1])
        p = pos[lo:hi]
        d = (p.unsqueeze(1) - p.unsqueeze(0)).pow(2).sum(-1)
        m = d < r * r
        if not loop:
            m &= ~torch.eye(hi - lo, dtype=torch.bool)
        rank = m.long().cumsum(1)
        m &= rank <= max_num_neighbors
        c, s = m.nonzero(as_tuple=True)
        srcs.append(s + lo)
        dsts.append(c + lo)
    if not srcs:
        return torch.zeros(2, 0, dtype=torch.long)
    return torch.stack([torch.cat(srcs), torch.cat(dsts)])


def gaussian_distance(edge_index: Tensor, pos: Tensor, sigma: float, dtype=torch.float32) -> Tensor:
    """reference transforms.py:50-58: exp(-d^2 / 2 sigma^2) / sqrt(2 pi sigma^2)."""
    row, col = edge_index[0], edge_index[1]
    sq_dist = (pos[row] - pos[col]).pow(2).sum(dim=-1)
    return (torch.exp(-sq_dist / (2 * sigma ** 2)) / math.sqrt(2 * math.pi * sigma ** 2)).to(dtype)


def to_csr(edge_index: Tensor, num_nodes: int):
    """torch_sparse.SparseTensor(row=edge_index[1], col=edge_index[0]) storage
    (reference datamodule.py:45 `ToSparseTensor`): stable sort on key
    row * N + col.  Returns (rowptr[N+1], col[E], perm[E]) with
    col = src[perm]; value = edge_weight[perm]."""
    src, dst = edge_index[0], edge_index[1]
    key = dst * num_nodes + src
    perm = torch.sort(key, stable=True).indices
    col = src[perm]
    rowptr = torch.zeros(num_nodes + 1, dtype=torch.long)
    rowptr[1:] = torch.bincount(dst, minlength=num_nodes).cumsum(0)
    return rowptr, col, perm


def to_csc(rowptr: Tensor, col: Tensor, num_nodes: int):
    """Transpose index for the backward pass: edges grouped by src, ordered by
    CSR slot.  Returns (colptr[N+1], row[E], csr2csc[E]): CSC slot q holds the
    CSR slot csr2csc[q] whose dst is row[q]."""
    E = col.numel()
    deg = rowptr[1:] - rowptr[:-1]
    row_of_slot = torch.repeat_interleave(torch.arange(num_nodes), deg)
    csr2csc = torch.sort(col, stable=True).indices
    colptr = torch.zeros(num_nodes + 1, dtype=torch.long)
    colptr[1:] = torch.bincount(col, minlength=num_nodes).cumsum(0)
    assert csr2csc.numel() == E
    return colptr, row_of_slot[csr2csc], csr2csc


# ----------------------------------------------------------------------------
# Model shells mirroring the reference's L1 modules (gat.py:17-59, gin.py:17-35,
# drgnet.py:16-69) plus the GCN / SAGE graph classifiers BASELINE.json adds.
# ----------------------------------------------------------------------------


class GAT(nn.Module):
    def __init__(self, input_features, hiddden_channels, num_classes, heads, dropout, num_st_seed_points=None):
        super().__init__()
        assert num_st_seed_points is None
        assert all(d % heads == 0 for d in hiddden_channels)
        self.in_proj = nn.Linear(input_features, hiddden_channels[0])
        self.convs = nn.ModuleList(
            [GATConv(d1, d2 // heads, heads=heads, dropout=dropout) for d1, d2 in pairwise(hiddden_channels)]
        )
        self.out_proj = nn.Linear(hiddden_channels[-1], num_classes)

    def forward(self, x, edge_index, batch, num_graphs=None):
        x = self.in_proj(x)
        for conv in self.convs:
            x = F.elu(conv(x, edge_index))
        x = global_mean_pool(x, batch, num_graphs)
        return self.out_proj(x)


class GIN(nn.Module):
    def __init__(self, input_features, hidden_channels, num_classes, dropout, pooling="mean"):
        super().__init__()
        self.in_proj = nn.Linear(input_features, hidden_channels[0])
        self.convs = nn.ModuleList(
            [GINConv(MLP([d1, d2, d2], act="ELU", dropout=dropout)) for d1, d2 in pairwise(hidden_channels)]
        )
        self.pooling = pooling
        mult = 2 if pooling == "mean+max" else 1
        self.out_proj = nn.Linear(mult * hidden_channels[-1], num_classes)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x, edge_index, batch, num_graphs=None):
        x = self.in_proj(x)
        for conv in self.convs:
            x = F.elu(conv(x, edge_index))
            x = self.dropout(x)
        if self.pooling == "mean+max":
            x = torch.cat([global_mean_pool(x, batch, num_graphs), global_max_pool(x, batch, num_graphs)], dim=1)
        else:
            x = global_mean_pool(x, batch, num_graphs)
        return self.out_proj(x)


class SAGE(nn.Module):
    """BASELINE.json config 3 (SAGE variant): 3 x SAGEConv(mean) + ELU, mean||max pooling."""

    def __init__(self, input_features, hidden_channels, num_classes, aggr="mean", pooling="mean+max"):
        super().__init__()
        dims = [input_features] + list(hidden_channels)
        self.convs = nn.ModuleList([SAGEConv(d1, d2, aggr=aggr) for d1, d2 in pairwise(dims)])
        self.pooling = pooling
        mult = 2 if pooling == "mean+max" else 1
        self.out_proj = nn.Linear(mult * hidden_channels[-1], num_classes)

    def forward(self, x, edge_index, batch, num_graphs=None):
        for conv in self.convs:
            x = F.elu(conv(x, edge_index))
        if self.pooling == "mean+max":
            x = torch.cat([global_mean_pool(x, batch, num_graphs), global_max_pool(x, batch, num_graphs)], dim=1)
        else:
            x = global_mean_pool(x, batch, num_graphs)
        return self.out_proj(x)


class GCN(nn.Module):
    """BASELINE.json config 1 / 4: n x (GCNConv + ELU) -> global_mean_pool -> Linear."""

    def __init__(self, input_features, hidden_channels, num_classes):
        super().__init__()
        dims = [input_features] + list(hidden_channels)
        self.convs = nn.ModuleList([GCNConv(d1, d2) for d1, d2 in pairwise(dims)])
        self.out_proj = nn.Linear(hidden_channels[-1], num_classes)

    def forward(self, x, edge_index, batch, num_graphs=None, edge_weight=None):
        for conv in self.convs:
            x = F.elu(conv(x, edge_index, edge_weight))
        x = global_mean_pool(x, batch, num_graphs)
        return self.out_proj(x)


class DRGNet(nn.Module):
    """reference drgnet.py:16-69."""

    def __init__(self, input_features, gnn_hidden_dim, num_layers, sortpool_k, num_classes, conv_hidden_dims=(16, 32)):
        super().__init__()
        gnn_dims = [input_features] + [gnn_hidden_dim] * num_layers
        self.graph_convs = nn.ModuleList([GraphConv(a, b) for a, b in pairwise(gnn_dims)])
        self.graph_convs.append(GraphConv(gnn_hidden_dim, 1))
        total_latent_dim = gnn_hidden_dim * num_layers + 1
        self.sortpool_k = sortpool_k
        kernel_size = 5
        self.conv1 = nn.Conv1d(1, conv_hidden_dims[0], kernel_size=total_latent_dim, stride=total_latent_dim)
        self.max_pool = nn.MaxPool1d(2, 2)
        self.conv2 = nn.Conv1d(conv_hidden_dims[0], conv_hidden_dims[1], kernel_size=kernel_size, stride=1)
        dense_dim = int((sortpool_k - 2) / 2 + 1)
        dense_dim = (dense_dim - kernel_size + 1) * conv_hidden_dims[1]
        self.mlp = MLP([dense_dim, 128, num_classes], dropout=0.5, norm=None, act=F.elu)

    def forward(self, x, edge_index, batch, edge_weight=None, num_graphs=None):
        xs = []
        for graph_conv in self.graph_convs:
            x = F.elu(graph_conv(x, edge_index, edge_weight))
            xs.append(x)
        x_cat = torch.cat(xs, dim=1)
        x = sort_aggregation(x_cat, batch, self.sortpool_k, num_graphs).unsqueeze(1)
        x = F.elu(self.conv1(x))
        x = self.max_pool(x)
        x = F.elu(self.conv2(x))
        x = x.view(x.size(0), -1)
        return self.mlp(x)


def training_step(model: nn.Module, optimizer, x, edge_index, batch, y, num_graphs, loss_type="MSE",
                  num_classes=5, edge_weight=None):
    """reference models/base.py:196-201 (+ optimiser step, base.py:162-170) and
    the regression clamp of gat.py:94-95."""
    optimizer.zero_grad(set_to_none=True)
    if edge_weight is not None:
        logits = model(x, edge_index, batch, edge_weight=edge_weight, num_graphs=num_graphs)
    else:
        logits = model(x, edge_index, batch, num_graphs=num_graphs)
    if loss_type == "MSE":
        logits = torch.clamp(logits.squeeze(1), min=0, max=num_classes - 1)
        loss = F.mse_loss(logits, y.to(logits.dtype))
    else:
        loss = F.cross_entropy(logits, y)
    loss.backward()
    optimizer.step()
    return loss.detach()
