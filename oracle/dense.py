"""Independent dense-matrix formulations of the same layers (TEST INFRASTRUCTURE).

Second opinion for oracle/pyg_semantics.py: every layer is re-derived from its
textbook matrix formula on a dense [N, N] adjacency so that a slip in the
gather/scatter restatement (self-loop handling, normalisation side, softmax
axis) shows up as a disagreement.  Small graphs only (O(N^2) memory).
A[i, j] = multiplicity/weight of edge j -> i.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def dense_adj(edge_index, num_nodes, edge_weight=None, dtype=torch.float64):
    A = torch.zeros(num_nodes, num_nodes, dtype=dtype)
    w = torch.ones(edge_index.size(1), dtype=dtype) if edge_weight is None else edge_weight.to(dtype)
    A.index_put_((edge_index[1], edge_index[0]), w, accumulate=True)
    return A


def gcn_dense(x, edge_index, W, b, edge_weight=None, improved=False):
    """D^-1/2 (A' ) D^-1/2 X W^T + b where A' = A with its diagonal replaced by
    (existing loop weight, else fill)."""
    n = x.size(0)
    src, dst = edge_index
    w = torch.ones(src.numel(), dtype=x.dtype) if edge_weight is None else edge_weight.to(x.dtype)
    nl = src != dst
    A = torch.zeros(n, n, dtype=x.dtype)
    A.index_put_((dst[nl], src[nl]), w[nl], accumulate=True)
    diag = torch.full((n,), 2.0 if improved else 1.0, dtype=x.dtype)
    for e in torch.nonzero(~nl).flatten().tolist():  # last existing loop wins
        diag[src[e]] = w[e]
    A = A + torch.diag(diag)
    deg = A.sum(dim=1)
    dinv = torch.where(deg > 0, deg.pow(-0.5), torch.zeros_like(deg))
    out = (dinv[:, None] * A * dinv[None, :]) @ (x @ W.t())
    return out if b is None else out + b


def gat_dense(x, edge_index, W, att_src, att_dst, b, heads, negative_slope=0.2):
    """Masked dense softmax over [N, N, H]; multi-edges counted with multiplicity."""
    n = x.size(0)
    H = heads
    C = W.size(0) // H
    h = (x @ W.t()).view(n, H, C)
    a_s = (h * att_src.view(1, H, C)).sum(-1)  # [N, H]
    a_d = (h * att_dst.view(1, H, C)).sum(-1)
    src, dst = edge_index
    nl = src != dst
    M = torch.zeros(n, n, dtype=x.dtype)
    M.index_put_((dst[nl], src[nl]), torch.ones(int(nl.sum()), dtype=x.dtype), accumulate=True)
    M = M + torch.eye(n, dtype=x.dtype)  # exactly one self loop per node
    e = F.leaky_relu(a_d.unsqueeze(1) + a_s.unsqueeze(0), negative_slope)  # [i, j, H]
    e = e.masked_fill((M == 0).unsqueeze(-1), float("-inf"))
    m = e.max(dim=1, keepdim=True).values
    p = (e - m).exp() * M.unsqueeze(-1)
    alpha = p / (p.sum(dim=1, keepdim=True) + 1e-16)
    out = torch.einsum("ijh,jhc->ihc", alpha, h).reshape(n, H * C)
    return out if b is None else out + b


def aggregate_dense(x, edge_index, aggr, edge_weight=None):
    n = x.size(0)
    A = dense_adj(edge_index, n, edge_weight, x.dtype)
    if aggr in ("sum", "add"):
        return A @ x
    if aggr == "mean":
        cnt = dense_adj(edge_index, n, None, x.dtype).sum(1).clamp(min=1)
        return (A @ x) / cnt[:, None]
    if aggr == "max":
        mask = dense_adj(edge_index, n, None, x.dtype) > 0
        vals = x.unsqueeze(0).expand(n, n, -1).masked_fill(~mask.unsqueeze(-1), float("-inf"))
        out = vals.max(dim=1).values
        return torch.where(torch.isinf(out), torch.zeros_like(out), out)
    raise ValueError(aggr)


def pool_dense(x, batch, num_graphs, reduce):
    onehot = F.one_hot(batch, num_graphs).to(x.dtype)  # [N, B]
    if reduce in ("sum", "add"):
        return onehot.t() @ x
    if reduce == "mean":
        return (onehot.t() @ x) / onehot.sum(0).clamp(min=1)[:, None]
    if reduce == "max":
        vals = x.unsqueeze(0).expand(num_graphs, -1, -1).masked_fill(~(onehot.t().bool()).unsqueeze(-1), float("-inf"))
        out = vals.max(dim=1).values
        return torch.where(torch.isinf(out), torch.zeros_like(out), out)
    raise ValueError(reduce)
