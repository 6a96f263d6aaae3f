"""The reference's L1 modules on the lgk layers: same constructor arguments, attribute names (state_dict keys)
and `forward(x, edge_index, batch)` signature as reference src/lesion_gnn/models/gat.py:17-59 and
gin.py:17-35, plus the GCN / GraphSAGE graph classifiers BASELINE.json's configs 1, 3 and 4 name.

`edge_index` may be the `[2, E]` int64 tensor or a `CSRGraph` (the reference's `adj_t`, gat.py:87-90); `batch`
may be the int64 graph-id vector or a `Segments` object.  The in_proj / out_proj transforms are
`lesion_gnn_b200.nn.Linear` (parameter names `weight`, `bias` as in torch.nn.Linear, torch's default init).
"""
from __future__ import annotations

import math
from itertools import pairwise

import os

import torch
import torch.nn as nn
from torch import Tensor

from ._lib import ACT_ELU
from .nn import (GATConv, GCNConv, GINConv, GraphConv, MLP, PointNetConv, SAGEConv, SortAggregation, Linear, elu, fps,
                 global_max_pool, global_mean_max_pool, global_mean_pool, linear, radius)


_FUSED_ELU = os.environ.get("LGK_FUSED_ELU", "1") != "0"  # profiling knob: 0 = separate lgk_elu_fwd launch after every GAT layer

class _TorchInitLinear(Linear):
    """torch.nn.Linear's initialisation (the reference's in_proj / out_proj are torch.nn.Linear)."""

    def reset_parameters(self):
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1 / math.sqrt(self.in_channels)
            nn.init.uniform_(self.bias, -bound, bound)


def _readout(x: Tensor, batch, pooling: str) -> Tensor:
    if pooling == "mean":
        return global_mean_pool(x, batch)
    if pooling == "max":
        return global_max_pool(x, batch)
    if pooling == "mean+max":
        return global_mean_max_pool(x, batch)
    raise ValueError(pooling)


class GAT(nn.Module):
    """reference models/gat.py:17-59 (num_st_seed_points=None branch)."""

    def __init__(self, input_features: int, hiddden_channels: list[int], num_classes: int, heads: int, dropout: float,
                 num_st_seed_points: int | None = None):
        super().__init__()
        if num_st_seed_points is not None:
            raise NotImplementedError("SetTransformerAggregation is outside the hot path (SURVEY.md 2)")
        assert all(d % heads == 0 for d in hiddden_channels)
        self.in_proj = _TorchInitLinear(input_features, hiddden_channels[0])
        self.convs = nn.ModuleList(
            [GATConv(d1, d2 // heads, heads=heads, dropout=dropout, fused_elu=_FUSED_ELU) for d1, d2 in pairwise(hiddden_channels)]
        )  # fused_elu: the F.elu of gat.py:51 is written by the attention kernel (one launch less per layer)
        self.st = None
        self.out_proj = _TorchInitLinear(hiddden_channels[-1], num_classes)

    def forward(self, x: Tensor, edge_index, batch) -> Tensor:
        x = self.in_proj(x)
        for conv in self.convs:
            x = conv(x, edge_index) if conv.fused_elu else elu(conv(x, edge_index))  # F.elu(GATConv(x)) of the reference
        x = global_mean_pool(x, batch)
        return self.out_proj(x)


class GIN(nn.Module):
    """reference models/gin.py:17-35; pooling='mean+max' is BASELINE.json config 3's read-out."""

    def __init__(self, input_features: int, hidden_channels: list[int], num_classes: int, dropout: float,
                 pooling: str = "mean"):
        super().__init__()
        assert all(d > 0 for d in hidden_channels)
        self.in_proj = _TorchInitLinear(input_features, hidden_channels[0])
        self.convs = nn.ModuleList(
            [GINConv(MLP([d1, d2, d2], act="ELU", dropout=dropout)) for d1, d2 in pairwise(hidden_channels)]
        )
        self.pooling = pooling
        mult = 2 if pooling == "mean+max" else 1
        self.out_proj = _TorchInitLinear(mult * hidden_channels[-1], num_classes)
        from .nn.functional import ActDropout
        self.dropout = ActDropout(ACT_ELU, dropout)  # `F.elu(...)` then `self.dropout(...)` of gin.py:31-32 (no parameters)

    def forward(self, x: Tensor, edge_index, batch) -> Tensor:
        x = self.in_proj(x)
        for conv in self.convs:  # F.elu + nn.Dropout of gin.py:31-32 as one launch
            x = self.dropout(conv(x, edge_index))
        return self.out_proj(_readout(x, batch, self.pooling))


class SAGE(nn.Module):
    """BASELINE.json config 3, GraphSAGE variant: n x (SAGEConv + ELU) -> mean||max -> Linear."""

    def __init__(self, input_features: int, hidden_channels: list[int], num_classes: int, aggr: str = "mean",
                 pooling: str = "mean+max"):
        super().__init__()
        dims = [input_features] + list(hidden_channels)
        self.convs = nn.ModuleList([SAGEConv(d1, d2, aggr=aggr) for d1, d2 in pairwise(dims)])
        self.pooling = pooling
        mult = 2 if pooling == "mean+max" else 1
        self.out_proj = _TorchInitLinear(mult * hidden_channels[-1], num_classes)

    def forward(self, x: Tensor, edge_index, batch) -> Tensor:
        for conv in self.convs:
            x = elu(conv(x, edge_index))
        return self.out_proj(_readout(x, batch, self.pooling))


class GCN(nn.Module):
    """BASELINE.json configs 1 and 4: n x (GCNConv + ELU) -> global_mean_pool -> Linear."""

    def __init__(self, input_features: int, hidden_channels: list[int], num_classes: int):
        super().__init__()
        dims = [input_features] + list(hidden_channels)
        self.convs = nn.ModuleList([GCNConv(d1, d2) for d1, d2 in pairwise(dims)])
        self.out_proj = _TorchInitLinear(hidden_channels[-1], num_classes)

    def forward(self, x: Tensor, edge_index, batch, edge_weight=None) -> Tensor:
        for conv in self.convs:
            x = elu(conv(x, edge_index, edge_weight))
        return self.out_proj(global_mean_pool(x, batch))


class DRGNet(nn.Module):
    """reference models/drgnet.py:16-69: `num_layers` GraphConv (+ one GraphConv to a single channel) with ELU, the
    layer outputs concatenated, SortAggregation(k) read-out, then the reference's small 1-D CNN and MLP head.  The
    message passing and the read-out run on the lgk kernels; Conv1d / MaxPool1d act on the [B, 1, k * D] read-out
    and stay torch modules (they are not message passing: SURVEY.md 8f row 2)."""

    def __init__(self, input_features: int, gnn_hidden_dim: int, num_layers: int, sortpool_k: int, num_classes: int,
                 conv_hidden_dims=(16, 32)):
        super().__init__()
        gnn_dims = [input_features] + [gnn_hidden_dim] * num_layers
        self.graph_convs = nn.ModuleList(GraphConv(a, b) for a, b in pairwise(gnn_dims))
        self.graph_convs.append(GraphConv(gnn_hidden_dim, 1))
        total_latent_dim = gnn_hidden_dim * num_layers + 1
        self.sort_pool = SortAggregation(sortpool_k)
        kernel_size = 5
        self.conv1 = nn.Conv1d(1, conv_hidden_dims[0], kernel_size=total_latent_dim, stride=total_latent_dim)
        self.max_pool = nn.MaxPool1d(2, 2)
        self.conv2 = nn.Conv1d(conv_hidden_dims[0], conv_hidden_dims[1], kernel_size=kernel_size, stride=1)
        dense_dim = int((sortpool_k - 2) / 2 + 1)
        dense_dim = (dense_dim - kernel_size + 1) * conv_hidden_dims[1]
        self.mlp = MLP([dense_dim, 128, num_classes], dropout=0.5, norm=None, act="ELU")

    def forward(self, x: Tensor, edge_index, batch, edge_weight=None) -> Tensor:
        xs = []
        for graph_conv in self.graph_convs:
            x = elu(graph_conv(x, edge_index, edge_weight))
            xs.append(x)
        x = self.sort_pool(torch.cat(xs, dim=1), batch)  # [B, k * D]
        # conv1 has kernel = stride = D over the k sorted rows: a D -> C1 transform of every row; conv2 (kernel 5,
        # stride 1) is a transform of the unfolded windows.  Both run on the fp32 lgk transform kernels (cuDNN
        # convolutions default to TF32, which breaks the 1e-5 bar); parameters keep Conv1d's names and shapes.
        B, k = x.size(0), self.sort_pool.k
        c1, c2 = self.conv1, self.conv2
        x = linear(x.reshape(B * k, -1), c1.weight.reshape(c1.out_channels, -1), c1.bias, act=ACT_ELU)
        x = self.max_pool(x.reshape(B, k, -1).transpose(1, 2))  # [B, C1, k // 2]
        ks = c2.kernel_size[0]
        win = x.unfold(2, ks, 1).permute(0, 2, 1, 3)  # [B, L, C1, ks]
        L = win.size(1)
        x = linear(win.reshape(B * L, -1), c2.weight.reshape(c2.out_channels, -1), c2.bias, act=ACT_ELU)
        x = x.reshape(B, L, -1).transpose(1, 2)  # [B, C2, L], the layout the reference flattens
        return self.mlp(x.reshape(B, -1))


class SAModule(nn.Module):
    """reference models/pointnet.py:13-29: farthest point sampling, radius query, PointNetConv (max of an edge MLP)"""

    def __init__(self, ratio: float, radius_: float, nn_: nn.Module, random_start: bool = True):
        super().__init__()
        self.ratio, self.r, self.random_start = ratio, radius_, random_start
        self.conv = PointNetConv(nn_, add_self_loops=False)

    def forward(self, x: Tensor, pos: Tensor, batch: Tensor):
        pos = pos.to(x.dtype)
        idx = fps(pos, batch, ratio=self.ratio, random_start=self.random_start)
        row, col = radius(pos, pos[idx], self.r, batch, batch[idx], max_num_neighbors=64)
        edge_index = torch.stack([col, row], dim=0)
        x_dst = None if x is None else x[idx]
        x = self.conv((x, x_dst), (pos, pos[idx]), edge_index)
        return x, pos[idx], batch[idx]


class GlobalSAModule(nn.Module):
    """reference models/pointnet.py:32-40"""

    def __init__(self, nn_: nn.Module):
        super().__init__()
        self.nn = nn_

    def forward(self, x: Tensor, pos: Tensor, batch: Tensor, size=None) -> Tensor:
        x = self.nn(torch.cat([x, pos], dim=1))
        return global_max_pool(x, batch, size)


class PointNet(nn.Module):
    """reference models/pointnet.py:43-59 on the lgk kernels.  `radii` (the reference hard-codes 0.2 / 0.4, i.e. unit-cube
    coordinates) and `random_start` (torch_cluster's default True; False for reproducible sampling) are extra arguments."""

    def __init__(self, input_features: int, pos_dim: int, num_classes: int, radii=(0.2, 0.4), random_start: bool = True):
        super().__init__()
        self.sa1_module = SAModule(0.5, radii[0], MLP([input_features + pos_dim, 64, 64, 128]), random_start)
        self.sa2_module = SAModule(0.25, radii[1], MLP([128 + pos_dim, 128, 128, 256]), random_start)
        self.sa3_module = GlobalSAModule(MLP([256 + pos_dim, 256, 512, 1024]))
        self.mlp = MLP([1024, 512, 256, num_classes], dropout=0.5, norm=None)

    def forward(self, x: Tensor, pos: Tensor, batch: Tensor) -> Tensor:
        size = int(batch[-1].item()) + 1 if batch.numel() else 0
        x, pos, batch = self.sa1_module(x, pos, batch)
        x, pos, batch = self.sa2_module(x, pos, batch)
        x = self.sa3_module(x, pos, batch, size)
        return self.mlp(x).log_softmax(dim=-1)
