"""CPU restatement of the PointNet++ operators the reference's models/pointnet.py uses.

TEST INFRASTRUCTURE (see oracle/__init__.py) -- parity unpinned: torch-cluster 1.6.3 / torch-geometric 2.5.1 are absent
(oracle/PROVENANCE.md).  Plain loops / torch ops on the CPU, dtype-generic:
  fps            torch_cluster/fps.py: fps -> fps_cpu / fps_cuda (models/pointnet.py:22)
  radius         torch_cluster/radius.py: radius (models/pointnet.py:23)
  PointNetConv   torch_geometric/nn/conv/point_conv.py (models/pointnet.py:18,26)
  SAModule, GlobalSAModule, PointNet   the reference itself: models/pointnet.py:13-59
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor

from . import pyg_semantics as O


def fps(pos: Tensor, batch: Tensor, ratio: float, random_start: bool = False, start=None) -> Tensor:
    """per graph: ceil(ratio * n) samples (computed in the dtype of pos, as torch_cluster does); start at the first node
    (or `start[g]`), then repeatedly the node with the largest distance to the chosen set, ties to the lower index"""
    out = []
    B = int(batch.max()) + 1 if batch.numel() else 0
    for g in range(B):
        idx = (batch == g).nonzero().flatten()
        n = idx.numel()
        if n == 0:
            continue
        m = int((torch.tensor(float(n), dtype=pos.dtype) * torch.tensor(ratio, dtype=pos.dtype)).ceil())
        p = pos[idx]
        cur = 0 if start is None else int(start[g])
        dist = torch.full((n,), float("inf"), dtype=pos.dtype)
        for _ in range(m):
            out.append(int(idx[cur]))
            d = (p - p[cur]).pow(2).sum(-1)
            dist = torch.minimum(dist, d)
            cur = int(torch.argmax(dist))  # first maximum
            best = dist.max()
            cur = int((dist == best).nonzero()[0])
    return torch.tensor(out, dtype=torch.long)


def radius(x: Tensor, y: Tensor, r: float, batch_x: Tensor, batch_y: Tensor, max_num_neighbors: int = 32) -> Tensor:
    """[2, E]: row 0 = query index into y, row 1 = point index into x; strict <, ascending point index, capped"""
    rows, cols = [], []
    r2 = torch.tensor(r, dtype=x.dtype) ** 2
    for i in range(y.size(0)):
        cand = (batch_x == batch_y[i]).nonzero().flatten()
        d = (x[cand] - y[i]).pow(2).sum(-1)
        hit = cand[d < r2][:max_num_neighbors]
        rows += [i] * hit.numel()
        cols += hit.tolist()
    return torch.tensor([rows, cols], dtype=torch.long).reshape(2, -1)


class PointNetConv(nn.Module):
    def __init__(self, local_nn=None, global_nn=None, add_self_loops=False):
        super().__init__()
        assert not add_self_loops
        self.local_nn, self.global_nn = local_nn, global_nn

    def forward(self, x, pos, edge_index):
        x_src = x[0] if isinstance(x, tuple) else x
        pos_src, pos_dst = pos if isinstance(pos, tuple) else (pos, pos)
        src, dst = edge_index[0], edge_index[1]
        msg = pos_src[src] - pos_dst[dst]
        if x_src is not None:
            msg = torch.cat([x_src[src], msg], dim=1)
        if self.local_nn is not None:
            msg = self.local_nn(msg)
        out = O.scatter(msg, dst, pos_dst.size(0), "max")
        return self.global_nn(out) if self.global_nn is not None else out


class SAModule(nn.Module):
    """reference models/pointnet.py:13-29"""

    def __init__(self, ratio, r, nn_):
        super().__init__()
        self.ratio, self.r = ratio, r
        self.conv = PointNetConv(nn_, add_self_loops=False)

    def forward(self, x, pos, batch):
        pos = pos.to(x.dtype)
        idx = fps(pos, batch, self.ratio)
        row, col = radius(pos, pos[idx], self.r, batch, batch[idx], max_num_neighbors=64)
        edge_index = torch.stack([col, row], dim=0)
        x = self.conv((x, x[idx]), (pos, pos[idx]), edge_index)
        return x, pos[idx], batch[idx]


class GlobalSAModule(nn.Module):
    """reference models/pointnet.py:32-40"""

    def __init__(self, nn_):
        super().__init__()
        self.nn = nn_

    def forward(self, x, pos, batch, num_graphs=None):
        x = self.nn(torch.cat([x, pos], dim=1))
        return O.global_max_pool(x, batch, num_graphs)


class PointNet(nn.Module):
    """reference models/pointnet.py:43-59 (radii are constructor arguments here: the reference's 0.2 / 0.4 assume unit-cube
    coordinates)"""

    def __init__(self, input_features, pos_dim, num_classes, radii=(0.2, 0.4)):
        super().__init__()
        self.sa1_module = SAModule(0.5, radii[0], O.MLP([input_features + pos_dim, 64, 64, 128], act="relu"))
        self.sa2_module = SAModule(0.25, radii[1], O.MLP([128 + pos_dim, 128, 128, 256], act="relu"))
        self.sa3_module = GlobalSAModule(O.MLP([256 + pos_dim, 256, 512, 1024], act="relu"))
        self.mlp = O.MLP([1024, 512, 256, num_classes], dropout=0.5, norm=None, act="relu")

    def forward(self, x, pos, batch, num_graphs=None):
        x, pos, batch = self.sa1_module(x, pos, batch)
        x, pos, batch = self.sa2_module(x, pos, batch)
        x = self.sa3_module(x, pos, batch, num_graphs)
        return self.mlp(x).log_softmax(dim=-1)
