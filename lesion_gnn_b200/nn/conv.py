"""Message-passing layers with PyG 2.5.1's constructor arguments, parameter names and forward signatures.

Drop-in for `torch_geometric.nn.{GCNConv, GATConv, GINConv, SAGEConv, GraphConv, MLP}` at the reference's call
sites (models/gat.py:31,51; gin.py:23,31; drgnet.py:31,33,55): `conv(x, edge_index[, edge_weight])` where
`edge_index` is the `[2, E]` int64 tensor or a `CSRGraph` (the `adj_t` stand-in).  state_dict keys match PyG's
(`lin.weight`, `att_src`, `att_dst`, `bias`; `lin_l.*`, `lin_r.weight`; `lin_rel.*`, `lin_root.weight`;
`nn.lins.*`, `nn.norms.*.module.*`, `eps`).  CUDA only: inputs on the CPU raise.
"""
from __future__ import annotations

import os

import math
from itertools import pairwise
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as TF
from torch import Tensor

from .._lib import ACT_ELU, ACT_NONE, ACT_RELU, SELF_ADD, SELF_NONE, SELF_REPLACE
from ..compile import OpaqueModule
from ..graph import CSRGraph, as_graph
from . import functional as F

# The attention-logit halves can come out of the transform's epilogue (lgk_linear_gat_fwd) instead of a separate
# lgk_gat_alpha launch over the stored h.  bf16 activations: always (the logits must not see h's rounding: 4x smaller
# gradient error, tests/test_gpu_bf16.py).  fp32: on by default since the main chain of the captured step runs on a
# high-priority stream (train.TrainStep): headline step 0.3115 -> 0.3091 ms, 53 -> 50 launches (B200, 500 steps, same box,
# A/B/A/B).  Before that the separate launch was the faster one (0.3267 vs 0.3312 ms: the transform's 4-warp epilogue is
# fully exposed, one tile per CTA); LGK_GAT_FUSED_ALPHA=0 brings it back.
_FUSED_ALPHA_F32 = os.environ.get("LGK_GAT_FUSED_ALPHA", "1") != "0"
# profiling knob: 0 = GCNConv applies the source-side d^-1/2 inside the aggregation (per-entry weights) instead of in the
# transform's epilogue
_GCN_FUSED_SCALE = os.environ.get("LGK_GCN_FUSED_SCALE", "1") != "0"
_seed_counter = [0]


def _next_seed() -> int:
    """seed for in-kernel attention dropout: derived from torch's global seed and a call counter"""
    _seed_counter[0] += 1
    return (torch.initial_seed() * 0x9E3779B1 + _seed_counter[0] * 0x85EBCA6B) & 0xFFFFFFFFFFFFFFFF


# A captured CUDA graph replays the launch arguments it recorded, so a host seed alone would freeze the attention-dropout
# mask of every replay.  While a clock is installed (train.TrainStep installs the optimiser's device step counter), the
# kernels mix its current value into the seed at run time: forward and backward of one step agree, the next step differs.
_DROPOUT_CLOCK = [None]


class dropout_clock:
    """context manager: `with dropout_clock(step_counter_int64_device_tensor): ...`"""

    def __init__(self, counter: Optional[Tensor]):
        if counter is not None and (counter.dtype != torch.int64 or not counter.is_cuda or counter.numel() < 1):
            raise TypeError("the dropout clock is a CUDA int64 tensor")
        self.counter = counter

    def __enter__(self):
        self._old = _DROPOUT_CLOCK[0]
        _DROPOUT_CLOCK[0] = self.counter
        return self

    def __exit__(self, *exc):
        _DROPOUT_CLOCK[0] = self._old
        return False


def _glorot(t: Tensor) -> None:
    a = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
    with torch.no_grad():
        t.uniform_(-a, a)


class Linear(OpaqueModule, nn.Module):
    """torch_geometric.nn.dense.linear.Linear: weight [out, in]; runs on the lgk transform kernels."""

    def __init__(self, in_channels: int, out_channels: int, bias: bool = True, weight_initializer: Optional[str] = None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels))
        self.bias = nn.Parameter(torch.empty(out_channels)) if bias else None
        self.weight_initializer = weight_initializer
        self.reset_parameters()
        self._register_opaque()

    def reset_parameters(self):
        if self.weight_initializer == "glorot":
            _glorot(self.weight)
        else:  # PyG default: kaiming_uniform(a=sqrt(5)), i.e. torch.nn.Linear's
            nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            if self.weight_initializer == "glorot":
                nn.init.zeros_(self.bias)
            else:
                bound = 1 / math.sqrt(self.in_channels) if self.in_channels > 0 else 0
                nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x: Tensor, act: int = ACT_NONE) -> Tensor:
        if act != ACT_NONE:  # fused-activation form: used by MLP, itself an opaque layer
            return F.linear(x, self.weight, self.bias, act=act)
        return self._dispatch(x)

    def _forward_eager(self, x: Tensor) -> Tensor:
        return F.linear(x, self.weight, self.bias)

    def _fake(self, x: Tensor) -> Tensor:
        return x.new_empty(tuple(x.shape[:-1]) + (self.out_channels,))


class GCNConv(OpaqueModule, nn.Module):
    """D^-1/2 (A + I) D^-1/2 X W^T + b.  Transform first, then one fused aggregation in which
    the symmetric normalisation is applied on the fly from a per-node deg^-1/2 vector (no [E] weight array)."""

    def __init__(self, in_channels: int, out_channels: int, improved: bool = False, cached: bool = False,
                 add_self_loops: bool = True, normalize: bool = True, bias: bool = True):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.improved, self.cached, self.add_self_loops, self.normalize = improved, cached, add_self_loops, normalize
        self.lin = Linear(in_channels, out_channels, bias=False, weight_initializer="glorot")
        self.bias = nn.Parameter(torch.zeros(out_channels)) if bias else None
        self._register_opaque()

    def forward(self, x: Tensor, edge_index, edge_weight: Optional[Tensor] = None) -> Tensor:
        return self._dispatch(x, edge_index, edge_weight)

    def _fake(self, x: Tensor, edge_index=None, edge_weight=None) -> Tensor:
        return x.new_empty((x.size(0), self.out_channels))

    def _forward_eager(self, x: Tensor, edge_index, edge_weight: Optional[Tensor] = None) -> Tensor:
        # a graph whose build is still running on its own stream (the first layer of a captured step): transform first, it
        # does not read the graph and overlaps the build; as_graph joins the build stream
        pending = isinstance(edge_index, CSRGraph) and edge_index._ready is not None
        if (not pending and self.normalize and edge_weight is None and self.lin.bias is None and x.dim() == 2
                and x.dtype == torch.float32 and _GCN_FUSED_SCALE):
            g = as_graph(edge_index, x.size(0), edge_weight)
            if g.value is None:
                # unweighted graph: the source-side d^-1/2 rides in the transform's epilogue (free), the aggregation runs
                # without per-entry weights
                dinv, selfw = g.gcn_norm(False, self.add_self_loops)
                return F.gcn_conv(x, self.lin.weight, self.bias, g, dinv, selfw, 1.0,
                                  SELF_REPLACE if self.add_self_loops else SELF_NONE)
        h = self.lin(x)
        g = as_graph(edge_index, x.size(0), edge_weight)
        if not self.normalize:
            return F.aggregate(h, g, "sum", use_value=g.value is not None, bias=self.bias)
        # PyG 2.5.1's gcn_norm adds the remaining self loops BEFORE it defaults edge_weight to ones, so without
        # edge weights the added loops weigh 1 even when improved=True (fill_value=2 only reaches weighted graphs)
        improved = self.improved and g.value is not None
        dinv, selfw = g.gcn_norm(improved, self.add_self_loops)
        return F.aggregate(h, g, "sum", use_value=g.value is not None, pre=dinv, post=dinv, selfw=selfw,
                           self_fill=2.0 if improved else 1.0,
                           self_mode=SELF_REPLACE if self.add_self_loops else SELF_NONE, bias=self.bias)


class GATConv(OpaqueModule, nn.Module):
    """Multi-head graph attention (reference models/gat.py:31,51): transform, then ONE kernel
    for leaky_relu + per-destination softmax + attention dropout + weighted aggregation + bias."""

    def __init__(self, in_channels: int, out_channels: int, heads: int = 1, concat: bool = True,
                 negative_slope: float = 0.2, dropout: float = 0.0, add_self_loops: bool = True,
                 edge_dim: Optional[int] = None, fill_value="mean", bias: bool = True, fused_elu: bool = False):
        """fused_elu (not a PyG argument): the layer returns elu(output) -- the `F.elu(conv(x))` of the reference's
        models/gat.py:51 written by the attention kernel itself; concat=True only."""
        super().__init__()
        if fused_elu and not concat:
            raise ValueError("fused_elu needs concat=True (the head mean comes before the activation otherwise)")
        self.fused_elu = fused_elu
        if edge_dim is not None:
            raise NotImplementedError("edge features (edge_dim) are not on the reference's path")
        if not isinstance(in_channels, int):
            raise NotImplementedError("bipartite (tuple) in_channels is not on the reference's path")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.concat, self.negative_slope, self.dropout = concat, negative_slope, dropout
        self.add_self_loops = add_self_loops
        self.lin = Linear(in_channels, heads * out_channels, bias=False, weight_initializer="glorot")
        self.att_src = nn.Parameter(torch.empty(1, heads, out_channels))
        self.att_dst = nn.Parameter(torch.empty(1, heads, out_channels))
        if bias:
            self.bias = nn.Parameter(torch.zeros(heads * out_channels if concat else out_channels))
        else:
            self.bias = None
        _glorot(self.att_src)
        _glorot(self.att_dst)
        self._register_opaque()

    def forward(self, x: Tensor, edge_index, edge_attr=None, size=None, return_attention_weights=None) -> Tensor:
        if edge_attr is not None or return_attention_weights:
            raise NotImplementedError("edge_attr / return_attention_weights are not on the reference's path")
        return self._dispatch(x, edge_index)

    def _fake(self, x: Tensor, edge_index=None) -> Tensor:
        return x.new_empty((x.size(0), self.heads * self.out_channels if self.concat else self.out_channels))

    def _forward_eager(self, x: Tensor, edge_index) -> Tensor:
        # before as_graph: the transform does not read the graph, whose build may still be running on its own stream
        # (as_graph joins it).  The attention-logit halves fall out of the transform's epilogue.
        alpha = None
        if self.lin.bias is None and x.dim() == 2 and (x.dtype == torch.bfloat16 or _FUSED_ALPHA_F32):
            h, a_src, a_dst = F.linear_gat(x, self.lin.weight, self.att_src, self.att_dst, self.heads, self.out_channels)
            alpha = (a_src, a_dst)
        else:
            h = self.lin(x)
        g = as_graph(edge_index, x.size(0))
        p = self.dropout if self.training else 0.0
        seed = _next_seed() if p > 0 else 0
        fused_bias = self.bias if self.concat else None
        out = F.gat_attention(h, self.att_src, self.att_dst, fused_bias, g, self.heads, self.out_channels,
                              self.negative_slope, self.add_self_loops, p, seed, fused_elu=self.fused_elu,
                              seed_dev=_DROPOUT_CLOCK[0] if p > 0 else None, alpha=alpha)
        if not self.concat:
            out = out.view(-1, self.heads, self.out_channels).mean(dim=1)
            if self.bias is not None:
                out = out + self.bias
        return out


class BatchNorm(nn.Module):
    """torch_geometric.nn.norm.BatchNorm (BatchNorm1d kept in `.module` so checkpoints interchange)."""

    def __init__(self, in_channels: int, eps: float = 1e-5, momentum: float = 0.1, affine: bool = True):
        super().__init__()
        self.module = nn.BatchNorm1d(in_channels, eps, momentum, affine)

    def forward(self, x: Tensor) -> Tensor:
        return self.module(x)


class MLP(OpaqueModule, nn.Module):
    """torch_geometric.nn.models.MLP as the reference uses it: gin.py:23 `MLP([d1, d2, d2], act="ELU", dropout=p)`
    and drgnet.py:48 `MLP([dense, 128, C], dropout=0.5, norm=None, act=F.elu)`."""

    def __init__(self, channel_list, act="relu", dropout: float = 0.0, norm: Optional[str] = "batch_norm",
                 plain_last: bool = True, bias: bool = True):
        super().__init__()
        channel_list = list(channel_list)
        self.channel_list, self.plain_last = channel_list, plain_last
        n_lin = len(channel_list) - 1
        self.dropout = [dropout] * n_lin
        if plain_last:
            self.dropout[-1] = 0.0
        if act in ("ELU", "elu") or act is TF.elu:
            self.act = ACT_ELU
        elif act in ("relu", "ReLU") or act is TF.relu:  # PyG's default (reference models/pointnet.py:48-52)
            self.act = ACT_RELU
        else:
            raise NotImplementedError("only the ELU and ReLU activations are on the reference's path")
        self.lins = nn.ModuleList(Linear(a, b, bias=bias) for a, b in pairwise(channel_list))
        hidden = channel_list[1:-1] if plain_last else channel_list[1:]
        self.norms = nn.ModuleList(BatchNorm(d) if norm is not None else nn.Identity() for d in hidden)
        self._register_opaque()

    def forward(self, x: Tensor) -> Tensor:
        return self._dispatch(x)

    def _fake(self, x: Tensor) -> Tensor:
        return x.new_empty((x.size(0), self.channel_list[-1]))

    def _forward_eager(self, x: Tensor) -> Tensor:
        for i, (lin, norm) in enumerate(zip(self.lins, self.norms)):
            p = self.dropout[i] if self.training else 0.0
            if isinstance(norm, nn.Identity):
                if self.act == ACT_ELU and p == 0:
                    x = lin(x, act=ACT_ELU)  # ELU fused into the transform's epilogue
                else:
                    x = F.act_dropout(lin(x), self.act, p, _next_seed() if p > 0 else 0, _DROPOUT_CLOCK[0] if p > 0 else None)
            else:  # BatchNorm1d + activation + Dropout: lgk_bn_act_fwd / bwd (no ATen kernel inside the MLP)
                x = F.bn_act(lin(x), norm.module, self.act, p, _next_seed() if p > 0 else 0, _DROPOUT_CLOCK[0] if p > 0 else None)
        if self.plain_last:
            x = self.lins[-1](x)
            p = self.dropout[-1] if self.training else 0.0
            if p > 0:
                x = F.act_dropout(x, ACT_NONE, p, _next_seed(), _DROPOUT_CLOCK[0])
        return x


class GINConv(OpaqueModule, nn.Module):
    """nn((1 + eps) x_i + sum_j x_j) (reference models/gin.py:23,31): the (1+eps) x_i term is an
    extra self entry inside the aggregation kernel."""

    def __init__(self, nn: nn.Module, eps: float = 0.0, train_eps: bool = False):
        super().__init__()
        self.nn = nn
        self.initial_eps = eps
        if train_eps:
            self.eps = torch.nn.Parameter(torch.tensor([eps], dtype=torch.float32))
        else:
            self.register_buffer("eps", torch.tensor([eps], dtype=torch.float32))
        self._eps_host = float(eps)
        self.train_eps = train_eps
        self._register_opaque()

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
        if not self.train_eps:
            self._eps_host = float(self.eps.detach().cpu())

    def forward(self, x: Tensor, edge_index, size=None) -> Tensor:
        return self._dispatch(x, edge_index)

    def _fake(self, x: Tensor, edge_index=None) -> Tensor:
        if hasattr(self.nn, "_fake"):
            return self.nn._fake(x)
        last = [m for m in self.nn.modules() if isinstance(m, (Linear, nn.Linear))][-1]
        return x.new_empty((x.size(0), last.weight.size(0)))

    def _forward_eager(self, x: Tensor, edge_index) -> Tensor:
        g = as_graph(edge_index, x.size(0))
        if self.train_eps:
            out = F.aggregate(x, g, "sum") + (1 + self.eps) * x
        else:
            out = F.aggregate(x, g, "sum", self_fill=1.0 + self._eps_host, self_mode=SELF_ADD)
        return self.nn(out)


class SAGEConv(OpaqueModule, nn.Module):
    """W_l aggr_j x_j + b_l + W_r x_i: aggregate at input width, then both products in one
    accumulating transform."""

    def __init__(self, in_channels: int, out_channels: int, aggr: str = "mean", normalize: bool = False,
                 root_weight: bool = True, project: bool = False, bias: bool = True):
        super().__init__()
        if project:
            raise NotImplementedError("project=True is not on the reference's path")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.aggr, self.normalize, self.root_weight = aggr, normalize, root_weight
        self.lin_l = Linear(in_channels, out_channels, bias=bias)
        self.lin_r = Linear(in_channels, out_channels, bias=False) if root_weight else None
        self._register_opaque()

    def forward(self, x: Tensor, edge_index, size=None) -> Tensor:
        return self._dispatch(x, edge_index)

    def _fake(self, x: Tensor, edge_index=None) -> Tensor:
        return x.new_empty((x.size(0), self.out_channels))

    def _forward_eager(self, x: Tensor, edge_index) -> Tensor:
        g = as_graph(edge_index, x.size(0))
        agg = F.aggregate(x, g, self.aggr)
        if self.lin_r is not None:
            out = F.linear(agg, self.lin_l.weight, self.lin_l.bias, x2=x, weight2=self.lin_r.weight)
        else:
            out = self.lin_l(agg)
        if self.normalize:
            out = TF.normalize(out, p=2.0, dim=-1)
        return out


class GraphConv(OpaqueModule, nn.Module):
    """W_rel sum_j e_ji x_j + b + W_root x_i (reference models/drgnet.py:31,33,55)."""

    def __init__(self, in_channels: int, out_channels: int, aggr: str = "add", bias: bool = True):
        super().__init__()
        self.in_channels, self.out_channels, self.aggr = in_channels, out_channels, aggr
        self.lin_rel = Linear(in_channels, out_channels, bias=bias)
        self.lin_root = Linear(in_channels, out_channels, bias=False)
        self._register_opaque()

    def forward(self, x: Tensor, edge_index, edge_weight: Optional[Tensor] = None, size=None) -> Tensor:
        return self._dispatch(x, edge_index, edge_weight)

    def _fake(self, x: Tensor, edge_index=None, edge_weight=None) -> Tensor:
        return x.new_empty((x.size(0), self.out_channels))

    def _forward_eager(self, x: Tensor, edge_index, edge_weight: Optional[Tensor] = None) -> Tensor:
        g = as_graph(edge_index, x.size(0), edge_weight)
        if g.value is not None and self.aggr not in ("add", "sum"):
            raise NotImplementedError("edge weights are defined for the sum aggregation only")
        agg = F.aggregate(x, g, self.aggr, use_value=g.value is not None)
        return F.linear(agg, self.lin_rel.weight, self.lin_rel.bias, x2=x, weight2=self.lin_root.weight)
