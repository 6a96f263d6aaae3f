"""Autograd wrappers: one torch.autograd.Function per fused op, saving only what the backward kernels need.

Forward kernels walk the CSR (rows = destinations), backward kernels walk the CSC (rows = sources), so no
pass uses atomics and every result is bit-reproducible run to run.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import Tensor

from .. import compile as _compile
from .. import ops
from .._lib import ACT_ELU, ACT_NONE, SELF_ADD, SELF_NONE, SELF_REPLACE  # noqa: F401
from ..graph import CSRGraph


# ---- gradients written straight into the parameters' .grad buffers --------------------------------------------
# With train.FlatParams every parameter's .grad is a view into one flat bucket.  Autograd would materialise each
# weight / bias / attention-vector gradient and then launch one `add` kernel per parameter to accumulate it; inside
# `direct_param_grads()` the backward kernels instead ADD into p.grad themselves (their second reduction stage has
# an accumulate flag) and hand autograd None for that input.  Off by default: plain autograd semantics.
_DIRECT = [False]


class direct_param_grads:
    def __enter__(self):
        self._old = _DIRECT[0]
        _DIRECT[0] = True
        return self

    def __exit__(self, *exc):
        _DIRECT[0] = self._old
        return False


def _bucket(p) -> Optional[Tensor]:
    """p.grad when gradients may be accumulated in place: a leaf parameter whose fp32, contiguous .grad is bound"""
    if not _DIRECT[0] or p is None or not isinstance(p, Tensor) or not p.is_leaf or not p.requires_grad:
        return None
    g = p.grad
    if g is None or g.dtype != torch.float32 or not g.is_contiguous() or g.shape != p.shape:
        return None
    return g


# ---- parameter-gradient kernels on a side stream -----------------------------------------------------------------
# Weight / bias / attention-vector gradients are leaves of the backward graph: nothing downstream waits for them
# until the optimiser.  Inside `side_stream(s)` (and only when they are written straight into the bucket, see above)
# they are enqueued on `s`, forked from the point of the main stream where their inputs are ready, so they overlap
# with the main chain (dX of this layer, the next layer's attention backward, ...) -- each of these launches fills
# only part of the GPU.  `join()` makes the main stream wait for them; tensors they read are kept alive until then,
# so the caching allocator cannot hand their memory to a later main-stream allocation.
_SIDE = [None]


class side_stream:
    """`stream`: one stream or a list of them.  With several, the kernels are dealt round-robin -- the parameter-gradient
    chain of the headline step (three slab reductions, four column sums, three attention-vector reductions and the
    in_proj dW) is longer than the main backward chain when it runs on a single stream -- except that all kernels adding
    into the same gradient buffer (`key`) stay on one stream, i.e. ordered."""

    def __init__(self, stream):
        # None: no side streams, every kernel stays on the current stream (per-kernel timing, debugging)
        self.streams = list(stream) if isinstance(stream, (list, tuple)) else ([] if stream is None else [stream])
        self.keep, self._next, self._by_key, self._used = [], 0, {}, []

    def __enter__(self):
        self._old = _SIDE[0]
        _SIDE[0] = self
        return self

    def pick(self, key=None):
        if key is not None and key in self._by_key:
            return self._by_key[key]
        s = self.streams[self._next % len(self.streams)]
        self._next += 1
        if key is not None:
            self._by_key[key] = s
        if s not in self._used:
            self._used.append(s)
        return s

    def join(self):
        cur = torch.cuda.current_stream()
        for s in self._used:  # only streams forked since the last join (a wait on an idle stream breaks graph capture)
            cur.wait_stream(s)
        self.keep.clear()
        self._by_key.clear()
        self._used.clear()

    def __exit__(self, *exc):
        self.join()
        _SIDE[0] = self._old
        return False


def _on_side(fn, *keep, key=None):
    sc = _SIDE[0]
    if sc is None or not sc.streams:
        return fn()
    stream = sc.pick(key)
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        fn()
    sc.keep.extend(t for t in keep if t is not None)
    return None


def _need_csc(g: CSRGraph):
    if g.colptr is None:
        raise RuntimeError("this CSRGraph was built without the CSC transpose index; backward needs with_csc=True")


class _Aggregate(torch.autograd.Function):
    """out = post * sum_p w_p pre[j_p] x[j_p] (+ self term) (+ bias), or mean / max over the row."""

    @staticmethod
    def forward(ctx, x, bias, g: CSRGraph, reduce, use_value, pre, post, selfw, self_fill, self_mode):
        ew = g.value if use_value else None
        ctx.g, ctx.reduce, ctx.ew = g, reduce, ew
        ctx.pre, ctx.post, ctx.selfw, ctx.self_fill, ctx.self_mode = pre, post, selfw, self_fill, self_mode
        ctx.has_bias = bias is not None
        ctx.bias_ref = bias
        x = x.contiguous()
        if reduce == "max":
            out, arg = ops.spmm(g.rowptr, g.col, x, "max", bias=bias, want_arg=True)
            ctx.arg = arg
            return out
        if reduce == "mean":
            ctx.inv_deg = g.inv_degree()  # memoised on the graph: three small launches once per batch, not per layer
            if ew is None and pre is None and post is None and self_mode == SELF_NONE and ops.blocked_eligible(x, "sum", None):
                # windowed kernel: mean = sum scaled by the memoised 1 / max(deg, 1) on the way out -- its post-scale variant is
                # 7 % faster than its mean variant, which forms the degree from the row pointer and divides per row
                return ops.spmm(g.rowptr, g.col, x, "sum", post=ctx.inv_deg, bias=bias, quads=g.quads("csr"))
        return ops.spmm(g.rowptr, g.col, x, reduce, ew=ew, pre=pre, post=post, selfw=selfw, self_fill=self_fill,
                        self_mode=self_mode, bias=bias, quads=g.quads("csr") if ops.blocked_eligible(x, reduce, ew) else None)

    @staticmethod
    def backward(ctx, dout):
        g = ctx.g
        _need_csc(g)
        dout = dout.contiguous()
        dx = dbias = None
        if ctx.needs_input_grad[0]:
            if ctx.reduce == "max":
                dx = ops.spmm_max_bwd(g.colptr, g.row, g.csr2csc, ctx.arg, dout)
            elif ctx.reduce == "mean":
                # d x_j = sum_{i : j -> i} dout_i / max(deg_i, 1)
                dx = ops.spmm(g.colptr, g.row, dout, "sum", pre=ctx.inv_deg,
                              quads=g.quads("csc") if ops.blocked_eligible(dout, "sum", None) else None)
            else:
                # transpose: scale on the gathered side <-> scale on the output side
                dx = ops.spmm(g.colptr, g.row, dout, "sum", emap=g.csr2csc if ctx.ew is not None else None, ew=ctx.ew,
                              pre=ctx.post, post=ctx.pre, selfw=ctx.selfw, self_fill=ctx.self_fill, self_mode=ctx.self_mode,
                              quads=g.quads("csc") if ops.blocked_eligible(dout, "sum", ctx.ew) else None)
        if ctx.has_bias and ctx.needs_input_grad[1]:
            d32 = dout  # the column-sum kernel reads fp32 or bf16 and accumulates in fp32
            gb = _bucket(ctx.bias_ref)
            if gb is not None:
                _on_side(lambda: ops.colsum(d32, out=gb), d32, key=gb.data_ptr())
            else:
                dbias = ops.colsum(d32)
        return dx, dbias, None, None, None, None, None, None, None, None


def aggregate(x: Tensor, g: CSRGraph, reduce: str = "sum", *, use_value: bool = False, pre: Optional[Tensor] = None,
              post: Optional[Tensor] = None, selfw: Optional[Tensor] = None, self_fill: float = 1.0,
              self_mode: int = SELF_NONE, bias: Optional[Tensor] = None) -> Tensor:
    reduce = {"add": "sum"}.get(reduce, reduce)
    if reduce not in ("sum", "mean", "max"):
        raise ValueError(f"unsupported aggregation {reduce!r}")
    return _Aggregate.apply(x, bias, g, reduce, use_value, pre, post, selfw, float(self_fill), int(self_mode))


class _GcnConv(torch.autograd.Function):
    """GCNConv on an UNWEIGHTED graph as two launches: out = D^-1/2 (A + c I) D^-1/2 (x W^T) + b.

    The source-side factor d_j^-1/2 is applied by the TRANSFORM's epilogue (h' = dinv * (x W^T), lgk_linear_fwd's
    rowscale: free), so the aggregation runs without per-entry weights -- the cheaper path of both aggregation kernels
    (config-4 shape: 449 -> ~420 us) -- and only scales its output rows.  The backward is that of the unfused form:
    dh = D^-1/2 (A + c I)^T D^-1/2 dout (one aggregation over the CSC), then the plain transform gradients."""

    @staticmethod
    def forward(ctx, x, weight, bias, g: CSRGraph, dinv, selfw, self_fill, self_mode):
        x = x.contiguous()
        w = weight.contiguous()
        hs = ops.linear_fwd(x, w, None, dinv)
        out = ops.spmm(g.rowptr, g.col, hs, "sum", post=dinv, selfw=selfw, self_fill=self_fill, self_mode=self_mode, bias=bias,
                       quads=g.quads("csr") if ops.blocked_eligible(hs, "sum", None) else None)
        ctx.save_for_backward(x, w)
        ctx.g, ctx.dinv, ctx.selfw, ctx.self_fill, ctx.self_mode = g, dinv, selfw, self_fill, self_mode
        ctx.refs = (weight, bias)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, w = ctx.saved_tensors
        g = ctx.g
        _need_csc(g)
        dout = dout.contiguous()
        p_w, p_b = ctx.refs
        ni = ctx.needs_input_grad
        dx = dw = db = None
        if p_b is not None and ni[2]:
            gb = _bucket(p_b)
            if gb is not None:
                _on_side(lambda: ops.colsum(dout, out=gb), dout, key=gb.data_ptr())
            else:
                db = ops.colsum(dout)
        if ni[0] or ni[1]:
            dh = ops.spmm(g.colptr, g.row, dout, "sum", pre=ctx.dinv, post=ctx.dinv, selfw=ctx.selfw, self_fill=ctx.self_fill,
                          self_mode=ctx.self_mode, quads=g.quads("csc") if ops.blocked_eligible(dout, "sum", None) else None)
            if ni[1]:
                gw = _bucket(p_w)
                if gw is not None:
                    _on_side(lambda: ops.linear_bwd_w(dh, x, want_dw=True, want_db=False, dw_out=gw), dh, x, key=gw.data_ptr())
                else:
                    dw = ops.linear_bwd_w(dh, x, want_dw=True, want_db=False)[0]
            if ni[0]:
                dx = ops.linear_bwd_x(dh, w)
        return dx, dw, db, None, None, None, None, None


def gcn_conv(x: Tensor, weight: Tensor, bias: Optional[Tensor], g: CSRGraph, dinv: Tensor, selfw: Optional[Tensor],
             self_fill: float, self_mode: int) -> Tensor:
    return _GcnConv.apply(x, weight, bias, g, dinv, selfw, float(self_fill), int(self_mode))


class _GatAttention(torch.autograd.Function):
    """h [N, H*C] -> edge-softmax attention output [N, H*C] (+ bias)."""

    @staticmethod
    def forward(ctx, h, att_src, att_dst, bias, g: CSRGraph, heads, chan, negative_slope, add_self_loops, dropout_p, seed,
                fused_elu=False, seed_dev=None, alpha=None):
        h = h.contiguous()
        a_s = att_src.detach().reshape(-1).contiguous()
        a_d = att_dst.detach().reshape(-1).contiguous()
        # alpha: (a_src, a_dst) already formed by the transform's epilogue (linear_gat)
        a_src, a_dst = alpha if alpha is not None else ops.gat_alpha(h, a_s, a_d, heads, chan)
        res = ops.gat_fwd(g.rowptr, g.col, h, a_src, a_dst, bias, heads, chan, negative_slope, add_self_loops, dropout_p,
                          seed, with_elu=fused_elu, seed_dev=seed_dev,
                          quads=g.quads("csr") if ops.gat_blocked_eligible(h, chan, dropout_p) else None)
        ctx.seed_dev = seed_dev
        out, rowmax, rowsum = res[:3]
        y = res[3] if fused_elu else h.new_empty(0)
        ctx.fused_elu = fused_elu
        ctx.save_for_backward(h, a_src, a_dst, rowmax, rowsum, out, a_s, a_d, bias if bias is not None else h.new_empty(0), y)
        ctx.g, ctx.cfg = g, (heads, chan, negative_slope, add_self_loops, dropout_p, seed)
        ctx.has_bias = bias is not None
        ctx.att_shape = att_src.shape
        ctx.refs = (att_src, att_dst, bias)
        return y if fused_elu else out

    @staticmethod
    def backward(ctx, dout):
        h, a_src, a_dst, rowmax, rowsum, out, a_s, a_d, bias, y = ctx.saved_tensors
        heads, chan, slope, add_self, p, seed = ctx.cfg
        g = ctx.g
        _need_csc(g)
        dout = dout.contiguous()
        p_src, p_dst, p_bias = ctx.refs
        res = ops.gat_bwd(g, h, a_src, a_dst, rowmax, rowsum, out, bias if ctx.has_bias else None,
                          a_s, a_d, dout, heads, chan, slope, add_self, p, seed, through_elu=ctx.fused_elu,
                          seed_dev=ctx.seed_dev)
        dh, da_src, da_dst = res[:3]
        if ctx.fused_elu:
            dout = res[3]  # gradient w.r.t. the pre-activation output (formed by the dst-side kernel), for db below
        g_src, g_dst = _bucket(p_src), _bucket(p_dst)
        if g_src is not None and g_dst is not None:
            _on_side(lambda: ops.gat_datt(h, da_src, da_dst, heads, chan, g_src.view(-1), g_dst.view(-1)), h, da_src, da_dst, key=g_src.data_ptr())
            datt_src = datt_dst = None
        else:
            datt_src, datt_dst = ops.gat_datt(h, da_src, da_dst, heads, chan)
            datt_src, datt_dst = datt_src.view(ctx.att_shape), datt_dst.view(ctx.att_shape)
        dbias = None
        if ctx.has_bias and ctx.needs_input_grad[3]:
            d32 = dout  # fp32 or bf16 in, fp32 sums out
            gb = _bucket(p_bias)
            if gb is not None:
                _on_side(lambda: ops.colsum(d32, out=gb), d32, key=gb.data_ptr())
            else:
                dbias = ops.colsum(d32)
        return (dh, datt_src, datt_dst, dbias, None, None, None, None, None, None, None, None, None, None)


def gat_attention(h, att_src, att_dst, bias, g, heads, chan, negative_slope=0.2, add_self_loops=True, dropout_p=0.0,
                  seed=0, fused_elu=False, seed_dev: Optional[Tensor] = None, alpha=None) -> Tensor:
    """fused_elu: return elu(attention output + bias) from the same launch (the bias must be the fused one).
    seed_dev: int64 device counter mixed into the dropout seed by the kernels (see conv.dropout_clock)"""
    return _GatAttention.apply(h, att_src, att_dst, bias, g, int(heads), int(chan), float(negative_slope),
                               bool(add_self_loops), float(dropout_p), int(seed), bool(fused_elu), seed_dev, alpha)


class _Linear(torch.autograd.Function):
    """out = act(rowscale * x W^T + bias) [+ second product x2 W2^T fused by accumulation]."""

    @staticmethod
    def forward(ctx, x, weight, bias, rowscale, act, x2, weight2, gat=None):
        x = x.contiguous()
        w = weight.contiguous()
        if gat is not None:
            # GATConv.lin: h = x W^T and the attention-logit halves from the transform's fp32 accumulators (no pass over h;
            # in bf16 mode the logits do not see h's rounding).  a_src / a_dst are by-products: their gradient reaches h
            # and the attention vectors inside the attention kernels' backward.
            att_src, att_dst, heads, chan = gat
            if bias is not None or rowscale is not None or act != ACT_NONE or x2 is not None:
                raise ValueError("the GAT transform takes no bias / rowscale / activation / second operand")
            out, a_src, a_dst = ops.linear_gat_fwd(x, w, att_src.detach().reshape(-1).contiguous(),
                                                   att_dst.detach().reshape(-1).contiguous(), heads, chan)
            ctx.save_for_backward(x, w, x.new_empty(0), x.new_empty(0))
            ctx.two, ctx.act, ctx.has_bias, ctx.has_rs = False, act, False, False
            ctx.refs = (weight, None, None)
            ctx.mark_non_differentiable(a_src, a_dst)
            ctx.set_materialize_grads(False)  # no zero tensors for the two by-products in the backward
            ctx.gat = True
            return out, a_src, a_dst
        ctx.gat = False
        if x2 is not None:
            # out = x W^T + bias + x2 W2^T  (SAGEConv / GraphConv root term): second kernel accumulates in place
            if rowscale is not None or act != ACT_NONE:
                raise ValueError("the fused two-operand form takes no rowscale/activation")
            x2 = x2.contiguous()
            w2 = weight2.contiguous()
            out = ops.linear_fwd(x, w, bias)
            out = ops.linear_fwd(x2, w2, out=out)
            ctx.save_for_backward(x, w, x2, w2)
        else:
            out = ops.linear_fwd(x, w, bias, rowscale, act)
            ctx.save_for_backward(x, w, out if act != ACT_NONE else x.new_empty(0), rowscale if rowscale is not None else x.new_empty(0))
        ctx.two, ctx.act, ctx.has_bias, ctx.has_rs = x2 is not None, act, bias is not None, rowscale is not None
        ctx.refs = (weight, bias, weight2)
        return out

    @staticmethod
    def backward(ctx, dout, *unused):
        dout = dout.contiguous()
        ni = ctx.needs_input_grad
        p_w, p_b, p_w2 = ctx.refs

        def grads_w(dy, xin, want_dw, want_db, pw, pb):
            """dW / db, added straight into the .grad buckets when both are bound (see direct_param_grads)"""
            if not (want_dw or want_db):
                return None, None
            gw, gb = (_bucket(pw) if want_dw else None), (_bucket(pb) if want_db else None)
            if (not want_dw or gw is not None) and (not want_db or gb is not None):
                _on_side(lambda: ops.linear_bwd_w(dy, xin, want_dw=want_dw, want_db=want_db, dw_out=gw, db_out=gb), dy, xin,
                         key=(gw if gw is not None else gb).data_ptr())
                return None, None
            return ops.linear_bwd_w(dy, xin, want_dw=want_dw, want_db=want_db)

        if ctx.two:
            x, w, x2, w2 = ctx.saved_tensors
            dx = ops.linear_bwd_x(dout, w) if ni[0] else None
            dw, db = grads_w(dout, x, ni[1], ctx.has_bias and ni[2], p_w, p_b)
            dx2 = ops.linear_bwd_x(dout, w2) if ni[5] else None
            dw2 = grads_w(dout, x2, ni[6], False, p_w2, None)[0]
            return dx, dw, db, None, None, dx2, dw2, None
        x, w, out, rs = ctx.saved_tensors
        if ctx.act == ACT_ELU:
            dout = ops.elu_bwd(out, dout)
        elif ctx.act != ACT_NONE:
            raise NotImplementedError
        want_db = ctx.has_bias and ni[2]
        if ctx.has_rs:
            # the bias sits outside the row scale: db = colsum(dout), dW uses the scaled dout
            db = None
            if want_db:
                gb, d0 = _bucket(p_b), dout
                if gb is not None:
                    _on_side(lambda: ops.colsum(d0, out=gb), d0, key=gb.data_ptr())
                else:
                    db = ops.colsum(d0)
            dout = (dout * rs.unsqueeze(1)).to(dout.dtype)
            dw = grads_w(dout, x, ni[1], False, p_w, None)[0]
        else:
            dw, db = grads_w(dout, x, ni[1], want_db, p_w, p_b)  # one pass: db falls out of the dy tiles of dW
        dx = ops.linear_bwd_x(dout, w) if ni[0] else None
        return dx, dw, db, None, None, None, None, None


def linear_gat(x: Tensor, weight: Tensor, att_src: Tensor, att_dst: Tensor, heads: int, chan: int):
    """-> (h [N, H*C], a_src [N, H], a_dst [N, H]): GATConv's transform with the attention-logit halves computed in its
    epilogue (ops.linear_gat_fwd); differentiable in x and weight through h"""
    return _Linear.apply(x, weight, None, None, ACT_NONE, None, None, (att_src, att_dst, int(heads), int(chan)))


def _linear_fake(x, weight, bias=None):
    return x.new_empty(tuple(x.shape[:-1]) + (weight.size(0),))


_LINEAR_OPS = {act: _compile.opaque_function((lambda act: lambda x, w, b=None: linear(x, w, b, act=act))(act), _linear_fake)
               for act in (ACT_NONE, ACT_ELU)}


def linear(x: Tensor, weight: Tensor, bias: Optional[Tensor] = None, *, rowscale: Optional[Tensor] = None,
           act: int = ACT_NONE, x2: Optional[Tensor] = None, weight2: Optional[Tensor] = None) -> Tensor:
    if _compile.is_compiling() and rowscale is None and x2 is None and act in _LINEAR_OPS:
        return _LINEAR_OPS[act](x, weight, bias)
    if x.dim() != 2:
        lead = x.shape[:-1]
        return linear(x.reshape(-1, x.size(-1)), weight, bias, rowscale=rowscale, act=act).reshape(*lead, weight.size(0))
    return _Linear.apply(x, weight, bias, rowscale, int(act), x2, weight2)


class _Elu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = ops.elu_fwd(x.contiguous())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return ops.elu_bwd(y, dy.contiguous())


def _elu_eager(x: Tensor) -> Tensor:
    return _Elu.apply(x)


elu = _compile.opaque_function(_elu_eager, lambda x: torch.empty_like(x))


class _BnAct(torch.autograd.Function):
    """dropout(act(batch_norm(x))) -- BatchNorm1d + ELU + Dropout of PyG's MLP (reference models/gin.py:23) in three
    launches; saves x and the [d] statistics only (the mask is a function of (seed, device clock, element))."""

    @staticmethod
    def forward(ctx, x, weight, bias, bn, act, dropout_p, seed, seed_dev):
        x = x.contiguous()
        training = bn.training or bn.running_mean is None
        if bn.momentum is None:
            raise NotImplementedError("BatchNorm with momentum=None (cumulative average) is not on the reference's path")
        track = bn.track_running_stats and bn.training
        y, mean, invstd = ops.bn_act_fwd(x, weight, bias, bn.running_mean if (track or not training) else None,
                                         bn.running_var if (track or not training) else None,
                                         bn.num_batches_tracked if track else None, bn.eps, bn.momentum, training, act,
                                         dropout_p, seed, seed_dev)
        ctx.save_for_backward(x, mean, invstd, weight if weight is not None else x.new_empty(0), bias if bias is not None else x.new_empty(0))
        ctx.cfg = (training, act, dropout_p, seed, seed_dev, weight is not None, bias is not None)
        ctx.refs = (weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mean, invstd, w, b = ctx.saved_tensors
        training, act, p, seed, seed_dev, has_w, has_b = ctx.cfg
        pw, pb = ctx.refs
        want_dw, want_db = has_w and ctx.needs_input_grad[1], has_b and ctx.needs_input_grad[2]
        gw, gb = (_bucket(pw) if want_dw else None), (_bucket(pb) if want_db else None)
        direct = (not want_dw or gw is not None) and (not want_db or gb is not None) and (want_dw or want_db)
        dx, dw, db = ops.bn_act_bwd(x, dy.contiguous(), w if has_w else None, b if has_b else None, mean, invstd, training, act, p, seed,
                                    seed_dev, want_dw=want_dw, want_db=want_db, dw_out=gw if direct else None,
                                    db_out=gb if direct else None)
        return dx, dw, db, None, None, None, None, None


def bn_act(x: Tensor, bn: "torch.nn.BatchNorm1d", act: int = ACT_ELU, dropout_p: float = 0.0, seed: int = 0,
           seed_dev: Optional[Tensor] = None) -> Tensor:
    """dropout(act(bn(x))) with torch.nn.BatchNorm1d semantics (training / eval, running statistics, momentum)"""
    if x.dim() != 2:
        raise NotImplementedError("bn_act takes [N, d] node features")
    return _BnAct.apply(x, bn.weight, bn.bias, bn, int(act), float(dropout_p), int(seed), seed_dev)


class _ActDropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act, dropout_p, seed, seed_dev):
        x = x.contiguous()
        ctx.save_for_backward(x)
        ctx.cfg = (act, dropout_p, seed, seed_dev)
        return ops.act_dropout_fwd(x, act, dropout_p, seed, seed_dev)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        act, p, seed, seed_dev = ctx.cfg
        return ops.act_dropout_bwd(x, dy.contiguous(), act, p, seed, seed_dev), None, None, None, None


def act_dropout(x: Tensor, act: int = ACT_ELU, dropout_p: float = 0.0, seed: int = 0, seed_dev: Optional[Tensor] = None) -> Tensor:
    """dropout(act(x)) in one launch (the F.elu + nn.Dropout after every GIN layer, reference models/gin.py:31-32)"""
    return _ActDropout.apply(x, int(act), float(dropout_p), int(seed), seed_dev)


class ActDropout(_compile.OpaqueModule, torch.nn.Module):
    """nn.Sequential(activation, nn.Dropout(p)) as one launch and, under torch.compile, one opaque operator"""

    def __init__(self, act: int = ACT_ELU, p: float = 0.0):
        super().__init__()
        self.act, self.p = int(act), float(p)
        self._register_opaque()

    def forward(self, x: Tensor) -> Tensor:
        return self._dispatch(x)

    def _fake(self, x: Tensor) -> Tensor:
        return torch.empty_like(x)

    def _forward_eager(self, x: Tensor) -> Tensor:
        from .conv import _DROPOUT_CLOCK, _next_seed
        p = self.p if self.training else 0.0
        return act_dropout(x, self.act, p, _next_seed() if p > 0 else 0, _DROPOUT_CLOCK[0] if p > 0 else None)


class _Pool(torch.autograd.Function):
    """segment reduce over graphs; mode in {'sum','mean','max','mean+max'}"""

    @staticmethod
    def forward(ctx, x, ptr, mode):
        x = x.contiguous()
        want_sum = mode in ("sum", "mean", "mean+max")
        want_max = mode in ("max", "mean+max")
        mean = mode in ("mean", "mean+max")
        out_sum, out_max, arg = ops.pool_fwd(x, ptr, want_sum, mean, want_max)
        ctx.ptr, ctx.arg, ctx.mode, ctx.mean = ptr, arg, mode, mean
        ctx.shape, ctx.dtype, ctx.device = x.shape, x.dtype, x.device
        if mode == "mean+max":
            return torch.cat([out_sum, out_max], dim=1)
        return out_sum if want_sum else out_max

    @staticmethod
    def backward(ctx, dout):
        n, f = ctx.shape
        dout = dout.contiguous()
        dsum = dmax = None
        if ctx.mode == "mean+max":
            dsum, dmax = dout[:, :f].contiguous(), dout[:, f:].contiguous()
        elif ctx.mode == "max":
            dmax = dout
        else:
            dsum = dout
        dx = ops.pool_bwd(dsum, ctx.mean, dmax, ctx.arg, ctx.ptr, n, f, ctx.dtype, ctx.device)
        return dx, None, None


def segment_pool(x: Tensor, ptr: Tensor, mode: str) -> Tensor:
    return _Pool.apply(x, ptr, mode)


class _SortPool(torch.autograd.Function):
    """SortAggregation(k): saves only the [B, k] selection"""

    @staticmethod
    def forward(ctx, x, ptr, k):
        out, sel = ops.sort_pool_fwd(x.contiguous(), ptr, k)
        ctx.save_for_backward(sel)
        ctx.shape = x.shape
        return out

    @staticmethod
    def backward(ctx, dy):
        (sel,) = ctx.saved_tensors
        n, f = ctx.shape
        return ops.sort_pool_bwd(dy.contiguous(), sel, n, f), None, None


def sort_pool(x: Tensor, ptr: Tensor, k: int) -> Tensor:
    return _SortPool.apply(x, ptr, int(k))
