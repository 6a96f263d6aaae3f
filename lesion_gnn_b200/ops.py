"""Tensor-level wrappers over the C ABI (no autograd; see nn/functional.py for that).

Every function checks that its tensors live on a CUDA device, are contiguous and have the dtype the kernel
expects, allocates outputs / workspaces with PyTorch's caching allocator, and enqueues on the current stream.
Nothing here synchronises, so all of it can be captured into a CUDA graph.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
from torch import Tensor

from . import _lib
from ._lib import F32, BF16, SUM, MEAN, MAX, SELF_NONE, SELF_REPLACE, SELF_ADD  # noqa: F401

_REDUCE = {"sum": SUM, "add": SUM, "mean": MEAN, "max": MAX}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _dt(t: Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"lesion_gnn_b200 kernels take float32 or bfloat16 features, got {t.dtype}")


def _chk(t: Optional[Tensor], dtype=None, name="tensor") -> Optional[int]:
    """validate and return the device pointer (None -> NULL)"""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: lesion_gnn_b200 has no CPU path")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}")
    return t.data_ptr() if t.numel() > 0 else None


def _ws(nbytes: int, device) -> Tensor:
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def launch_count() -> int:
    return int(_lib.lib().query("lgk_launch_count"))


def reset_launch_count() -> None:
    _lib.lib().query("lgk_reset_launch_count")


# ---------------------------------------------------------------------------------------------- structure


# set by csr_build(..., dst_sorted=True): int32 device scalar, 1 when the last sorted build found its contract violated
last_sorted_violation: Optional[Tensor] = None


def csr_build(edge_index: Tensor, num_nodes: int, with_csc: bool = True, dst_sorted: bool = False):
    """-> rowptr, col, perm, colptr, row, csr2csc (int32); the last three are None without with_csc.
    dst_sorted: the caller guarantees edge_index[1] is non-decreasing (kNN / radius graphs, collated kNN samples): the
    boundary-detection build lgk_csr_build_sorted runs instead of the general bucket sort (same outputs)."""
    L = _lib.lib()
    if edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError("edge_index must have shape [2, E]")
    ei = _chk(edge_index, torch.int64, "edge_index")
    dev = edge_index.device
    E, N = edge_index.size(1), int(num_nodes)
    i32 = dict(dtype=torch.int32, device=dev)
    rowptr, col, perm = torch.empty(N + 1, **i32), torch.empty(E, **i32), torch.empty(E, **i32)
    if with_csc:
        colptr, row, csr2csc = torch.empty(N + 1, **i32), torch.empty(E, **i32), torch.empty(E, **i32)
    else:
        colptr = row = csr2csc = None
    nbytes = L.query("lgk_csr_workspace_bytes", N, E)
    ws = _ws(nbytes, dev)
    if dst_sorted:
        global last_sorted_violation
        last_sorted_violation = torch.empty(1, **i32)
        L.call("lgk_csr_build_sorted", ei, E, N, rowptr.data_ptr(), _chk(col), _chk(perm),
               None if colptr is None else colptr.data_ptr(), _chk(row), _chk(csr2csc), last_sorted_violation.data_ptr(),
               ws.data_ptr(), ws.numel(), _stream())
        return rowptr, col, perm, colptr, row, csr2csc
    L.call("lgk_csr_build", ei, E, N, rowptr.data_ptr(), _chk(col), _chk(perm),
           None if colptr is None else colptr.data_ptr(), _chk(row), _chk(csr2csc),
           ws.data_ptr(), ws.numel(), _stream())
    return rowptr, col, perm, colptr, row, csr2csc


def batch_ptr(batch: Tensor, num_graphs: int) -> Tensor:
    L = _lib.lib()
    b = _chk(batch, torch.int64, "batch")
    ptr = torch.empty(int(num_graphs) + 1, dtype=torch.int32, device=batch.device)
    L.call("lgk_batch_ptr", b, batch.numel(), int(num_graphs), ptr.data_ptr(), _stream())
    return ptr


def gather_f32(src: Tensor, index: Tensor) -> Tensor:
    L = _lib.lib()
    out = torch.empty(index.numel(), dtype=torch.float32, device=src.device)
    L.call("lgk_gather_f32", _chk(src, torch.float32, "src"), _chk(index, torch.int32, "index"), _chk(out),
           index.numel(), _stream())
    return out


def knn_graph_raw(pos: Tensor, ptr: Tensor, k: int, loop: bool, num_edges: Optional[int] = None) -> Tensor:
    """kNN edges for a collated batch; `ptr` is the int32 graph pointer.  num_edges, when the caller knows it
    from host-side graph sizes, avoids the one device->host read of eptr[-1]."""
    L = _lib.lib()
    if pos.dim() != 2:
        raise ValueError("pos must be [N, dim]")
    if pos.dtype not in (torch.float32, torch.float64):
        raise TypeError("pos must be float32 or float64")
    B, N = ptr.numel() - 1, pos.size(0)
    eptr = torch.empty(B + 1, dtype=torch.int64, device=pos.device)
    L.call("lgk_knn_edge_ptr", _chk(ptr, torch.int32, "ptr"), B, int(k), int(loop), eptr.data_ptr(), _stream())
    if num_edges is None:
        num_edges = int(eptr[-1].item())
    out = torch.empty(2, num_edges, dtype=torch.int64, device=pos.device)
    L.call("lgk_knn_graph", _chk(pos, None, "pos"), 0 if pos.dtype == torch.float32 else 1, pos.size(1),
           ptr.data_ptr(), eptr.data_ptr(), B, N, int(k), int(loop), _chk(out), num_edges, _stream())
    return out


def radius_graph_raw(pos: Tensor, ptr: Tensor, r: float, loop: bool, max_num_neighbors: int) -> Tensor:
    """radius-graph edges for a collated batch (two passes: count, exclusive scan, fill).  The edge count depends on
    the data, so this reads one int64 back from the device -- graph construction is outside the captured step."""
    L = _lib.lib()
    if pos.dim() != 2:
        raise ValueError("pos must be [N, dim]")
    if pos.dtype not in (torch.float32, torch.float64):
        raise TypeError("pos must be float32 or float64")
    B, N = ptr.numel() - 1, pos.size(0)
    dt = 0 if pos.dtype == torch.float32 else 1
    counts = torch.empty(N, dtype=torch.int32, device=pos.device)
    L.call("lgk_radius_graph", _chk(pos, None, "pos"), dt, pos.size(1), _chk(ptr, torch.int32, "ptr"), B, N, float(r), int(loop),
           int(max_num_neighbors), _chk(counts), None, None, 0, _stream())
    incl = torch.cumsum(counts, 0, dtype=torch.int64)
    num_edges = int(incl[-1].item()) if N > 0 else 0
    out = torch.empty(2, num_edges, dtype=torch.int64, device=pos.device)
    if num_edges > 0:
        offsets = (incl - counts).contiguous()
        L.call("lgk_radius_graph", _chk(pos, None, "pos"), dt, pos.size(1), _chk(ptr, torch.int32, "ptr"), B, N, float(r), int(loop),
               int(max_num_neighbors), None, _chk(offsets, torch.int64, "offsets"), _chk(out), num_edges, _stream())
    return out


def gaussian_distance(pos: Tensor, edge_index: Tensor, sigma: float) -> Tensor:
    L = _lib.lib()
    E = edge_index.size(1)
    out = torch.empty(E, dtype=torch.float32, device=pos.device)
    L.call("lgk_gaussian_distance", _chk(pos, None, "pos"), 0 if pos.dtype == torch.float32 else 1, pos.size(1),
           _chk(edge_index, torch.int64, "edge_index"), E, float(sigma), _chk(out), _stream())
    return out


# -------------------------------------------------------------------------------------------- aggregation


class Quads:
    """Row-blocked form of a CSR / CSC (lgk_quad_build): merged neighbour lists of four consecutive rows with
    membership masks, plus the source-id range of every chunk of 64 quads.  Index work, built once per batch."""

    __slots__ = ("qlen", "qent", "cmin", "cmax", "qmeta")

    def __init__(self, qlen, qent, cmin, cmax, qmeta):
        self.qlen, self.qent, self.cmin, self.cmax, self.qmeta = qlen, qent, cmin, cmax, qmeta


def quad_build(ptr: Tensor, idx: Tensor) -> Quads:
    L = _lib.lib()
    n = ptr.numel() - 1
    dev = ptr.device
    i32 = dict(dtype=torch.int32, device=dev)
    nq = (n + 3) // 4
    nc = int(L.query("lgk_quad_chunks", n))
    q = Quads(torch.empty(2 * max(nq, 1), **i32), torch.empty(max(idx.numel(), 1), **i32), torch.empty(max(nc, 1), **i32),
              torch.empty(max(nc, 1), **i32), torch.empty(4, **i32))
    L.call("lgk_quad_build", _chk(ptr, torch.int32, "ptr"), _chk(idx, torch.int32, "idx"), n, q.qlen.data_ptr(),
           q.qent.data_ptr(), q.cmin.data_ptr(), q.cmax.data_ptr(), q.qmeta.data_ptr(), _stream())
    return q


# rows x features from which the windowed, row-blocked aggregation kernel takes over (tuning / test knob)
_WIN_MIN_ELEMS = int(os.environ.get("LGK_WIN_MIN_ELEMS", str(1 << 23)))


def blocked_eligible(x: Tensor, reduce: str, ew) -> bool:
    return (ew is None and _REDUCE[reduce] in (SUM, MEAN) and x.dtype == torch.float32 and x.dim() == 2
            and x.size(1) % 4 == 0 and x.numel() >= _WIN_MIN_ELEMS and x.size(0) < (1 << 28) and x.numel() // 4 < (1 << 31))


def spmm(ptr: Tensor, idx: Tensor, x: Tensor, reduce: str = "sum", *, emap: Optional[Tensor] = None,
         ew: Optional[Tensor] = None, pre: Optional[Tensor] = None, post: Optional[Tensor] = None,
         selfw: Optional[Tensor] = None, self_fill: float = 1.0, self_mode: int = SELF_NONE,
         bias: Optional[Tensor] = None, want_arg: bool = False, quads: Optional[Quads] = None):
    """quads: the row-blocked form of (ptr, idx); when given (and the call is eligible: see blocked_eligible) the
    windowed kernel lgk_spmm_blocked runs instead of lgk_spmm -- same result up to fp32 summation order"""
    L = _lib.lib()
    n, f = x.shape
    out = torch.empty_like(x)
    if quads is not None and not want_arg and blocked_eligible(x, reduce, ew):
        f32 = torch.float32
        L.call("lgk_spmm_blocked", _chk(ptr, torch.int32, "ptr"), _chk(quads.qlen, torch.int32), _chk(quads.qent, torch.int32),
               _chk(quads.cmin, torch.int32), _chk(quads.cmax, torch.int32), _chk(quads.qmeta, torch.int32), _chk(pre, f32, "pre"), _chk(post, f32, "post"),
               _chk(selfw, f32, "selfw"), float(self_fill), int(self_mode), _chk(x, None, "x"), _chk(bias, f32, "bias"),
               _chk(out), n, f, _REDUCE[reduce], _dt(x), _stream())
        return out
    arg = torch.empty(n, f, dtype=torch.int32, device=x.device) if (want_arg and _REDUCE[reduce] == MAX) else None
    f32 = torch.float32
    L.call("lgk_spmm", _chk(ptr, torch.int32, "ptr"), _chk(idx, torch.int32, "idx"), _chk(emap, torch.int32, "emap"),
           _chk(ew, f32, "ew"), _chk(pre, f32, "pre"), _chk(post, f32, "post"), _chk(selfw, f32, "selfw"),
           float(self_fill), int(self_mode), _chk(x, None, "x"), _chk(bias, f32, "bias"), _chk(out), _chk(arg),
           n, f, _REDUCE[reduce], _dt(x), _stream())
    return (out, arg) if want_arg else out


def spmm_max_bwd(colptr: Tensor, row: Tensor, csr2csc: Tensor, arg: Tensor, dy: Tensor) -> Tensor:
    L = _lib.lib()
    dx = torch.empty_like(dy)
    L.call("lgk_spmm_max_bwd", _chk(colptr, torch.int32), _chk(row, torch.int32), _chk(csr2csc, torch.int32),
           _chk(arg, torch.int32), _chk(dy, None, "dy"), _chk(dx), dy.size(0), dy.size(1), _dt(dy), _stream())
    return dx


def gcn_norm(rowptr: Tensor, col: Tensor, ew: Optional[Tensor], fill: float, add_self_loops: bool):
    L = _lib.lib()
    n = rowptr.numel() - 1
    dinv = torch.empty(n, dtype=torch.float32, device=rowptr.device)
    selfw = torch.empty(n, dtype=torch.float32, device=rowptr.device) if (ew is not None and add_self_loops) else None
    L.call("lgk_gcn_norm", _chk(rowptr, torch.int32), _chk(col, torch.int32), _chk(ew, torch.float32, "ew"), float(fill),
           int(add_self_loops), _chk(dinv), _chk(selfw), n, _stream())
    return dinv, selfw


# ---------------------------------------------------------------------------------------------------- GAT


def gat_alpha(h: Tensor, att_src: Tensor, att_dst: Tensor, heads: int, chan: int):
    L = _lib.lib()
    n = h.size(0)
    a_src = torch.empty(n, heads, dtype=torch.float32, device=h.device)
    a_dst = torch.empty_like(a_src)
    L.call("lgk_gat_alpha", _chk(h, None, "h"), _chk(att_src, torch.float32, "att_src"), _chk(att_dst, torch.float32, "att_dst"),
           _chk(a_src), _chk(a_dst), n, heads, chan, _dt(h), _stream())
    return a_src, a_dst


def gat_blocked_eligible(h: Tensor, chan: int, dropout_p: float) -> bool:
    """the windowed GAT forward takes over on large graphs: fp32, no attention dropout, a 128-byte slice inside one head"""
    return (h.dtype == torch.float32 and dropout_p == 0.0 and chan % 32 == 0 and h.numel() >= _WIN_MIN_ELEMS
            and h.size(0) < (1 << 28) and h.numel() // 4 < (1 << 31))


def gat_fwd(rowptr, col, h, a_src, a_dst, bias, heads, chan, negative_slope, add_self_loops, dropout_p=0.0, seed=0,
            with_elu=False, seed_dev: Optional[Tensor] = None, quads: Optional[Quads] = None):
    """-> out, rowmax, rowsum [, elu(out)].  seed_dev: int64 device counter mixed into the dropout seed at run time.
    quads: the row-blocked form of (rowptr, col); when given and gat_blocked_eligible, lgk_gat_fwd_blocked runs (same result
    up to fp32 summation order)"""
    L = _lib.lib()
    n = h.size(0)
    out = torch.empty_like(h)
    rowmax = torch.empty(n, heads, dtype=torch.float32, device=h.device)
    rowsum = torch.empty_like(rowmax)
    if quads is not None and gat_blocked_eligible(h, chan, dropout_p):
        y = torch.empty_like(h) if with_elu else None
        ws = _ws(L.query("lgk_gat_blocked_workspace_bytes", n, heads), h.device)
        L.call("lgk_gat_fwd_blocked", _chk(rowptr, torch.int32), _chk(quads.qlen, torch.int32), _chk(quads.qent), _chk(quads.cmin, torch.int32),
               _chk(quads.cmax, torch.int32), _chk(quads.qmeta, torch.int32), _chk(h, torch.float32, "h"), _chk(a_src, torch.float32),
               _chk(a_dst, torch.float32), _chk(bias, torch.float32, "bias"), _chk(out), _chk(y), _chk(rowmax), _chk(rowsum), n,
               heads, chan, float(negative_slope), int(add_self_loops), ws.data_ptr(), ws.numel(), F32, _stream())
        return (out, rowmax, rowsum, y) if with_elu else (out, rowmax, rowsum)
    if with_elu:
        y = torch.empty_like(h)
        L.call("lgk_gat_fwd_elu", _chk(rowptr, torch.int32), _chk(col, torch.int32), _chk(h, None, "h"), _chk(a_src, torch.float32),
               _chk(a_dst, torch.float32), _chk(bias, torch.float32, "bias"), _chk(out), _chk(y), _chk(rowmax), _chk(rowsum), n,
               heads, chan, float(negative_slope), int(add_self_loops), float(dropout_p), int(seed), _chk(seed_dev, torch.int64, "seed_dev"),
               _dt(h), _stream())
        return out, rowmax, rowsum, y
    L.call("lgk_gat_fwd", _chk(rowptr, torch.int32), _chk(col, torch.int32), _chk(h, None, "h"), _chk(a_src, torch.float32),
           _chk(a_dst, torch.float32), _chk(bias, torch.float32, "bias"), _chk(out), _chk(rowmax), _chk(rowsum), n, heads, chan,
           float(negative_slope), int(add_self_loops), float(dropout_p), int(seed), _chk(seed_dev, torch.int64, "seed_dev"),
           _dt(h), _stream())
    return out, rowmax, rowsum


def gat_bwd(graph, h, a_src, a_dst, rowmax, rowsum, out, bias, att_src, att_dst, dout, heads, chan, negative_slope,
            add_self_loops, dropout_p=0.0, seed=0, through_elu=False, seed_dev: Optional[Tensor] = None):
    """-> dh, da_src [N,H], da_dst [N,H] (the attention-vector gradients follow from gat_datt).
    through_elu: `dout` is the gradient of elu(out) (layer run with gat_fwd(with_elu=True)); the gradient of out is
    formed by the dst-side kernel and returned as a fourth value"""
    L = _lib.lib()
    n, dev = h.size(0), h.device
    f32 = dict(dtype=torch.float32, device=dev)
    stats = torch.empty(n, heads, 4, **f32)
    da_dst = torch.empty(n, heads, **f32)
    da_src = torch.empty(n, heads, **f32)
    dh = torch.empty_like(h)
    dt, st = _dt(h), _stream()
    sd = _chk(seed_dev, torch.int64, "seed_dev")
    if through_elu:
        dy, dout = dout, torch.empty_like(h)
        L.call("lgk_gat_bwd_dst_elu", _chk(graph.rowptr), _chk(graph.col), _chk(h), _chk(a_src), _chk(a_dst), _chk(rowmax),
               _chk(rowsum), _chk(out, None, "out"), _chk(bias, torch.float32), _chk(dy, h.dtype, "dy"), _chk(dout), _chk(stats),
               _chk(da_dst), n, heads, chan, float(negative_slope), int(add_self_loops), float(dropout_p), int(seed), sd, dt, st)
    else:
        L.call("lgk_gat_bwd_dst", _chk(graph.rowptr), _chk(graph.col), _chk(h), _chk(a_src), _chk(a_dst), _chk(rowmax),
               _chk(rowsum), _chk(out, None, "out"), _chk(bias, torch.float32), _chk(dout, h.dtype, "dout"), _chk(stats),
               _chk(da_dst), n, heads, chan, float(negative_slope), int(add_self_loops), float(dropout_p), int(seed), sd, dt, st)
    L.call("lgk_gat_bwd_src", _chk(graph.colptr), _chk(graph.row), _chk(graph.csr2csc), _chk(h), _chk(a_src), _chk(a_dst),
           _chk(rowmax), _chk(rowsum), _chk(stats), _chk(da_dst), _chk(att_src, torch.float32), _chk(att_dst, torch.float32),
           _chk(dout), _chk(dh), _chk(da_src), n, heads, chan, float(negative_slope), int(add_self_loops),
           float(dropout_p), int(seed), sd, dt, st)
    return (dh, da_src, da_dst, dout) if through_elu else (dh, da_src, da_dst)


def gat_datt(h, da_src, da_dst, heads, chan, datt_src_out=None, datt_dst_out=None):
    """datt_src / datt_dst [H*C] = sum_n da[n,h] h[n,h,:]; with the *_out buffers given the sums are ADDED into them
    (gradient bucket) and (None, None) is returned"""
    L = _lib.lib()
    n, dev = h.size(0), h.device
    direct = datt_src_out is not None
    datt_src = datt_src_out if direct else torch.empty(heads * chan, dtype=torch.float32, device=dev)
    datt_dst = datt_dst_out if direct else torch.empty(heads * chan, dtype=torch.float32, device=dev)
    ws = _ws(L.query("lgk_gat_datt_workspace_bytes", n, heads, chan), dev)
    L.call("lgk_gat_datt", _chk(h), _chk(da_src), _chk(da_dst), _chk(datt_src, torch.float32, "datt_src"),
           _chk(datt_dst, torch.float32, "datt_dst"), n, heads, chan, int(direct), ws.data_ptr(), ws.numel(), _dt(h), _stream())
    return (None, None) if direct else (datt_src, datt_dst)


# ------------------------------------------------------------------------------------------------ pooling


def pool_fwd(x: Tensor, ptr: Tensor, want_sum: bool, mean: bool, want_max: bool):
    """-> (out_sum|None, out_max|None, arg|None)"""
    L = _lib.lib()
    n, f = x.shape
    b = ptr.numel() - 1
    out_sum = torch.empty(b, f, dtype=x.dtype, device=x.device) if want_sum else None
    out_max = torch.empty(b, f, dtype=x.dtype, device=x.device) if want_max else None
    arg = torch.empty(b, f, dtype=torch.int32, device=x.device) if want_max else None
    ws = _ws(L.query("lgk_pool_workspace_bytes", n, b, f), x.device)
    L.call("lgk_pool_fwd", _chk(x, None, "x"), _chk(ptr, torch.int32, "ptr"), _chk(out_sum), int(mean), _chk(out_max), _chk(arg),
           n, b, f, ws.data_ptr(), ws.numel(), _dt(x), _stream())
    return out_sum, out_max, arg


def pool_bwd(dsum: Optional[Tensor], mean: bool, dmax: Optional[Tensor], arg: Optional[Tensor], ptr: Tensor,
             num_nodes: int, feat: int, dtype, device) -> Tensor:
    L = _lib.lib()
    dx = torch.empty(num_nodes, feat, dtype=dtype, device=device)
    b = ptr.numel() - 1
    L.call("lgk_pool_bwd", _chk(dsum, dtype, "dsum"), int(mean), _chk(dmax, dtype, "dmax"), _chk(arg, torch.int32), _chk(ptr, torch.int32),
           _chk(dx), num_nodes, b, feat, _dt(dx), _stream())
    return dx


def sort_pool_fwd(x: Tensor, ptr: Tensor, k: int):
    """SortAggregation(k) -> (out [B, k*F], sel [B, k] int32 node index or -1)"""
    L = _lib.lib()
    n, f = x.shape
    b = ptr.numel() - 1
    out = torch.empty(b, k * f, dtype=x.dtype, device=x.device)
    sel = torch.empty(b, k, dtype=torch.int32, device=x.device)
    L.call("lgk_sort_pool_fwd", _chk(x, None, "x"), _chk(ptr, torch.int32, "ptr"), _chk(out), _chk(sel), b, f, int(k), _dt(x), _stream())
    return out, sel


def sort_pool_bwd(dy: Tensor, sel: Tensor, num_nodes: int, feat: int) -> Tensor:
    L = _lib.lib()
    b, k = sel.shape
    dx = torch.empty(num_nodes, feat, dtype=dy.dtype, device=dy.device)
    L.call("lgk_sort_pool_bwd", _chk(dy, None, "dy"), _chk(sel, torch.int32, "sel"), _chk(dx), num_nodes, b, feat, k, _dt(dy), _stream())
    return dx


# ------------------------------------------------------------------------------------------------- linear


def linear_fwd(x: Tensor, weight: Tensor, bias: Optional[Tensor] = None, rowscale: Optional[Tensor] = None,
               act: int = 0, out: Optional[Tensor] = None) -> Tensor:
    """out = act(rowscale * x W^T + bias [+ out]);  passing `out` accumulates into it in place."""
    L = _lib.lib()
    n, k = x.shape
    m = weight.size(0)
    accumulate = out is not None
    if out is None:
        out = torch.empty(n, m, dtype=x.dtype, device=x.device)
    L.call("lgk_linear_fwd", _chk(x, None, "x"), _chk(weight, torch.float32, "weight"), _chk(bias, torch.float32, "bias"),
           _chk(rowscale, torch.float32, "rowscale"), _chk(out, x.dtype, "out"), n, k, m, int(act), int(accumulate), _dt(x), _stream())
    return out


def linear_gat_fwd(x: Tensor, weight: Tensor, att_src: Tensor, att_dst: Tensor, heads: int, chan: int):
    """-> (h = x W^T [n, H*C] in x.dtype, a_src [n, H] f32, a_dst [n, H] f32): include/lgk.h lgk_linear_gat_fwd"""
    n, k = x.shape
    if weight.size(0) != heads * chan:
        raise ValueError("weight must have heads * chan rows")
    L = _lib.lib()
    if x.dtype == torch.bfloat16 and not L.query("lgk_linear_gat_fused", n, k, int(heads), int(chan), BF16):
        # a head width the epilogues do not cover (chan does not divide 64): keep the promise that the logits never see
        # h's bf16 rounding -- transform in fp32, logit halves from it, then round h.  Off the reference's shapes.
        h32, a_src, a_dst = linear_gat_fwd(x.float(), weight, att_src, att_dst, heads, chan)
        return h32.to(torch.bfloat16), a_src, a_dst
    out = torch.empty(n, heads * chan, dtype=x.dtype, device=x.device)
    a_src = torch.empty(n, heads, dtype=torch.float32, device=x.device)
    a_dst = torch.empty_like(a_src)
    L.call("lgk_linear_gat_fwd", _chk(x, None, "x"), _chk(weight, torch.float32, "weight"), _chk(att_src, torch.float32, "att_src"),
                    _chk(att_dst, torch.float32, "att_dst"), _chk(out), _chk(a_src), _chk(a_dst), n, k, int(heads), int(chan), _dt(x),
                    _stream())
    return out, a_src, a_dst


def linear_bwd_x(dy: Tensor, weight: Tensor, out: Optional[Tensor] = None) -> Tensor:
    """dx = dy W (+ out).  Large aligned products (the hidden transforms of the 1e5..1e6-node configs) go through the
    tensor-core transform as dy (W^T)^T: dX is a forward product against the transposed weight, a [K, M] copy that
    costs nothing next to N x M; the FFMA kernel lgk_linear_bwd_x keeps the small and unaligned ones."""
    L = _lib.lib()
    n, m = dy.shape
    k = weight.size(1)
    if n >= 32768 and m >= 32 and m % 4 == 0 and k % 4 == 0 and dy.dtype == torch.float32:
        return linear_fwd(dy, weight.t().contiguous(), out=out)
    accumulate = out is not None
    if out is None:
        out = torch.empty(n, k, dtype=dy.dtype, device=dy.device)
    L.call("lgk_linear_bwd_x", _chk(dy, None, "dy"), _chk(weight, torch.float32, "weight"), _chk(out, dy.dtype),
           n, k, m, int(accumulate), _dt(dy), _stream())
    return out


def linear_bwd_w(dy: Tensor, x: Optional[Tensor], want_dw: bool = True, want_db: bool = False,
                 dw_out: Optional[Tensor] = None, db_out: Optional[Tensor] = None):
    """-> (dW [M,K] | None, db [M] | None).  dw_out / db_out: ADD the gradients into these buffers (the flat gradient
    bucket of train.FlatParams) instead of returning fresh tensors -- both must be given for the outputs wanted."""
    L = _lib.lib()
    n, m = dy.shape
    k = x.size(1) if x is not None else 1
    dev = dy.device
    accumulate = (dw_out is not None) or (db_out is not None)
    if accumulate and ((want_dw and dw_out is None) or (want_db and db_out is None)):
        raise ValueError("accumulating form needs an output buffer for every requested gradient")
    dw = (dw_out if accumulate else torch.empty(m, k, dtype=torch.float32, device=dev)) if want_dw else None
    db = (db_out if accumulate else torch.empty(m, dtype=torch.float32, device=dev)) if want_db else None
    ws = _ws(L.query("lgk_linear_bwd_w_workspace_bytes", n, k, m), dev)
    L.call("lgk_linear_bwd_w", _chk(dy, None, "dy"), _chk(x, dy.dtype, "x") if want_dw else None,
           _chk(dw, torch.float32, "dw"), _chk(db, torch.float32, "db"),
           n, k, m, int(accumulate), ws.data_ptr(), ws.numel(), _dt(dy), _stream())
    return (None, None) if accumulate else (dw, db)


def colsum(dy: Tensor, out: Optional[Tensor] = None) -> Optional[Tensor]:
    return linear_bwd_w(dy, None, want_dw=False, want_db=True, db_out=out)[1]


# -------------------------------------------------------------------------------------------- elementwise


def elu_fwd(x: Tensor) -> Tensor:
    L = _lib.lib()
    y = torch.empty_like(x)
    L.call("lgk_elu_fwd", _chk(x, None, "x"), _chk(y), x.numel(), _dt(x), _stream())
    return y


def elu_bwd(y: Tensor, dy: Tensor) -> Tensor:
    L = _lib.lib()
    dx = torch.empty_like(y)
    L.call("lgk_elu_bwd", _chk(y, None, "y"), _chk(dy, y.dtype, "dy"), _chk(dx), y.numel(), _dt(y), _stream())
    return dx


def adam_step(param: Tensor, grad: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, *, lr: float, betas=(0.9, 0.999),
              eps: float = 1e-8, weight_decay: float = 0.0, step: int = 1, step_dev: Optional[Tensor] = None,
              grad_scale: float = 1.0, adamw: bool = False, zero_grad: bool = False, advance_step: bool = False) -> None:
    """zero_grad: the kernel clears every gradient it has consumed (the next step needs no fill launch);
    advance_step: step_dev is int64[2] = {steps completed, 0} and the kernel itself moves it on by one"""
    if advance_step and (step_dev is None or step_dev.numel() < 2):
        raise ValueError("advance_step needs a two-element device step counter")
    L = _lib.lib()
    f32 = torch.float32
    L.call("lgk_adam_step", _chk(param, f32, "param"), _chk(grad, f32, "grad"), _chk(exp_avg, f32), _chk(exp_avg_sq, f32),
           param.numel(), float(lr), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), int(step),
           _chk(step_dev, torch.int64, "step_dev"), float(grad_scale), int(adamw), int(zero_grad), int(advance_step), _stream())


def dp_allreduce_adam(peer_base_dev: int, rank: int, world: int, param: Tensor, exp_avg: Tensor, exp_avg_sq: Tensor, *, lr: float,
                      betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0, step_dev: Tensor, epoch_dev: Tensor,
                      adamw: bool = False, advance_step: bool = False) -> None:
    """gradient all-reduce over peer memory fused with Adam (include/lgk.h lgk_dp_allreduce_adam); peer_base_dev: address of the
    device array of the ranks' symmetric buffers; advance_step: the kernel itself moves step_dev[0] on by one"""
    f32 = torch.float32
    _lib.lib().call("lgk_dp_allreduce_adam", int(peer_base_dev), int(rank), int(world), param.numel(), _chk(param, f32, "param"),
                    _chk(exp_avg, f32), _chk(exp_avg_sq, f32), float(lr), float(betas[0]), float(betas[1]), float(eps),
                    float(weight_decay), _chk(step_dev, torch.int64, "step_dev"), _chk(epoch_dev, torch.int64, "epoch_dev"),
                    int(adamw), int(advance_step), _stream())


def clamped_regression_loss(logits: Tensor, y: Tensor, lo: float, hi: float, kind: int, want_grad: bool = True,
                            grad_scale: float = 1.0):
    """-> (loss [] f32, grad_scale * d loss / d logits [n] f32 | None): MSE (kind 0) / SmoothL1 (kind 1) of clamp(logits, lo, hi)
    against int64 labels, forward and gradient in one launch (reference models/base.py:196-199, gat.py:94-95)"""
    L = _lib.lib()
    n = logits.numel()
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dl = torch.empty(n, dtype=torch.float32, device=logits.device) if want_grad else None
    L.call("lgk_clamped_regression_loss", _chk(logits, torch.float32, "logits"), _chk(y, torch.int64, "y"), n,
           float(lo), float(hi), int(kind), float(grad_scale), loss.data_ptr(), _chk(dl, torch.float32, "dlogits"), _stream())
    return loss, dl


def sgd_step(param: Tensor, grad: Tensor, momentum_buf: Optional[Tensor] = None, *, lr: float, momentum: float = 0.0,
             weight_decay: float = 0.0, nesterov: bool = False, grad_scale: float = 1.0, zero_grad: bool = False) -> None:
    """torch.optim.SGD over flat buffers (reference models/base.py:167-168), one launch"""
    L = _lib.lib()
    f32 = torch.float32
    L.call("lgk_sgd_step", _chk(param, f32, "param"), _chk(grad, f32, "grad"), _chk(momentum_buf, f32, "momentum_buf"),
           param.numel(), float(lr), float(momentum), float(weight_decay), int(nesterov), float(grad_scale), int(zero_grad), _stream())


def cross_entropy_loss(logits: Tensor, y: Tensor, class_weight: Optional[Tensor] = None, want_grad: bool = True,
                       grad_scale: float = 1.0):
    """-> (loss [] f32, grad_scale * d loss / d logits [n, C] | None): nn.CrossEntropyLoss(weight=class_weight), forward and
    gradient in one launch (reference models/base.py:93-94)"""
    L = _lib.lib()
    n, c = logits.shape
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    dl = torch.empty(n, c, dtype=torch.float32, device=logits.device) if want_grad else None
    L.call("lgk_cross_entropy_loss", _chk(logits, torch.float32, "logits"), _chk(y, torch.int64, "y"),
           _chk(class_weight, torch.float32, "class_weight"), n, c, float(grad_scale), loss.data_ptr(),
           _chk(dl, torch.float32, "dlogits"), _stream())
    return loss, dl


def gather_rows(src: Tensor, index: Tensor, out: Optional[Tensor] = None) -> Tensor:
    """out[i] = src[index[i]] for a [R, ...] tensor of any 4-byte-multiple row size (device-side collation by node index)"""
    L = _lib.lib()
    n = index.numel()
    if out is None:
        out = torch.empty((n,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    row_bytes = src[0].numel() * src.element_size() if src.size(0) else 0
    L.call("lgk_gather_rows", _chk(src, None, "src"), _chk(index, torch.int64, "index"), _chk(out, src.dtype, "out"), n, row_bytes,
           src.size(0), _stream())
    return out


def bn_act_fwd(x: Tensor, weight, bias, running_mean, running_var, num_batches_tracked, eps: float, momentum: float,
               training: bool, act: int, dropout_p: float = 0.0, seed: int = 0, seed_dev: Optional[Tensor] = None):
    """-> y, save_mean, save_invstd: dropout(act(batch_norm(x))) in three launches (two in eval mode)"""
    L = _lib.lib()
    n, d = x.shape
    f32 = torch.float32
    y = torch.empty_like(x)
    mean = torch.empty(d, dtype=f32, device=x.device)
    invstd = torch.empty(d, dtype=f32, device=x.device)
    ws = _ws(L.query("lgk_bn_workspace_bytes", n, d), x.device)
    L.call("lgk_bn_act_fwd", _chk(x, None, "x"), _chk(weight, f32, "weight"), _chk(bias, f32, "bias"), _chk(running_mean, f32),
           _chk(running_var, f32), _chk(num_batches_tracked, torch.int64), _chk(y), _chk(mean), _chk(invstd), n, d, float(eps),
           float(momentum), int(training), int(act), float(dropout_p), int(seed), _chk(seed_dev, torch.int64), ws.data_ptr(),
           ws.numel(), _dt(x), _stream())
    return y, mean, invstd


def bn_act_bwd(x: Tensor, dy: Tensor, weight, bias, mean: Tensor, invstd: Tensor, training: bool, act: int, dropout_p: float = 0.0,
               seed: int = 0, seed_dev: Optional[Tensor] = None, want_dw: bool = True, want_db: bool = True,
               dw_out: Optional[Tensor] = None, db_out: Optional[Tensor] = None):
    """-> dx, dweight | None, dbias | None; dw_out / db_out: ADD into these (gradient bucket) and return None for them"""
    L = _lib.lib()
    n, d = x.shape
    f32 = torch.float32
    dx = torch.empty_like(x)
    accumulate = dw_out is not None or db_out is not None
    dw = (dw_out if accumulate else torch.empty(d, dtype=f32, device=x.device)) if want_dw else None
    db = (db_out if accumulate else torch.empty(d, dtype=f32, device=x.device)) if want_db else None
    ws = _ws(L.query("lgk_bn_workspace_bytes", n, d), x.device)
    L.call("lgk_bn_act_bwd", _chk(x, None, "x"), _chk(dy, x.dtype, "dy"), _chk(weight, f32), _chk(bias, f32), _chk(mean, f32), _chk(invstd, f32),
           _chk(dx), _chk(dw, f32, "dweight"), _chk(db, f32, "dbias"), int(accumulate), n, d, int(training), int(act), float(dropout_p),
           int(seed), _chk(seed_dev, torch.int64), ws.data_ptr(), ws.numel(), _dt(x), _stream())
    return dx, (None if accumulate else dw), (None if accumulate else db)


def act_dropout_fwd(x: Tensor, act: int, dropout_p: float, seed: int = 0, seed_dev: Optional[Tensor] = None) -> Tensor:
    L = _lib.lib()
    y = torch.empty_like(x)
    L.call("lgk_act_dropout_fwd", _chk(x, None, "x"), _chk(y), x.numel(), int(act), float(dropout_p), int(seed),
           _chk(seed_dev, torch.int64), _dt(x), _stream())
    return y


def act_dropout_bwd(x: Tensor, dy: Tensor, act: int, dropout_p: float, seed: int = 0, seed_dev: Optional[Tensor] = None) -> Tensor:
    L = _lib.lib()
    dx = torch.empty_like(x)
    L.call("lgk_act_dropout_bwd", _chk(x, None, "x"), _chk(dy, x.dtype, "dy"), _chk(dx), x.numel(), int(act),
           float(dropout_p), int(seed), _chk(seed_dev, torch.int64), _dt(x), _stream())
    return dx
