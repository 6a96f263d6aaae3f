"""GPU parity of the drop-in layers (PyG signatures) against the fp64 oracle and the committed golden vectors:
transform kernels, GATConv (fused edge softmax), GCNConv, GINConv, SAGEConv, GraphConv."""
import os

import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_BF16, TOL_F32, assert_close, assert_parity, random_graph, row_rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "layers.pt")


def _run(layer_gpu, layer_ref, x, args_ref, dev, tol=TOL_F32, what="", tol_grad=None, fp32_floor=False):
    """same weights, same inputs; compares output, dx and every parameter gradient, whole-tensor and per-row.
    fp32_floor: the bound is max(tol, 4 x the error of the oracle itself run in fp32 on the CPU) -- tests/util.assert_parity"""
    import copy
    layer_gpu.load_state_dict({k: v.float() for k, v in layer_ref.state_dict().items()})
    layer_gpu = layer_gpu.to(dev)
    xr = x.clone().requires_grad_(True)
    want = layer_ref(xr, *args_ref)
    gen = torch.Generator().manual_seed(17)
    dy = torch.randn(want.shape, generator=gen, dtype=torch.float64)
    want.backward(dy)
    want32 = x32 = grads32 = None
    if fp32_floor:
        ref32 = copy.deepcopy(layer_ref).float()
        ref32.zero_grad()
        x32 = x.float().requires_grad_(True)
        want32 = ref32(x32, *[a if a.dtype == torch.int64 else a.float() for a in args_ref])
        want32.backward(dy.float())
        grads32 = {k: p.grad for k, p in ref32.named_parameters()}
    xg = x.float().to(dev).requires_grad_(True)
    args_gpu = [a.to(dev) if a.dtype == torch.int64 else a.float().to(dev) for a in args_ref]
    got = layer_gpu(xg, *args_gpu)
    got.backward(dy.float().to(dev))

    def check(a, b, b32, t, name):
        if fp32_floor:
            assert_parity(a, b, b32, name, t)
        else:
            assert_close(a, b, t, name)
            assert row_rel_err(a, b) <= 10 * t, f"{name}: per-row error {row_rel_err(a, b):.3e}"

    check(got, want, want32, tol, what + " out")
    check(xg.grad, xr.grad, None if x32 is None else x32.grad, tol_grad or tol, what + " dx")
    ref_grads = {k: p.grad for k, p in layer_ref.named_parameters()}
    scale = max(float(g.abs().max()) for g in ref_grads.values() if g is not None)
    for k, p in layer_gpu.named_parameters():
        if ref_grads[k] is not None:
            assert p.grad is not None, k
            if float(ref_grads[k].abs().max()) < 1e-9 * scale:
                # analytically zero gradient (a bias in front of BatchNorm): only rounding noise on either side
                assert float(p.grad.abs().max()) < 1e-4 * scale, k
                continue
            check(p.grad.view_as(ref_grads[k]), ref_grads[k], None if grads32 is None else grads32[k], tol_grad or tol, f"{what} d{k}")


@pytest.mark.parametrize("n,k,m", [(1, 1, 1), (7, 3, 5), (300, 16, 64), (1000, 64, 64), (513, 1025, 64), (2000, 128, 128),
                                   (4096, 64, 5), (100, 64, 1), (129, 20, 130),
                                   # TMA-fed tensor-core kernel: K not a multiple of the 32-wide block, partial row and
                                   # column tiles, and n >= 32768 where dX runs as a forward product against W^T
                                   (600, 36, 70), (33000, 100, 36), (40000, 64, 64)])
def test_linear_fwd_bwd(dev, n, k, m):
    from lesion_gnn_b200.nn import functional as F
    gen = torch.Generator().manual_seed(n + k + m)
    x = torch.randn(n, k, generator=gen, dtype=torch.float64)
    w = torch.randn(m, k, generator=gen, dtype=torch.float64) / k ** 0.5
    b = torch.randn(m, generator=gen, dtype=torch.float64)
    dy = torch.randn(n, m, generator=gen, dtype=torch.float64)
    xr, wr, br = (t.clone().requires_grad_(True) for t in (x, w, b))
    want = torch.nn.functional.elu(torch.nn.functional.linear(xr, wr, br))
    want.backward(dy)
    xg, wg, bg = (t.float().to(dev).requires_grad_(True) for t in (x, w, b))
    got = F.linear(xg, wg, bg, act=1)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, "linear out")
    assert_close(xg.grad, xr.grad, TOL_F32, "linear dx")
    assert_close(wg.grad, wr.grad, TOL_F32, "linear dW")
    assert_close(bg.grad, br.grad, TOL_F32, "linear db")


@pytest.mark.parametrize("n", [300, 5000])
@pytest.mark.parametrize("k", [64, 36, 1025])
@pytest.mark.parametrize("heads,chan", [(4, 16), (2, 32), (2, 64), (1, 64), (8, 8), (16, 4), (3, 20), (1, 260)])
def test_gat_transform_with_logit_halves(dev, n, k, heads, chan):
    """lgk_linear_gat_fwd: h = x W^T and a_src / a_dst = <h[n, h, :], att[h, :]> from the transform's accumulators (tensor-core
    epilogue for tall aligned products, FFMA epilogue otherwise, separate kernel for head widths neither covers)"""
    from lesion_gnn_b200 import ops
    gen = torch.Generator().manual_seed(n + k + heads * chan)
    x = torch.randn(n, k, generator=gen, dtype=torch.float64)
    w = torch.randn(heads * chan, k, generator=gen, dtype=torch.float64) / k ** 0.5
    a_s, a_d = (torch.randn(heads * chan, generator=gen, dtype=torch.float64) for _ in range(2))
    h = x @ w.t()
    want_s = (h.view(n, heads, chan) * a_s.view(1, heads, chan)).sum(-1)
    want_d = (h.view(n, heads, chan) * a_d.view(1, heads, chan)).sum(-1)
    got_h, got_s, got_d = ops.linear_gat_fwd(x.float().to(dev), w.float().to(dev), a_s.float().to(dev), a_d.float().to(dev), heads, chan)
    assert_close(got_h, h, TOL_F32, "h")
    assert_close(got_s, want_s, TOL_F32, "a_src")
    assert_close(got_d, want_d, TOL_F32, "a_dst")


def test_linear_two_operand_form(dev):
    from lesion_gnn_b200.nn import functional as F
    gen = torch.Generator().manual_seed(0)
    x, x2 = (torch.randn(700, 48, generator=gen, dtype=torch.float64) for _ in range(2))
    w, w2 = (torch.randn(32, 48, generator=gen, dtype=torch.float64) for _ in range(2))
    b = torch.randn(32, generator=gen, dtype=torch.float64)
    ts = [t.clone().requires_grad_(True) for t in (x, w, b, x2, w2)]
    want = ts[0] @ ts[1].t() + ts[2] + ts[3] @ ts[4].t()
    want.backward(torch.ones_like(want))
    tg = [t.float().to(dev).requires_grad_(True) for t in (x, w, b, x2, w2)]
    got = F.linear(tg[0], tg[1], tg[2], x2=tg[3], weight2=tg[4])
    got.backward(torch.ones_like(got))
    assert_close(got, want, TOL_F32)
    for a, r in zip(tg, ts):
        assert_close(a.grad, r.grad, TOL_F32)


@pytest.mark.parametrize("heads,chan", [(1, 4), (2, 64), (4, 16), (4, 32), (8, 8), (3, 20), (1, 260)])
@pytest.mark.parametrize("add_self_loops", [True, False])
def test_gatconv(dev, heads, chan, add_self_loops):
    import lesion_gnn_b200 as lg
    n, fin = 257, 24
    ei = random_graph(n, 1800, heads * chan)
    ei = ei[:, ei[1] != 11]  # node 11: no in-edges (only its added self loop, or nothing at all)
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(n, fin, generator=gen, dtype=torch.float64)
    torch.manual_seed(2)
    ref = O.GATConv(fin, chan, heads=heads, add_self_loops=add_self_loops).double()
    ref.bias.data.normal_()
    _run(lg.nn.GATConv(fin, chan, heads=heads, add_self_loops=add_self_loops), ref, x, [ei], dev,
         what=f"GAT H={heads} C={chan} loops={add_self_loops}")


def test_gatconv_mean_heads(dev):
    import lesion_gnn_b200 as lg
    ei = random_graph(100, 600, 0)
    x = torch.randn(100, 12, dtype=torch.float64)
    torch.manual_seed(0)
    ref = O.GATConv(12, 8, heads=3, concat=False).double()
    _run(lg.nn.GATConv(12, 8, heads=3, concat=False), ref, x, [ei], dev, what="GAT concat=False")


def test_gatconv_large_logits_are_stable(dev):
    """online softmax must survive logits of +-80 (exp overflow territory without the running max).  With
    near-one-hot attention the gradients are differences of nearly equal terms, so fp32 itself (the oracle run in
    fp32 included) is only good to ~1e-3 there: forward keeps the 1e-5 bar, backward gets the fp32 oracle's own
    error (x4) as its bar."""
    import lesion_gnn_b200 as lg
    ei = random_graph(64, 700, 4)
    x = torch.randn(64, 8, dtype=torch.float64) * 30
    torch.manual_seed(1)
    ref = O.GATConv(8, 4, heads=2).double()
    ref.att_src.data.mul_(5)
    ref.att_dst.data.mul_(5)
    _run(lg.nn.GATConv(8, 4, heads=2), ref, x, [ei], dev, tol=TOL_F32, what="GAT large logits", fp32_floor=True)


def test_gat_attention_dropout_is_consistent_and_unbiased(dev):
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.nn import functional as F
    n, H, C = 2000, 2, 8
    ei = random_graph(n, 30000, 3)
    g = lg.CSRGraph.from_edge_index(ei.to(dev), n)
    h = torch.randn(n, H * C, device=dev)
    att = torch.randn(1, H, C, device=dev)
    base = F.gat_attention(h, att, att, None, g, H, C)
    acc = torch.zeros_like(base)
    reps = 200
    for s in range(reps):
        acc += F.gat_attention(h, att, att, None, g, H, C, dropout_p=0.35, seed=1000 + s)
    # E[dropout(alpha)] = alpha
    assert float((acc / reps - base).abs().mean() / base.abs().mean()) < 0.05
    a = F.gat_attention(h, att, att, None, g, H, C, dropout_p=0.35, seed=5)
    b = F.gat_attention(h, att, att, None, g, H, C, dropout_p=0.35, seed=5)
    assert torch.equal(a, b) and not torch.equal(a, base)
    # backward uses the same mask as forward: finite-difference check of d<out, r>/dh along a random direction
    hd = h.double()
    r = torch.randn_like(base)
    hh = h.clone().requires_grad_(True)
    F.gat_attention(hh, att, att, None, g, H, C, dropout_p=0.35, seed=5).mul(r).sum().backward()
    d = torch.randn_like(h)
    eps = 1e-2
    fp = F.gat_attention((hd + eps * d).float(), att, att, None, g, H, C, dropout_p=0.35, seed=5).double().mul(r).sum()
    fm = F.gat_attention((hd - eps * d).float(), att, att, None, g, H, C, dropout_p=0.35, seed=5).double().mul(r).sum()
    fd = float((fp - fm) / (2 * eps))
    an = float((hh.grad.double() * d).sum())
    assert abs(fd - an) <= 2e-2 * max(abs(fd), abs(an), 1.0), (fd, an)


def _gat_attention_with_mask(h, att_src, att_dst, rowptr, col, H, C, keep_edge, keep_self, p, slope=0.2):
    """fp64 restatement of GATConv's attention (oracle order: drop stored loops, add one loop per node, leaky_relu,
    softmax per destination, dropout with a GIVEN keep mask, weighted sum) on the CSR's own edge order"""
    n = rowptr.numel() - 1
    dst = torch.repeat_interleave(torch.arange(n), rowptr[1:] - rowptr[:-1])
    src = col
    nl = src != dst
    src_all = torch.cat([src[nl], torch.arange(n)])
    dst_all = torch.cat([dst[nl], torch.arange(n)])
    keep = torch.cat([keep_edge[nl], keep_self]).to(h.dtype) / (1.0 - p)          # [E', H]
    hv = h.view(n, H, C)
    a_s = (hv * att_src.view(1, H, C)).sum(-1)
    a_d = (hv * att_dst.view(1, H, C)).sum(-1)
    e = torch.nn.functional.leaky_relu(a_s[src_all] + a_d[dst_all], slope)
    alpha = O.softmax(e, dst_all, n) * keep
    msg = hv[src_all] * alpha.unsqueeze(-1)
    return torch.zeros(n, H, C, dtype=h.dtype).index_add_(0, dst_all, msg).view(n, H * C)


@pytest.mark.parametrize("H,C", [(2, 8), (4, 16), (2, 64)])
@pytest.mark.parametrize("clock", [None, 7, 8])
def test_gat_attention_dropout_matches_the_oracle_given_the_mask(dev, H, C, clock):
    """attention dropout (reference configs/config.py:63: p = 0.35) EXACTLY: the kernels' keep decision is a pure function
    of (seed, device step counter, CSR slot, head) restated on the host (tests/util.attention_dropout_keep); with that
    mask the fp64 oracle must reproduce the forward and every gradient at 1e-5 -- so forward and both backward passes
    use one and the same mask, with and without the device-side clock (captured graphs freeze the host seed)"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.nn import functional as F
    from tests.util import attention_dropout_keep
    n, p, seed = 900, 0.35, 12345
    ei = random_graph(n, 7000, 4)
    g = lg.CSRGraph.from_edge_index(ei.to(dev), n)
    rowptr, col = g.rowptr.cpu().long(), g.col.cpu().long()
    E = col.numel()
    gen = torch.Generator().manual_seed(3)
    h = torch.randn(n, H * C, generator=gen, dtype=torch.float64)
    a_src = torch.randn(1, H, C, generator=gen, dtype=torch.float64) * 0.5
    a_dst = torch.randn(1, H, C, generator=gen, dtype=torch.float64) * 0.5
    dy = torch.randn(n, H * C, generator=gen, dtype=torch.float64)
    heads = torch.arange(H)
    keep_edge = attention_dropout_keep(seed, clock, torch.arange(E).repeat_interleave(H), heads.repeat(E), p).view(E, H)
    keep_self = attention_dropout_keep(seed, clock, (torch.arange(n) | 0x80000000).repeat_interleave(H), heads.repeat(n), p).view(n, H)
    assert 0.55 < float(keep_edge.float().mean()) < 0.75
    hr, sr, dr = h.clone().requires_grad_(True), a_src.clone().requires_grad_(True), a_dst.clone().requires_grad_(True)
    want = _gat_attention_with_mask(hr, sr, dr, rowptr, col, H, C, keep_edge, keep_self, p)
    want.backward(dy)
    hg = h.float().to(dev).requires_grad_(True)
    sg, dg = a_src.float().to(dev).requires_grad_(True), a_dst.float().to(dev).requires_grad_(True)
    clock_dev = None if clock is None else torch.tensor([clock, 0], dtype=torch.int64, device=dev)
    got = F.gat_attention(hg, sg, dg, None, g, H, C, dropout_p=p, seed=seed, seed_dev=clock_dev)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, "out")
    assert_close(hg.grad, hr.grad, TOL_F32, "dh")
    assert_close(sg.grad, sr.grad, TOL_F32, "datt_src")
    assert_close(dg.grad, dr.grad, TOL_F32, "datt_dst")


@pytest.mark.parametrize("fin,fout", [(16, 64), (64, 64), (128, 128), (7, 5)])
@pytest.mark.parametrize("weighted", [False, True])
def test_gcnconv(dev, fin, fout, weighted):
    import lesion_gnn_b200 as lg
    n = 400
    ei = torch.cat([random_graph(n, 3000, fin), torch.tensor([[5, 5, 8], [5, 5, 8]])], 1)
    gen = torch.Generator().manual_seed(fin)
    x = torch.randn(n, fin, generator=gen, dtype=torch.float64)
    args = [ei] + ([torch.rand(ei.size(1), generator=gen, dtype=torch.float64) + 0.1] if weighted else [])
    torch.manual_seed(3)
    ref = O.GCNConv(fin, fout).double()
    ref.bias.data.normal_()
    _run(lg.nn.GCNConv(fin, fout), ref, x, args, dev, what=f"GCN {fin}->{fout} w={weighted}")


@pytest.mark.parametrize("opts", [dict(add_self_loops=False), dict(normalize=False), dict(improved=True), dict(bias=False)])
def test_gcnconv_options(dev, opts):
    import lesion_gnn_b200 as lg
    ei = random_graph(150, 900, 2)
    x = torch.randn(150, 12, dtype=torch.float64)
    torch.manual_seed(4)
    ref = O.GCNConv(12, 8, **opts).double()
    _run(lg.nn.GCNConv(12, 8, **opts), ref, x, [ei], dev, what=f"GCN {opts}")


@pytest.mark.parametrize("aggr", ["mean", "max", "sum"])
def test_sageconv(dev, aggr):
    import lesion_gnn_b200 as lg
    ei = random_graph(500, 4000, 6)
    x = torch.randn(500, 64, dtype=torch.float64)
    torch.manual_seed(5)
    ref = O.SAGEConv(64, 48, aggr=aggr).double()
    _run(lg.nn.SAGEConv(64, 48, aggr=aggr), ref, x, [ei], dev, what=f"SAGE {aggr}")


@pytest.mark.parametrize("weighted", [False, True])
def test_graphconv(dev, weighted):
    import lesion_gnn_b200 as lg
    ei = random_graph(300, 2500, 7)
    gen = torch.Generator().manual_seed(7)
    x = torch.randn(300, 110, generator=gen, dtype=torch.float64)
    args = [ei] + ([torch.rand(ei.size(1), generator=gen, dtype=torch.float64)] if weighted else [])
    torch.manual_seed(6)
    _run(lg.nn.GraphConv(110, 1), O.GraphConv(110, 1).double(), x, args, dev, what="GraphConv 110->1")
    _run(lg.nn.GraphConv(110, 110), O.GraphConv(110, 110).double(), x, args, dev, what="GraphConv 110->110")


@pytest.mark.parametrize("norm", [None, "batch_norm"])
def test_ginconv(dev, norm):
    import lesion_gnn_b200 as lg
    ei = random_graph(600, 5000, 8)
    x = torch.randn(600, 64, dtype=torch.float64)
    torch.manual_seed(7)
    ref = O.GINConv(O.MLP([64, 64, 64], act="ELU", dropout=0.0, norm=norm)).double()
    got = lg.nn.GINConv(lg.nn.MLP([64, 64, 64], act="ELU", dropout=0.0, norm=norm))
    _run(got, ref, x, [ei], dev, tol=TOL_F32, what=f"GIN norm={norm}", fp32_floor=True)


@pytest.mark.parametrize("n,d", [(2, 4), (777, 64), (5000, 100), (60000, 64)])
@pytest.mark.parametrize("affine", [True, False])
def test_bn_act_matches_torch_batchnorm_elu(dev, n, d, affine):
    """lgk_bn_act_fwd / bwd (BatchNorm1d + ELU of PyG's MLP, reference models/gin.py:23) against torch.nn.BatchNorm1d + F.elu in
    fp64: training mode (batch statistics, running statistics and num_batches_tracked updated), then eval mode"""
    from lesion_gnn_b200.nn import functional as F
    from lesion_gnn_b200._lib import ACT_ELU
    gen = torch.Generator().manual_seed(n + d)
    x = torch.randn(n, d, generator=gen, dtype=torch.float64) * 3 + 7        # a large mean: cancellation in naive variance
    dy = torch.randn(n, d, generator=gen, dtype=torch.float64)
    ref = torch.nn.BatchNorm1d(d, affine=affine).double()
    got = torch.nn.BatchNorm1d(d, affine=affine).to(dev)
    if affine:
        with torch.no_grad():
            ref.weight.copy_(torch.rand(d, generator=gen, dtype=torch.float64) + 0.5)
            ref.bias.copy_(torch.randn(d, generator=gen, dtype=torch.float64))
            got.weight.copy_(ref.weight.float())
            got.bias.copy_(ref.bias.float())
    for mode in ("train", "train", "eval"):
        ref.train(mode == "train"), got.train(mode == "train")
        ref.zero_grad(), got.zero_grad()
        state_before = {k: v.clone() for k, v in ref.state_dict().items()}
        xr = x.clone().requires_grad_(True)
        want = torch.nn.functional.elu(ref(xr))
        want.backward(dy)
        xg = x.float().to(dev).requires_grad_(True)
        out = F.bn_act(xg, got, ACT_ELU)
        out.backward(dy.float().to(dev))
        assert_close(out, want, TOL_F32, f"{mode} y")
        # dx of batch statistics is a difference of nearly equal terms (exactly zero for n = 2): its bar is what torch's
        # own fp32 BatchNorm loses on these inputs (x4), never below 1e-5
        ref32 = torch.nn.BatchNorm1d(d, affine=affine)
        ref32.load_state_dict({k: v.float() if v.is_floating_point() else v for k, v in state_before.items()})
        ref32.train(mode == "train")
        x32 = x.float().requires_grad_(True)
        torch.nn.functional.elu(ref32(x32)).backward(dy.float())
        assert_parity(xg.grad, xr.grad, x32.grad, f"{mode} dx")
        if affine:
            assert_close(got.weight.grad, ref.weight.grad, TOL_F32, f"{mode} dweight")
            assert_close(got.bias.grad, ref.bias.grad, TOL_F32, f"{mode} dbias")
        assert_close(got.running_mean, ref.running_mean, TOL_F32, "running_mean")
        assert_close(got.running_var, ref.running_var, TOL_F32, "running_var")
        assert int(got.num_batches_tracked) == int(ref.num_batches_tracked)


def test_act_dropout_masks_are_consistent(dev):
    """dropout(elu(x)) in one launch: kept elements are elu(x) / (1 - p), the backward zeroes exactly the dropped ones, the
    mask is a function of (seed, device clock) and changes with either"""
    from lesion_gnn_b200.nn import functional as F
    from lesion_gnn_b200._lib import ACT_ELU
    x = torch.randn(4000, 64, device=dev)
    p = 0.35
    clock = torch.tensor([3, 0], dtype=torch.int64, device=dev)
    xg = x.clone().requires_grad_(True)
    y = F.act_dropout(xg, ACT_ELU, p, 11, clock)
    y.backward(torch.ones_like(y))
    e = torch.nn.functional.elu(x)
    kept = y != 0
    assert 0.63 < float(kept.float().mean()) < 0.67
    assert_close(y[kept], (e / (1 - p))[kept], TOL_F32, "kept values")
    de = torch.where(x > 0, torch.ones_like(x), e + 1) / (1 - p)
    assert_close(xg.grad[kept], de[kept], TOL_F32, "kept gradients")
    assert float(xg.grad[~kept].abs().max()) == 0.0
    assert torch.equal(y, F.act_dropout(x, ACT_ELU, p, 11, clock))
    clock[0] += 1
    assert not torch.equal(y, F.act_dropout(x, ACT_ELU, p, 11, clock))
    assert torch.equal(F.act_dropout(x, ACT_ELU, 0.0), e) or float((F.act_dropout(x, ACT_ELU, 0.0) - e).abs().max()) < 1e-6


def test_state_dict_keys_match_pyg_names():
    import lesion_gnn_b200 as lg
    assert set(lg.nn.GATConv(8, 4, heads=2).state_dict()) == {"att_src", "att_dst", "bias", "lin.weight"}
    assert set(lg.nn.GCNConv(8, 4).state_dict()) == {"bias", "lin.weight"}
    assert set(lg.nn.SAGEConv(8, 4).state_dict()) == {"lin_l.weight", "lin_l.bias", "lin_r.weight"}
    assert set(lg.nn.GraphConv(8, 4).state_dict()) == {"lin_rel.weight", "lin_rel.bias", "lin_root.weight"}
    gin = lg.nn.GINConv(lg.nn.MLP([8, 4, 4]))
    assert {"eps", "nn.lins.0.weight", "nn.lins.1.bias", "nn.norms.0.module.running_mean"} <= set(gin.state_dict())


def test_golden_vectors(dev):
    """committed fixtures (tests/golden/make_golden.py): two-node, star with stored + duplicate edges, path with
    an isolated node over two graphs, mixed-size kNN batch incl. a single-node graph"""
    import lesion_gnn_b200 as lg
    gold = torch.load(GOLD)
    makers = {
        "gcn": (lambda: lg.nn.GCNConv(16, 8), False), "gcn_w": (lambda: lg.nn.GCNConv(16, 8), True),
        "gat": (lambda: lg.nn.GATConv(16, 4, heads=2), False),
        "gin": (lambda: lg.nn.GINConv(lg.nn.MLP([16, 8, 8], act="ELU", norm=None)), False),
        "sage_mean": (lambda: lg.nn.SAGEConv(16, 8, aggr="mean"), False),
        "sage_max": (lambda: lg.nn.SAGEConv(16, 8, aggr="max"), False),
        "graphconv_w": (lambda: lg.nn.GraphConv(16, 8), True),
    }
    for name, case in gold["cases"].items():
        ei, n, x = case["edge_index"].to(dev), case["num_nodes"], case["x"]
        g = lg.CSRGraph.from_edge_index(ei, n)
        for key in ("rowptr", "col", "perm", "colptr", "row", "csr2csc"):
            assert torch.equal(getattr(g, key).cpu().long(), case["csr"][key]), (name, key)
        for lname, (make, weighted) in makers.items():
            rec = case[lname]
            layer = make()
            layer.load_state_dict({k: v.float() for k, v in rec["state"].items()})
            layer = layer.to(dev)
            xg = x.float().to(dev).requires_grad_(True)
            args = [ei] + ([case["edge_weight"].float().to(dev)] if weighted else [])
            out = layer(xg, *args)
            out.backward(rec["dy"].float().to(dev))
            assert_close(out, rec["out"], TOL_F32, f"{name}/{lname} out")
            assert_close(xg.grad, rec["dx"], TOL_F32, f"{name}/{lname} dx")
            for k, p in layer.named_parameters():
                if k in rec["dparams"]:
                    assert_close(p.grad.view_as(rec["dparams"][k]), rec["dparams"][k], TOL_F32, f"{name}/{lname} d{k}")
        B = int(case["batch"].max()) + 1
        for red, fn in (("mean", lg.nn.global_mean_pool), ("max", lg.nn.global_max_pool), ("sum", lg.nn.global_add_pool)):
            rec = case["pool_" + red]
            xg = x.float().to(dev).requires_grad_(True)
            out = fn(xg, case["batch"].to(dev), B)
            out.backward(rec["dy"].float().to(dev))
            assert_close(out, rec["out"], TOL_F32, f"{name}/pool_{red}")
            assert_close(xg.grad, rec["dx"], TOL_F32, f"{name}/pool_{red} dx")


def test_layers_accept_adj_t_sparse_tensors(dev):
    """reference models/gat.py:87-90 hands the models `adj_t` (a torch_sparse.SparseTensor, rows = destinations) when
    compile=False: a torch sparse CSR tensor, a sparse COO tensor and any object with SparseTensor's `.csr()` work"""
    import lesion_gnn_b200 as lg
    from tests.util import random_graph
    n = 300
    ei = random_graph(n, 2500, 4, duplicates=False).to(dev)
    w = torch.rand(ei.size(1), device=dev) + 0.1
    x = torch.randn(n, 32, device=dev)
    conv = lg.nn.GraphConv(32, 16).to(dev)
    want = conv(x, ei, w)
    g = lg.CSRGraph.from_edge_index(ei, n, w)
    adj_csr = torch.sparse_csr_tensor(g.rowptr.long(), g.col.long(), g.value, size=(n, n))

    class FakeSparseTensor:  # torch_sparse.SparseTensor's accessor
        def csr(self):
            return g.rowptr.long(), g.col.long(), g.value

    for adj in (adj_csr, adj_csr.to_sparse_coo(), FakeSparseTensor()):
        got = conv(x, adj)
        assert torch.equal(got, want)
    gcn = lg.nn.GCNConv(32, 8).to(dev)
    unweighted = torch.sparse_csr_tensor(g.rowptr.long(), g.col.long(), torch.ones_like(g.value), size=(n, n))
    assert_close(gcn(x, unweighted), gcn(x, ei, torch.ones_like(w)), 1e-6, "gcn adj_t")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_narrow_and_small_weight_products(dev, dtype):
    """the products the tiled kernel is the wrong shape for (csrc/linear.cu): forward with <= 8 outputs (a warp per row; row
    scale, bias, accumulate, activation), its dX (outer-product kernel) and the dX against a weight of at most 128 x 128 (weight
    and a tile of dy rows in shared memory) -- aligned and unaligned reductions, column counts that leave lanes idle, ragged
    last tiles, the accumulating forms"""
    from lesion_gnn_b200 import ops
    gen = torch.Generator().manual_seed(21)
    tol = TOL_F32 if dtype == torch.float32 else TOL_BF16

    def dev_act(t):
        return t.float().to(dev).to(dtype)

    def rounded(t):  # what the kernel sees
        return t.float().to(dtype).double()

    for n, k, m in [(256, 128, 1), (4096, 64, 5), (37, 1025, 8), (1, 4, 1), (300, 30, 3)]:
        x = torch.randn(n, k, generator=gen, dtype=torch.float64)
        w = torch.randn(m, k, generator=gen, dtype=torch.float64) / k ** 0.5
        b = torch.randn(m, generator=gen, dtype=torch.float64)
        rs = torch.rand(n, generator=gen, dtype=torch.float64) + 0.5
        base = torch.randn(n, m, generator=gen, dtype=torch.float64)
        xw = rounded(x) @ w.float().double().t()
        got = ops.linear_fwd(dev_act(x), w.float().to(dev), b.float().to(dev), rs.float().to(dev), act=1)
        assert_close(got, torch.nn.functional.elu(rs.float().double()[:, None] * xw + b.float().double()), tol, f"narrow fwd {n}x{k}->{m}")
        out = dev_act(base).clone()
        ops.linear_fwd(dev_act(x), w.float().to(dev), out=out)
        assert_close(out, rounded(base) + xw, tol, f"narrow fwd accumulate {n}x{k}->{m}")
        dy = torch.randn(n, m, generator=gen, dtype=torch.float64)
        want_dx = rounded(dy) @ w.float().double()
        assert_close(ops.linear_bwd_x(dev_act(dy), w.float().to(dev)), want_dx, tol, f"narrow dX {n}x{m}->{k}")
        acc = dev_act(x).clone()
        ops.linear_bwd_x(dev_act(dy), w.float().to(dev), out=acc)
        assert_close(acc, rounded(x) + want_dx, tol, f"narrow dX accumulate {n}x{m}->{k}")
    for n, k, m in [(11447, 64, 64), (1000, 128, 128), (333, 16, 128), (70, 20, 12), (5, 128, 16), (4097, 100, 36)]:
        dy = torch.randn(n, m, generator=gen, dtype=torch.float64)
        w = torch.randn(m, k, generator=gen, dtype=torch.float64) / m ** 0.5
        base = torch.randn(n, k, generator=gen, dtype=torch.float64)
        want_dx = rounded(dy) @ w.float().double()
        got = ops.linear_bwd_x(dev_act(dy), w.float().to(dev))
        assert_close(got, want_dx, tol, f"small-weight dX {n}x{m}->{k}")
        if dtype == torch.float32:  # every element on its own: a misplaced tile would hide under max-normalisation
            assert float((got.double().cpu() - want_dx).abs().max()) <= 1e-5 * float(want_dx.abs().max())
        acc = dev_act(base).clone()
        ops.linear_bwd_x(dev_act(dy), w.float().to(dev), out=acc)
        assert_close(acc, rounded(base) + want_dx, tol, f"small-weight dX accumulate {n}x{m}->{k}")


def test_linear_tensor_core_epilogue_options(dev):
    """row scale, bias, accumulate-into-out and ELU of the tensor-core transform (aligned rows -> linear_tma.cu,
    1025-float rows -> linear_tc.cu), against fp64"""
    from lesion_gnn_b200 import ops
    gen = torch.Generator().manual_seed(9)
    for n, k, m in [(777, 64, 48), (900, 1025, 64)]:
        x = torch.randn(n, k, generator=gen, dtype=torch.float64)
        w = torch.randn(m, k, generator=gen, dtype=torch.float64) / k ** 0.5
        b = torch.randn(m, generator=gen, dtype=torch.float64)
        rs = torch.rand(n, generator=gen, dtype=torch.float64) + 0.5
        base = torch.randn(n, m, generator=gen, dtype=torch.float64)
        want = torch.nn.functional.elu(rs[:, None] * (x @ w.t()) + b)
        got = ops.linear_fwd(x.float().to(dev), w.float().to(dev), b.float().to(dev), rs.float().to(dev), act=1)
        assert_close(got, want, TOL_F32, f"rowscale+bias+elu {n}x{k}->{m}")
        out = base.float().to(dev).clone()
        ops.linear_fwd(x.float().to(dev), w.float().to(dev), out=out)
        assert_close(out, base + x @ w.t(), TOL_F32, f"accumulate {n}x{k}->{m}")


@pytest.mark.parametrize("n,k,m", [(11583, 1025, 64), (2048, 256, 128), (5000, 300, 100), (3001, 257, 40),
                                   (2500, 1024, 7), (4100, 513, 64)])
def test_wide_weight_gradient_on_tensor_cores(dev, n, k, m):
    """dW of a wide transform (grad_w_tc.cu: nodes as the MMA K dimension, both operands MN-major as they lie in
    memory): ragged last node block, last feature tile of one column, m not a multiple of 32, accumulate form"""
    from lesion_gnn_b200 import ops
    gen = torch.Generator().manual_seed(n + k + m)
    x = torch.randn(n, k, generator=gen, dtype=torch.float64)
    dy = torch.randn(n, m, generator=gen, dtype=torch.float64)
    want_w, want_b = dy.t() @ x, dy.sum(0)
    xg, dyg = x.float().to(dev), dy.float().to(dev)
    dw, db = ops.linear_bwd_w(dyg, xg, want_dw=True, want_db=True)
    assert_close(dw, want_w, TOL_F32, "dW")
    assert_close(db, want_b, TOL_F32, "db")
    # every element on its own, not only the largest: a misplaced tile would hide under max-normalisation
    assert float((dw.double().cpu() - want_w).abs().max()) <= 1e-5 * float(want_w.abs().max())
    base_w = torch.randn(m, k, generator=gen).to(dev)
    base_b = torch.randn(m, generator=gen).to(dev)
    acc_w, acc_b = base_w.clone(), base_b.clone()
    ops.linear_bwd_w(dyg, xg, want_dw=True, want_db=True, dw_out=acc_w, db_out=acc_b)
    assert_close(acc_w, base_w.double().cpu() + want_w, TOL_F32, "dW accumulate")
    assert_close(acc_b, base_b.double().cpu() + want_b, TOL_F32, "db accumulate")
    assert torch.equal(ops.linear_bwd_w(dyg, xg)[0], dw)   # deterministic


def test_gat_fused_elu_equals_separate_activation(dev):
    """GATConv(fused_elu=True) (the attention kernel writes elu(out) itself) against elu(GATConv(x)): bit-identical
    output, gradients within the fp32 bar"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.nn import functional as F
    from tests.util import knn_batch
    hb = knn_batch(12, "ddr", 32, 6, seed=9)
    torch.manual_seed(4)
    a = lg.nn.GATConv(32, 16, heads=4).to(dev)
    b = lg.nn.GATConv(32, 16, heads=4, fused_elu=True).to(dev)
    b.load_state_dict(a.state_dict())
    ei = hb.edge_index.to(dev)
    xa = hb.x.to(dev).requires_grad_(True)
    xb = hb.x.to(dev).requires_grad_(True)
    ya, yb = F.elu(a(xa, ei)), b(xb, ei)
    assert torch.equal(ya, yb)
    dy = torch.randn_like(ya)
    ya.backward(dy)
    yb.backward(dy)
    assert_close(xb.grad, xa.grad, TOL_F32, "dx")
    for (k, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert_close(pb.grad, pa.grad, TOL_F32, "d" + k)


def test_cluster_reduced_weight_gradient_opt_in(dev):
    """LGK_DW_CLUSTER=1 (csrc/linear.cu bwd_w_cluster_kernel: slab partials reduced through distributed shared memory, one
    launch): the knob is read once per process, so the check runs in a child process -- dW and db against fp64, plain and
    accumulating, several cluster counts, unaligned widths, and bitwise run-to-run determinism"""
    import os
    import subprocess
    import sys
    code = r"""
import torch
from lesion_gnn_b200 import ops
dev = torch.device("cuda")
g = torch.Generator().manual_seed(0)
for n, k, m in [(5, 3, 2), (300, 64, 64), (513, 60, 33), (11447, 64, 64), (40000, 64, 1), (2000, 17, 64)]:
    x = torch.randn(n, k, generator=g, dtype=torch.float64); dy = torch.randn(n, m, generator=g, dtype=torch.float64)
    xd, dyd = x.float().to(dev), dy.float().to(dev)
    ops.reset_launch_count()
    dw, db = ops.linear_bwd_w(dyd, xd, want_dw=True, want_db=True)
    assert ops.launch_count() == 1, ops.launch_count()
    dw2, db2 = ops.linear_bwd_w(dyd, xd, want_dw=True, want_db=True)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)
    want_w, want_b = dy.t() @ x, dy.sum(0)
    assert float((dw.double().cpu() - want_w).abs().max() / want_w.abs().max()) < 1e-5, (n, k, m)
    assert float((db.double().cpu() - want_b).abs().max() / want_b.abs().max().clamp(min=1e-30)) < 1e-5, (n, k, m)
    accw, accb = torch.ones(m, k, device=dev), torch.ones(m, device=dev)
    ops.linear_bwd_w(dyd, xd, want_dw=True, want_db=True, dw_out=accw, db_out=accb)
    assert float((accw.double().cpu() - 1 - want_w).abs().max() / want_w.abs().max()) < 1e-5
    only_w, none_b = ops.linear_bwd_w(dyd, xd, want_dw=True, want_db=False)
    assert none_b is None and torch.equal(only_w, dw)
print("ok")
"""
    env = dict(os.environ, LGK_DW_CLUSTER="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "ok" in res.stdout, res.stdout + res.stderr
