"""GPU parity of the windowed, row-blocked aggregation (lgk_quad_build + lgk_spmm_blocked, csrc/spmm_win.cu):
the quad structure bit-exact against a host restatement of the four-way merge, the aggregation against the fp64
oracle (<= 1e-5) and against lgk_spmm on the same graph, for graphs that exercise every path of the producer:
windowed chunks (raster / block-diagonal batches), gaps between graphs, backward jumps, chunks too wide for the
ring (global-gather path), multigraph duplicates, stored self loops, empty rows, ragged sizes."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_F32, assert_close, random_graph

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _force_blocked(monkeypatch):
    from lesion_gnn_b200 import ops
    monkeypatch.setattr(ops, "_WIN_MIN_ELEMS", 0)


def _graph(ei, n, dev):
    from lesion_gnn_b200 import CSRGraph
    return CSRGraph.from_edge_index(ei.to(dev), n)


def _host_quads(rowptr, col, n):
    """four-way merge of the sorted rows of every quad: [(source, mask)] per quad, and the per-chunk source range"""
    rowptr, col = rowptr.tolist(), col.tolist()
    quads = []
    for g in range((n + 3) // 4):
        rows = [col[rowptr[r]:rowptr[r + 1]] if r < n else [] for r in range(4 * g, 4 * g + 4)]
        pos = [0, 0, 0, 0]
        ent = []
        while True:
            heads = [rows[k][pos[k]] if pos[k] < len(rows[k]) else None for k in range(4)]
            live = [h for h in heads if h is not None]
            if not live:
                break
            m = min(live)
            mask = 0
            for k in range(4):
                if heads[k] == m:
                    mask |= 1 << k
                    pos[k] += 1
            ent.append((m, mask))
        quads.append(ent)
    return quads


def _band_graph(n, k, seed, width=40):  # noqa: E302
    """every node i receives k edges from nodes within +-width (raster-like band), duplicates possible"""
    g = torch.Generator().manual_seed(seed)
    dst = torch.arange(n).repeat_interleave(k)
    off = torch.randint(-width, width + 1, (n * k,), generator=g)
    src = (dst + off).clamp(0, n - 1)
    return torch.stack([src, dst])


def _anti_diagonal_graph(n, k, seed, block=300):
    """destination block b reads from source block (B-1-b): every chunk's window lies BELOW the previous one"""
    g = torch.Generator().manual_seed(seed)
    nb = (n + block - 1) // block
    dst = torch.arange(n).repeat_interleave(k)
    b = dst // block
    src_b = nb - 1 - b
    src = (src_b * block + torch.randint(0, block, (n * k,), generator=g)).clamp(0, n - 1)
    return torch.stack([src, dst])


GRAPHS = {
    "band": lambda: (_band_graph(3001, 9, 1), 3001),                                # narrow windows: 256-byte slices
    "band_mid": lambda: (_band_graph(9000, 6, 7, width=500), 9000),                 # too wide for those: 128-byte slices
    "band_few_wide": lambda: (torch.cat([_band_graph(40000, 5, 8), torch.tensor([[39000, 5], [17, 20001], [3, 20002]]).t()], 1), 40000),
    "band_wide_direct": lambda: (_band_graph(9000, 6, 2, width=4000), 9000),       # window > ring: global gathers
    "random_direct": lambda: (random_graph(6000, 40000, 3), 6000),
    "anti_diagonal": lambda: (_anti_diagonal_graph(2500, 7, 4), 2500),             # backward jumps, drains
    "tiny": lambda: (random_graph(3, 7, 5), 3),
    "no_edges": lambda: (torch.zeros(2, 0, dtype=torch.long), 37),
    "sparse_rows": lambda: (random_graph(4000, 900, 6), 4000),                     # most rows and some chunks empty
}


def _knn_batch_graph(seed):
    from lesion_gnn_b200 import data
    hb = data.synth_batch(40, "ddr", 4, seed=seed)
    return O.knn_graph(hb.pos, 6, hb.batch, loop=True), hb.num_nodes


@pytest.mark.parametrize("name", sorted(GRAPHS) + ["knn_batch"])
def test_quad_structure_is_the_four_way_merge(dev, name):
    ei, n = _knn_batch_graph(11) if name == "knn_batch" else GRAPHS[name]()
    g = _graph(ei, n, dev)
    for which, ptr, idx in (("csr", g.rowptr, g.col), ("csc", g.colptr, g.row)):
        q = g.quads(which)
        ptr_h, idx_h = ptr.cpu(), idx.cpu()
        want = _host_quads(ptr_h, idx_h, n)
        qhdr = q.qlen.cpu().view(-1, 2)
        qlen, qent = qhdr[:, 0].tolist(), q.qent.cpu()
        los, his = [], []
        for gq, ent in enumerate(want):
            assert qlen[gq] == len(ent), (which, gq)
            start = int(ptr_h[min(4 * gq, n)])
            words = qent[start:start + len(ent)].tolist()
            got = [((w & 0xFFFFFFFF) & 0x0FFFFFFF, ((w & 0xFFFFFFFF) >> 28) & 0xF) for w in words]
            assert got == ent, (which, gq)
            selfc = sum(min(255, sum(1 for m, k in ent if m == 4 * gq + r and (k >> r) & 1)) << (8 * r) for r in range(4))
            assert int(qhdr[gq, 1]) & 0xFFFFFFFF == selfc, (which, gq)
        for c in range((len(want) + 15) // 16):
            cols = [m for ent in want[16 * c:16 * c + 16] for m, _ in ent]
            los.append(min(cols) if cols else 2 ** 31 - 1)
            his.append(max(cols) if cols else -1)
        nc = len(los)
        assert q.cmin.cpu().tolist()[:nc] == los and q.cmax.cpu().tolist()[:nc] == his, which
        widths = [h - l + 1 for l, h in zip(los, his) if h >= l]
        meta = q.qmeta.cpu().tolist()
        assert meta[0] == (max(widths) if widths else 0) and meta[1] == sum(w > 700 for w in widths), (which, meta)


@pytest.mark.parametrize("feat", [4, 32, 64, 100, 128, 132])
@pytest.mark.parametrize("name", sorted(GRAPHS) + ["knn_batch"])
def test_blocked_sum_matches_oracle_and_lgk_spmm(dev, name, feat):
    from lesion_gnn_b200 import ops
    ei, n = _knn_batch_graph(12) if name == "knn_batch" else GRAPHS[name]()
    gen = torch.Generator().manual_seed(feat)
    x = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    want = O.propagate(x, ei, n, "sum")
    g = _graph(ei, n, dev)
    xd = x.float().to(dev)
    ops.reset_launch_count()
    got = ops.spmm(g.rowptr, g.col, xd, "sum", quads=g.quads("csr"))
    assert_close(got, want, TOL_F32, f"{name} F={feat} fwd")
    old = ops.spmm(g.rowptr, g.col, xd, "sum")
    assert_close(got, old.double(), 2e-6, f"{name} F={feat} vs lgk_spmm")
    # transpose direction on the CSC
    want_t = O.propagate(x, ei.flip(0), n, "sum")
    got_t = ops.spmm(g.colptr, g.row, xd, "sum", quads=g.quads("csc"))
    assert_close(got_t, want_t, TOL_F32, f"{name} F={feat} transpose")
    again = ops.spmm(g.rowptr, g.col, xd, "sum", quads=g.quads("csr"))
    assert torch.equal(got, again), "run-to-run difference"


@pytest.mark.parametrize("name", ["band", "random_direct", "knn_batch", "anti_diagonal"])
def test_blocked_mean_and_bias(dev, name):
    from lesion_gnn_b200 import ops
    ei, n = _knn_batch_graph(13) if name == "knn_batch" else GRAPHS[name]()
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(n, 64, generator=gen, dtype=torch.float64)
    b = torch.randn(64, generator=gen, dtype=torch.float64)
    want = O.propagate(x, ei, n, "mean") + b
    g = _graph(ei, n, dev)
    got = ops.spmm(g.rowptr, g.col, x.float().to(dev), "mean", bias=b.float().to(dev), quads=g.quads("csr"))
    assert_close(got, want, TOL_F32, name)


@pytest.mark.parametrize("name", ["band", "random_direct", "knn_batch", "tiny"])
@pytest.mark.parametrize("mode", ["gcn", "gin"])
def test_blocked_self_loop_modes_through_the_layers(dev, name, mode):
    """GCNConv (pre / post scaling, SELF_REPLACE with stored, duplicated and missing self loops) and GINConv's
    SELF_ADD term, forward and backward, through the windowed kernel against the fp64 oracle"""
    import lesion_gnn_b200 as lg
    ei, n = _knn_batch_graph(14) if name == "knn_batch" else GRAPHS[name]()
    gen = torch.Generator().manual_seed(9)
    x = torch.randn(n, 32, generator=gen, dtype=torch.float64)
    dy = torch.randn(n, 32, generator=gen, dtype=torch.float64)
    if mode == "gcn":
        ref = O.GCNConv(32, 32).double()
        mod = lg.nn.GCNConv(32, 32)
    else:
        ref = O.GINConv(torch.nn.Linear(32, 32).double(), eps=0.25)
        mod = lg.nn.GINConv(lg.nn.Linear(32, 32), eps=0.25)
        mod.nn.weight.data.copy_(ref.nn.weight.data.float())
        mod.nn.bias.data.copy_(ref.nn.bias.data.float())
    if mode == "gcn":
        mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    xr = x.clone().requires_grad_(True)
    want = ref(xr, ei)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = mod(xg, _graph(ei, n, dev))
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, f"{mode} {name} fwd")
    assert_close(xg.grad, xr.grad, TOL_F32, f"{mode} {name} bwd")


def test_blocked_kernel_is_selected_by_size(dev, monkeypatch):
    """large inputs take the windowed kernel, small ones lgk_spmm -- both through F.aggregate"""
    from lesion_gnn_b200 import _lib, ops
    from lesion_gnn_b200.nn import functional as F
    calls = []
    L = _lib.lib()
    monkeypatch.setattr(L, "hook", lambda phase, name, args: calls.append(name) if phase == "begin" else None)
    ei, n = GRAPHS["band"]()
    g = _graph(ei, n, dev)
    x = torch.randn(n, 64, device=dev)
    monkeypatch.setattr(ops, "_WIN_MIN_ELEMS", 1 << 40)
    F.aggregate(x, g, "sum")
    assert "lgk_spmm" in calls and "lgk_spmm_blocked" not in calls
    calls.clear()
    monkeypatch.setattr(ops, "_WIN_MIN_ELEMS", 0)
    F.aggregate(x, g, "sum")
    assert "lgk_spmm_blocked" in calls and "lgk_spmm" not in calls


@pytest.mark.parametrize("heads,chan", [(1, 32), (4, 32), (2, 64), (3, 96)])
@pytest.mark.parametrize("add_self_loops", [True, False])
@pytest.mark.parametrize("name", sorted(GRAPHS) + ["knn_batch"])
def test_blocked_gat_forward(dev, name, heads, chan, add_self_loops):
    """lgk_gat_fwd_blocked (edge soft-max + aggregation on the windowed, row-blocked structure) against the fp64 oracle's
    GATConv and against lgk_gat_fwd: output, fused ELU output and the soft-max statistics the backward reads -- on every
    producer path (windowed, gaps, backward jumps, global-gather chunks), multigraph duplicates, stored self loops, rows
    without in-edges"""
    from lesion_gnn_b200 import ops
    ei, n = _knn_batch_graph(13) if name == "knn_batch" else GRAPHS[name]()
    fin = 12
    gen = torch.Generator().manual_seed(heads * chan + n)
    x = torch.randn(n, fin, generator=gen, dtype=torch.float64)
    torch.manual_seed(3)
    ref = O.GATConv(fin, chan, heads=heads, add_self_loops=add_self_loops).double()
    ref.bias.data.normal_()
    ref.att_src.data.mul_(3.0)  # logits of a few units: the soft-max is far from uniform
    ref.att_dst.data.mul_(3.0)
    want = ref(x, ei)
    g = _graph(ei, n, dev)
    w = ref.lin.weight.detach().float().to(dev)
    h = ops.linear_fwd(x.float().to(dev), w)
    a_s, a_d = ops.gat_alpha(h, ref.att_src.detach().float().reshape(-1).to(dev), ref.att_dst.detach().float().reshape(-1).to(dev), heads, chan)
    bias = ref.bias.detach().float().to(dev)
    base = ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, heads, chan, 0.2, add_self_loops, with_elu=True)
    q = g.quads("csr")
    ops.reset_launch_count()
    got = ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, heads, chan, 0.2, add_self_loops, with_elu=True, quads=q)
    assert ops.launch_count() == 2  # the transpose of a_src and the windowed kernel: the blocked path really ran
    assert_close(got[0], want, TOL_F32, "out vs oracle")
    assert_close(got[0], base[0], TOL_F32, "out vs lgk_gat_fwd")
    assert_close(got[3], torch.nn.functional.elu(want), TOL_F32, "elu(out) vs oracle")
    # statistics: rows with at least one entry (others hold -inf / 0 in both kernels)
    live = torch.isfinite(base[1])
    assert torch.equal(torch.isfinite(got[1]), live)
    assert torch.equal(got[1][live], base[1][live]), "the row maximum is exact in any order"
    assert_close(got[2], base[2], TOL_F32, "rowsum")
    again = ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, heads, chan, 0.2, add_self_loops, with_elu=True, quads=g.quads("csr"))
    assert all(torch.equal(a, b) for a, b in zip(got, again)), "deterministic run to run"


def test_blocked_gat_through_the_layer_backward(dev):
    """GATConv with the blocked forward (forced by the size threshold) and the ordinary backward kernels, which read the
    blocked kernel's statistics: output and all gradients against the fp64 oracle"""
    import lesion_gnn_b200 as lg
    ei, n = _knn_batch_graph(14)
    x = torch.randn(n, 20, generator=torch.Generator().manual_seed(1), dtype=torch.float64)
    torch.manual_seed(4)
    ref = O.GATConv(20, 32, heads=2).double()
    ref.bias.data.normal_()
    mod = lg.nn.GATConv(20, 32, heads=2)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    xr = x.clone().requires_grad_(True)
    want = ref(xr, ei)
    dy = torch.randn(want.shape, generator=torch.Generator().manual_seed(2), dtype=torch.float64)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    from lesion_gnn_b200 import ops
    ops.reset_launch_count()
    got = mod(xg, ei.to(dev))
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, "out")
    assert_close(xg.grad, xr.grad, TOL_F32, "dx")
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        assert_close(p.grad.view_as(pr.grad), pr.grad, TOL_F32, "d" + k)
