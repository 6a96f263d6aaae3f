"""GPU parity, floating point (<= 1e-5 relative in fp32, 1e-2 in bf16): fused gather + segment-reduce
aggregation, its CSC backward, gcn_norm, global pooling -- through the C ABI against the fp64 oracle."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_BF16, TOL_F32, assert_close, random_graph

pytestmark = pytest.mark.gpu

FEATS = [1, 3, 4, 8, 16, 20, 64, 128, 132, 260, 1025]


def _graph(ei, n, dev, ew=None):
    from lesion_gnn_b200 import CSRGraph
    return CSRGraph.from_edge_index(ei.to(dev), n, None if ew is None else ew.to(dev))


@pytest.mark.parametrize("reduce", ["sum", "mean", "max"])
@pytest.mark.parametrize("feat", FEATS)
def test_aggregate_fwd_bwd(dev, reduce, feat):
    from lesion_gnn_b200.nn import functional as F
    n = 300
    ei = random_graph(n, 2000, feat)
    ei = ei[:, ei[1] != 7]  # node 7 has no in-edges: empty row
    gen = torch.Generator().manual_seed(feat)
    x = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    dy = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    want = O.propagate(xr, ei, n, reduce)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = F.aggregate(xg, _graph(ei, n, dev), reduce)
    got.backward(dy.float().to(dev))
    assert float(got[7].detach().abs().max()) == 0.0
    assert_close(got, want, TOL_F32, f"{reduce} fwd F={feat}")
    assert_close(xg.grad, xr.grad, TOL_F32, f"{reduce} bwd F={feat}")


@pytest.mark.parametrize("feat", [4, 16, 64, 100])
def test_weighted_sum_fwd_bwd(dev, feat):
    from lesion_gnn_b200.nn import functional as F
    n = 500
    ei = random_graph(n, 4000, 1)
    gen = torch.Generator().manual_seed(2)
    ew = torch.rand(ei.size(1), generator=gen, dtype=torch.float64) + 0.1
    x = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    dy = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    want = O.propagate(xr, ei, n, "sum", ew)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = F.aggregate(xg, _graph(ei, n, dev, ew.float()), "sum", use_value=True)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, "weighted fwd")
    assert_close(xg.grad, xr.grad, TOL_F32, "weighted bwd")


@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("improved", [False, True])
def test_gcn_norm_matches_oracle(dev, weighted, improved):
    """dinv and the self-loop weights of add_remaining_self_loops, incl. stored and duplicate loops"""
    n = 200
    ei = random_graph(n, 1500, 5)
    loops = torch.tensor([[3, 3, 9], [3, 3, 9]])  # node 3 has two stored loops: the last one wins
    ei = torch.cat([ei, loops], 1)
    gen = torch.Generator().manual_seed(6)
    ew = (torch.rand(ei.size(1), generator=gen, dtype=torch.float64) + 0.1) if weighted else None
    g = _graph(ei, n, dev, None if ew is None else ew.float())
    # (PyG quirk restated by the oracle: fill_value=2 of improved=True only reaches graphs WITH edge weights)
    fill = 2.0 if (improved and weighted) else 1.0
    dinv, selfw = g.gcn_norm(fill == 2.0, True)
    ei2, w2 = O.gcn_norm(ei, ew, n, improved, True, torch.float64)
    # recover deg^-1/2 from the oracle: weight of the loop edge (i,i) = dinv_i^2 * selfw_i
    loop_w = torch.full((n,), fill, dtype=torch.float64)
    if ew is not None:
        for e in range(ei.size(1)):
            if ei[0, e] == ei[1, e]:
                loop_w[ei[0, e]] = ew[e]
    want_dinv = (w2[-n:] / loop_w).sqrt()
    assert_close(dinv, want_dinv, TOL_F32, "dinv")
    if weighted:
        assert_close(selfw, loop_w, 1e-7, "self-loop weights")


@pytest.mark.parametrize("mode", ["mean", "max", "sum", "mean+max"])
@pytest.mark.parametrize("feat", [1, 5, 16, 64, 128, 1024, 1028])
def test_global_pool_fwd_bwd(dev, mode, feat):
    import lesion_gnn_b200 as lg
    counts = [40, 0, 1, 333, 0, 7, 90]
    B = len(counts)
    n = sum(counts)
    batch = torch.repeat_interleave(torch.arange(B), torch.tensor(counts))
    gen = torch.Generator().manual_seed(feat)
    x = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    if mode == "mean+max":
        want = torch.cat([O.global_mean_pool(xr, batch, B), O.global_max_pool(xr, batch, B)], 1)
        fn = lg.nn.global_mean_max_pool
    else:
        want = O.scatter(xr, batch, B, mode)
        fn = {"mean": lg.nn.global_mean_pool, "max": lg.nn.global_max_pool, "sum": lg.nn.global_add_pool}[mode]
    dy = torch.randn(want.shape, generator=gen, dtype=torch.float64)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = fn(xg, batch.to(dev), B)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, f"pool {mode} fwd")
    assert_close(xg.grad, xr.grad, TOL_F32, f"pool {mode} bwd")
    assert float(got[1].detach().abs().max()) == 0.0 and float(got[4].detach().abs().max()) == 0.0  # empty graphs -> 0


def test_global_pool_size_none_and_single_graph(dev):
    import lesion_gnn_b200 as lg
    x = torch.randn(50, 8)
    batch = torch.sort(torch.randint(0, 4, (50,))).values
    got = lg.nn.global_mean_pool(x.to(dev), batch.to(dev))
    assert_close(got, O.global_mean_pool(x.double(), batch), TOL_F32)
    got = lg.nn.global_max_pool(x.to(dev), None)
    assert_close(got, x.double().max(0, keepdim=True).values, TOL_F32)


def test_pool_long_graph_uses_split_path(dev):
    """one 200k-node graph among small ones: the split + combine stage"""
    import lesion_gnn_b200 as lg
    counts = [200000, 10, 3]
    batch = torch.repeat_interleave(torch.arange(3), torch.tensor(counts))
    gen = torch.Generator().manual_seed(0)
    x = torch.randn(sum(counts), 32, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    want = torch.cat([O.global_mean_pool(xr, batch, 3), O.global_max_pool(xr, batch, 3)], 1)
    dy = torch.randn(want.shape, generator=gen, dtype=torch.float64)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = lg.nn.global_mean_max_pool(xg, batch.to(dev), 3)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32)
    assert_close(xg.grad, xr.grad, TOL_F32)


@pytest.mark.parametrize("reduce", ["sum", "mean", "max"])
def test_aggregate_bf16(dev, reduce):
    from lesion_gnn_b200 import ops
    n = 400
    ei = random_graph(n, 3000, 8)
    x = torch.randn(n, 64).to(torch.bfloat16)
    g = _graph(ei, n, dev)
    got = ops.spmm(g.rowptr, g.col, x.to(dev), reduce)
    assert got.dtype == torch.bfloat16
    assert_close(got.float(), O.propagate(x.double(), ei, n, reduce), TOL_BF16, f"bf16 {reduce}")


def test_aggregate_is_bitwise_deterministic(dev):
    from lesion_gnn_b200.nn import functional as F
    n = 5000
    ei = random_graph(n, 80000, 3)
    g = _graph(ei, n, dev)
    x = torch.randn(n, 64, device=dev, requires_grad=True)
    outs, grads = [], []
    for _ in range(3):
        x.grad = None
        y = F.aggregate(x, g, "sum")
        y.backward(torch.ones_like(y))
        outs.append(y.detach().clone())
        grads.append(x.grad.clone())
    assert all(torch.equal(outs[0], o) for o in outs) and all(torch.equal(grads[0], gr) for gr in grads)


@pytest.mark.parametrize("feat", [8, 128, 5])
def test_max_aggregation_ties_go_to_the_first_slot(dev, feat):
    """torch-scatter's rule (what PyG's scatter(reduce='max') uses for gradients on CUDA): among equal maxima the FIRST
    entry of the destination's list wins.  Integer-valued features give many ties; rows of every length and alignment."""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import ops
    n = 500
    ei = random_graph(n, 9000, 17)
    x = torch.randint(0, 3, (n, feat), generator=torch.Generator().manual_seed(1)).float()
    g = lg.CSRGraph.from_edge_index(ei.to(dev), n)
    out, arg = ops.spmm(g.rowptr, g.col, x.to(dev), "max", want_arg=True)
    rowptr, col = g.rowptr.cpu().long(), g.col.cpu().long()
    want_out = torch.zeros(n, feat)
    want_arg = torch.full((n, feat), -1, dtype=torch.long)
    for i in range(n):
        lo, hi = int(rowptr[i]), int(rowptr[i + 1])
        if hi > lo:
            vals = x[col[lo:hi]]                       # [deg, F]
            m, _ = vals.max(0)
            first = (vals == m).float().argmax(0)      # first position attaining the maximum
            want_out[i] = m
            want_arg[i] = lo + first
    assert torch.equal(out.cpu(), want_out)
    assert torch.equal(arg.cpu().long(), want_arg)
