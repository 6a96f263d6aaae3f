"""CPU tests of the oracle itself: the reference's only known answers (GaussianDistance), the gather/scatter
restatement against the independent dense formulation, gradcheck, kNN against scipy's cKDTree, CSR against
scipy.sparse, and reproduction of the committed golden vectors."""
import math
import os

import numpy as np
import pytest
import torch

from oracle import dense as D
from oracle import pyg_semantics as O
from tests.util import random_graph

GOLD = os.path.join(os.path.dirname(__file__), "golden", "layers.pt")


def test_gaussian_distance_reference_kat():
    """/root/reference/test/test_transforms.py:8-38: two nodes at unit distance."""
    pos = torch.tensor([[0.0, 0.0], [1.0, 0.0]])
    ei = torch.tensor([[0, 1], [1, 0]])
    for sigma, want in ((1.0, 0.2420), (0.5, 0.1080), (2.0, 0.1760)):
        w = O.gaussian_distance(ei, pos, sigma)
        assert torch.allclose(w, torch.full((2,), want), rtol=1e-3, atol=1e-3)
        assert math.isclose(float(w[0]), math.exp(-1 / (2 * sigma ** 2)) / math.sqrt(2 * math.pi * sigma ** 2), rel_tol=1e-6)


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_layers_match_dense_formulation(seed):
    n, e = 23, 90
    ei = random_graph(n, e, seed)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 12, generator=g, dtype=torch.float64)
    ew = torch.rand(ei.size(1), generator=g, dtype=torch.float64) + 0.05
    torch.manual_seed(seed)
    gcn = O.GCNConv(12, 8).double()
    gcn.bias.data.normal_()
    for w in (None, ew):
        got = gcn(x, ei, w)
        want = D.gcn_dense(x, ei, gcn.lin.weight, gcn.bias, w)
        assert torch.allclose(got, want, atol=1e-12)
    gat = O.GATConv(12, 4, heads=3).double()
    gat.bias.data.normal_()
    want = D.gat_dense(x, ei, gat.lin.weight, gat.att_src, gat.att_dst, gat.bias, 3)
    assert torch.allclose(gat(x, ei), want, atol=1e-12)
    for aggr in ("sum", "mean", "max"):
        w = ew if aggr == "sum" else None
        assert torch.allclose(O.propagate(x, ei, n, aggr, w), D.aggregate_dense(x, ei, aggr, w), atol=1e-12)
    batch = torch.sort(torch.randint(0, 5, (n,), generator=g)).values
    for red in ("sum", "mean", "max"):
        assert torch.allclose(O.scatter(x, batch, 6, red), D.pool_dense(x, batch, 6, red), atol=1e-12)


def test_gradcheck_gat_and_gcn():
    ei = random_graph(7, 20, 3)
    x = torch.randn(7, 4, dtype=torch.float64, requires_grad=True)
    torch.manual_seed(0)
    gat = O.GATConv(4, 2, heads=2).double()
    gcn = O.GCNConv(4, 3).double()
    assert torch.autograd.gradcheck(lambda t: gat(t, ei), (x,), atol=1e-6)
    assert torch.autograd.gradcheck(lambda t: gcn(t, ei), (x,), atol=1e-6)


def test_empty_segments_and_isolated_nodes():
    x = torch.randn(4, 3, dtype=torch.float64)
    batch = torch.tensor([0, 0, 2, 2])
    assert torch.equal(O.global_max_pool(x, batch, 4)[1], torch.zeros(3, dtype=torch.float64))
    assert torch.equal(O.global_mean_pool(x, batch, 4)[3], torch.zeros(3, dtype=torch.float64))
    ei = torch.tensor([[0], [1]])
    assert torch.equal(O.propagate(x, ei, 4, "max")[0], torch.zeros(3, dtype=torch.float64))
    assert torch.equal(O.propagate(x, ei, 4, "mean")[1], x[0])


def test_knn_against_ckdtree():
    from scipy.spatial import cKDTree
    g = torch.Generator().manual_seed(5)
    counts = [50, 3, 1, 120]
    pos = torch.rand(sum(counts), 2, generator=g, dtype=torch.float64) * 1536
    batch = torch.repeat_interleave(torch.arange(len(counts)), torch.tensor(counts))
    for loop in (True, False):
        k = 6
        ei = O.knn_graph(pos, k, batch, loop=loop)
        lo = 0
        want_src, want_dst = [], []
        for n in counts:
            p = pos[lo:lo + n].numpy()
            kk = min(k if loop else k + 1, n)
            _, idx = cKDTree(p).query(p, k=kk)
            idx = np.asarray(idx).reshape(n, kk)
            for i in range(n):
                nb = [j for j in idx[i] if loop or j != i][: k]
                want_src += [j + lo for j in nb]
                want_dst += [i + lo] * len(nb)
            lo += n
        assert ei[0].tolist() == want_src and ei[1].tolist() == want_dst


def test_csr_against_scipy():
    import scipy.sparse as sp
    n = 40
    ei = random_graph(n, 300, 11, duplicates=False)
    rowptr, col, perm = O.to_csr(ei, n)
    m = sp.csr_matrix((np.arange(1, ei.size(1) + 1), (ei[1].numpy(), ei[0].numpy())), shape=(n, n))
    m.sort_indices()
    assert rowptr.tolist() == m.indptr.tolist() and col.tolist() == m.indices.tolist()
    assert (perm + 1).tolist() == m.data.tolist()
    colptr, row, csr2csc = O.to_csc(rowptr, col, n)
    mt = m.tocsc()
    mt.sort_indices()
    assert colptr.tolist() == mt.indptr.tolist() and row.tolist() == mt.indices.tolist()
    assert torch.equal(col[csr2csc], torch.repeat_interleave(torch.arange(n), colptr[1:] - colptr[:-1]))


def test_golden_vectors_reproduce():
    gold = torch.load(GOLD)
    from tests.golden.make_golden import run_layer
    for name, case in gold["cases"].items():
        ei, n, x, ew = case["edge_index"], case["num_nodes"], case["x"], case["edge_weight"]
        rowptr, col, perm = O.to_csr(ei, n)
        assert torch.equal(rowptr, case["csr"]["rowptr"]) and torch.equal(col, case["csr"]["col"])
        again = run_layer(lambda: O.GATConv(16, 4, heads=2), x, ei)
        assert torch.equal(again["out"], case["gat"]["out"]) and torch.equal(again["dx"], case["gat"]["dx"])
        again = run_layer(lambda: O.GCNConv(16, 8), x, ei, ew)
        assert torch.equal(again["out"], case["gcn_w"]["out"])


def test_real_pyg_if_present():
    """PARITY PIN, opportunistic: when a box has torch_geometric, the oracle must equal it."""
    pyg = pytest.importorskip("torch_geometric")
    from torch_geometric.nn import GATConv, GCNConv, SAGEConv, global_max_pool, global_mean_pool
    ei = random_graph(30, 150, 4)
    x = torch.randn(30, 16)
    for ours, theirs in ((O.GCNConv(16, 8), GCNConv(16, 8)), (O.GATConv(16, 4, heads=2), GATConv(16, 4, heads=2)),
                         (O.SAGEConv(16, 8), SAGEConv(16, 8))):
        ours.load_state_dict(theirs.state_dict())
        assert torch.allclose(ours(x, ei), theirs(x, ei), atol=1e-6)
    batch = torch.sort(torch.randint(0, 4, (30,))).values
    assert torch.allclose(O.global_mean_pool(x, batch), global_mean_pool(x, batch), atol=1e-6)
    assert torch.allclose(O.global_max_pool(x, batch), global_max_pool(x, batch), atol=1e-6)
    del pyg
