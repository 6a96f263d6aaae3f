"""GPU parity, integer work (bit-exact bar): CSR/CSC builder, graph segment pointer, kNN graph, through the
C ABI against the oracle."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import random_graph

pytestmark = pytest.mark.gpu


def _check_csr(ei, n, dev):
    from lesion_gnn_b200 import ops
    rowptr, col, perm, colptr, row, csr2csc = ops.csr_build(ei.to(dev), n, with_csc=True)
    r_rowptr, r_col, r_perm = O.to_csr(ei, n)
    r_colptr, r_row, r_c2c = O.to_csc(r_rowptr, r_col, n)
    for got, want, name in ((rowptr, r_rowptr, "rowptr"), (col, r_col, "col"), (perm, r_perm, "perm"),
                            (colptr, r_colptr, "colptr"), (row, r_row, "row"), (csr2csc, r_c2c, "csr2csc")):
        assert got.dtype == torch.int32
        assert torch.equal(got.cpu().long(), want), name


@pytest.mark.parametrize("n,e,seed", [(1, 1, 0), (2, 2, 1), (50, 400, 2), (1000, 8000, 3), (3200, 25600, 4),
                                      (5000, 100, 5), (20000, 200000, 6)])
def test_csr_random_multigraphs(dev, n, e, seed):
    _check_csr(random_graph(n, e, seed), n, dev)


def test_csr_empty_and_edgeless(dev):
    _check_csr(torch.zeros(2, 0, dtype=torch.long), 7, dev)
    _check_csr(torch.zeros(2, 0, dtype=torch.long), 0, dev)


def test_csr_big_rows(dev):
    """rows beyond the rank-by-counting limit (64), the shared-memory bitonic limit (4096) and with duplicates"""
    g = torch.Generator().manual_seed(0)
    n = 9000
    parts = [torch.stack([torch.randint(0, n, (100,), generator=g), torch.full((100,), 3)]),       # 100 -> node 3
             torch.stack([torch.randint(0, n, (3000,), generator=g), torch.full((3000,), 17)]),    # 3000 -> node 17
             torch.stack([torch.randint(0, n, (6000,), generator=g), torch.full((6000,), 8999)]),  # 6000 -> last node
             torch.stack([torch.full((5000,), 42), torch.randint(0, n, (5000,), generator=g)]),    # hub as source
             random_graph(n, 20000, 1)]
    ei = torch.cat(parts, dim=1)
    ei = ei[:, torch.randperm(ei.size(1), generator=g)].contiguous()
    _check_csr(ei, n, dev)


def test_csr_knn_config1(dev):
    from tests.util import knn_batch
    hb = knn_batch(32, 100, 16, 8, seed=1234)
    assert hb.edge_index.size(1) == 25600
    _check_csr(hb.edge_index, hb.num_nodes, dev)


def test_csr_without_csc(dev):
    from lesion_gnn_b200 import ops
    ei = random_graph(100, 700, 9)
    rowptr, col, perm, colptr, row, c2c = ops.csr_build(ei.to(dev), 100, with_csc=False)
    assert colptr is None and torch.equal(col.cpu().long(), O.to_csr(ei, 100)[1])


def test_csr_is_deterministic(dev):
    from lesion_gnn_b200 import ops
    ei = random_graph(4000, 60000, 12).to(dev)
    a = ops.csr_build(ei, 4000)
    for _ in range(3):
        b = ops.csr_build(ei, 4000)
        assert all(torch.equal(x, y) for x, y in zip(a, b))


def test_csr_built_on_a_side_stream_joins_on_first_use(dev):
    """CSRGraph.from_edge_index(stream=...) forks the build; as_graph joins it before a layer reads it"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.graph import as_graph
    ei = random_graph(3000, 40000, 21).to(dev)
    ref = lg.CSRGraph.from_edge_index(ei, 3000)
    side = torch.cuda.Stream()
    for _ in range(3):
        g = lg.CSRGraph.from_edge_index(ei, 3000, stream=side)
        assert g._ready is not None
        g2 = as_graph(g, 3000)
        assert g2 is g and g._ready is None
        x = torch.ones(3000, 8, device=dev)
        out = lg.ops.spmm(g.rowptr, g.col, x, "sum")
        assert torch.equal(out[:, 0].cpu().long(), (ref.rowptr[1:] - ref.rowptr[:-1]).cpu().long())
        assert torch.equal(g.col, ref.col) and torch.equal(g.csr2csc, ref.csr2csc)


@pytest.mark.parametrize("counts", [[3, 0, 5, 1], [0, 0, 4], [7], [1] * 50, [2, 3, 0, 0]])
def test_batch_ptr_with_empty_graphs(dev, counts):
    from lesion_gnn_b200 import ops
    batch = torch.repeat_interleave(torch.arange(len(counts)), torch.tensor(counts))
    ptr = ops.batch_ptr(batch.to(dev), len(counts))
    assert ptr.cpu().long().tolist() == O.batch_to_ptr(batch, len(counts)).tolist()


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("loop", [True, False])
@pytest.mark.parametrize("k", [1, 6, 8, 16, 33])
def test_knn_graph_bit_exact(dev, dtype, loop, k):
    from lesion_gnn_b200 import transforms
    g = torch.Generator().manual_seed(k)
    counts = [100, 1, 2, 7, 300, 45, 600]
    pos = (torch.rand(sum(counts), 2, generator=g, dtype=torch.float64) * 1536).to(dtype)
    batch = torch.repeat_interleave(torch.arange(len(counts)), torch.tensor(counts))
    want = O.knn_graph(pos, k, batch, loop=loop)
    got = transforms.knn_graph(pos.to(dev), k, batch.to(dev), loop=loop)
    assert got.dtype == torch.int64 and torch.equal(got.cpu(), want)
    got2 = transforms.knn_graph(pos.to(dev), k, batch.to(dev), loop=loop,
                                num_edges=transforms.knn_num_edges(counts, k, loop))
    assert torch.equal(got2.cpu(), want)


def test_knn_single_graph_and_ties(dev):
    from lesion_gnn_b200 import transforms
    # a regular grid has many exact distance ties: lower index must win, as in the oracle's stable sort
    xs, ys = torch.meshgrid(torch.arange(6.0), torch.arange(6.0), indexing="ij")
    pos = torch.stack([xs.flatten(), ys.flatten()], 1).double()
    want = O.knn_graph(pos, 5, None, loop=False)
    got = transforms.knn_graph(pos.to(dev), 5, None, loop=False)
    assert torch.equal(got.cpu(), want)


def test_gaussian_distance(dev):
    from lesion_gnn_b200 import transforms
    g = torch.Generator().manual_seed(3)
    pos = torch.rand(200, 2, generator=g, dtype=torch.float64) * 20
    ei = random_graph(200, 1500, 3)
    for sigma in (0.5, 1.0, 7.0):
        got = transforms.GaussianDistance(sigma)(pos.to(dev), ei.to(dev))
        want = O.gaussian_distance(ei, pos, sigma)
        assert torch.allclose(got.cpu(), want, rtol=1e-6, atol=1e-30)
    # the reference's own known answers (/root/reference/test/test_transforms.py:20,29,38)
    p2 = torch.tensor([[0.0, 0.0], [1.0, 0.0]], device=dev)
    e2 = torch.tensor([[0, 1], [1, 0]], device=dev)
    for sigma, want in ((1.0, 0.2420), (0.5, 0.1080), (2.0, 0.1760)):
        assert torch.allclose(transforms.GaussianDistance(sigma)(p2, e2).cpu(), torch.full((2,), want), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("r,loop,cap", [(150.0, False, 32), (150.0, True, 32), (400.0, False, 8), (5.0, False, 32), (4000.0, True, 32)])
def test_radius_graph_bit_exact(dev, dtype, r, loop, cap):
    """RadiusGraph (reference scripts/sweep.py:113-117) on the device == the oracle's torch_cluster restatement"""
    from lesion_gnn_b200 import data, transforms
    hb = data.synth_batch(9, "ddr", 4, seed=31)
    pos = hb.pos.to(dtype)
    ref = O.radius_graph(pos, r, hb.batch, loop=loop, max_num_neighbors=cap)
    got = transforms.radius_graph(pos.to(dev), r, hb.batch.to(dev), loop=loop, max_num_neighbors=cap)
    assert got.dtype == torch.int64 and torch.equal(got.cpu(), ref)


def test_radius_graph_empty_and_single(dev):
    from lesion_gnn_b200 import transforms
    pos = torch.tensor([[0.0, 0.0], [3.0, 4.0], [100.0, 100.0]], dtype=torch.float64)
    got = transforms.radius_graph(pos.to(dev), 5.0)  # |p0 - p1| = 5 exactly: strict <, no edge
    assert got.shape == (2, 0)
    got = transforms.radius_graph(pos.to(dev), 5.0001)
    assert torch.equal(got.cpu(), torch.tensor([[1, 0], [0, 1]]))


def _sorted_case(n, e, seed, hub=None):
    """random multigraph (self loops, duplicates, empty rows) with its edges grouped by destination, sources in random
    order inside a row; hub: (node, degree) gets that many extra in-edges"""
    ei = random_graph(n, e, seed)
    if hub is not None:
        g = torch.Generator().manual_seed(seed + 1)
        extra = torch.stack([torch.randint(0, n, (hub[1],), generator=g), torch.full((hub[1],), hub[0], dtype=torch.long)])
        ei = torch.cat([ei, extra], 1)
    ei = ei[:, ei[1] != min(3, n - 1)] if n > 1 else ei   # an empty row in the middle
    order = torch.sort(ei[1], stable=True).indices
    return ei[:, order].contiguous()


@pytest.mark.parametrize("n,e,hub", [(1, 5, None), (50, 0, None), (300, 2500, None), (5000, 80000, None), (2000, 9000, (17, 6000)),
                                     (129, 6000, None), (100000, 1600000, None)])
@pytest.mark.parametrize("with_csc", [True, False])
def test_sorted_input_build_equals_general_build(dev, n, e, hub, with_csc):
    """lgk_csr_build_sorted (boundary detection + 32-bit stable ranks) against lgk_csr_build on the same destination-sorted
    edges: all six index arrays bit for bit, incl. a 6000-wide hub row (cooperative ranking), windows too wide to stage
    (129 nodes x 6000 edges), empty rows and an edgeless graph"""
    from lesion_gnn_b200 import ops
    ei = _sorted_case(n, e, seed=n + e).to(dev)
    want = ops.csr_build(ei, n, with_csc)
    got = ops.csr_build(ei, n, with_csc, dst_sorted=True)
    assert int(ops.last_sorted_violation) == 0
    for a, b, name in zip(got, want, ("rowptr", "col", "perm", "colptr", "row", "csr2csc")):
        assert (a is None) == (b is None), name
        if a is not None:
            assert torch.equal(a, b), name


def test_sorted_input_build_reports_a_violated_contract(dev):
    from lesion_gnn_b200 import ops
    ei = random_graph(400, 3000, 5).to(dev)  # not sorted by destination
    ops.csr_build(ei, 400, True, dst_sorted=True)
    torch.cuda.synchronize()
    assert int(ops.last_sorted_violation) == 1
    bad = _sorted_case(400, 3000, 6).to(dev)
    bad[0, 7] = 400  # source out of range
    ops.csr_build(bad, 400, True, dst_sorted=True)
    torch.cuda.synchronize()
    assert int(ops.last_sorted_violation) == 1


def test_knn_graph_feeds_the_sorted_build(dev):
    """knn_graph / radius_graph emit edges grouped by destination: the transform pipeline (KNNGraph -> ToSparseTensor) takes
    the sorted build and gets the general build's graph"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops, transforms
    hb = data.synth_batch(40, "ddr", 8, seed=3)
    d = transforms.Data(x=hb.x.to(dev), pos=hb.pos.float().to(dev), batch=hb.batch.to(dev))
    lg.graph._SORTED_MIN_EDGES = 0  # small graphs normally keep the general build (faster below ~1 M edges)
    d = transforms.KNNGraph(k=6, loop=True)(d)
    assert d.edges_grouped_by_dst and bool((d.edge_index[1][1:] >= d.edge_index[1][:-1]).all())
    ei = d.edge_index.clone()
    d = transforms.ToSparseTensor(attr=None)(d)
    assert int(ops.last_sorted_violation) == 0
    lg.graph._SORTED_MIN_EDGES = 1 << 20
    want = lg.CSRGraph.from_edge_index(ei, hb.num_nodes)
    for name in ("rowptr", "col", "perm", "colptr", "row", "csr2csc"):
        assert torch.equal(getattr(d.adj_t, name), getattr(want, name)), name
