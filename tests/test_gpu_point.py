"""GPU parity of the PointNet++ operators and the unsorted segment reduce (reference
models/pointnet.py:13-59, datasets/nodes/lesions.py:88-93) against oracle/pointnet.py: index work bit-exact,
floating point at 1e-5."""
import pytest
import torch

from oracle import pointnet as OP
from oracle import pyg_semantics as O
from tests.util import TOL_F32, assert_close

pytestmark = pytest.mark.gpu


def _points(num_graphs, seed, dtype=torch.float32, dim=2):
    from lesion_gnn_b200 import data
    hb = data.synth_batch(num_graphs, "ddr", 4, seed=seed)
    pos = (hb.pos / 1536.0).to(dtype)
    if dim == 3:
        g = torch.Generator().manual_seed(seed)
        pos = torch.cat([pos, torch.rand(pos.size(0), 1, generator=g, dtype=torch.float64).to(dtype)], 1)
    return pos, hb.batch, hb


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("ratio", [0.5, 0.25, 1.0])
def test_fps_bit_exact(dev, dtype, ratio):
    import lesion_gnn_b200 as lg
    pos, batch, _ = _points(17, 3, dtype)
    want = OP.fps(pos, batch, ratio)
    got = lg.nn.fps(pos.to(dev), batch.to(dev), ratio=ratio, random_start=False)
    assert torch.equal(got.cpu(), want)
    # a random start still returns ceil(ratio * n) distinct nodes of every graph
    rnd = lg.nn.fps(pos.to(dev), batch.to(dev), ratio=ratio, random_start=True).cpu()
    assert rnd.numel() == want.numel() and rnd.unique().numel() == rnd.numel()
    assert torch.equal(batch[rnd], batch[want])


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("r,cap", [(0.1, 64), (0.3, 8), (2.0, 64)])
def test_radius_bipartite_bit_exact(dev, dtype, r, cap):
    import lesion_gnn_b200 as lg
    pos, batch, _ = _points(9, 5, dtype, dim=3)
    idx = OP.fps(pos, batch, 0.5)
    want = OP.radius(pos, pos[idx], r, batch, batch[idx], max_num_neighbors=cap)
    got = lg.nn.radius(pos.to(dev), pos[idx].to(dev), r, batch.to(dev), batch[idx].to(dev), max_num_neighbors=cap)
    assert torch.equal(got.cpu(), want)


@pytest.mark.parametrize("reduce", ["sum", "mean", "max"])
@pytest.mark.parametrize("n,feat,labels", [(5000, 37, 60), (40000, 1025, 33), (10, 4, 20)])
def test_scatter_on_unsorted_labels(dev, reduce, n, feat, labels):
    """reference datasets/nodes/lesions.py:93 `scatter(features [H*W, C], cc.flatten(), 0, reduce)`: labels in pixel order
    (unsorted), some labels empty"""
    import lesion_gnn_b200 as lg
    gen = torch.Generator().manual_seed(n + feat)
    index = torch.randint(0, labels, (n,), generator=gen)
    index[index == 3] = 4  # label 3 stays empty
    x = torch.randn(n, feat, generator=gen, dtype=torch.float64)
    dy = torch.randn(labels, feat, generator=gen, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    want = O.scatter(xr, index, labels, reduce)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = lg.nn.scatter(xg, index.to(dev), 0, labels, reduce)
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, f"{reduce} fwd")
    assert_close(xg.grad, xr.grad, TOL_F32, f"{reduce} bwd")
    assert float(got[3].abs().max()) == 0.0
    assert torch.equal(lg.nn.scatter(xg.detach(), index.to(dev), 0, None, reduce)[: int(index.max()) + 1],
                       got.detach()[: int(index.max()) + 1])


def test_pointnetconv_forward_backward(dev):
    import lesion_gnn_b200 as lg
    pos, batch, hb = _points(12, 8)
    gen = torch.Generator().manual_seed(2)
    x = torch.randn(pos.size(0), 6, generator=gen, dtype=torch.float64)
    idx = OP.fps(pos, batch, 0.5)
    row, col = OP.radius(pos, pos[idx], 0.25, batch, batch[idx], max_num_neighbors=64)
    ei = torch.stack([col, row])
    torch.manual_seed(4)
    ref = OP.PointNetConv(O.MLP([6 + 2, 16, 32], act="relu", norm=None), add_self_loops=False).double()
    mod = lg.nn.PointNetConv(lg.nn.MLP([6 + 2, 16, 32], norm=None), add_self_loops=False)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    xr = x.clone().requires_grad_(True)
    want = ref((xr, xr[idx]), (pos.double(), pos[idx].double()), ei)
    dy = torch.randn(want.shape, generator=gen, dtype=torch.float64)
    want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True)
    got = mod((xg, xg[idx.to(dev)]), (pos.to(dev), pos[idx].to(dev)), ei.to(dev))
    got.backward(dy.float().to(dev))
    assert_close(got, want, TOL_F32, "out")
    assert_close(xg.grad, xr.grad, TOL_F32, "dx")
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        assert_close(p.grad, pr.grad, TOL_F32, "d" + k)


def test_pointnet_model_forward(dev):
    """reference models/pointnet.py:43-59 end to end (eval mode, deterministic sampling): fps, radius, two PointNetConv set
    abstractions, the global one, the MLP head and log_softmax"""
    import lesion_gnn_b200 as lg
    pos, batch, hb = _points(6, 11)
    gen = torch.Generator().manual_seed(3)
    x = torch.randn(pos.size(0), 5, generator=gen, dtype=torch.float64)
    torch.manual_seed(6)
    ref = OP.PointNet(5, 2, 5, radii=(0.25, 0.5)).double().eval()
    mod = lg.models.PointNet(5, 2, 5, radii=(0.25, 0.5), random_start=False)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev).eval()
    want = ref(x, pos.double(), batch, num_graphs=6)
    got = mod(x.float().to(dev), pos.to(dev), batch.to(dev))
    assert got.shape == want.shape
    assert_close(got, want, 2e-5, "log-probabilities")
