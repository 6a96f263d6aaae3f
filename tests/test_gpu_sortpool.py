"""GPU parity of SortAggregation (reference models/drgnet.py:37,59) and of the whole DRGNet (drgnet.py:16-69) against
the oracle: selection bit-exact (it is index work), values and gradients within the fp32 tolerance."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_F32, assert_close, knn_batch, assert_parity

pytestmark = pytest.mark.gpu


def _case(counts, feat, seed, ties=False):
    g = torch.Generator().manual_seed(seed)
    n = sum(counts)
    x = torch.randn(n, feat, generator=g)
    if ties:  # repeated keys: the stable order (lower node index first) decides
        x[:, -1] = torch.randint(0, 3, (n,), generator=g).float()
    batch = torch.repeat_interleave(torch.arange(len(counts)), torch.tensor(counts))
    return x, batch


@pytest.mark.parametrize("counts,feat,k,ties", [
    ([5, 9, 1, 30], 7, 4, False),
    ([5, 9, 1, 30], 8, 12, False),       # k larger than most graphs: zero padding
    ([3, 0, 6, 0], 5, 4, False),         # empty graphs
    ([40, 17, 64], 193, 30, False),      # DRGNet width 3 * 64 + 1
    ([50, 20, 33], 6, 10, True),
    ([600, 2, 300], 33, 50, True),
    ([9000], 3, 16, False),              # slab larger than the shared-memory key cache
])
def test_sort_pool_matches_oracle(dev, counts, feat, k, ties):
    import lesion_gnn_b200 as lg
    x, batch = _case(counts, feat, seed=len(counts) * 100 + feat, ties=ties)
    B = len(counts)
    xr = x.double().requires_grad_(True)
    ref = O.sort_aggregation(xr, batch, k, B)
    w = torch.randn(ref.shape, generator=torch.Generator().manual_seed(3), dtype=torch.float64)
    (ref * w).sum().backward()
    xg = x.to(dev).requires_grad_(True)
    seg = lg.nn.Segments.from_batch(batch.to(dev), B)
    out = lg.nn.SortAggregation(k)(xg, seg)
    (out * w.float().to(dev)).sum().backward()
    assert out.shape == (B, k * feat)
    assert torch.equal(out.cpu(), ref.detach().float()), "selected rows differ from the oracle (pure copy: must be bit-exact)"
    assert torch.equal(xg.grad.cpu(), xr.grad.float()), "gradient routing differs from the oracle"


def test_sort_pool_accepts_batch_vector_and_is_deterministic(dev):
    import lesion_gnn_b200 as lg
    x, batch = _case([12, 7, 25], 16, seed=2)
    pool = lg.nn.SortAggregation(8)
    a = pool(x.to(dev), batch.to(dev))
    b = pool(x.to(dev), lg.nn.Segments.from_batch(batch.to(dev), 3))
    assert torch.equal(a, b) and torch.equal(a.cpu(), O.sort_aggregation(x, batch, 8, 3))


def test_sort_pool_refuses_cpu(dev):
    import lesion_gnn_b200 as lg
    with pytest.raises(RuntimeError):
        lg.nn.SortAggregation(4)(torch.randn(5, 3), torch.zeros(5, dtype=torch.long))


def test_drgnet_forward_backward(dev):
    """reference drgnet.py:16-69 with GaussianDistance edge weights (drgnet.py:55), CE loss, eval mode (no dropout)"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(12, "ddr", 40, 6, seed=77)
    ew = O.gaussian_distance(hb.edge_index, hb.pos, sigma=100.0)
    torch.manual_seed(5)
    ref = O.DRGNet(40, 32, 3, 10, 5).double().eval()
    mod = lg.models.DRGNet(40, 32, 3, 10, 5)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev).eval()
    import copy
    ref32 = copy.deepcopy(ref).float()

    def run_ref(model, dt):
        x = hb.x.detach().clone().to(dt).requires_grad_(True)
        o = model(x, hb.edge_index, hb.batch, edge_weight=ew.to(dt), num_graphs=hb.num_graphs)
        torch.nn.functional.cross_entropy(o, hb.y).backward()
        return x, o

    xr, out_r = run_ref(ref, torch.float64)
    x32, out_32 = run_ref(ref32, torch.float32)
    xg = hb.x.detach().clone().to(dev).requires_grad_(True)
    out = mod(xg, hb.edge_index.to(dev), hb.batch.to(dev), ew.float().to(dev))
    torch.nn.functional.cross_entropy(out, hb.y.to(dev)).backward()
    # the stated 1e-5 bar, or 4x what the oracle itself loses in fp32 on these inputs (tests/util.assert_parity)
    assert_parity(out, out_r, out_32, "logits")
    assert_parity(xg.grad, xr.grad, x32.grad, "dx")
    refg, ref32g = dict(ref.named_parameters()), dict(ref32.named_parameters())
    for k, p in mod.named_parameters():
        if refg[k].grad is not None and float(refg[k].grad.abs().max()) > 0:
            assert_parity(p.grad.view_as(refg[k].grad), refg[k].grad, ref32g[k].grad, "d" + k)
