"""bf16 mode (BASELINE.json north_star: "1e-2 in bf16"): activations stored as bf16, parameters and every accumulator
fp32.  Each layer, the read-outs and whole models run forward AND backward on bf16 tensors and are compared with the
fp64 oracle fed the SAME bf16-rounded inputs and upstream gradients, at the north_star's 1e-2 in its whole-tensor metric;
the per-row L2 metric (a sanity bound: rows of small magnitude must not be arbitrarily wrong) is held at 5e-2 -- one bf16
rounding of a stored row is up to 2^-8 of each element and a layer stores two or three rows on the way."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_BF16, knn_batch, random_graph, rel_err, row_rel_err

pytestmark = pytest.mark.gpu
ROW_BF16 = 5e-2


def _distinct(n, f, seed):
    """[n, f] fp64 of bf16-representable values, all distinct within a column: arg-max decisions cannot tie (bf16 has 8
    mantissa bits: rounded Gaussians tie all the time, and a tie is a routing convention, not arithmetic)"""
    gen = torch.Generator().manual_seed(seed)
    mags = torch.arange(0x3B00, 0x3B00 + n, dtype=torch.int16).view(torch.bfloat16).double()
    cols = [mags[torch.randperm(n, generator=gen)] * (torch.randint(0, 2, (n,), generator=gen) * 2 - 1) for _ in range(f)]
    return torch.stack(cols, 1)


def _r(t):
    """round to bf16, keep fp64 storage: what the oracle sees is exactly what the GPU reads"""
    return t.to(torch.bfloat16).double()


class _RoundSTE(torch.autograd.Function):
    """y = bf16(x) stored, dx = bf16(dy) stored: what any implementation that keeps this tensor in bf16 does to it"""

    @staticmethod
    def forward(ctx, x):
        return _r(x)

    @staticmethod
    def backward(ctx, g):
        return _r(g)


def _emulate_bf16_storage(ref):
    """a copy of the fp64 oracle module whose transforms, normalisations and layers round their outputs (and the
    gradients flowing back through them) to bf16 -- the storage points of a bf16 pipeline, all arithmetic still fp64.
    Its distance from the plain fp64 oracle is the error that bf16 STORAGE alone causes on these inputs: the floor
    under any bf16 implementation, however exact its arithmetic."""
    import copy
    emu = copy.deepcopy(ref)
    kinds = (O.PygLinear, O.PygBatchNorm, torch.nn.Linear, O.GATConv, O.GCNConv, O.SAGEConv, O.GINConv, O.GraphConv)
    for m in emu.modules():
        if isinstance(m, kinds):
            m.register_forward_hook(lambda mod, inp, out: _RoundSTE.apply(out))
    return emu


def _check(got, want, what, tol=TOL_BF16, row=ROW_BF16, floor=None, mult=4.0):
    """1e-2 (whole tensor) / 5e-2 (per row); where bf16 storage itself cannot hold that on these inputs (cancellation in
    the attention-vector gradients, depth), the bound is `mult` x the storage floor measured by _emulate_bf16_storage"""
    e, er = rel_err(got, want), row_rel_err(got, want)
    f = fr = 0.0
    if floor is not None:
        f, fr = rel_err(floor, want), row_rel_err(floor, want)
    assert e <= max(tol, mult * f), f"{what}: relative error {e:.3e} > max({tol:.0e}, {mult} x storage floor {f:.3e})"
    assert er <= max(row, mult * fr), f"{what}: per-row error {er:.3e} > max({row:.0e}, {mult} x storage floor {fr:.3e})"


def _run(layer_gpu, layer_ref, x, args_ref, dev, what, seed=17):
    layer_gpu.load_state_dict({k: v.float() for k, v in layer_ref.state_dict().items()})
    layer_gpu = layer_gpu.to(dev)
    x = _r(x)
    xr = x.clone().requires_grad_(True)
    want = layer_ref(xr, *args_ref)
    gen = torch.Generator().manual_seed(seed)
    dy = _r(torch.randn(want.shape, generator=gen, dtype=torch.float64))
    want.backward(dy)
    emu = _emulate_bf16_storage(layer_ref)
    emu.zero_grad()
    xe = x.clone().requires_grad_(True)
    floor = emu(xe, *args_ref)
    floor.backward(dy)
    emu_grads = {k: p.grad for k, p in emu.named_parameters()}
    xg = x.to(torch.bfloat16).to(dev).requires_grad_(True)
    args_gpu = [a.to(dev) if a.dtype == torch.int64 else a.float().to(dev) for a in args_ref]
    got = layer_gpu(xg, *args_gpu)
    assert got.dtype == torch.bfloat16, f"{what}: output is {got.dtype}"
    got.backward(dy.to(torch.bfloat16).to(dev))
    assert xg.grad.dtype == torch.bfloat16
    _check(got, want, what + " out", floor=floor)
    _check(xg.grad, xr.grad, what + " dx", floor=xe.grad)
    ref_grads = {k: p.grad for k, p in layer_ref.named_parameters()}
    scale = max(float(g.abs().max()) for g in ref_grads.values() if g is not None)
    for k, p in layer_gpu.named_parameters():
        if ref_grads[k] is None:
            continue
        assert p.grad is not None and p.grad.dtype == torch.float32, k
        if float(ref_grads[k].abs().max()) < 1e-9 * scale:  # analytically zero (a bias in front of BatchNorm)
            assert float(p.grad.abs().max()) < 2e-2 * scale, k
            continue
        _check(p.grad.view_as(ref_grads[k]), ref_grads[k], f"{what} d{k}", floor=emu_grads[k])


@pytest.mark.parametrize("n,k,m", [(7, 3, 5), (1000, 64, 64), (513, 1025, 64), (2000, 128, 128), (129, 20, 130), (40000, 64, 64)])
def test_linear_bf16(dev, n, k, m):
    from lesion_gnn_b200.nn import functional as F
    gen = torch.Generator().manual_seed(n + k + m)
    x = _r(torch.randn(n, k, generator=gen, dtype=torch.float64))
    w = torch.randn(m, k, generator=gen, dtype=torch.float64).float().double() / k ** 0.5
    w = w.float().double()
    b = torch.randn(m, generator=gen, dtype=torch.float64).float().double()
    dy = _r(torch.randn(n, m, generator=gen, dtype=torch.float64))
    xr, wr, br = (t.clone().requires_grad_(True) for t in (x, w, b))
    want = torch.nn.functional.elu(torch.nn.functional.linear(xr, wr, br))
    want.backward(dy)
    xg = x.to(torch.bfloat16).to(dev).requires_grad_(True)
    wg, bg = (t.float().to(dev).requires_grad_(True) for t in (w, b))
    got = F.linear(xg, wg, bg, act=1)
    assert got.dtype == torch.bfloat16
    got.backward(dy.to(torch.bfloat16).to(dev))
    _check(got, want, "linear out")
    _check(xg.grad, xr.grad, "linear dx")
    _check(wg.grad, wr.grad, "linear dW")
    _check(bg.grad, br.grad, "linear db")


@pytest.mark.parametrize("heads,chan", [(4, 16), (2, 64), (3, 20)])
def test_gatconv_bf16(dev, heads, chan):
    import lesion_gnn_b200 as lg
    n, fin = 257, 24
    ei = random_graph(n, 1800, heads * chan)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1), dtype=torch.float64)
    torch.manual_seed(2)
    ref = O.GATConv(fin, chan, heads=heads).double()
    ref.bias.data.normal_()
    _run(lg.nn.GATConv(fin, chan, heads=heads), ref, x, [ei], dev, f"GAT bf16 H={heads} C={chan}")


def test_gcnconv_bf16(dev):
    import lesion_gnn_b200 as lg
    n = 300
    ei = random_graph(n, 2400, 3, self_loops=False)
    x = torch.randn(n, 32, generator=torch.Generator().manual_seed(4), dtype=torch.float64)
    torch.manual_seed(5)
    ref = O.GCNConv(32, 48).double()
    ref.bias.data.normal_()
    _run(lg.nn.GCNConv(32, 48), ref, x, [ei], dev, "GCN bf16")


@pytest.mark.parametrize("aggr", ["mean", "max", "sum"])
def test_sageconv_bf16(dev, aggr):
    import lesion_gnn_b200 as lg
    n = 300
    ei = random_graph(n, 2400, 6)
    x = _distinct(n, 32, 7) if aggr == "max" else torch.randn(n, 32, generator=torch.Generator().manual_seed(7), dtype=torch.float64)
    torch.manual_seed(8)
    ref = O.SAGEConv(32, 40, aggr=aggr).double()
    _run(lg.nn.SAGEConv(32, 40, aggr=aggr), ref, x, [ei], dev, f"SAGE bf16 {aggr}")


def test_graphconv_bf16(dev):
    import lesion_gnn_b200 as lg
    n = 300
    ei = random_graph(n, 2000, 9)
    x = torch.randn(n, 24, generator=torch.Generator().manual_seed(10), dtype=torch.float64)
    torch.manual_seed(11)
    ref = O.GraphConv(24, 24).double()
    _run(lg.nn.GraphConv(24, 24), ref, x, [ei], dev, "GraphConv bf16")


def test_ginconv_with_batchnorm_bf16(dev):
    """GINConv(MLP([d, d, d])): Linear -> BatchNorm1d (batch statistics) -> ELU -> Linear, every stored activation bf16"""
    import lesion_gnn_b200 as lg
    n = 400
    ei = random_graph(n, 3000, 12)
    x = torch.randn(n, 32, generator=torch.Generator().manual_seed(13), dtype=torch.float64)
    torch.manual_seed(14)
    ref = O.GINConv(O.MLP([32, 32, 32], act="elu")).double().train()
    gpu = lg.nn.GINConv(lg.nn.MLP([32, 32, 32], act="ELU")).train()
    _run(gpu, ref, x, [ei], dev, "GIN bf16")


@pytest.mark.parametrize("reduce", ["mean", "max", "sum"])
def test_global_pool_bf16(dev, reduce):
    import lesion_gnn_b200 as lg
    hb = knn_batch(9, "ddr", 48, 6, seed=3)
    x = _distinct(hb.x.size(0), 48, 5) if reduce == "max" else _r(hb.x.double())
    xr = x.clone().requires_grad_(True)
    want = {"mean": O.global_mean_pool, "max": O.global_max_pool, "sum": O.global_add_pool}[reduce](xr, hb.batch, hb.num_graphs)
    dy = _r(torch.randn(want.shape, generator=torch.Generator().manual_seed(2), dtype=torch.float64))
    want.backward(dy)
    xg = x.to(torch.bfloat16).to(dev).requires_grad_(True)
    got = {"mean": lg.nn.global_mean_pool, "max": lg.nn.global_max_pool, "sum": lg.nn.global_add_pool}[reduce](xg, hb.batch.to(dev))
    assert got.dtype == torch.bfloat16
    got.backward(dy.to(torch.bfloat16).to(dev))
    _check(got, want, f"pool {reduce}")
    _check(xg.grad, xr.grad, f"pool {reduce} dx")


def _model_bf16(make_gpu, make_ref, hb, dev, loss_type):
    """whole model, forward + loss + all gradients at 1e-2 / 5e-2 per row, or 4x (the multiple tests/util.assert_parity uses for fp32) the bf16 storage floor of the same
    model on the same inputs where depth and cancellation push the floor itself above the bar"""
    import lesion_gnn_b200 as lg
    torch.manual_seed(11)
    ref = make_ref().double()
    mod = make_gpu()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    x = _r(hb.x.double())

    def run_ref(model):
        model.zero_grad()
        xr = x.clone().requires_grad_(True)
        out = model(xr, hb.edge_index, hb.batch, num_graphs=hb.num_graphs)
        if loss_type == "MSE":
            loss = torch.nn.functional.mse_loss(torch.clamp(out.squeeze(1), 0, 4), hb.y.double())
        else:
            loss = torch.nn.functional.cross_entropy(out, hb.y)
        loss.backward()
        return xr, out, loss

    xr, out_r, loss_r = run_ref(ref)
    emu = _emulate_bf16_storage(ref)
    xe, out_e, loss_e = run_ref(emu)
    xg = x.to(torch.bfloat16).to(dev).requires_grad_(True)
    out = mod(xg, hb.edge_index.to(dev), hb.batch.to(dev))
    assert out.dtype == torch.bfloat16
    loss = lg.train.loss_fn(out.float(), hb.y.to(dev), loss_type, 5)
    loss.backward()
    # [B, out]: one row for the L2 metric (single logits near zero have no meaningful relative error of their own)
    _check(out.reshape(1, -1), out_r.reshape(1, -1), "logits", floor=out_e.reshape(1, -1))
    _check(loss, loss_r, "loss", floor=loss_e)
    _check(xg.grad, xr.grad, "dx", floor=xe.grad)
    refg, emug = dict(ref.named_parameters()), dict(emu.named_parameters())
    for k, p in mod.named_parameters():
        if refg[k].grad is not None and float(refg[k].grad.abs().max()) > 0:
            _check(p.grad.view_as(refg[k].grad), refg[k].grad, "d" + k, floor=emug[k].grad)


def test_gat_model_bf16(dev):
    """BASELINE.json config 2 shape in bf16: GAT(1025, [64]*4, heads=4) on 64 DDR-shaped graphs"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(64, "ddr", 1025, 6, seed=1234)
    _model_bf16(lambda: lg.models.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0),
                lambda: O.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0), hb, dev, "MSE")


def test_gcn_model_bf16(dev):
    import lesion_gnn_b200 as lg
    hb = knn_batch(32, 100, 16, 8, seed=1234)
    _model_bf16(lambda: lg.models.GCN(16, [64, 64], 5), lambda: O.GCN(16, [64, 64], 5), hb, dev, "CE")


def test_gin_model_bf16(dev):
    """BASELINE.json config 3 (GIN, BatchNorm in train mode, mean read-out) in bf16"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(64, 100, 64, 8, seed=5)
    _model_bf16(lambda: lg.models.GIN(64, [64] * 3, 5, dropout=0.0).train(), lambda: O.GIN(64, [64] * 3, 5, dropout=0.0).train(), hb, dev, "CE")


def test_training_step_bf16(dev):
    """TrainStep over a DeviceBatch whose features are bf16: captured, replayed, loss falls, parameters stay fp32"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(32, "ddr", 64, 6, seed=9)
    hb.x = hb.x.to(torch.bfloat16)
    torch.manual_seed(3)
    model = lg.models.GAT(64, [32, 32, 32], 1, heads=4, dropout=0.0).to(dev).train()
    opt = train.FlatAdam(model, lr=3e-3)
    step = train.TrainStep(model, opt, "MSE", 5, use_cuda_graph=True)
    db = train.DeviceBatch(hb, dev)
    assert db.x.dtype == torch.bfloat16
    db.copy_from(hb)
    losses = [float(step(db)) for _ in range(30)]
    assert all(l == l for l in losses) and losses[-1] < losses[0]
    assert all(p.dtype == torch.float32 for p in model.parameters())
