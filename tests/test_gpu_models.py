"""GPU parity of whole models (reference models/gat.py:48-59, gin.py:28-35 and the GCN / SAGE classifiers of
BASELINE.json) and of the training step (models/base.py:196-201 + Adam) against the oracle; size-independent
properties at BASELINE.json's full sizes."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_F32, assert_close, knn_batch, rel_err

pytestmark = pytest.mark.gpu


def _model_parity(make_gpu, make_ref, hb, dev, loss_type, tol, edge_weight=None):
    import lesion_gnn_b200 as lg
    torch.manual_seed(11)
    ref = make_ref().double()
    mod = make_gpu()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    xr = hb.x.double().requires_grad_(True)
    kw = {} if edge_weight is None else {"edge_weight": edge_weight.double()}
    out_r = ref(xr, hb.edge_index, hb.batch, num_graphs=hb.num_graphs, **kw)
    if loss_type == "MSE":
        loss_r = torch.nn.functional.mse_loss(torch.clamp(out_r.squeeze(1), 0, 4), hb.y.double())
    else:
        loss_r = torch.nn.functional.cross_entropy(out_r, hb.y)
    loss_r.backward()
    xg = hb.x.to(dev).requires_grad_(True)
    args = [] if edge_weight is None else [edge_weight.float().to(dev)]
    out = mod(xg, hb.edge_index.to(dev), hb.batch.to(dev), *args)
    loss = lg.train.loss_fn(out, hb.y.to(dev), loss_type, 5)
    loss.backward()
    assert_close(out, out_r, tol, "logits")
    assert_close(loss, loss_r, tol, "loss")
    assert_close(xg.grad, xr.grad, tol, "dx")
    refg = dict(ref.named_parameters())
    for k, p in mod.named_parameters():
        if refg[k].grad is not None and float(refg[k].grad.abs().max()) > 0:
            assert_close(p.grad.view_as(refg[k].grad), refg[k].grad, tol, "d" + k)


def test_gcn_config1(dev):
    """BASELINE.json config 1: 2-layer GCN, 32 graphs x 100 nodes, 16-d, kNN k=8"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(32, 100, 16, 8, seed=1234)
    _model_parity(lambda: lg.models.GCN(16, [64, 64], 5), lambda: O.GCN(16, [64, 64], 5), hb, dev, "CE", TOL_F32)


def test_gat_config2_small(dev):
    """BASELINE.json config 2 shape (3 GAT layers, 4 heads, 64-d, DDR-shaped graphs, k=6 loop=True), 24 graphs"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(24, "ddr", 1025, 6, seed=1234)
    _model_parity(lambda: lg.models.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0),
                  lambda: O.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0), hb, dev, "MSE", 2e-5)


def test_gin_and_sage_config3_small(dev):
    import lesion_gnn_b200 as lg
    hb = knn_batch(40, 100, 64, 8, seed=5)
    # eval(): BatchNorm running statistics, no dropout -- BN batch statistics are checked in test_gpu_layers
    for train_mode in (False, True):
        torch.manual_seed(11)
        ref = O.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max").double().train(train_mode)
        mod = lg.models.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max")
        mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
        mod = mod.to(dev).train(train_mode)
        want = ref(hb.x.double(), hb.edge_index, hb.batch, num_graphs=40)
        got = mod(hb.x.to(dev), hb.edge_index.to(dev), hb.batch.to(dev))
        assert_close(got, want, 5e-5, f"GIN logits train={train_mode}")
    _model_parity(lambda: lg.models.SAGE(64, [64] * 3, 5), lambda: O.SAGE(64, [64] * 3, 5), hb, dev, "CE", TOL_F32)


def test_gcn_with_gaussian_edge_weights(dev):
    import lesion_gnn_b200 as lg
    hb = knn_batch(6, 50, 16, 6, seed=8, loop=False)
    ew = O.gaussian_distance(hb.edge_index, hb.pos, 200.0, torch.float64) * 500
    _model_parity(lambda: lg.models.GCN(16, [32, 32], 5), lambda: O.GCN(16, [32, 32], 5), hb, dev, "CE", TOL_F32, ew)


def test_csrgraph_input_equals_edge_index_input(dev):
    """`adj_t`-style CSRGraph and Segments inputs give the same result as plain tensors (reference gat.py:87-90)"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.nn.pool import Segments
    hb = knn_batch(10, "ddr", 32, 6, seed=3)
    torch.manual_seed(0)
    mod = lg.models.GAT(32, [32, 32], 5, heads=2, dropout=0.0).to(dev)
    x, ei, b = hb.x.to(dev), hb.edge_index.to(dev), hb.batch.to(dev)
    a = mod(x, ei, b)
    c = mod(x, lg.CSRGraph.from_edge_index(ei, hb.num_nodes), Segments.from_batch(b, hb.num_graphs))
    assert torch.equal(a, c)


@pytest.mark.parametrize("loss_type", ["MSE", "SmoothL1"])
def test_clamped_regression_loss_matches_torch(dev, loss_type):
    """loss + gradient in one launch (lgk_clamped_regression_loss) vs the reference's criterion on clamped logits
    (models/base.py:196-199, gat.py:94-95): values below, inside, above and exactly on the clamp bounds"""
    from lesion_gnn_b200 import train
    torch.manual_seed(3)
    n = 777
    x = (torch.randn(n, 1, dtype=torch.float64) * 3 + 2)
    x[:6, 0] = torch.tensor([0.0, 4.0, -0.0, 4.0000001, -1e-7, 2.5])
    y = torch.randint(0, 5, (n,))
    xr = x.clone().requires_grad_(True)
    c = torch.clamp(xr.squeeze(1), 0, 4)
    want = (torch.nn.functional.mse_loss if loss_type == "MSE" else torch.nn.functional.smooth_l1_loss)(c, y.double())
    (want * 0.7).backward()
    xg = x.float().to(dev).requires_grad_(True)
    got = train.loss_fn(xg, y.to(dev), loss_type, 5)
    (got * 0.7).backward()
    assert_close(got, want, TOL_F32, "loss")
    # the clamp mask is decided on the fp32 logits: compare against the fp64 reference evaluated at those values
    xr2 = x.float().double().requires_grad_(True)
    c2 = torch.clamp(xr2.squeeze(1), 0, 4)
    w2 = (torch.nn.functional.mse_loss if loss_type == "MSE" else torch.nn.functional.smooth_l1_loss)(c2, y.double())
    (w2 * 0.7).backward()
    assert_close(xg.grad, xr2.grad, TOL_F32, "dlogits")
    assert torch.equal(train.loss_fn(xg, y.to(dev), loss_type, 5), got.detach())   # deterministic


def test_adam_matches_torch(dev):
    from lesion_gnn_b200 import ops
    torch.manual_seed(0)
    for adamw, wd in ((False, 0.0), (False, 0.01), (True, 0.05)):
        p = torch.randn(1000, dtype=torch.float64)
        ref = p.clone().requires_grad_(True)
        opt = (torch.optim.AdamW if adamw else torch.optim.Adam)([ref], lr=1e-2, weight_decay=wd)
        pg = p.float().to(dev)
        m, v = torch.zeros_like(pg), torch.zeros_like(pg)
        step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        for step in range(1, 6):
            g = torch.randn(1000, dtype=torch.float64)
            ref.grad = g.clone()
            opt.step()
            if step % 2:
                ops.adam_step(pg, g.float().to(dev), m, v, lr=1e-2, weight_decay=wd, step=step, adamw=adamw)
                step_dev.add_(1)
            else:  # device-resident step counter (CUDA-graph mode)
                step_dev.add_(1)
                ops.adam_step(pg, g.float().to(dev), m, v, lr=1e-2, weight_decay=wd, step=0, step_dev=step_dev, adamw=adamw)
        assert_close(pg, ref, 1e-5, f"adam adamw={adamw} wd={wd}")
        # the kernel advances the counter itself (int64[2] = {steps completed, scratch}) and clears the gradients
        ref2 = p.clone().requires_grad_(True)
        opt2 = (torch.optim.AdamW if adamw else torch.optim.Adam)([ref2], lr=1e-2, weight_decay=wd)
        pg2 = p.float().to(dev)
        m2, v2 = torch.zeros_like(pg2), torch.zeros_like(pg2)
        counter = torch.zeros(2, dtype=torch.int64, device=dev)
        for step in range(1, 6):
            g = torch.randn(1000, dtype=torch.float64)
            ref2.grad = g.clone()
            opt2.step()
            gg = g.float().to(dev)
            ops.adam_step(pg2, gg, m2, v2, lr=1e-2, weight_decay=wd, step=0, step_dev=counter, adamw=adamw,
                          zero_grad=True, advance_step=True)
            assert counter.tolist() == [step, 0] and float(gg.abs().max()) == 0.0
        assert_close(pg2, ref2, 1e-5, f"adam (self-advancing) adamw={adamw} wd={wd}")


@pytest.mark.parametrize("use_graph", [False, True])
def test_train_step_matches_oracle_training(dev, use_graph):
    """5 optimisation steps (fwd + loss + bwd + Adam): eager and CUDA-graph replay vs the oracle's training_step"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(16, "ddr", 40, 6, seed=2)
    torch.manual_seed(5)
    ref = O.GAT(40, [32, 32, 32], 1, heads=4, dropout=0.0).double()
    mod = lg.models.GAT(40, [32, 32, 32], 1, heads=4, dropout=0.0)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    opt_r = torch.optim.Adam(ref.parameters(), lr=1e-3)
    opt = train.FlatAdam(mod, lr=1e-3)
    step = train.TrainStep(mod, opt, "MSE", 5, use_cuda_graph=use_graph)
    db = train.DeviceBatch(hb, dev)
    hb_p = hb.pin()
    for it in range(5):
        loss_r = O.training_step(ref, opt_r, hb.x.double(), hb.edge_index, hb.batch, hb.y, hb.num_graphs, "MSE", 5)
        db.copy_from(hb_p)
        loss = step(db)
        assert rel_err(loss, loss_r) < 1e-4, (it, float(loss), float(loss_r))
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        assert_close(p, pr, 1e-4, "param " + k)


def test_full_size_properties_config4(dev):
    """BASELINE.json config 4 at 16 graphs x 10k nodes x 128-d, k=16 (oracle too slow here): linearity of the
    aggregation, <A x, y> = <x, A^T y> (forward CSR vs backward CSC), and CSR structural invariants."""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops, transforms
    from lesion_gnn_b200.nn.pool import Segments
    G, n, k, F = 16, 10000, 16, 128
    hb = data.synth_batch(G, n, F, seed=1234, layout="raster")
    pos, batch = hb.pos.to(dev), hb.batch.to(dev)
    seg = Segments.from_batch(batch, G)
    ei = transforms.knn_graph(pos, k, seg, loop=True, num_edges=G * n * k)
    assert ei.shape == (2, G * n * k)
    g = lg.CSRGraph.from_edge_index(ei, G * n)
    deg = g.rowptr[1:] - g.rowptr[:-1]
    assert int(deg.min()) == k and int(deg.max()) == k and int(g.rowptr[-1]) == G * n * k
    col = g.col.view(G * n, k)
    assert bool((col[:, 1:] > col[:, :-1]).all())                       # sorted, distinct neighbours per row
    assert bool((col // n == torch.arange(G * n, device=dev).unsqueeze(1) // n).all())  # block diagonal
    assert bool((col == torch.arange(G * n, device=dev).unsqueeze(1)).any(1).all())     # loop=True: self edge
    assert torch.equal(torch.sort(g.perm.long()).values, torch.arange(G * n * k, device=dev))
    assert torch.equal(ei[0][g.perm.long()], g.col.long())
    assert torch.equal(g.col[g.csr2csc.long()].long(), torch.repeat_interleave(
        torch.arange(G * n, device=dev), (g.colptr[1:] - g.colptr[:-1]).long()))
    x, y = hb.x.to(dev), torch.randn(G * n, F, device=dev)
    dinv, _ = g.gcn_norm(False, True)
    kw = dict(pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
    ax = ops.spmm(g.rowptr, g.col, x, "sum", **kw)
    ay = ops.spmm(g.rowptr, g.col, y, "sum", **kw)
    axy = ops.spmm(g.rowptr, g.col, 2 * x - 3 * y, "sum", **kw)
    assert rel_err(axy, 2 * ax.double() - 3 * ay.double()) < 1e-5
    aty = ops.spmm(g.colptr, g.row, y, "sum", **kw)
    lhs, rhs = float((ax.double() * y.double()).sum()), float((x.double() * aty.double()).sum())
    assert abs(lhs - rhs) <= 1e-6 * max(abs(lhs), abs(rhs))
    # pooling: mean over equal-size graphs of a constant column is that constant; checksum of sums
    s = lg.nn.global_add_pool(x, seg)
    assert rel_err(s.double().sum(0), x.double().sum(0)) < 1e-6
