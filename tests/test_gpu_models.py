"""GPU parity of whole models (reference models/gat.py:48-59, gin.py:28-35 and the GCN / SAGE classifiers of
BASELINE.json) and of the training step (models/base.py:196-201 + Adam) against the oracle; size-independent
properties at BASELINE.json's full sizes."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import TOL_BF16, TOL_F32, assert_close, assert_parity, knn_batch, rel_err, row_rel_err

pytestmark = pytest.mark.gpu


ROW_TOL = 2e-5  # per-row L2 metric (stricter than the north_star's whole-tensor metric): twice the whole-tensor bar


def _model_parity(make_gpu, make_ref, hb, dev, loss_type, tol=TOL_F32, edge_weight=None, train_mode=None, route_max=False):
    """forward, loss and EVERY gradient (inputs and parameters) of a whole model against the fp64 oracle at the stated
    1e-5 bar (whole-tensor metric) and 2e-5 per row; where fp32 cannot hold it the bound is 4x the fp32 oracle's own
    error on the same inputs (tests/util.assert_parity).  route_max: the oracle's max read-out routes its gradient
    through the arg-max the CUDA forward chose (tests/util.routed_max_pool: near-ties are decisions, not arithmetic)."""
    import contextlib
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import ops
    from tests.util import routed_max_pool
    torch.manual_seed(11)
    ref = make_ref().double()
    ref32 = make_ref().float()
    ref32.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = make_gpu()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    if train_mode is not None:
        ref.train(train_mode), ref32.train(train_mode), mod.train(train_mode)

    # CUDA first: its read-out arg-max is what the oracle's gradient is routed through (route_max)
    args_seen = []
    real_pool_fwd = ops.pool_fwd

    def spy(*a, **k):
        res = real_pool_fwd(*a, **k)
        args_seen.append(res[2])
        return res

    ops.pool_fwd = spy
    try:
        xg = hb.x.detach().clone().to(dev).requires_grad_(True)
        args = [] if edge_weight is None else [edge_weight.float().to(dev)]
        out = mod(xg, hb.edge_index.to(dev), hb.batch.to(dev), *args)
    finally:
        ops.pool_fwd = real_pool_fwd
    loss = lg.train.loss_fn(out, hb.y.to(dev), loss_type, 5)
    loss.backward()

    def run_ref(model, dt):
        x = hb.x.detach().clone().to(dt).requires_grad_(True)
        kw = {} if edge_weight is None else {"edge_weight": edge_weight.to(dt)}
        ctx = routed_max_pool(args_seen[-1]) if route_max else contextlib.nullcontext()
        with ctx:
            out = model(x, hb.edge_index, hb.batch, num_graphs=hb.num_graphs, **kw)
        if loss_type == "MSE":
            loss = torch.nn.functional.mse_loss(torch.clamp(out.squeeze(1), 0, 4), hb.y.to(dt))
        else:
            loss = torch.nn.functional.cross_entropy(out, hb.y)
        loss.backward()
        return x, out, loss

    xr, out_r, loss_r = run_ref(ref, torch.float64)
    x32, out_32, loss_32 = run_ref(ref32, torch.float32)
    kw = dict(tol=tol, row_tol=ROW_TOL)
    assert_parity(out, out_r, out_32, "logits", **kw)
    assert_parity(loss, loss_r, loss_32, "loss", **kw)
    assert_parity(xg.grad, xr.grad, x32.grad, "dx", **kw)
    refg, ref32g = dict(ref.named_parameters()), dict(ref32.named_parameters())
    scale = max(float(g.grad.abs().max()) for g in refg.values() if g.grad is not None)
    for k, p in mod.named_parameters():
        if refg[k].grad is None:
            continue
        if float(refg[k].grad.abs().max()) < 1e-9 * scale:  # analytically zero (a bias in front of BatchNorm)
            assert float(p.grad.abs().max()) < 1e-4 * scale, k
            continue
        assert_parity(p.grad.view_as(refg[k].grad), refg[k].grad, ref32g[k].grad, "d" + k, **kw)


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
                  lambda: O.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0), hb, dev, "MSE")


def test_gat_config2_benched_size(dev):
    """the benched workload itself (bench.py: GAT(1025, [64]*4, heads=4), 256 DDR-shaped graphs, k=6 loop=True, MSE on the
    clamped scalar): forward, loss and all gradients against the fp64 oracle at 1e-5"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(256, "ddr", 1025, 6, seed=1234)
    _model_parity(lambda: lg.models.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0),
                  lambda: O.GAT(1025, [64] * 4, 1, heads=4, dropout=0.0), hb, dev, "MSE")


def test_gat_reference_operating_point(dev):
    """the reference's own configuration (configs/config.py:49,53-65): hiddden_channels=[128]*4, heads=2, k=6"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(64, "ddr", 1025, 6, seed=77)
    _model_parity(lambda: lg.models.GAT(1025, [128] * 4, 1, heads=2, dropout=0.0),
                  lambda: O.GAT(1025, [128] * 4, 1, heads=2, dropout=0.0), hb, dev, "MSE")


@pytest.mark.parametrize("train_mode", [False, True])
def test_gin_config3(dev, train_mode):
    """BASELINE.json config 3, GIN with mean||max read-out at 512 graphs x 100 nodes (k=8): forward AND backward of the
    whole model, in eval mode (running statistics) and in train mode (per-batch BatchNorm statistics, gin.py:23)"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(512, 100, 64, 8, seed=5)
    _model_parity(lambda: lg.models.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max"),
                  lambda: O.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max"), hb, dev, "CE", train_mode=train_mode,
                  route_max=True)
    # the reference's own read-out (gin.py:33: global_mean_pool), no arg-max decisions anywhere
    _model_parity(lambda: lg.models.GIN(64, [64] * 4, 5, dropout=0.0), lambda: O.GIN(64, [64] * 4, 5, dropout=0.0), hb, dev, "CE",
                  train_mode=train_mode)


def test_sage_config3(dev):
    """BASELINE.json config 3, GraphSAGE variant at 512 graphs x 100 nodes"""
    import lesion_gnn_b200 as lg
    hb = knn_batch(512, 100, 64, 8, seed=6)
    _model_parity(lambda: lg.models.SAGE(64, [64] * 3, 5), lambda: O.SAGE(64, [64] * 3, 5), hb, dev, "CE", route_max=True)


def test_gcn_config4_one_graph(dev):
    """BASELINE.json config 4 at G=1: a 10 000-node raster superpixel graph, 128-d, k=16, 2-layer GCN -- through the
    windowed aggregation kernel (forced) and through lgk_spmm, both against the fp64 oracle"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops
    hb = data.synth_batch(1, 10000, 128, seed=1234, layout="raster")
    hb.edge_index = O.knn_graph(hb.pos, 16, hb.batch, loop=True)
    old = ops._WIN_MIN_ELEMS
    try:
        for thresh in (0, 1 << 40):
            ops._WIN_MIN_ELEMS = thresh
            lg.graph.clear_cache()
            _model_parity(lambda: lg.models.GCN(128, [128, 128], 5), lambda: O.GCN(128, [128, 128], 5), hb, dev, "CE")
    finally:
        ops._WIN_MIN_ELEMS = old


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


def test_graph_replays_draw_fresh_attention_dropout_masks(dev):
    """the reference trains with dropout=0.35 (configs/config.py:63).  A captured CUDA graph replays the launch arguments
    it recorded, so the dropout seed must come from device memory: with lr = 0 (parameters frozen) consecutive replays
    of the same batch must still give different losses, and the eager step likewise"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(12, "ddr", 24, 6, seed=5)
    torch.manual_seed(1)
    for graphed in (True, False):
        # 5 logits + cross-entropy: no clamp that could hide the masks (the regression loss clamps logits to [0, 4])
        mod = lg.models.GAT(24, [32, 32, 32], 5, heads=4, dropout=0.35).to(dev).train()
        opt = train.FlatAdam(mod, lr=0.0)
        step = train.TrainStep(mod, opt, "CE", 5, use_cuda_graph=graphed)
        db = train.DeviceBatch(hb, dev)
        db.copy_from(hb)
        losses = [float(step(db)) for _ in range(4)]
        assert len(set(losses)) == 4, (graphed, losses)
        mod.eval()
        ev = train.TrainStep(mod, train.FlatAdam(mod, lr=0.0), "CE", 5, use_cuda_graph=graphed)
        l0, l1 = float(ev(db)), float(ev(db))
        assert l0 == l1, "eval mode must not drop anything"


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


@pytest.mark.parametrize("momentum,nesterov,wd", [(0.0, False, 0.0), (0.0, False, 2e-6), (0.9, False, 0.01), (0.9, True, 0.0)])
def test_sgd_matches_torch(dev, momentum, nesterov, wd):
    """torch.optim.SGD (reference models/base.py:167-168) as one launch over the flat buffers"""
    from lesion_gnn_b200 import ops
    torch.manual_seed(0)
    p = torch.randn(1000, dtype=torch.float64)
    ref = p.clone().requires_grad_(True)
    opt = torch.optim.SGD([ref], lr=1e-2, momentum=momentum, nesterov=nesterov, weight_decay=wd)
    pg = p.float().to(dev)
    buf = torch.zeros_like(pg) if momentum else None
    for _ in range(5):
        g = torch.randn(1000, dtype=torch.float64)
        ref.grad = g.clone()
        opt.step()
        gg = (2 * g).float().to(dev)
        ops.sgd_step(pg, gg, buf, lr=1e-2, momentum=momentum, weight_decay=wd, nesterov=nesterov, grad_scale=0.5, zero_grad=True)
        assert float(gg.abs().max()) == 0.0
    assert_close(pg, ref, 1e-6, "sgd")


@pytest.mark.parametrize("weighted", [False, True])
def test_cross_entropy_loss_matches_torch(dev, weighted):
    """nn.CrossEntropyLoss(weight=class_weights) (reference models/base.py:93-94): loss and gradient in one launch"""
    from lesion_gnn_b200 import train
    torch.manual_seed(4)
    n, C = 1300, 5
    x = torch.randn(n, C, dtype=torch.float64) * 4
    y = torch.randint(0, C, (n,))
    w = torch.tensor([0.3, 1.0, 2.5, 0.7, 4.0], dtype=torch.float64) if weighted else None
    xr = x.clone().requires_grad_(True)
    want = torch.nn.functional.cross_entropy(xr, y, weight=w)
    (want * 0.7).backward()
    xg = x.float().to(dev).requires_grad_(True)
    got = train.loss_fn(xg, y.to(dev), "CE", C, class_weights=None if w is None else w.float())
    (got * 0.7).backward()
    # rows whose softmax has saturated carry gradients of a few ulp: the per-row bound is what torch's own fp32
    # cross-entropy loses on the same logits (assert_parity), the whole-tensor bound is the stated 1e-5
    x32 = x.float().requires_grad_(True)
    want32 = torch.nn.functional.cross_entropy(x32, y, weight=None if w is None else w.float())
    (want32 * 0.7).backward()
    assert_parity(got, want, want32, "loss")
    assert_parity(xg.grad, xr.grad, x32.grad, "dlogits")


def test_train_step_sgd_weighted_ce_matches_oracle(dev):
    """the reference's other optimiser / loss branch: SGD + class-weighted cross-entropy, 3 steps, graph replay"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(16, "ddr", 24, 6, seed=9)
    torch.manual_seed(6)
    w = torch.tensor([0.5, 1.0, 2.0, 1.5, 3.0])
    ref = O.GAT(24, [32, 32], 5, heads=2, dropout=0.0).double()
    mod = lg.models.GAT(24, [32, 32], 5, heads=2, dropout=0.0)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    opt_r = torch.optim.SGD(ref.parameters(), lr=5e-2, weight_decay=2e-6)
    opt = train.make_optimizer(mod, "sgd", lr=5e-2, weight_decay=2e-6)
    step = train.TrainStep(mod, opt, "CE", 5, use_cuda_graph=True, class_weights=w)
    db = train.DeviceBatch(hb, dev)
    db.copy_from(hb)
    for it in range(3):
        opt_r.zero_grad()
        out = ref(hb.x.double(), hb.edge_index, hb.batch, num_graphs=hb.num_graphs)
        loss_r = torch.nn.functional.cross_entropy(out, hb.y, weight=w.double())
        loss_r.backward()
        opt_r.step()
        loss = step(db)
        assert rel_err(loss, loss_r) < 1e-5, (it, float(loss), float(loss_r))
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        assert_close(p, pr, 1e-5, "param " + k)


@pytest.mark.parametrize("use_graph", [False, True])
def test_train_step_matches_oracle_training(dev, use_graph):
    """5 optimisation steps (fwd + loss + bwd + Adam): eager and CUDA-graph replay vs the oracle's training_step"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(16, "ddr", 40, 6, seed=2)
    torch.manual_seed(5)
    ref = O.GAT(40, [32, 32, 32], 1, heads=4, dropout=0.0).double()
    ref32 = O.GAT(40, [32, 32, 32], 1, heads=4, dropout=0.0).float()
    ref32.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = lg.models.GAT(40, [32, 32, 32], 1, heads=4, dropout=0.0)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev)
    opt_r = torch.optim.Adam(ref.parameters(), lr=1e-3)
    opt_32 = torch.optim.Adam(ref32.parameters(), lr=1e-3)
    opt = train.FlatAdam(mod, lr=1e-3)
    step = train.TrainStep(mod, opt, "MSE", 5, use_cuda_graph=use_graph)
    db = train.DeviceBatch(hb, dev)
    hb_p = hb.pin()
    for it in range(5):
        loss_r = O.training_step(ref, opt_r, hb.x.double(), hb.edge_index, hb.batch, hb.y, hb.num_graphs, "MSE", 5)
        loss_32 = O.training_step(ref32, opt_32, hb.x, hb.edge_index, hb.batch, hb.y, hb.num_graphs, "MSE", 5)
        db.copy_from(hb_p)
        loss = step(db)
        # Adam divides by sqrt(v): early steps amplify fp32 rounding of tiny gradients, so the bound is what the oracle
        # itself loses when it trains in fp32 (x4), never below the stated 1e-5
        assert_parity(loss, loss_r, loss_32, f"loss at step {it}")
    for (k, p), (_, pr), (_, p32) in zip(mod.named_parameters(), ref.named_parameters(), ref32.named_parameters()):
        assert_parity(p, pr, p32, "param " + k)


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
