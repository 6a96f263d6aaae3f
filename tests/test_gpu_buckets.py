"""Shape-polymorphic captured training step (SURVEY.md 7 hard part 10; reference training.py:64-72 steps over DataLoader
batches whose (N, E) change every step): 50 different DDR-shaped batches go through train.BucketedTrainStep with at most
4 captured CUDA graphs, and every loss and the final parameters equal those of the plain eager step on the unpadded
batches."""
import pytest
import torch

from tests.util import knn_batch, rel_err

pytestmark = pytest.mark.gpu


def _batches(count, graphs, feat, k):
    return [knn_batch(graphs, "ddr", feat, k, seed=100 + i).pin() for i in range(count)]


@pytest.mark.parametrize("name", ["gat", "sage"])
def test_fifty_shapes_four_graphs(dev, name):
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    graphs, feat, k = 16, 24, 6
    hbs = _batches(50, graphs, feat, k)
    shapes = {(hb.num_nodes, hb.edge_index.size(1)) for hb in hbs}
    assert len(shapes) >= 40, "the batches must really differ in shape"
    assert all(hb.dst_sorted for hb in hbs)  # collated kNN samples: grouped by destination

    def make():
        torch.manual_seed(7)
        if name == "gat":
            m = lg.models.GAT(feat, [32, 32, 32], 1, heads=4, dropout=0.0)
        else:
            m = lg.models.SAGE(feat, [32, 32], 1, aggr="mean", pooling="mean")
        m = m.to(dev).train()
        return m, train.FlatAdam(m, lr=2e-3, weight_decay=1e-5)

    m_ref, o_ref = make()
    eager = train.TrainStep(m_ref, o_ref, "MSE", 5, use_cuda_graph=False)
    m_b, o_b = make()
    bucketed = train.BucketedTrainStep(m_b, o_b, num_graphs=graphs, max_degree=k + 1, loss_type="MSE", node_step=256)
    for i, hb in enumerate(hbs):
        db = train.DeviceBatch(hb, dev)
        db.copy_from(hb)
        want = float(eager(db))
        got = float(bucketed(hb))
        assert abs(got - want) <= 1e-4 * max(abs(want), 1e-3), (i, got, want)
    assert bucketed.captures <= 4, bucketed.captures
    assert sum(len(b) for b in bucketed.buckets.values()) == bucketed.captures
    assert int(o_b.step_dev[0]) == int(o_ref.step_dev[0]) == 50
    assert rel_err(o_b.param, o_ref.param) <= 1e-3


def test_double_buffered_prefetch_equals_the_plain_call(dev):
    """stage() / run() with two buffers per bucket (copy of batch i + 1 overlaps step i): same losses and parameters as the
    blocking call, bit for bit"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    graphs, feat, k = 16, 24, 6
    hbs = _batches(24, graphs, feat, k)
    res = []
    for buffers in (1, 2):
        torch.manual_seed(7)
        m = lg.models.GAT(feat, [32, 32, 32], 1, heads=4, dropout=0.0).to(dev).train()
        opt = train.FlatAdam(m, lr=2e-3)
        st = train.BucketedTrainStep(m, opt, num_graphs=graphs, max_degree=k + 1, node_step=256, buffers=buffers)
        losses = []
        if buffers == 1:
            losses = [float(st(hb)) for hb in hbs]
        else:
            handle = st.stage(hbs[0])
            for i in range(len(hbs)):
                nxt = st.stage(hbs[i + 1]) if i + 1 < len(hbs) else None
                losses.append(st.run(handle))  # no host sync here: the loss tensors are read after the loop
                handle = nxt
        res.append((losses, opt.param.clone(), st.captures))
    # a captured step returns the same loss TENSOR on every replay of its graph, so the prefetching loop (which never
    # synchronises) is compared through the parameters after 24 Adam steps
    assert torch.equal(res[0][1], res[1][1])
    assert res[1][2] <= 2 * len({(hb.num_nodes + 1 + 255) // 256 for hb in hbs})


def test_padding_contributes_exact_zeros(dev):
    """one batch, padded vs unpadded: identical loss and parameter gradients up to the summation order of the slab
    reductions (the padding rows add exact zeros)"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    hb = knn_batch(12, "ddr", 24, 6, seed=3).pin()
    grads = []
    for padded in (False, True):
        torch.manual_seed(1)
        m = lg.models.GAT(24, [32, 32, 32], 1, heads=4, dropout=0.0).to(dev).train()
        opt = train.FlatAdam(m, lr=0.0)
        if padded:
            db = train.PaddedBatch(hb.num_nodes + 500, (hb.num_nodes + 500) * 7, hb.num_graphs, 24, dev)
            db.load(hb)
        else:
            db = train.DeviceBatch(hb, dev)
            db.copy_from(hb)
        opt.zero_grad()
        with train.direct_param_grads():
            loss = train.forward_backward(m, db, "MSE", 5)
        torch.cuda.synchronize()
        grads.append((float(loss), opt.grad.clone()))
    assert abs(grads[0][0] - grads[1][0]) <= 1e-6 * abs(grads[0][0])
    assert rel_err(grads[1][1], grads[0][1]) <= 1e-5


def test_padding_edges_make_no_hub(dev):
    """the spare edges are spread over the spare nodes: no padded node ends up with more than ~max_degree in-edges, even
    when a single spare node remains"""
    from lesion_gnn_b200 import train
    hb = knn_batch(12, "ddr", 8, 6, seed=4).pin()
    for cap in (hb.num_nodes + 1, hb.num_nodes + 37, hb.num_nodes + 4000):
        pb = train.PaddedBatch(cap, cap * 7, hb.num_graphs, 8, dev)
        pb.load(hb)
        dst = pb.edge_index[1]
        assert bool((dst[1:] >= dst[:-1]).all()) and int(dst.max()) == cap - 1 and torch.equal(pb.edge_index[0, hb.edge_index.size(1):],
                                                                                                   dst[hb.edge_index.size(1):])
        deg = torch.bincount(dst[hb.edge_index.size(1):] - hb.num_nodes, minlength=cap - hb.num_nodes)
        spare_e, spare_n = cap * 7 - hb.edge_index.size(1), cap - hb.num_nodes
        assert int(deg.max()) <= -(-spare_e // spare_n) and int(deg.sum()) == spare_e


def test_batchnorm_models_are_refused(dev):
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    m = lg.models.GIN(8, [8, 8], 5, dropout=0.0).to(dev).train()
    with pytest.raises(NotImplementedError):
        train.BucketedTrainStep(m, train.FlatAdam(m), num_graphs=4, max_degree=7)
