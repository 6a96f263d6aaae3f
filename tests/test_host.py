"""Host-side logic on CPU: synthetic data shapes, data-parallel sharding, flat parameter buckets, and the
world_size-2 gradient exchange over gloo (the N>1 path of SURVEY.md 8e without GPUs)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from lesion_gnn_b200 import data, train, transforms
from oracle import pyg_semantics as O


def test_synth_batch_shapes_and_determinism():
    a = data.synth_batch(32, 100, 16)
    b = data.synth_batch(32, 100, 16)
    assert a.x.shape == (3200, 16) and a.pos.dtype == torch.float64 and a.batch.dtype == torch.int64
    assert torch.equal(a.x, b.x) and torch.equal(a.pos, b.pos) and torch.equal(a.y, b.y)
    assert bool((a.batch[1:] >= a.batch[:-1]).all())
    d = data.synth_batch(256, "ddr", 8)
    assert min(d.counts) >= 1 and max(d.counts) <= 600 and 20 < sum(d.counts) / 256 < 80


def test_knn_num_edges_matches_oracle():
    hb = data.synth_batch(6, "ddr", 4, seed=3)
    for loop in (True, False):
        ei = O.knn_graph(hb.pos, 6, hb.batch, loop=loop)
        assert ei.size(1) == transforms.knn_num_edges(hb.counts, 6, loop)


def test_shard_counts_balanced_and_complete():
    hb = data.synth_batch(256, "ddr", 4)
    for w in (1, 2, 4, 8):
        sh = data.shard_counts(hb.counts, w)
        assert sh[0][0] == 0 and sh[-1][1] == 256 and all(a[1] == b[0] for a, b in zip(sh, sh[1:]))
        loads = [sum(hb.counts[a:b]) for a, b in sh]
        assert max(loads) - min(loads) <= 600
    # more ranks than graphs: empty shards are allowed, nothing is lost
    sh = data.shard_counts([5, 5], 4)
    assert sum(b - a for a, b in sh) == 2


def test_take_graphs_rebases_indices():
    hb = data.synth_batch(5, "ddr", 4, seed=9)
    hb.edge_index = O.knn_graph(hb.pos, 3, hb.batch, loop=True)
    sub = data.take_graphs(hb, 2, 4)
    n = sum(hb.counts[2:4])
    assert sub.num_nodes == n and sub.num_graphs == 2 and int(sub.batch.min()) == 0
    assert int(sub.edge_index.min()) >= 0 and int(sub.edge_index.max()) < n
    assert torch.equal(sub.edge_index, O.knn_graph(sub.pos, 3, sub.batch, loop=True))


def test_flat_params_are_views_and_grads_accumulate_in_bucket():
    torch.manual_seed(0)
    m = O.GCN(8, [16, 16], 5)
    want = {k: v.clone() for k, v in m.state_dict().items()}
    fp = train.FlatParams(m)
    for k, v in m.state_dict().items():
        assert torch.equal(v, want[k])
    fp.zero_grad()
    hb = data.synth_batch(3, 10, 8)
    ei = O.knn_graph(hb.pos, 3, hb.batch, loop=True)
    torch.nn.functional.cross_entropy(m(hb.x, ei, hb.batch, num_graphs=3), hb.y).backward()
    assert float(fp.grad.abs().sum()) > 0
    for p, o in zip(fp.params, fp.offsets):
        assert p.grad.data_ptr() == fp.grad.data_ptr() + 4 * o and o % 32 == 0
        assert p.data_ptr() == fp.param.data_ptr() + 4 * o


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dp_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    hb = data.synth_batch(9, "ddr", 8, seed=21)
    hb.edge_index = O.knn_graph(hb.pos, 4, hb.batch, loop=True)
    lo, hi = data.shard_counts(hb.counts, world)[rank]
    sub = data.take_graphs(hb, lo, hi)
    torch.manual_seed(3)
    m = O.GCN(8, [16, 16], 5).float()
    fp = train.FlatParams(m)
    fp.zero_grad()
    loss = torch.nn.functional.cross_entropy(m(sub.x, sub.edge_index, sub.batch, num_graphs=sub.num_graphs), sub.y)
    (loss * train.dp_loss_weight(sub.num_graphs, hb.num_graphs, world)).backward()
    fp.all_reduce()
    q.put((rank, (fp.grad / world).clone()))
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_gradient_equals_full_batch_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_dp_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process gradient on the whole batch
    hb = data.synth_batch(9, "ddr", 8, seed=21)
    hb.edge_index = O.knn_graph(hb.pos, 4, hb.batch, loop=True)
    torch.manual_seed(3)
    m = O.GCN(8, [16, 16], 5).float()
    fp = train.FlatParams(m)
    fp.zero_grad()
    torch.nn.functional.cross_entropy(m(hb.x, hb.edge_index, hb.batch, num_graphs=9), hb.y).backward()
    for r in range(world):
        assert torch.allclose(got[r], fp.grad, rtol=1e-4, atol=1e-6)
    assert torch.equal(got[0], got[1])


def test_collate_matches_pyg_batching():
    """PyG collation (reference datasets/datamodule.py:64-69): cat, cumulative node offsets on edge_index, batch vector,
    ptr; splitting the collated batch again (take_graphs) returns every graph"""
    import torch
    from lesion_gnn_b200 import data
    from oracle import pyg_semantics as O
    g = torch.Generator().manual_seed(3)
    samples = []
    for n in (5, 1, 9, 4):
        pos = torch.rand(n, 2, generator=g, dtype=torch.float64) * 100
        ei = O.knn_graph(pos, 3, None, loop=True)
        samples.append(data.GraphSample(x=torch.randn(n, 6, generator=g), pos=pos, edge_index=ei,
                                        y=torch.randint(0, 5, (1,), generator=g), edge_weight=torch.rand(ei.size(1), generator=g)))
    hb = data.collate(samples)
    assert hb.counts == [5, 1, 9, 4] and hb.num_graphs == 4 and hb.num_nodes == 19
    assert torch.equal(hb.batch, torch.tensor([0] * 5 + [1] + [2] * 9 + [3] * 4))
    assert torch.equal(data.ptr_from_counts(hb.counts), torch.tensor([0, 5, 6, 15, 19]))
    assert torch.equal(hb.y, torch.cat([s.y for s in samples]))
    off = 0
    for i, s in enumerate(samples):
        sub = data.take_graphs(hb, i, i + 1)
        assert torch.equal(sub.x, s.x) and torch.equal(sub.pos, s.pos)
        assert torch.equal(sub.edge_index, s.edge_index) and torch.equal(sub.edge_weight, s.edge_weight)
        lo, hi = off, off + s.x.size(0)
        m = (hb.edge_index[1] >= lo) & (hb.edge_index[1] < hi)
        assert bool(((hb.edge_index[0][m] >= lo) & (hb.edge_index[0][m] < hi)).all()), "an edge crosses graphs"
        off = hi
    # batching == building the kNN graph of the whole batch at once (what the device pipeline does)
    assert torch.equal(hb.edge_index, O.knn_graph(hb.pos, 3, hb.batch, loop=True))
    import pytest
    with pytest.raises(ValueError):
        data.collate([data.GraphSample(x=torch.zeros(2, 3), edge_index=torch.tensor([[0], [5]]))])
