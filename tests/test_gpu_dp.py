"""The data-parallel optimiser step without NCCL (csrc/dp.cu, train.FlatAdam.step_dp): the gradient
all-reduce over NVLink peer memory fused with Adam.  One GPU: a single-rank bucket reproduces lgk_adam_step bit for bit,
eagerly and inside the captured training step.  Two GPUs (skipped on a one-GPU box): two ranks with different shards end
with bit-identical parameters, equal to the NCCL all-reduce + lgk_adam_step path bit for bit, over many graph replays."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(lg, dev, seed=5):
    from lesion_gnn_b200.nn import conv
    torch.manual_seed(seed)
    conv._seed_counter[0] = 0  # dropout seeds are a function of torch's seed and a call counter: same masks in every run
    return lg.models.GAT(24, [32, 32, 32], 1, heads=4, dropout=0.35).to(dev).train()


@pytest.mark.parametrize("adamw", [False, True])
def test_single_rank_bucket_equals_adam_step(dev, adamw):
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    ma, mb = _model(lg, dev), _model(lg, dev)
    a = train.FlatAdam(ma, lr=1e-2, weight_decay=1e-3, adamw=adamw)
    b = train.FlatAdam(mb, lr=1e-2, weight_decay=1e-3, adamw=adamw, peer_bucket=train.PeerBucket())
    assert b.grad.data_ptr() % 16 == 0 and b.peer_bucket.world == 1
    gen = torch.Generator(device=dev).manual_seed(1)
    for step in range(4):
        g = torch.randn(a.numel, device=dev, generator=gen)
        a.zero_grad()
        b.zero_grad()
        a.grad.copy_(g)
        b.grad.copy_(g)
        a.step()
        b.step_dp()
        assert torch.equal(a.param, b.param), step
        assert torch.equal(a.exp_avg_sq, b.exp_avg_sq)
        assert float(b.grad.abs().max()) == 0.0  # released and cleared for the next backward
    assert int(b.peer_bucket.epoch_dev[0]) == 4 and int(b.step_dev[0]) == 4


def test_single_rank_bucket_inside_the_captured_step(dev):
    """TrainStep with world_size 1 never takes the data-parallel path; force it with a one-rank bucket and compare with
    the ordinary captured step over replays (warm-up rewinds the optimiser's step count, not the bucket's epoch)"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    from tests.util import knn_batch
    hb = knn_batch(12, "ddr", 24, 6, seed=3)
    res = []
    for fused in (False, True):
        model = _model(lg, dev)
        opt = train.FlatAdam(model, lr=1e-2, peer_bucket=train.PeerBucket() if fused else None)
        step = train.TrainStep(model, opt, "MSE", 5, use_cuda_graph=True, world_size=1)
        if fused:
            step.world_size, step._fused_dp = 2, True  # one rank standing in for the group: mean over `world` = 1 bucket
            opt.peer_bucket.world = 1
        db = train.DeviceBatch(hb, dev)
        db.copy_from(hb)
        losses = [float(step(db)) for _ in range(5)]
        assert len(set(losses)) == 5  # dropout masks move with the step count
        res.append((losses, opt.param.clone(), int(opt.step_dev[0])))
    assert res[0][0] == res[1][0]
    assert torch.equal(res[0][1], res[1][1])
    assert res[0][2] == res[1][2] == 5


def _worker(rank, world, port, fused, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import train
    from tests.util import knn_batch
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    try:
        model = _model(lg, dev)
        bucket = train.PeerBucket(dist.group.WORLD) if fused else None
        opt = train.FlatAdam(model, lr=1e-2, weight_decay=1e-4, peer_bucket=bucket)
        step = train.TrainStep(model, opt, "MSE", 5, use_cuda_graph=True, world_size=world)
        assert step._fused_dp == fused
        hb = knn_batch(10 + 3 * rank, "ddr", 24, 6, seed=7 + rank)  # different shards (and sizes) per rank
        db = train.DeviceBatch(hb, dev)
        db.copy_from(hb)
        db.loss_weight = train.dp_loss_weight(db.num_graphs, 10 + 13, world)
        losses = [float(step(db)) for _ in range(20)]
        torch.cuda.synchronize()
        torch.save({"param": opt.param.cpu(), "losses": losses, "steps": int(opt.step_dev[0])}, os.path.join(out_dir, f"{int(fused)}_{rank}.pt"))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_ranks_fused_equals_nccl(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    for fused, port in ((False, 29611), (True, 29612)):
        mp.spawn(_worker, args=(2, port, fused, str(tmp_path)), nprocs=2, join=True)
    r = {(f, k): torch.load(tmp_path / f"{f}_{k}.pt") for f in (0, 1) for k in (0, 1)}
    for f in (0, 1):  # replicas stay bit-identical
        assert torch.equal(r[(f, 0)]["param"], r[(f, 1)]["param"]), f
        assert r[(f, 0)]["steps"] == 20
    # two ranks: a + b is the same in either order, so the peer-memory sum equals NCCL's bit for bit
    assert torch.equal(r[(0, 0)]["param"], r[(1, 0)]["param"])
    assert r[(0, 0)]["losses"] == r[(1, 0)]["losses"]
