import sys, torch
sys.path.insert(0, '.')
import lesion_gnn_b200 as lg
from lesion_gnn_b200 import data, transforms, ops
from lesion_gnn_b200.nn.pool import Segments
dev = torch.device('cuda')
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / reps
for graphs, n, k in ((128, 10000, 16), (256, 67, 6)):
    hb = data.synth_batch(graphs, n if n != 67 else "ddr", 4, seed=data.SEED, layout="raster" if n != 67 else "uniform")
    N = hb.num_nodes
    seg = Segments.from_batch(hb.batch.to(dev), graphs)
    ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True)
    E = ei.size(1)
    a = ops.csr_build(ei, N, True); b = ops.csr_build(ei, N, True, dst_sorted=True)
    same = all(torch.equal(x, y) for x, y in zip(a, b))
    tg, ts = t(lambda: ops.csr_build(ei, N, True)), t(lambda: ops.csr_build(ei, N, True, dst_sorted=True))
    byts = 28 * E + 8 * (N + 1)
    print(f"N={N} E={E} same={same} violation={int(ops.last_sorted_violation)} general {tg:.1f} us ({byts/tg/1e3/6551:.3f}) sorted {ts:.1f} us ({byts/ts/1e3/6551:.3f})")
