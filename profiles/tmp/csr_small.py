import sys, torch
sys.path.insert(0, '.')
import lesion_gnn_b200 as lg
from lesion_gnn_b200 import data, transforms, ops
from lesion_gnn_b200.nn.pool import Segments
dev = torch.device('cuda')
hb = data.synth_batch(256, "ddr", 4, seed=data.SEED)
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), 256)
ei = transforms.knn_graph(hb.pos.float().to(dev), 6, seg, loop=True)
def t(fn, reps=50):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (reps * 10)
print("N", N, "E", ei.size(1))
print("general in-graph us", round(t(lambda: ops.csr_build(ei, N, True)), 2))
print("sorted  in-graph us", round(t(lambda: ops.csr_build(ei, N, True, dst_sorted=True)), 2))
