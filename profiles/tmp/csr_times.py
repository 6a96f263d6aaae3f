import sys, torch
sys.path.insert(0, '.')
import lesion_gnn_b200 as lg
from lesion_gnn_b200 import data, transforms, ops
from lesion_gnn_b200.nn.pool import Segments
dev = torch.device('cuda')
graphs, n, k = 128, 10000, 16
hb = data.synth_batch(graphs, n, 4, seed=data.SEED, layout="raster")
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), graphs)
ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
torch.cuda.synchronize()
for _ in range(2):
    ops.csr_build(ei, N, True, dst_sorted=True)
torch.cuda.synchronize()
