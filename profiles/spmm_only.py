"""Aggregation kernels only, on the SpMM stress shape (graph cached under /tmp): the command profiled with ncu.
usage: python profiles/spmm_only.py [graphs] [reps] [feat]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
F = int(sys.argv[3]) if len(sys.argv) > 3 else 128
dev = torch.device("cuda:0")
n, k = 10000, 16
hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), graphs)
ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
g = lg.CSRGraph.from_edge_index(ei, N)
x = hb.x.to(dev)
dinv, _ = g.gcn_norm(False, True)
for _ in range(reps):
    ops.spmm(g.rowptr, g.col, x, "sum")
    ops.spmm(g.rowptr, g.col, x, "sum", pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
torch.cuda.synchronize()
print("ok", N, ei.size(1))
