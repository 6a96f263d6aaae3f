"""Runs each hot kernel a few times on the SpMM stress shape (BASELINE.json configs[3]: superpixel graphs of
10 000 nodes, 128-d, kNN k=16) and on the in_proj transform of configs[1] -- the command profiled with ncu
(B200_PROFILING.md).  usage: python profiles/stress_kernels.py [graphs] [reps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = torch.device("cuda:0")
n, k, F = 10000, 16, 128
hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), graphs)
ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
g = lg.CSRGraph.from_edge_index(ei, N)
x = hb.x.to(dev)
dinv, _ = g.gcn_norm(False, True)
att = torch.randn(128, device=dev) * 0.1
a_s, a_d = ops.gat_alpha(x, att, att, 4, 32)
xin = torch.randn(11583, 1025, device=dev)
w = torch.randn(64, 1025, device=dev)
dyin = torch.randn(11583, 64, device=dev)
for _ in range(reps):
    ops.spmm(g.rowptr, g.col, x, "sum")
    ops.spmm(g.rowptr, g.col, x, "sum", pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
    ops.spmm(g.colptr, g.row, x, "sum", pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
    ops.spmm(g.rowptr, g.col, x, "max", want_arg=True)
    ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, 4, 32, 0.2, True)
    ops.pool_fwd(x, seg.ptr, True, True, True)
    ops.linear_fwd(xin, w)
    ops.linear_bwd_w(dyin, xin)
    lg.CSRGraph.from_edge_index(ei, N)
torch.cuda.synchronize()
print("ok", N, ei.size(1))
