"""GAT soft-max + aggregate on the config-4 stress shape: lgk_gat_fwd vs lgk_gat_fwd_blocked.  usage: python profiles/gat_bench.py [graphs]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda")
n, k, F, H = 10000, 16, 128, 4
C = F // H
hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), graphs)
ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
E = ei.size(1)
g = lg.CSRGraph.from_edge_index(ei, N, dst_sorted=True)
x = hb.x.to(dev)
att = torch.randn(H * C, device=dev) * 0.1
a_s, a_d = ops.gat_alpha(x, att, att, H, C)
q = g.quads("csr")


def t(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / reps


nbytes = 4 * (2 * N * F + 4 * N * H) + 4 * E + 4 * (N + 1)
peak = 6551.0
for name, fn in (("lgk_gat_fwd", lambda: ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True)),
                 ("lgk_gat_fwd_blocked", lambda: ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True, quads=q))):
    us = t(fn)
    print(json.dumps({"kernel": name, "graphs": graphs, "us": round(us, 1), "GBps": round(nbytes / us / 1e3, 1), "frac": round(nbytes / us / 1e3 / peak, 4)}))
a = ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True)
b = ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True, quads=q)
print(json.dumps({"max_rel_diff": float((a[0] - b[0]).abs().max() / a[0].abs().max())}))
