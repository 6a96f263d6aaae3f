"""the three GAT kernels on the headline shape (256 DDR graphs, H = 4, C = 16), timed inside a CUDA graph of 20 calls"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

dev = torch.device("cuda")
H, C = 4, 16
hb = data.synth_batch(256, "ddr", H * C, seed=data.SEED)
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), 256)
ei = transforms.knn_graph(hb.pos.float().to(dev), 6, seg, loop=True)
g = lg.CSRGraph.from_edge_index(ei, N)
h = hb.x.to(dev)
att = torch.randn(H * C, device=dev) * 0.3
a_s, a_d = ops.gat_alpha(h, att, att, H, C)
bias = torch.randn(H * C, device=dev)
dout = torch.randn(N, H * C, device=dev)


def t(fn, reps=50):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for _ in range(20):
            fn()
    gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        gr.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (reps * 20)


for p in (0.0, 0.35):
    out, rmax, rsum, y = ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, H, C, 0.2, True, p, 7, with_elu=True)
    print(f"dropout {p}: gat_fwd_elu {t(lambda: ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, H, C, 0.2, True, p, 7, with_elu=True)):.2f} us", end="  ")
    print(f"gat_bwd (dst + src) {t(lambda: ops.gat_bwd(g, h, a_s, a_d, rmax, rsum, out, bias, att, att, dout, H, C, 0.2, True, p, 7, through_elu=True)):.2f} us")
