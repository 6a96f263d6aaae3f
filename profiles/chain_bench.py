"""the kernels of the headline step's DEPENDENT CHAIN (256 DDR graphs, F_in 1025, GAT 3 x 64, 4 heads), each timed alone
inside a CUDA graph of 20 calls -- what one link costs when nothing else runs beside it.  One JSON line per kernel."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

dev = torch.device("cuda")
H, C, FIN = 4, 16, 1025
F = H * C
hb = data.synth_batch(256, "ddr", FIN, seed=data.SEED)
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), 256)
ei = transforms.knn_graph(hb.pos.float().to(dev), 6, seg, loop=True)
g = lg.CSRGraph.from_edge_index(ei, N)
x = hb.x.to(dev)
gen = torch.Generator(device=dev).manual_seed(1)
w_in = torch.randn(F, FIN, device=dev, generator=gen) * 0.03
b_in = torch.randn(F, device=dev, generator=gen)
w = torch.randn(F, F, device=dev, generator=gen) * 0.1
att = torch.randn(F, device=dev, generator=gen) * 0.3
bias = torch.randn(F, device=dev, generator=gen)
h = torch.randn(N, F, device=dev, generator=gen)
dout = torch.randn(N, F, device=dev, generator=gen)
w_out = torch.randn(1, 2 * F, device=dev, generator=gen) * 0.1
a_s, a_d = ops.gat_alpha(h, att, att, H, C)
out, rmax, rsum, y_elu = ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, H, C, 0.2, True, 0.35, 7, with_elu=True)
pooled_sum, pooled_max, arg = ops.pool_fwd(h, seg.ptr, True, True, True)
pooled = torch.cat([pooled_sum, pooled_max], 1)
dpool = torch.randn(256, F, device=dev, generator=gen)
yb = hb.y.to(dev).long()
logit0 = pooled[:, 0].contiguous()
dlogit = dpool[:, :1].contiguous()


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
    return round(e0.elapsed_time(e1) * 1000 / (reps * 20), 2)


res = {"shape": {"nodes": N, "edges": int(ei.size(1)), "F_in": FIN, "F": F, "heads": H}}
res["in_proj_fwd (elu, bias)"] = t(lambda: ops.linear_fwd(x, w_in, b_in, act=1))
res["layer transform + logit halves"] = t(lambda: ops.linear_gat_fwd(h, w, att, att, H, C))
res["layer transform"] = t(lambda: ops.linear_fwd(h, w))
res["gat_alpha"] = t(lambda: ops.gat_alpha(h, att, att, H, C))
res["gat_fwd_elu p=0.35"] = t(lambda: ops.gat_fwd(g.rowptr, g.col, h, a_s, a_d, bias, H, C, 0.2, True, 0.35, 7, with_elu=True))
res["gat_bwd dst+src p=0.35"] = t(lambda: ops.gat_bwd(g, h, a_s, a_d, rmax, rsum, out, bias, att, att, dout, H, C, 0.2, True, 0.35, 7,
                                                      through_elu=True))
res["layer dX (64 x 64)"] = t(lambda: ops.linear_bwd_x(dout, w))
res["layer dW + db (side stream)"] = t(lambda: ops.linear_bwd_w(dout, h, want_dw=True, want_db=True))
res["pool_fwd mean+max"] = t(lambda: ops.pool_fwd(h, seg.ptr, True, True, True))
res["out_proj fwd"] = t(lambda: ops.linear_fwd(pooled, w_out))
res["loss"] = t(lambda: ops.clamped_regression_loss(logit0, yb, 0.0, 4.0, 0))
res["out_proj dX"] = t(lambda: ops.linear_bwd_x(dlogit, w_out))
res["pool_bwd"] = t(lambda: ops.pool_bwd(dpool, True, dpool, arg, seg.ptr, N, F, torch.float32, dev))
res["in_proj dW + db"] = t(lambda: ops.linear_bwd_w(dout, x, want_dw=True, want_db=True))
print(json.dumps(res))
