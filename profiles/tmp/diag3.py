import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import lesion_gnn_b200 as lg
from lesion_gnn_b200 import ops
from oracle import pyg_semantics as O
from tests.util import knn_batch, rel_err
dev = torch.device("cuda:0")
hb = knn_batch(512, 100, 64, 8, seed=5)
torch.manual_seed(11)
ref = O.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max").double().train(True)
mod = lg.models.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max")
mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}); mod = mod.to(dev).train(True)
store = {}
mod.dropout.register_forward_hook(lambda m, i, o: store.__setitem__("x", o.detach()))
mod(hb.x.to(dev), hb.edge_index.to(dev), hb.batch.to(dev))
X = store["x"]                      # [N, 64] activations entering the read-out
Xd = X.double().cpu().requires_grad_(True)
want = torch.cat([O.global_mean_pool(Xd, hb.batch, 512), O.global_max_pool(Xd, hb.batch, 512)], 1)
dy = torch.randn_like(want); want.backward(dy)
Xg = X.clone().requires_grad_(True)
got = lg.nn.global_mean_max_pool(Xg, hb.batch.to(dev)); got.backward(dy.float().to(dev))
print("pool fwd", f"{rel_err(got, want):.1e}", "bwd", f"{rel_err(Xg.grad, Xd.grad):.1e}")
# argmax comparison
xv = X.cpu().view(512, 100, 64)
am = xv.argmax(1)                                   # first max index per (graph, channel)
top2 = xv.topk(2, dim=1).values
gap = (top2[:, 0] - top2[:, 1])
print("exact ties among top-2:", int((gap == 0).sum()), "of", gap.numel(), " min gap", float(gap.min()))
tied = (gap == 0)
if tied.any():
    g, c = tied.nonzero()[0].tolist()
    col = xv[g, :, c]
    print("example tied channel values (top 5):", col.topk(5).values.tolist(), "count equal to max:", int((col == col.max()).sum()))
_, _, arg = ops.pool_fwd(X, lg.nn.Segments.from_batch(hb.batch.to(dev), 512).ptr, True, True, True)
arg = arg.cpu().long() - (torch.arange(512) * 100).unsqueeze(1)
print("argmax mismatches vs torch.argmax:", int((arg != am).sum()))
