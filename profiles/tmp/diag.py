import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import lesion_gnn_b200 as lg
from lesion_gnn_b200 import ops
from lesion_gnn_b200.nn import functional as F
from oracle import pyg_semantics as O
from tests.util import knn_batch, rel_err, random_graph
dev = torch.device("cuda:0")

def layer_check(name, mk_gpu, mk_ref, x, ei):
    torch.manual_seed(7)
    ref = mk_ref().double(); mod = mk_gpu()
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}); mod = mod.to(dev)
    xr = x.double().requires_grad_(True); want = ref(xr, ei)
    dy = torch.randn(want.shape, dtype=torch.float64); want.backward(dy)
    xg = x.float().to(dev).requires_grad_(True); got = mod(xg, ei.to(dev)); got.backward(dy.float().to(dev))
    errs = {"out": rel_err(got, want), "dx": rel_err(xg.grad, xr.grad)}
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        if pr.grad is not None and float(pr.grad.abs().max()) > 0: errs["d" + k] = rel_err(p.grad, pr.grad)
    print(name, {k: f"{v:.1e}" for k, v in errs.items()}, flush=True)

hb = knn_batch(512, 100, 64, 8, seed=5)
x, ei = hb.x, hb.edge_index
layer_check("GINConv(Linear) n=51200", lambda: lg.nn.GINConv(lg.nn.Linear(64, 64)), lambda: O.GINConv(O.PygLinear(64, 64)), x, ei)
layer_check("GINConv(MLP norm=None)", lambda: lg.nn.GINConv(lg.nn.MLP([64, 64, 64], norm=None)), lambda: O.GINConv(O.MLP([64, 64, 64], norm=None)), x, ei)
layer_check("GINConv(MLP bn)", lambda: lg.nn.GINConv(lg.nn.MLP([64, 64, 64])), lambda: O.GINConv(O.MLP([64, 64, 64])), x, ei)
layer_check("SAGEConv", lambda: lg.nn.SAGEConv(64, 64), lambda: O.SAGEConv(64, 64), x, ei)
layer_check("GCNConv", lambda: lg.nn.GCNConv(64, 64), lambda: O.GCNConv(64, 64), x, ei)
# plain linear at that size
class LinR(torch.nn.Module):
    def __init__(s): super().__init__(); s.l = O.PygLinear(64, 64)
    def forward(s, x, ei): return s.l(x)
class LinG(torch.nn.Module):
    def __init__(s): super().__init__(); s.l = lg.nn.Linear(64, 64)
    def forward(s, x, ei): return s.l(x)
layer_check("Linear 64->64 n=51200", LinG, LinR, x, ei)
# aggregate SELF_ADD fwd/bwd
g = lg.CSRGraph.from_edge_index(ei.to(dev), hb.num_nodes)
xr = x.double().requires_grad_(True)
want = O.propagate(xr, ei, hb.num_nodes, "sum") + 1.25 * xr
dy = torch.randn_like(want); want.backward(dy)
xg = x.to(dev).requires_grad_(True)
got = F.aggregate(xg, g, "sum", self_fill=1.25, self_mode=ops.SELF_ADD); got.backward(dy.float().to(dev))
print("aggregate SELF_ADD", f"{rel_err(got, want):.1e} {rel_err(xg.grad, xr.grad):.1e}", flush=True)

# dropout FD check over shapes
def fd_check(n, E, H, C, clock, scale):
    ei = random_graph(n, E, 4)
    g = lg.CSRGraph.from_edge_index(ei.to(dev), n)
    torch.manual_seed(1)
    h = torch.randn(n, H * C, device=dev); att = torch.randn(1, H, C, device=dev) * scale
    kw = dict(dropout_p=0.35, seed=5, seed_dev=torch.tensor([7, 0], dtype=torch.int64, device=dev) if clock else None)
    r = torch.randn(n, H * C, device=dev)
    hh = h.clone().requires_grad_(True)
    F.gat_attention(hh, att, att, None, g, H, C, **kw).mul(r).sum().backward()
    d = torch.randn_like(h); hd = h.double()
    res = []
    for eps in (1e-2, 1e-3):
        fp = F.gat_attention((hd + eps * d).float(), att, att, None, g, H, C, **kw).double().mul(r).sum()
        fm = F.gat_attention((hd - eps * d).float(), att, att, None, g, H, C, **kw).double().mul(r).sum()
        res.append(float((fp - fm) / (2 * eps)))
    print(f"FD n={n} E={E} H={H} C={C} clock={clock} scale={scale}: fd={res} an={float((hh.grad.double() * d).sum()):.4f}", flush=True)
for args in [(2000, 30000, 2, 8, False, 1.0), (2000, 30000, 2, 8, True, 1.0), (1500, 12000, 4, 16, False, 1.0), (1500, 12000, 4, 16, True, 1.0),
             (1500, 12000, 4, 16, False, 0.2), (1500, 12000, 2, 8, False, 1.0), (2000, 30000, 4, 16, False, 1.0)]:
    fd_check(*args)
