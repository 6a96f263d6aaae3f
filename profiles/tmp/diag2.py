import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import lesion_gnn_b200 as lg
from oracle import pyg_semantics as O
from tests.util import knn_batch, rel_err
dev = torch.device("cuda:0")
hb = knn_batch(512, 100, 64, 8, seed=5)
for train_mode in (False, True):
    torch.manual_seed(11)
    ref = O.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max").double().train(train_mode)
    mod = lg.models.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max")
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()}); mod = mod.to(dev).train(train_mode)
    acts_r, acts_g = {}, {}
    def hook(store, name):
        def f(m, inp, out):
            store[name] = out
            if out.requires_grad: out.retain_grad()
        return f
    for name, m in ref.named_modules():
        if name and name.count(".") <= 3: m.register_forward_hook(hook(acts_r, name))
    for name, m in mod.named_modules():
        if name and name.count(".") <= 3: m.register_forward_hook(hook(acts_g, name))
    xr = hb.x.double().requires_grad_(True)
    out_r = ref(xr, hb.edge_index, hb.batch, num_graphs=hb.num_graphs)
    torch.nn.functional.cross_entropy(out_r, hb.y).backward()
    xg = hb.x.clone().to(dev).requires_grad_(True)
    out = mod(xg, hb.edge_index.to(dev), hb.batch.to(dev))
    lg.train.loss_fn(out, hb.y.to(dev), "CE", 5).backward()
    print("train", train_mode, "logits", f"{rel_err(out, out_r):.1e}", "dx", f"{rel_err(xg.grad, xr.grad):.1e}")
    for name in acts_r:
        if name in acts_g and isinstance(acts_r[name], torch.Tensor) and acts_g[name].shape == acts_r[name].shape:
            a, b = acts_g[name], acts_r[name]
            ge = f"{rel_err(a.grad, b.grad):.1e}" if (a.grad is not None and b.grad is not None) else "-"
            print(f"  {name:28s} act {rel_err(a, b):.1e}  grad {ge}  |act|max {float(b.abs().max()):.2e}")
    for (k, p), (_, pr) in zip(mod.named_parameters(), ref.named_parameters()):
        if pr.grad is not None and float(pr.grad.abs().max()) > 0: print(f"  d{k:30s} {rel_err(p.grad, pr.grad):.1e}")
