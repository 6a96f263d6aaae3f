import sys, torch
sys.path.insert(0, '.')
from lesion_gnn_b200 import ops
dev = torch.device('cuda')
n, k, H, C = 17200, 64, 4, 16
x = torch.randn(n, k, device=dev); w = torch.randn(H * C, k, device=dev) / 8
a_s = torch.randn(H * C, device=dev); a_d = torch.randn(H * C, device=dev)
flush = torch.empty(64 << 20, device=dev)
def t(fn, reps=50, graph=True):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000 / (reps * 20)
def sep():
    h = ops.linear_fwd(x, w)
    return ops.gat_alpha(h, a_s, a_d, H, C)
print('linear_fwd only        us', round(t(lambda: ops.linear_fwd(x, w)), 2))
print('linear_fwd + gat_alpha us', round(t(sep), 2))
print('linear_gat_fwd (fused) us', round(t(lambda: ops.linear_gat_fwd(x, w, a_s, a_d, H, C)), 2))
