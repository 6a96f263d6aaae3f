"""in_proj forward transform only (the command profiled with ncu).  usage: python profiles/linear_only.py [n] [k] [m]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lesion_gnn_b200 import ops  # noqa: E402

n, k, m = (int(a) for a in (sys.argv[1:4] + ["11583", "1025", "64"][len(sys.argv) - 1:]))
dev = torch.device("cuda:0")
x = torch.randn(n, k, device=dev)
w = torch.randn(m, k, device=dev) / k ** 0.5
b = torch.randn(m, device=dev)
dy = torch.randn(n, m, device=dev)
for _ in range(3):
    ops.linear_fwd(x, w, b)
    ops.linear_bwd_w(dy, x, True, True)
torch.cuda.synchronize()
print("ok")
