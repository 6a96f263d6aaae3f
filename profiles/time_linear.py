"""Times the node-feature transform entry points (forward, dX, dW) on the headline shapes, GPU-side (CUDA graph of
launches over rotating inputs larger than L2).  usage: python profiles/time_linear.py   (LGK_LINEAR_TC=0: FFMA)"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lesion_gnn_b200 import ops  # noqa: E402
from profiles.gputime import graph_time_us  # noqa: E402

dev = torch.device("cuda:0")
shapes = [(11583, 1025, 64), (11583, 64, 64), (409600, 64, 64), (1280000, 128, 128), (100000, 1025, 128)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
for (n, k, m) in shapes:
    nv = max(2, min(8, int(300e6 / (4 * n * (k + m))) + 1))
    nv = int(os.environ.get("NV", nv))   # NV=1: inputs stay in L2 (is DRAM the limiter?)
    xs = [torch.randn(n, k, device=dev) for _ in range(nv)]
    dys = [torch.randn(n, m, device=dev) for _ in range(nv)]
    outs = [torch.empty(n, m, device=dev) for _ in range(nv)]
    w = torch.randn(m, k, device=dev) / k ** 0.5
    b = torch.randn(m, device=dev)
    ref = (xs[0].double() @ w.double().t() + b.double())
    out = ops.linear_fwd(xs[0], w, b)
    err = float((out.double() - ref).abs().max() / ref.abs().max())
    del ref
    res = {"n": n, "k": k, "m": m, "tc": os.environ.get("LGK_LINEAR_TC", "1"), "debug": os.environ.get("LGK_TC_DEBUG"), "fwd_rel_err": err}
    nbytes = 4 * (n * k + n * m + k * m)
    us = graph_time_us(lambda i: (lambda: ops.linear_fwd(xs[i], w, b)), nv)
    res["fwd_us"], res["fwd_GBs"] = round(us, 1), round(nbytes / us / 1e3, 1)
    if not os.environ.get("LGK_TC_DEBUG"):
        us = graph_time_us(lambda i: (lambda: ops.linear_bwd_x(dys[i], w)), nv)
        res["bwd_x_us"] = round(us, 1)
        us = graph_time_us(lambda i: (lambda: ops.linear_bwd_w(dys[i], xs[i], True, True)), nv)
        res["bwd_w_us"], res["bwd_w_GBs"] = round(us, 1), round(nbytes / us / 1e3, 1)
    print(json.dumps(res), flush=True)
    del xs, dys, outs
