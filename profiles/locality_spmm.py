"""How much of the aggregation time is neighbour locality?  Same N, F, degree; four neighbour patterns:
window (j = i-8..i+7: every gather an L1 hit), raster kNN (the stress shape), random inside 10k-node graphs
(L2 hits), random over all nodes (HBM).  usage: python profiles/locality_spmm.py [nodes] [feat]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import ops  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1280000
F = int(sys.argv[2]) if len(sys.argv) > 2 else 128
K = 16
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
dst = torch.arange(N, device=dev).repeat_interleave(K)
x = torch.randn(N, F, device=dev)


def timeit(fn, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


pats = {}
off = torch.arange(K, device=dev) - K // 2
pats["window16"] = (torch.arange(N, device=dev)[:, None] + off[None, :]).clamp(0, N - 1).reshape(-1)
blk = (torch.arange(N, device=dev) // 10000 * 10000)[:, None]
pats["random_in_10k_graph"] = (blk + torch.randint(0, 10000, (N, K), device=dev, generator=g)).clamp(max=N - 1).reshape(-1)
pats["random_all"] = torch.randint(0, N, (N * K,), device=dev, generator=g)
plain = 2 * N * F * 4 + 4 * N * K + 4 * (N + 1)
for name, src in pats.items():
    gr = lg.CSRGraph.from_edge_index(torch.stack([src, dst]), N)
    us = timeit(lambda: ops.spmm(gr.rowptr, gr.col, x, "sum"))
    xb = x.bfloat16()
    usb = timeit(lambda: ops.spmm(gr.rowptr, gr.col, xb, "sum"))
    print(json.dumps({"pattern": name, "N": N, "F": F, "sum_us": round(us, 1), "GBs": round(plain / us / 1e3, 1),
                      "clk_per_gather_per_sm": round(us * 1965 / (N * K / 148), 2), "sum_bf16_us": round(usb, 1)}), flush=True)
    del gr
