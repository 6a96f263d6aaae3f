"""Aggregation kernels on the SpMM stress shape (BASELINE.json configs[3]: `graphs` superpixel graphs x 10 000
nodes, 128-d, kNN k=16, raster order): lgk_spmm (register-landing L1 gathers) against lgk_spmm_blocked (row-blocked
CSR + TMA-staged source window), forward (CSR) and backward (CSC), plain sum and GCN-normalised, with a parity
check between the two.  One JSON line per measurement.
usage: python profiles/agg_bench.py [graphs=128] [iters=20] [feat=128]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 128
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
F = int(sys.argv[3]) if len(sys.argv) > 3 else 128
peak = 6551.0
if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")):
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
ops._WIN_MIN_ELEMS = 0
dev = torch.device("cuda:0")
n, k = 10000, 16
hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
N = hb.num_nodes
seg = Segments.from_batch(hb.batch.to(dev), graphs)
ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
E = ei.size(1)
g = lg.CSRGraph.from_edge_index(ei, N)
x = hb.x.to(dev)
dinv, _ = g.gcn_norm(False, True)


def timeit(fn, iters=iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


t_q = timeit(lambda: ops.quad_build(g.rowptr, g.col), 5, 1)
qc, qr = g.quads("csr"), g.quads("csc")
torch.cuda.synchronize()
merged = int(qc.qlen.view(-1, 2)[:, 0].sum().item())
merged_t = int(qr.qlen.view(-1, 2)[:, 0].sum().item())
band = (qc.cmax - qc.cmin + 1)[: (N // 4 + 15) // 16]
print(json.dumps({"what": "quad_build", "us": round(t_q * 1e3, 1), "N": N, "E": E, "merged_csr": merged, "merged_csc": merged_t,
                  "gathers_saved": round(E / merged, 3), "chunk_window_rows_max": int(band.max().item()),
                  "chunk_window_rows_median": int(band.float().median().item()), "qmeta": qc.qmeta.tolist(),
                  "slice": os.environ.get("LGK_WIN_SLICE", "auto")}), flush=True)

plain = 2 * N * F * 4 + 4 * E + 4 * (N + 1)
gcn = plain + 8 * N
kw = dict(pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
cases = [
    ("sum_fwd", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc), plain),
    ("gcn_fwd", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc, **kw), gcn),
    ("gcn_bwd", lambda q: ops.spmm(g.colptr, g.row, x, "sum", quads=q and qr, **kw), gcn),
    ("mean_fwd", lambda q: ops.spmm(g.rowptr, g.col, x, "mean", quads=q and qc), plain),
    ("sum_pre_only", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc, pre=dinv), plain + 4 * N),
    ("sum_post_only", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc, post=dinv), plain + 4 * N),
    ("sum_selfreplace_only", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc, self_mode=ops.SELF_REPLACE), plain),
    ("sum_selfadd_only", lambda q: ops.spmm(g.rowptr, g.col, x, "sum", quads=q and qc, self_mode=ops.SELF_ADD, self_fill=1.5), plain),
]
if os.environ.get("AGG_ONLY_BLOCKED"):
    for name, fn, nbytes in cases:
        ms = timeit(lambda: fn(True))
        print(json.dumps({"what": name, "kernel": "lgk_spmm_blocked", "us": round(ms * 1e3, 1), "frac": round(nbytes / (ms * 1e-3) / 1e9 / peak, 4)}), flush=True)
    sys.exit(0)
for name, fn, nbytes in cases:
    a = fn(None)
    b = fn(True)
    err = float((a - b).abs().max() / a.abs().max())
    for label, q in (("lgk_spmm", None), ("lgk_spmm_blocked", True)):
        ms = timeit(lambda: fn(q))
        gbs = nbytes / (ms * 1e-3) / 1e9
        print(json.dumps({"what": name, "kernel": label, "us": round(ms * 1e3, 1), "GBps": round(gbs, 1), "frac": round(gbs / peak, 4),
                          "graphs": graphs, "feat": F, "max_rel_diff_between_kernels": err}), flush=True)
