"""Tuning sweep of the aggregation launch geometry on the SpMM stress shape (BASELINE.json configs[3]).

    python profiles/sweep_spmm.py [graphs]          # parent: spawns one child per configuration
Each child reads LGK_SPMM_{THREADS,SLICE4,WAVES,MAX_BAND} (read once by liblgk), times the kernels with CUDA events
and prints one JSON line.  The graph is built once and cached under /tmp.
"""
import itertools
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(graphs: int):
    import torch
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops, transforms
    from lesion_gnn_b200.nn.pool import Segments
    dev = torch.device("cuda:0")
    n, k, F = 10000, 16, 128
    cache = f"/tmp/stress_{graphs}.pt"
    if os.path.exists(cache):
        d = torch.load(cache)
        ei, x = d["ei"].to(dev), d["x"].to(dev)
        N = x.size(0)
    else:
        hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
        N = hb.num_nodes
        seg = Segments.from_batch(hb.batch.to(dev), graphs)
        ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
        x = hb.x.to(dev)
        torch.save({"ei": ei.cpu(), "x": hb.x}, cache)
    g = lg.CSRGraph.from_edge_index(ei, N)
    E = ei.size(1)
    dinv, _ = g.gcn_norm(False, True)

    def t(fn, iters=10):
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

    plain = 2 * N * F * 4 + 4 * E + 4 * (N + 1)
    out = {"cfg": {k_: os.environ.get(k_) for k_ in ("LGK_SPMM_THREADS", "LGK_SPMM_SLICE4", "LGK_SPMM_WAVES", "LGK_SPMM_MAX_BAND", "LGK_SPMM_CARVEOUT")}}
    us = t(lambda: ops.spmm(g.rowptr, g.col, x, "sum"))
    out["sum_us"], out["sum_GBs"] = round(us, 1), round(plain / us / 1e3, 1)
    us = t(lambda: ops.spmm(g.rowptr, g.col, x, "sum", pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE))
    out["gcn_us"], out["gcn_GBs"] = round(us, 1), round((plain + 8 * N) / us / 1e3, 1)
    us = t(lambda: ops.spmm(g.colptr, g.row, x, "sum", pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE))
    out["gcnT_us"] = round(us, 1)
    us = t(lambda: ops.spmm(g.rowptr, g.col, x, "max", want_arg=True))
    out["max_us"] = round(us, 1)
    xb = x.bfloat16()
    us = t(lambda: ops.spmm(g.rowptr, g.col, xb, "sum"))
    out["sum_bf16_us"] = round(us, 1)
    x64 = x[:, :64].contiguous()
    us = t(lambda: ops.spmm(g.rowptr, g.col, x64, "sum"))
    out["sum_f64wide_us"] = round(us, 1)
    print(json.dumps(out), flush=True)


def main():
    graphs = int(sys.argv[1]) if len(sys.argv) > 1 else 128
    grid = []
    for carve in (-1, 25, 40, 60):
        grid.append((512, 32, 16, 2044, carve))
    for threads, slice4, waves, max_band, pf in grid:
        env = dict(os.environ, LGK_SPMM_THREADS=str(threads), LGK_SPMM_SLICE4=str(slice4), LGK_SPMM_WAVES=str(waves),
                   LGK_SPMM_MAX_BAND=str(max_band), LGK_SPMM_CARVEOUT=str(pf))
        r = subprocess.run([sys.executable, __file__, "--child", str(graphs)], env=env, capture_output=True, text=True)
        line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "ERR " + r.stderr[-300:]
        print(line, flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        child(int(sys.argv[2]))
    else:
        main()
