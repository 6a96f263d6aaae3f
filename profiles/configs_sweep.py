"""BASELINE.json configs 1, 3, 4, 5 on one B200: step time of the model of each config on the lgk layers
(CSR/CSC build + forward + loss + backward + Adam, CUDA-graph replay, inputs resident) and the graph-pipeline sweep
(kNN build, CSR+CSC build, mean+max pooling forward+backward) at 1e3..1e6 nodes.  The CPU oracle is timed beside the
two small configs.  One JSON line per measurement.  usage: python profiles/configs_sweep.py [--no-cpu]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import data, ops, train, transforms  # noqa: E402
from lesion_gnn_b200.nn.pool import Segments  # noqa: E402

dev = torch.device("cuda:0")
CPU = "--no-cpu" not in sys.argv


def device_batches(num_graphs, nodes, feat, k, count, layout="uniform"):
    out = []
    for i in range(count):
        hb = data.synth_batch(num_graphs, nodes, feat, seed=data.SEED + i, layout=layout)
        seg = Segments.from_batch(hb.batch.to(dev), hb.num_graphs)
        pos = hb.pos.to(dev) if hb.num_nodes < 500000 else hb.pos.float().to(dev)
        hb.edge_index = transforms.knn_graph(pos, k, seg, loop=True, num_edges=transforms.knn_num_edges(hb.counts, k, True)).cpu()
        out.append(hb)
    return out


def time_step(model, host, loss, steps):
    opt = train.FlatAdam(model, lr=1e-3)
    step = train.TrainStep(model, opt, loss, 5, use_cuda_graph=True)
    devb = [train.DeviceBatch(hb, dev) for hb in host]
    for hb, db in zip(host, devb):
        db.copy_from(hb)
    for i in range(2 * len(devb)):
        step(devb[i % len(devb)])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        step(devb[i % len(devb)])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def cpu_step(ref, hb, loss, iters=5):
    from oracle import pyg_semantics as O
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
    torch.set_num_threads(os.cpu_count() or 1)
    ts = []
    for i in range(iters + 2):
        t0 = time.perf_counter()
        O.training_step(ref, opt, hb.x, hb.edge_index, hb.batch, hb.y, hb.num_graphs, loss_type=loss)
        ts.append(time.perf_counter() - t0)
    return 1e3 * sorted(ts[2:])[len(ts[2:]) // 2]


def report(cfg, what, ms, graphs, hb, extra=None):
    d = {"config": cfg, "what": what, "ms_per_step": round(ms, 4), "graphs_per_s": round(graphs / ms * 1e3, 1),
         "graphs": graphs, "nodes": hb.num_nodes, "edges": int(hb.edge_index.size(1))}
    d.update(extra or {})
    print(json.dumps(d), flush=True)


torch.manual_seed(1234)
# config 1: 2-layer GCN, 32 graphs x 100 nodes, 16-d, k=8, CE
host = device_batches(32, 100, 16, 8, 4)
ms = time_step(lg.models.GCN(16, [64, 64], 5).to(dev), host, "CE", 500)
report(1, "GCN 16->64->64, B=32, n=100, k=8 (B200)", ms, 32, host[0])
if CPU:
    from oracle import pyg_semantics as O
    report(1, f"same, CPU oracle, {os.cpu_count()} threads", cpu_step(O.GCN(16, [64, 64], 5), host[0], "CE"), 32, host[0])

# config 3: GIN / GraphSAGE, mean+max read-out, 4096 graphs x 100 nodes, 64-d, k=8 (per-GPU share of the weak-scaling run)
host = device_batches(4096, 100, 64, 8, 2)
ms = time_step(lg.models.GIN(64, [64] * 4, 5, dropout=0.0, pooling="mean+max").to(dev), host, "CE", 50)
report(3, "GIN 64->[64]x4 (3 GINConv), mean||max, B=4096, n=100, k=8 (B200)", ms, 4096, host[0])
ms = time_step(lg.models.SAGE(64, [64] * 3, 5, aggr="mean", pooling="mean+max").to(dev), host, "CE", 50)
report(3, "SAGE 3 x SAGEConv(64,64,mean), mean||max, B=4096 (B200)", ms, 4096, host[0])
del host

# config 4: 2-layer GCN(128,128) on superpixel graphs of 10 000 nodes, k=16
for G in (1, 16, 128):
    host = device_batches(G, 10000, 128, 16, 1 if G == 128 else 2, layout="raster")
    ms = time_step(lg.models.GCN(128, [128, 128], 5).to(dev), host, "CE", 200 if G == 1 else 20)
    report(4, f"GCN 128->128->128, G={G} x 10 000 nodes, k=16 (B200)", ms, G, host[0])
    del host
    torch.cuda.empty_cache()

# config 5: graph pipeline sweep, graphs of 100 nodes, k=8, F=64
for N in (1000, 10000, 100000, 1000000):
    hb = data.synth_batch(N // 100, 100, 64, seed=data.SEED)
    pos = hb.pos.to(dev)
    batch = hb.batch.to(dev)
    seg = Segments.from_batch(batch, hb.num_graphs)
    E = transforms.knn_num_edges(hb.counts, 8, True)
    x = hb.x.to(dev).requires_grad_(True)

    def t(fn, iters=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters * 1e3

    us_knn = t(lambda: transforms.knn_graph(pos, 8, seg, loop=True, num_edges=E))
    ei = transforms.knn_graph(pos, 8, seg, loop=True, num_edges=E)
    us_csr = t(lambda: lg.CSRGraph.from_edge_index(ei, N))

    def pool_fb():
        x.grad = None
        lg.nn.global_mean_max_pool(x, seg).sum().backward()

    us_pool = t(pool_fb)
    print(json.dumps({"config": 5, "nodes": N, "edges": E, "knn_us": round(us_knn, 1), "knn_Medges_per_s": round(E / us_knn, 1),
                      "csr_csc_us": round(us_csr, 1), "csr_GBs": round((28 * E + 8 * (N + 1)) / us_csr / 1e3, 1),
                      "mean_max_pool_fwd_bwd_us": round(us_pool, 1), "pool_Mnodes_per_s": round(N / us_pool, 1)}), flush=True)
