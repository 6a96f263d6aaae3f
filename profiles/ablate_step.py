"""Where does the headline step's time go?  CUDA-graph replay time of the bench.py step with parts switched off
(no profiler: differences of whole-step times).  usage: python profiles/ablate_step.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import lesion_gnn_b200 as lg  # noqa: E402
from lesion_gnn_b200 import train  # noqa: E402

dev = torch.device("cuda:0")
host = [hb.pin() for hb in bench.make_host_batches(0, bench.POOL, with_edges_on=dev)]


def run(tag, tweak=None, loss="MSE", side=True, fwd_only=False):
    torch.manual_seed(1234)
    model = lg.models.GAT(bench.FEAT_IN, bench.HIDDEN, 1, heads=bench.HEADS, dropout=0.0).to(dev)
    if tweak:
        tweak(model)
    opt = train.FlatAdam(model, lr=1e-3)
    step = train.TrainStep(model, opt, loss, 5, use_cuda_graph=True)
    if not side:
        step._side = [torch.cuda.current_stream()]
        step._build = None
    if fwd_only:
        def body(db):
            with torch.no_grad():
                g = lg.CSRGraph.from_edge_index(db.edge_index, db.num_nodes, stream=step._build)
                seg = lg.nn.Segments.from_batch(db.batch, db.num_graphs)
                return model(db.x, g, seg).sum()
        step._body = body
    devb = [train.DeviceBatch(hb, dev) for hb in host]
    for hb, db in zip(host, devb):
        db.copy_from(hb)
    for i in range(2 * bench.POOL):
        step(devb[i % bench.POOL])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 300
    for i in range(n):
        step(devb[i % bench.POOL])
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({"variant": tag, "us_per_step": round(e0.elapsed_time(e1) / n * 1e3, 1)}), flush=True)


def freeze_in_proj(m):
    for p in m.in_proj.parameters():
        p.requires_grad_(False)


def freeze_all_but_out(m):
    for n, p in m.named_parameters():
        if not n.startswith("out_proj"):
            p.requires_grad_(False)


run("baseline")
run("in_proj frozen (no in_proj dW, db)", freeze_in_proj)
run("only out_proj trainable (no parameter-gradient kernels but out_proj's)", freeze_all_but_out)
run("forward only (CSR build + forward, no loss / backward / Adam)", fwd_only=True)
