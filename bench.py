#!/usr/bin/env python
"""bench.py -- graphs/sec fwd+bwd of the lesion-gnn message-passing hot path on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...   # the CPU oracle (PyG restatement) on the host cores

Workload (BASELINE.json configs[1], SURVEY.md 8d cfg 2): `GAT(1025, [64]*4, heads=4)` -- in_proj, 3 x (GATConv +
ELU), global_mean_pool, out_proj, MSE on the clamped scalar (reference models/gat.py:17-59,94-95) -- on 256
DDR-shaped synthetic lesion graphs per GPU (n ~ ceil(LogNormal(ln 30, 0.9)) in [1, 600], kNN k=6 loop=True), one
step = CSR/CSC build + forward + loss + backward + Adam (+ one NCCL all-reduce of the flat gradient bucket when
N > 1; weak scaling: 256 graphs per GPU).  A pool of distinct batches larger than L2 is rotated so no step
finds its inputs in cache.

One JSON line on stdout (rank 0).  `value`: inputs resident in HBM.  `e2e`: the same step driven from pinned HOST
buffers (x, edge_index, batch, y copied host->device inside the timed region, the loss read back).
`roofline`: the dominant kernel of the step, CUDA-event timed per launch in an instrumented pass of the same
steps.  `aggregation`: the fused aggregation kernels on the SpMM stress shape (BASELINE.json configs[3]).
`cpu_baseline`: the oracle on this box's host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOAD = "gat3x64h4_ddr_b256_k6"
GRAPHS_PER_GPU = 256
FEAT_IN, HIDDEN, HEADS, KNN = 1025, [64] * 4, 4, 6
POOL = 6  # distinct batches rotated: 6 x ~47 MB of node features > 126 MB L2
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the aggregation stress and cpu_baseline legs")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline sample")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return float(p["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def measured_traffic(name: str, tag: str):
    """ncu `dram__bytes_read.sum + dram__bytes_write.sum` per launch for an entry point, from the newest
    profiles/r*/traffic.json that has it (one `ncu --set full` capture per round); None if never captured"""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*", "traffic.json")), reverse=True):
        try:
            hit = json.load(open(path)).get(f"{name}|{tag}")
        except (OSError, ValueError):
            continue
        if hit:
            return int(hit["bytes"])
    return None


def make_host_batches(rank: int, count: int, with_edges_on=None):
    """`count` distinct synthetic batches for this rank; kNN edges are built on `with_edges_on` (a CUDA device,
    lgk_knn_graph) or by the oracle on the CPU (reference arm) -- graph construction is outside the timed step,
    as in the reference where it happens in the DataLoader (datasets/datamodule.py:43-48)."""
    from lesion_gnn_b200 import data
    out = []
    for i in range(count):
        hb = data.synth_batch(GRAPHS_PER_GPU, "ddr", FEAT_IN, seed=data.SEED + 1000 * rank + i)
        if with_edges_on is not None:
            from lesion_gnn_b200 import transforms
            from lesion_gnn_b200.nn.pool import Segments
            seg = Segments.from_batch(hb.batch.to(with_edges_on), hb.num_graphs)
            ei = transforms.knn_graph(hb.pos.to(with_edges_on), KNN, seg, loop=True,
                                      num_edges=transforms.knn_num_edges(hb.counts, KNN, True))
            hb.edge_index = ei.cpu()
        else:
            from oracle import pyg_semantics as O
            hb.edge_index = O.knn_graph(hb.pos, KNN, hb.batch, loop=True)
        out.append(hb)
    return out


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region through NVML (the same counters as the
    nvidia-smi clocks line of B200_PROFILING.md, polled every 5 ms so that sub-second runs still get samples)."""

    REASONS = (("hw_slowdown", 0x8), ("sw_power_cap", 0x4), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
               ("hw_power_brake", 0x80))

    def __init__(self, index: int):
        self.sm, self.bits, self.smax, self.err = [], 0, None, None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _loop(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                self.bits |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception as e:  # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(0.005)

    def start(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=1.0)
        sm = sorted(self.sm)
        out = {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.smax,
               "reasons": [n for n, b in self.REASONS if self.bits & b], "samples": len(sm)}
        if self.err:
            out["error"] = self.err
        return out


# ------------------------------------------------------------------------------------------------------
# algorithmic (compulsory) bytes per entry point: each distinct array touched once; fp32 = 4 B,
# int32 index = 4 B, int64 edge_index = 8 B.  E is supplied by the caller (it lives on the device).
# ------------------------------------------------------------------------------------------------------
def algorithmic_bytes(name: str, a: dict, num_edges: int) -> int:
    E = num_edges
    if name == "lgk_linear_fwd":
        return 4 * (a["n"] * a["k"] + a["n"] * a["m"] * (2 if a["accumulate"] else 1) + a["k"] * a["m"])
    if name in ("lgk_linear_bwd_x", "lgk_linear_bwd_w"):
        return 4 * (a["n"] * a["k"] + a["n"] * a["m"] + a["k"] * a["m"])
    if name == "lgk_spmm":
        n, f = a["n_rows"], a["feat"]
        extra = (4 * E if a["ew"] else 0) + (4 * E if a["emap"] else 0) + (4 * n if a["pre"] else 0) + (4 * n if a["post"] else 0)
        extra += 4 * n * f if a["arg_out"] else 0
        return 2 * n * f * 4 + 4 * E + 4 * (n + 1) + extra
    if name == "lgk_spmm_max_bwd":
        n, f = a["n_rows"], a["feat"]
        return 3 * n * f * 4 + 8 * E + 4 * (n + 1)
    if name in ("lgk_gat_fwd", "lgk_gat_bwd_dst", "lgk_gat_bwd_src", "lgk_gat_alpha", "lgk_gat_datt"):
        n, h, c = a["num_nodes"], a["heads"], a["chan"]
        nf, nh = n * h * c, n * h
        idx = 4 * E + 4 * (n + 1)
        return {"lgk_gat_fwd": 4 * (2 * nf + 4 * nh) + idx, "lgk_gat_bwd_dst": 4 * (3 * nf + 9 * nh) + idx,
                "lgk_gat_bwd_src": 4 * (3 * nf + 7 * nh) + idx, "lgk_gat_alpha": 4 * (nf + 2 * nh),
                "lgk_gat_datt": 4 * (nf + 2 * nh)}[name]
    if name in ("lgk_pool_fwd", "lgk_pool_bwd"):
        n, b, f = a["num_nodes"], a["num_graphs"], a["feat"]
        return 4 * (n * f + b * f) + 4 * (b + 1)
    if name == "lgk_elu_fwd":
        return 8 * a["n"]
    if name == "lgk_elu_bwd":
        return 12 * a["n"]
    if name == "lgk_csr_build":
        return 28 * a["num_edges"] + 8 * (a["num_nodes"] + 1)
    if name == "lgk_batch_ptr":
        return 8 * a["num_nodes"] + 4 * (a["num_graphs"] + 1)
    if name == "lgk_adam_step":
        return 28 * a["n"]
    if name == "lgk_gather_f32":
        return 12 * a["n"]
    return 0


class KernelTimer:
    """_lib hook: brackets every C-ABI call with CUDA events on the launching (current) stream."""

    def __init__(self, lib):
        self.lib, self.records, self.num_edges = lib, [], 0
        self._open = None

    def __call__(self, phase, name, args):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream())
        if phase == "begin":
            self._open = ev
        else:
            names = self.lib.protos[name][2]
            vals = {k: (v if isinstance(v, (int, float)) else (0 if v is None else 1)) for k, v in zip(names, args)}
            self.records.append((name, self._open, ev, algorithmic_bytes(name, vals, self.num_edges), _shape_tag(name, vals)))

    def summary(self) -> dict:
        torch.cuda.synchronize()
        agg = {}
        for name, e0, e1, nbytes, tag in self.records:
            d = agg.setdefault((name, tag), {"ms": 0.0, "bytes": 0, "calls": 0})
            d["ms"] += e0.elapsed_time(e1)
            d["bytes"] += nbytes
            d["calls"] += 1
        return agg


def _shape_tag(name: str, a: dict) -> str:
    # row counts (n, N, E) vary from batch to batch of the pool: group launches by the per-row shape only
    keys = [k for k in ("k", "m", "feat", "heads", "chan", "reduce") if k in a]
    if name == "lgk_linear_bwd_w" and not a.get("dw"):
        keys = ["m"]  # bias-gradient column sum only
    if name in ("lgk_linear_fwd", "lgk_linear_bwd_x", "lgk_linear_bwd_w") and a["n"] <= GRAPHS_PER_GPU:
        keys.append("n")  # per-graph (out_proj) products
    return ",".join(f"{k}={a[k]}" for k in keys)


# ------------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch.distributed as dist

    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import _lib, ops, train

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: lesion_gnn_b200 has no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    torch.manual_seed(1234)
    model = lg.models.GAT(FEAT_IN, HIDDEN, 1, heads=HEADS, dropout=0.0).to(dev)
    if world > 1:  # identical replicas
        for p in model.parameters():
            dist.broadcast(p.data, 0)
    opt = train.FlatAdam(model, lr=1e-3)
    step = train.TrainStep(model, opt, "MSE", 5, use_cuda_graph=True, process_group=None, world_size=world)

    host = [hb.pin() for hb in make_host_batches(rank, POOL, with_edges_on=dev)]
    devb = [train.DeviceBatch(hb, dev) for hb in host]
    for hb, db in zip(host, devb):
        db.copy_from(hb)
    total_graphs = GRAPHS_PER_GPU * world
    for db in devb:
        db.loss_weight = train.dp_loss_weight(db.num_graphs, total_graphs, world)
    resident = sum(hb.h2d_bytes() for hb in host)
    assert resident > L2_BYTES, "batch pool must exceed L2"

    # capture + warm-up (W untimed steps over the pool, at least one full rotation so every graph is captured)
    warm = max(args.warmup, POOL)
    for i in range(warm):
        step(devb[i % POOL])
    barrier()

    # ---- value: inputs resident in HBM --------------------------------------------------------------
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        step(devb[i % POOL])
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms)
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: host buffers -> device inside the timed region, loss read back every step -----------------------
    # Double buffered: while step i computes, the pinned-host -> device copy of step i+1's inputs runs on a copy
    # stream; every step's inputs are copied inside the timed region and every step's loss is read on the host.
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event() for _ in range(POOL)]
    consumed = [torch.cuda.Event() for _ in range(POOL)]

    def prefetch(i):
        s = i % POOL
        copy_stream.wait_event(consumed[s])  # the previous user of this device buffer has finished
        with torch.cuda.stream(copy_stream):
            devb[s].copy_from(host[s])
            ready[s].record(copy_stream)

    def e2e_steps(count):
        main = torch.cuda.current_stream()
        for s in range(POOL):
            consumed[s].record(main)
        prefetch(0)
        loss = 0.0
        for i in range(count):
            if i + 1 < count:
                prefetch(i + 1)
            main.wait_event(ready[i % POOL])
            out_dev = step(devb[i % POOL])
            consumed[i % POOL].record(main)
            loss = float(out_dev)  # device->host read of this step's loss (synchronises the host with the step)
        return loss

    e2e_steps(POOL)
    barrier()
    t0 = time.perf_counter()
    e0.record()
    loss_host = e2e_steps(args.steps)
    e1.record()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = torch.tensor([max(e0.elapsed_time(e1), wall_ms)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    e2e_ms_total = float(ms)
    h2d = sum(host[i % POOL].h2d_bytes() for i in range(args.steps)) / args.steps

    # ---- instrumented pass: per entry point CUDA-event times on the same steps (eager, same kernels) -----------
    eager = train.TrainStep(model, opt, "MSE", 5, use_cuda_graph=False, world_size=world)
    eager._side = eager._build = None  # one stream: every launch is timed alone, not beside the side-stream kernels
    L = _lib.lib()
    eager(devb[0])
    ops.reset_launch_count()
    eager(devb[0])
    launches_per_step = ops.launch_count()
    timer = KernelTimer(L)
    L.hook = timer
    reps = 2
    for r in range(reps):
        for i in range(POOL):
            timer.num_edges = host[i].edge_index.size(1)
            torch.cuda._sleep(20_000_000)  # ~10 ms of GPU idle spin: the host runs ahead, so the events bracket
            eager(devb[i])                 # back-to-back device execution instead of Python launch gaps
            torch.cuda.synchronize()
    L.hook = None
    agg = timer.summary()
    kern_total = sum(d["ms"] for d in agg.values())
    top = sorted(agg.items(), key=lambda kv: -kv[1]["ms"])
    peak, peak_src = peaks()
    breakdown = [{"kernel": n, "shape": tag, "calls_per_step": d["calls"] / (reps * POOL), "us_per_call": 1e3 * d["ms"] / d["calls"],
                  "share": d["ms"] / kern_total, "GBps": (d["bytes"] / d["calls"]) / (d["ms"] / d["calls"] * 1e-3) / 1e9 if d["bytes"] else None}
                 for (n, tag), d in top[:12]]
    (dom_name, dom_tag), dom = top[0]
    achieved = (dom["bytes"] / dom["calls"]) / (dom["ms"] / dom["calls"] * 1e-3) / 1e9
    roofline = {"kernel": dom_name, "shape": dom_tag, "bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                "peak_source": peak_src, "unit": "GB/s", "frac": round(achieved / peak, 4),
                "traffic": measured_traffic(dom_name, dom_tag),
                "bytes_per_launch": dom["bytes"] // dom["calls"], "us_per_launch": round(1e3 * dom["ms"] / dom["calls"], 2),
                "share_of_step_kernel_time": round(dom["ms"] / kern_total, 3)}

    out = None
    if rank == 0:
        ms_step = ms_total / args.steps
        out = {
            "metric": "graphs/sec fwd+bwd", "value": round(total_graphs * args.steps / (ms_total * 1e-3), 1), "unit": "graphs/s",
            "n_gpus": world, "steps": args.steps, "warmup": warm, "ms_per_step": round(ms_step, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "graphs_per_gpu": GRAPHS_PER_GPU, "global_graphs": total_graphs,
                       "nodes_per_step_rank0": sum(h.num_nodes for h in host) / POOL,
                       "edges_per_step_rank0": sum(h.edge_index.size(1) for h in host) / POOL,
                       "feat_in": FEAT_IN, "hidden": HIDDEN, "heads": HEADS, "knn_k": KNN, "loss": "MSE", "optimizer": "Adam",
                       "parallelism": f"dp{world}", "cuda_graph": True,
                       "l2_policy": f"rotating pool of {POOL} distinct batches, {resident / 2**20:.0f} MiB > L2 (126 MiB)"},
            "e2e": {"value": round(total_graphs * args.steps / (e2e_ms_total * 1e-3), 1), "unit": "graphs/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "ms_per_step": round(e2e_ms_total / args.steps, 4),
                    "last_loss": loss_host},
            "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
            "clocks": clocks, "roofline": roofline, "kernels": breakdown,
        }
    if not args.no_extras and rank == 0:
        out["aggregation"] = aggregation_stress(dev, peak, peak_src)
        out["cpu_baseline"] = cpu_baseline(args.cpu_seconds)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)


def _time_cuda(fn, iters: int, warmup: int = 3) -> float:
    """mean milliseconds per call, CUDA events on the current stream"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def aggregation_stress(dev, peak: float, peak_src: str, graphs: int = 128) -> dict:
    """BASELINE.json configs[3] / SURVEY.md 8d cfg 4: `graphs` superpixel graphs x 10 000 nodes, 128-d, kNN k=16
    (loop=True), GCN-normalised aggregation forward (CSR) and backward (CSC), plus mean pooling.  X = 655 MB at
    128 graphs, far beyond L2, so every launch streams from HBM."""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops, transforms
    from lesion_gnn_b200.nn.pool import Segments
    n, k, F = 10000, 16, 128
    hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
    N = hb.num_nodes
    seg = Segments.from_batch(hb.batch.to(dev), graphs)
    ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
    E = ei.size(1)
    t_csr = _time_cuda(lambda: lg.CSRGraph.from_edge_index(ei, N), 3, 1)
    g = lg.CSRGraph.from_edge_index(ei, N)
    x = hb.x.to(dev)
    dinv, _ = g.gcn_norm(False, True)
    kw = dict(pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
    res = {"shape": {"graphs": graphs, "nodes": N, "edges": E, "feat": F, "k": k}, "peak": peak, "peak_source": peak_src, "unit": "GB/s"}
    agg_bytes = 2 * N * F * 4 + 4 * E + 4 * (N + 1) + 8 * N

    def entry(ms, nbytes):
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"us": round(ms * 1e3, 1), "bytes": nbytes, "achieved": round(gbs, 1), "frac": round(gbs / peak, 4)}

    res["gcn_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", **kw), 20), agg_bytes)
    res["gcn_aggregate_bwd"] = entry(_time_cuda(lambda: ops.spmm(g.colptr, g.row, x, "sum", **kw), 20), agg_bytes)
    plain = 2 * N * F * 4 + 4 * E + 4 * (N + 1)
    res["sum_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum"), 20), plain)
    res["mean_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "mean"), 20), plain)
    res["max_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "max", want_arg=True), 20), plain + 4 * N * F)
    H, C = 4, 32
    att = torch.randn(H * C, device=dev) * 0.1
    a_s, a_d = ops.gat_alpha(x, att, att, H, C)
    gat_bytes = 4 * (2 * N * F + 4 * N * H) + 4 * E + 4 * (N + 1)
    res["gat_softmax_aggregate_fwd"] = entry(_time_cuda(lambda: ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True), 20), gat_bytes)
    pool_bytes = 4 * (N * F + graphs * F) + 4 * (graphs + 1)
    res["global_mean_pool_fwd"] = entry(_time_cuda(lambda: ops.pool_fwd(x, seg.ptr, True, True, False), 20), pool_bytes)
    res["global_mean_max_pool_fwd"] = entry(_time_cuda(lambda: ops.pool_fwd(x, seg.ptr, True, True, True), 20), pool_bytes + 8 * graphs * F)
    res["csr_csc_build"] = entry(t_csr, 28 * E + 8 * (N + 1))
    del x, g, ei
    torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------------------------------
# CPU legs: the oracle (op-for-op PyG restatement, oracle/pyg_semantics.py) on the host cores
# ------------------------------------------------------------------------------------------------------
def _oracle_stepper():
    from oracle import pyg_semantics as O
    torch.set_num_threads(os.cpu_count() or 1)
    torch.manual_seed(1234)
    model = O.GAT(FEAT_IN, HIDDEN, 1, heads=HEADS, dropout=0.0)
    optim = torch.optim.Adam(model.parameters(), lr=1e-3)
    batches = make_host_batches(0, 2)

    def one(i):
        hb = batches[i % len(batches)]
        return O.training_step(model, optim, hb.x, hb.edge_index, hb.batch, hb.y, hb.num_graphs, "MSE", 5)
    return one


def cpu_baseline(budget_s: float) -> dict:
    one = _oracle_stepper()
    one(0)
    t0 = time.perf_counter()
    n = 0
    while True:
        one(n)
        n += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or n >= 200:
            break
    return {"value": round(GRAPHS_PER_GPU * n / dt, 1), "unit": "graphs/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} full steps of {WORKLOAD} (256 graphs each, fwd+loss+bwd+Adam, fp32) in {dt:.1f} s after 1 warm-up",
            "ms_per_step": round(1e3 * dt / n, 2)}


def run_reference(args):
    """The reference's CPU path for this workload = the oracle (PyG is not installable here: SURVEY.md 8c), all
    host threads.  Under torchrun only rank 0 works."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    one = _oracle_stepper()
    warm = max(args.warmup, 1)
    for i in range(warm):
        one(i)
    steps = min(args.steps, 100)
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    dt = time.perf_counter() - t0
    v = round(GRAPHS_PER_GPU * steps / dt, 1)
    cores = torch.get_num_threads()
    print(json.dumps({
        "impl": "reference", "metric": "graphs/sec fwd+bwd", "value": v, "unit": "graphs/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": round(1e3 * dt / steps, 3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "graphs_per_step": GRAPHS_PER_GPU, "feat_in": FEAT_IN, "hidden": HIDDEN, "heads": HEADS,
                   "knn_k": KNN, "loss": "MSE", "optimizer": "Adam", "device": "cpu"},
        "cpu_baseline": {"value": v, "unit": "graphs/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} full steps of 256 graphs on {cores} host threads"},
        "e2e": {"value": v, "unit": "graphs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def _keep_stdout_for_the_json_line():
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner on the
    first collective when NCCL_DEBUG is set on the box), so file descriptor 1 is pointed at stderr for the whole run
    and `print` gets a private duplicate of the real stdout."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


if __name__ == "__main__":
    a = parse_args()
    _keep_stdout_for_the_json_line()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
