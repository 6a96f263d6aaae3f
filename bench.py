#!/usr/bin/env python
"""bench.py -- graphs/sec fwd+bwd of the lesion-gnn message-passing hot path on N B200s (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W             # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...    # the CPU oracle (PyG restatement) on the host cores
    python bench.py --workload gin_b4096 --scaling strong ...  # the other BASELINE.json configs (see WORKLOADS)

Default workload (BASELINE.json configs[1], SURVEY.md 8d cfg 2): `GAT(1025, [64]*4, heads=4, dropout=0.35)` --
in_proj, 3 x (GATConv + ELU), global_mean_pool, out_proj, MSE on the clamped scalar (reference models/gat.py:17-59,
94-95), Adam with weight_decay 2e-6 (configs/config.py:53-63), training mode -- on 256 DDR-shaped synthetic lesion
graphs per GPU (n ~ ceil(LogNormal(ln 30, 0.9)) in [1, 600], kNN k=6 loop=True).  One step = CSR/CSC build + forward +
loss + backward + optimiser (+ the gradient all-reduce when N > 1; weak scaling: 256 graphs per GPU).  A pool of
distinct batches larger than L2 is rotated so no step finds its inputs in cache.

One JSON line on stdout (rank 0).  `value`: inputs resident in HBM.  `e2e`: the same step driven from pinned HOST
buffers (x, edge_index, batch, y copied host->device inside the timed region, the loss read back every step).
`roofline`: the dominant entry point of the step, CUDA-event timed per launch in an instrumented pass of the same
steps.  `aggregation`: the fused aggregation kernels on the SpMM stress shape (BASELINE.json configs[3]).
`cpu_baseline` (N = 1 only): the oracle on this box's host cores on a bounded sample of the same workload.
Extra keys (never replacing the above): `parity_config` (the same model with dropout 0 / weight decay 0, the
configuration the parity tests check), `e2e_bf16` (features cross PCIe as bf16), `e2e_resident_dataset` (the dataset
lives in HBM; per step: gather by graph index + kNN + CSR on the device), `collective_us` (the gradient all-reduce alone).
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

L2_BYTES = 126 * 1024 * 1024

# name -> model / data of one data-parallel rank.  `graphs` is per GPU under weak scaling and global under strong.
WORKLOADS = {
    # BASELINE.json configs[1] -- the headline
    "gat3x64h4_ddr_b256_k6": dict(model="gat", feat=1025, hidden=[64] * 4, heads=4, out=1, graphs=256, nodes="ddr", k=6,
                                  layout="uniform", loss="MSE", dropout=0.35, weight_decay=2e-6, pool=6),
    # the reference's own operating point (configs/config.py:49,53-65): [128]*4, 2 heads, batch_size 10 000
    "gat4x128h2_ddr_b10000_k6": dict(model="gat", feat=1025, hidden=[128] * 4, heads=2, out=1, graphs=10000, nodes="ddr", k=6,
                                     layout="uniform", loss="MSE", dropout=0.35, weight_decay=2e-6, pool=2),
    # BASELINE.json configs[2]: GIN / GraphSAGE, mean||max read-out, 4096 graphs per step
    "gin_b4096": dict(model="gin", feat=64, hidden=[64] * 4, out=5, graphs=4096, nodes=100, k=8, layout="uniform", loss="CE",
                      dropout=0.35, weight_decay=2e-6, pool=3),
    "sage_b4096": dict(model="sage", feat=64, hidden=[64] * 3, out=5, graphs=4096, nodes=100, k=8, layout="uniform", loss="CE",
                       dropout=0.0, weight_decay=2e-6, pool=3),
    # BASELINE.json configs[3]: superpixel graphs, 2-layer GCN (the SpMM bandwidth stress as a whole training step)
    "gcn_cfg4": dict(model="gcn", feat=128, hidden=[128, 128], out=5, graphs=16, nodes=10000, k=16, layout="raster", loss="CE",
                     dropout=0.0, weight_decay=0.0, pool=2),
    # BASELINE.json configs[0]: the reference's CPU-runnable case
    "gcn_cfg1": dict(model="gcn", feat=16, hidden=[64, 64], out=5, graphs=32, nodes=100, k=8, layout="uniform", loss="CE",
                     dropout=0.0, weight_decay=0.0, pool=4),
}
DEFAULT_WORKLOAD = "gat3x64h4_ddr_b256_k6"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: `graphs` per GPU; strong: `graphs` in total, split over the GPUs")
    ap.add_argument("--no-extras", action="store_true", help="only value / e2e / roofline (skip the secondary legs)")
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


def graphs_per_rank(spec: dict, world: int, scaling: str) -> int:
    return spec["graphs"] if scaling == "weak" else max(1, spec["graphs"] // world)


def make_host_batches(spec: dict, graphs: int, rank: int, count: int, with_edges_on=None, world: int = 1):
    """`count` distinct synthetic batches for this rank; kNN edges are built on `with_edges_on` (a CUDA device,
    lgk_knn_graph) or by the oracle on the CPU (reference arm) -- for `value` and `e2e` graph construction is outside
    the timed step, as in the reference where it happens in the DataLoader (datasets/datamodule.py:43-48); the
    `e2e_resident_dataset` leg times it."""
    from lesion_gnn_b200 import data
    out = []
    for i in range(count):
        nodes, n_graphs = spec["nodes"], graphs
        if world > 1:
            # data parallel: ONE global batch of graphs x world graphs, split over the ranks by node count (data.shard_counts:
            # contiguous ranges of graphs, long-tailed sizes) -- every rank derives the same sizes from the shared seed and
            # generates only its own shard
            sizes = data.node_counts(graphs * world, spec["nodes"], torch.Generator().manual_seed(data.SEED + 7919 * (i + 1)))
            lo, hi = data.shard_counts(sizes, world)[rank]
            nodes, n_graphs = sizes[lo:hi], hi - lo
        hb = data.synth_batch(n_graphs, nodes, spec["feat"], seed=data.SEED + 1000 * rank + i, layout=spec["layout"],
                              num_classes=5)
        if with_edges_on is not None:
            from lesion_gnn_b200 import transforms
            from lesion_gnn_b200.nn.pool import Segments
            seg = Segments.from_batch(hb.batch.to(with_edges_on), hb.num_graphs)
            ei = transforms.knn_graph(hb.pos.to(with_edges_on), spec["k"], seg, loop=True,
                                      num_edges=transforms.knn_num_edges(hb.counts, spec["k"], True))
            hb.edge_index = ei.cpu()
        else:
            from oracle import pyg_semantics as O
            hb.edge_index = O.knn_graph(hb.pos, spec["k"], hb.batch, loop=True)
        out.append(hb)
    return out


def build_model(spec: dict, module, dropout=None):
    """`module`: lesion_gnn_b200.models or oracle.pyg_semantics -- same constructors"""
    p = spec["dropout"] if dropout is None else dropout
    if spec["model"] == "gat":
        return module.GAT(spec["feat"], spec["hidden"], spec["out"], heads=spec["heads"], dropout=p)
    if spec["model"] == "gin":
        return module.GIN(spec["feat"], spec["hidden"], spec["out"], dropout=p, pooling="mean+max")
    if spec["model"] == "sage":
        return module.SAGE(spec["feat"], spec["hidden"], spec["out"])
    if spec["model"] == "gcn":
        return module.GCN(spec["feat"], spec["hidden"], spec["out"])
    raise ValueError(spec["model"])


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
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(physical_index(index))
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


def physical_index(local: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        parts = vis.split(",")
        if local < len(parts) and parts[local].strip().isdigit():
            return int(parts[local])
    return local


def bind_to_gpu_numa(local: int) -> dict:
    """Pin this rank's host threads to the CPUs NVML reports as local to its GPU BEFORE any pinned buffer is allocated
    (first touch then places the staging buffers on the GPU's NUMA node): at N >= 4 every rank shared NUMA node 0 in
    round 1 and the per-rank host->device rate fell from 53 to 23 GB/s."""
    info = {"bound": False}
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(physical_index(local))
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = sorted(set(cpus) & allowed)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info = {"bound": True, "cpus": f"{cpus[0]}-{cpus[-1]}", "n_cpus": len(cpus)}
    except Exception as e:  # noqa: BLE001
        info["error"] = repr(e)[:120]
    return info


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
    if name in ("lgk_spmm", "lgk_spmm_blocked"):
        n, f = a["n_rows"], a["feat"]
        extra = (4 * E if a.get("ew") else 0) + (4 * E if a.get("emap") else 0) + (4 * n if a["pre"] else 0) + (4 * n if a["post"] else 0)
        extra += 4 * n * f if a.get("arg_out") else 0
        return 2 * n * f * 4 + 4 * E + 4 * (n + 1) + extra
    if name == "lgk_quad_build":
        return 8 * E + 5 * a["n_rows"]
    if name == "lgk_spmm_max_bwd":
        n, f = a["n_rows"], a["feat"]
        return 3 * n * f * 4 + 8 * E + 4 * (n + 1)
    if name in ("lgk_gat_fwd", "lgk_gat_fwd_elu", "lgk_gat_bwd_dst", "lgk_gat_bwd_dst_elu", "lgk_gat_bwd_src", "lgk_gat_alpha",
                "lgk_gat_datt"):
        n, h, c = a["num_nodes"], a["heads"], a["chan"]
        nf, nh = n * h * c, n * h
        idx = 4 * E + 4 * (n + 1)
        return {"lgk_gat_fwd": 4 * (2 * nf + 4 * nh) + idx, "lgk_gat_fwd_elu": 4 * (3 * nf + 4 * nh) + idx,
                "lgk_gat_bwd_dst": 4 * (3 * nf + 9 * nh) + idx, "lgk_gat_bwd_dst_elu": 4 * (4 * nf + 9 * nh) + idx,
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
    if name == "lgk_sgd_step":
        return 12 * a["n"]
    if name == "lgk_gather_f32":
        return 12 * a["n"]
    es = 2 if a.get("dtype") == 1 else 4  # activation element size (LGK_BF16 = 1)
    if name == "lgk_bn_act_fwd":    # x read for the statistics and again for the output, y written
        return (3 if a.get("training") else 2) * a["n"] * a["d"] * es
    if name == "lgk_bn_act_bwd":    # x and dy read for the sums and again for dx, dx written
        return 5 * a["n"] * a["d"] * es
    if name == "lgk_act_dropout_fwd":
        return 2 * a["n"] * es
    if name == "lgk_act_dropout_bwd":
        return 3 * a["n"] * es
    if name == "lgk_csr_build_sorted":
        return 28 * a["num_edges"] + 8 * (a["num_nodes"] + 1)
    if name == "lgk_linear_gat_fwd":
        m = a["heads"] * a["chan"]
        return (a["n"] * a["k"] + a["n"] * m) * es + 4 * a["k"] * m + 8 * a["n"] * a["heads"]
    return 0


class KernelTimer:
    """_lib hook: brackets every C-ABI call with CUDA events on the launching (current) stream."""

    def __init__(self, lib, small_n: int):
        self.lib, self.records, self.num_edges, self.small_n = lib, [], 0, small_n
        self._open = None

    def __call__(self, phase, name, args):
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream())
        if phase == "begin":
            self._open = ev
        else:
            names = self.lib.protos[name][2]
            vals = {k: (v if isinstance(v, (int, float)) else (0 if v is None else 1)) for k, v in zip(names, args)}
            self.records.append((name, self._open, ev, algorithmic_bytes(name, vals, self.num_edges), _shape_tag(name, vals, self.small_n)))

    def summary(self) -> dict:
        torch.cuda.synchronize()
        agg = {}
        for name, e0, e1, nbytes, tag in self.records:
            d = agg.setdefault((name, tag), {"ms": 0.0, "bytes": 0, "calls": 0})
            d["ms"] += e0.elapsed_time(e1)
            d["bytes"] += nbytes
            d["calls"] += 1
        return agg


def _shape_tag(name: str, a: dict, small_n: int) -> str:
    # row counts (n, N, E) vary from batch to batch of the pool: group launches by the per-row shape only
    keys = [k for k in ("k", "m", "feat", "heads", "chan", "reduce") if k in a]
    if name == "lgk_linear_bwd_w" and not a.get("dw"):
        keys = ["m"]  # bias-gradient column sum only
    if name in ("lgk_linear_fwd", "lgk_linear_bwd_x", "lgk_linear_bwd_w") and a["n"] <= small_n:
        keys.append("n")  # per-graph (out_proj) products
    return ",".join(f"{k}={a[k]}" for k in keys)


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


# ------------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------------
class E2EPipe:
    """Double-buffered host -> device feed: while step i computes, the pinned-host -> device copy of step i+1's inputs
    runs on a copy stream; every step's inputs are copied inside the timed region and every step's loss is read on
    the host.  `stage(slot)` does the copies of one slot (and whatever device-side preparation belongs to them)."""

    def __init__(self, n_slots, stage):
        self.n, self.stage = n_slots, stage
        self.copy_stream = torch.cuda.Stream()
        self.ready = [torch.cuda.Event() for _ in range(n_slots)]
        self.consumed = [torch.cuda.Event() for _ in range(n_slots)]

    def _prefetch(self, i):
        s = i % self.n
        self.copy_stream.wait_event(self.consumed[s])  # the previous user of this device buffer has finished
        with torch.cuda.stream(self.copy_stream):
            self.stage(s)
            self.ready[s].record(self.copy_stream)

    def run(self, count, step):
        main = torch.cuda.current_stream()
        for s in range(self.n):
            self.consumed[s].record(main)
        self._prefetch(0)
        loss = 0.0
        for i in range(count):
            if i + 1 < count:
                self._prefetch(i + 1)
            main.wait_event(self.ready[i % self.n])
            out_dev = step(i % self.n)
            self.consumed[i % self.n].record(main)
            loss = float(out_dev)  # device->host read of this step's loss (synchronises the host with the step)
        return loss


def run_b200(args):
    import torch.distributed as dist

    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import _lib, ops, train

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py --impl b200 needs a CUDA device: lesion_gnn_b200 has no CPU path")
    spec = WORKLOADS[args.workload]
    # the CPU leg runs first, at N = 1 only: under torchrun OMP_NUM_THREADS is 1 and the other ranks would spin
    cpu = None
    if world == 1 and not args.no_extras:
        cpu = cpu_baseline(spec, args.workload, args.cpu_seconds)
    numa = bind_to_gpu_numa(local)  # before any pinned allocation
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t)

    graphs = graphs_per_rank(spec, world, args.scaling)
    total_graphs = graphs * world
    POOL = spec["pool"]

    def make_step(dropout=None, weight_decay=None):
        torch.manual_seed(1234)
        model = build_model(spec, lg.models, dropout).to(dev).train()
        if world > 1:  # identical replicas
            for p in model.parameters():
                dist.broadcast(p.data, 0)
        # N > 1: gradients in a peer-visible bucket; the all-reduce is fused with Adam inside the captured step (csrc/dp.cu)
        bucket = train.PeerBucket(dist.group.WORLD) if world > 1 and train.DP_FUSED else None
        opt = train.FlatAdam(model, lr=1e-3, weight_decay=spec["weight_decay"] if weight_decay is None else weight_decay,
                             peer_bucket=bucket)
        return model, opt, train.TrainStep(model, opt, spec["loss"], 5, use_cuda_graph=True, process_group=None, world_size=world)

    model, opt, step = make_step()
    host = [hb.pin() for hb in make_host_batches(spec, graphs, rank, POOL, with_edges_on=dev, world=world)]
    devb = [train.DeviceBatch(hb, dev) for hb in host]
    for hb, db in zip(host, devb):
        db.copy_from(hb)
        db.loss_weight = train.dp_loss_weight(db.num_graphs, total_graphs, world)
    resident = sum(hb.h2d_bytes() for hb in host)
    l2_policy = (f"rotating pool of {POOL} distinct batches, {resident / 2**20:.0f} MiB > L2 (126 MiB)" if resident > L2_BYTES else
                 f"rotating pool of {POOL} distinct batches ({resident / 2**20:.0f} MiB) + a {2 * L2_BYTES >> 20} MiB L2 flush write between steps")
    flush = None if resident > L2_BYTES else torch.empty(2 * L2_BYTES // 4, dtype=torch.float32, device=dev)

    def run_steps(count, stepper, batches):
        for i in range(count):
            if flush is not None:
                flush.zero_()
            stepper(batches[i % len(batches)])

    def timed(count, stepper, batches):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        if flush is None:
            e0.record()
            run_steps(count, stepper, batches)
            e1.record()
            barrier()
            return max_over_ranks(e0.elapsed_time(e1))
        # with the L2 flush between steps only the steps themselves are timed (event pairs per step)
        evs = []
        for i in range(count):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            stepper(batches[i % len(batches)])
            b.record()
            evs.append((a, b))
        barrier()
        return max_over_ranks(sum(a.elapsed_time(b) for a, b in evs))

    # setup (not a warm-up step: the CUDA graph of every pool batch is captured once, training state restored),
    # then EXACTLY --warmup untimed steps and --steps timed ones
    for db in devb:
        step(db)
    barrier()
    run_steps(args.warmup, step, devb)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_total = timed(args.steps, step, devb)
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: host buffers -> device inside the timed region, loss read back every step -----------------------
    pipe = E2EPipe(POOL, lambda s: devb[s].copy_from(host[s]))
    pipe.run(POOL, lambda s: step(devb[s]))
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    loss_host = pipe.run(args.steps, lambda s: step(devb[s]))
    e1.record()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    e2e_ms_total = max_over_ranks(max(e0.elapsed_time(e1), wall_ms))
    h2d = sum(host[i % POOL].h2d_bytes() for i in range(args.steps)) / args.steps

    # ---- instrumented pass: per entry point CUDA-event times on the same steps (eager, same kernels) -----------
    eager = train.TrainStep(model, opt, spec["loss"], 5, use_cuda_graph=False, world_size=world)
    eager._side = eager._build = None  # one stream: every launch is timed alone, not beside the side-stream kernels
    L = _lib.lib()
    eager(devb[0])
    ops.reset_launch_count()
    eager(devb[0])
    launches_per_step = ops.launch_count()
    timer = KernelTimer(L, graphs)
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
        e2e_ms = e2e_ms_total / args.steps
        out = {
            "metric": "graphs/sec fwd+bwd", "value": round(total_graphs * args.steps / (ms_total * 1e-3), 1), "unit": "graphs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_step, 4),
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "graphs_per_gpu": graphs, "global_graphs": total_graphs,
                       "nodes_per_step_rank0": sum(h.num_nodes for h in host) / POOL,
                       "edges_per_step_rank0": sum(h.edge_index.size(1) for h in host) / POOL,
                       "model": spec["model"], "feat_in": spec["feat"], "hidden": spec["hidden"], "heads": spec.get("heads"),
                       "knn_k": spec["k"], "loss": spec["loss"], "optimizer": "Adam", "dropout": spec["dropout"],
                       "weight_decay": spec["weight_decay"], "train_mode": True,
                       "parallelism": f"dp{world}", "cuda_graph": True, "l2_policy": l2_policy,
                       "sharding": ("one global batch of graphs_per_gpu x n_gpus graphs per step, split over the ranks in contiguous "
                                    "ranges balanced by node count (data.shard_counts); graphs_per_gpu is the mean" if world > 1 else "none"),
                       "setup": f"{POOL} graph captures before the {args.warmup} warm-up steps (training state restored)"},
            "e2e": {"value": round(total_graphs * args.steps / (e2e_ms_total * 1e-3), 1), "unit": "graphs/s",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4, "ms_per_step": round(e2e_ms, 4),
                    "h2d_GBps_per_rank": round(h2d / (e2e_ms * 1e-3) / 1e9, 2), "numa": numa, "last_loss": loss_host},
            "gpu_launches": int(launches_per_step * args.steps), "gpu_launches_per_step": int(launches_per_step),
            "clocks": clocks, "roofline": roofline, "kernels": breakdown,
        }
    if not args.no_extras:
        extras = secondary_legs(args, spec, graphs, total_graphs, world, rank, dev, host, devb, make_step, timed, barrier,
                                max_over_ranks, opt)
        if rank == 0:
            out.update(extras)
            if world == 1:
                out["aggregation"] = aggregation_stress(dev, peak, peak_src)
                out["cpu_baseline"] = cpu
            else:
                out["cpu_baseline"] = {"skipped": "N > 1: the CPU leg is timed at N = 1 only (torchrun pins OMP_NUM_THREADS=1 and "
                                                  "the other ranks would spin in a barrier); see --impl reference"}
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(out), flush=True)


def secondary_legs(args, spec, graphs, total_graphs, world, rank, dev, host, devb, make_step, timed, barrier, max_over_ranks, opt):
    """Extra keys, each a separate measurement that never replaces value / e2e: see the module docstring."""
    import torch.distributed as dist

    from lesion_gnn_b200 import ops, train, transforms
    from lesion_gnn_b200.nn.pool import Segments
    POOL = len(devb)
    steps = max(20, min(args.steps, 300))
    out = {}

    def e2e_time(pipe, stepper, h2d_bytes):
        pipe.run(POOL, stepper)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        loss = pipe.run(steps, stepper)
        e1.record()
        barrier()
        ms = max_over_ranks(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
        return {"value": round(total_graphs * steps / (ms * 1e-3), 1), "unit": "graphs/s", "ms_per_step": round(ms / steps, 4),
                "steps": steps, "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": 4,
                "h2d_GBps_per_rank": round(h2d_bytes / (ms / steps * 1e-3) / 1e9, 2), "last_loss": loss}

    # ---- parity_config: the configuration the parity tests check (dropout 0, weight decay 0), resident inputs ----
    if spec["dropout"] != 0.0 or spec["weight_decay"] != 0.0:
        _, _, step0 = make_step(dropout=0.0, weight_decay=0.0)
        for db in devb:
            step0(db)
        ms = timed(steps, step0, devb)
        out["parity_config"] = {"dropout": 0.0, "weight_decay": 0.0, "value": round(total_graphs * steps / (ms * 1e-3), 1),
                                "unit": "graphs/s", "ms_per_step": round(ms / steps, 4), "steps": steps}
        del step0

    model, opt2, step = make_step()
    for db in devb:
        step(db)

    # ---- e2e_bf16: node features cross PCIe as bf16 and are widened on the device (1e-2 bar: inputs are rounded) ----
    x16_host = [hb.x.to(torch.bfloat16).pin_memory() for hb in host]
    x16_dev = [torch.empty_like(t, device=dev) for t in x16_host]

    def stage_bf16(s):
        hb, db = host[s], devb[s]
        x16_dev[s].copy_(x16_host[s], non_blocking=True)
        db.batch.copy_(hb.batch, non_blocking=True)
        db.y.copy_(hb.y, non_blocking=True)
        db.edge_index.copy_(hb.edge_index, non_blocking=True)
        db.x.copy_(x16_dev[s])  # bf16 -> fp32 on the device

    h2d16 = sum(x16_host[s].numel() * 2 + sum(t.numel() * t.element_size() for t in (host[s].batch, host[s].y, host[s].edge_index))
                for s in range(POOL)) / POOL
    out["e2e_bf16"] = e2e_time(E2EPipe(POOL, stage_bf16), lambda s: step(devb[s]), h2d16)
    out["e2e_bf16"]["note"] = "features rounded to bf16 on the host, fp32 arithmetic on the device; separate from the fp32 e2e"

    # ---- bf16_activations: the bf16 mode of the north_star (1e-2 bar, tests/test_gpu_bf16.py): every stored activation is
    # bf16, parameters / accumulators / optimiser fp32; resident and end to end (features cross PCIe as bf16 and stay bf16)
    import copy
    host16 = []
    for hb, x16 in zip(host, x16_host):
        h16 = copy.copy(hb)
        h16.x = x16
        host16.append(h16)
    devb16 = [train.DeviceBatch(hb, dev) for hb in host16]
    for hb, db in zip(host16, devb16):
        db.copy_from(hb)
        db.loss_weight = train.dp_loss_weight(db.num_graphs, total_graphs, world)
    _, _, step16 = make_step()
    for db in devb16:
        step16(db)
    ms = timed(steps, step16, devb16)
    leg = {"value": round(total_graphs * steps / (ms * 1e-3), 1), "unit": "graphs/s", "ms_per_step": round(ms / steps, 4), "steps": steps}

    def stage16(s):
        devb16[s].copy_from(host16[s])

    leg["e2e"] = e2e_time(E2EPipe(POOL, stage16), lambda s: step16(devb16[s]), h2d16)
    out["bf16_activations"] = leg
    del x16_host, x16_dev, devb16, step16

    # ---- e2e_distinct_batches: what a DataLoader does -- a NEW batch shape every step (reference training.py:64-72).
    # 50 distinct batches stream from pinned host memory through train.BucketedTrainStep: padded into a few (N, E) buckets,
    # one captured graph per bucket (models with BatchNorm in training mode are refused by it: GAT / SAGE / GCN only)
    # N = 1 only: under data parallelism the ranks would meet new buckets (warm-up + capture, i.e. extra collective calls) at
    # different steps
    if spec["model"] in ("gat", "sage", "gcn") and spec["graphs"] <= 1024 and world == 1:
        distinct = 50
        hosts50 = [hb.pin() for hb in make_host_batches(spec, graphs, rank + 64, distinct, with_edges_on=dev)]
        model50, opt50, _ = make_step()
        kdeg = spec["k"] + 1
        bstep = train.BucketedTrainStep(model50, opt50, num_graphs=graphs, max_degree=kdeg, loss_type=spec["loss"], node_step=1024,
                                        buffers=2, process_group=None, world_size=world)
        for _ in range(2):  # captures happen here (both buffers of every bucket), outside the timed region
            for pb_hb in hosts50:
                bstep(pb_hb)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        rounds = max(1, steps // distinct)
        seq = [h for _ in range(rounds) for h in hosts50]
        # every step's loss goes device -> pinned host inside the timed region; the host reads it one step LATER, so that it
        # can stage batch i + 2 while step i + 1 runs instead of idling on step i + 1's result (a replayed graph returns the
        # same loss tensor each time: the asynchronous copy is what keeps the value)
        loss_host = torch.zeros(2, dtype=torch.float32).pin_memory()
        loss_done = [torch.cuda.Event(), torch.cuda.Event()]
        last = None
        handle = bstep.stage(seq[0])
        for i in range(len(seq)):  # the copy of batch i + 1 runs on the copy stream while batch i trains
            nxt = bstep.stage(seq[i + 1]) if i + 1 < len(seq) else None
            loss50 = bstep.run(handle)
            loss_host[i & 1:(i & 1) + 1].copy_(loss50.reshape(1), non_blocking=True)
            loss_done[i & 1].record()
            if i > 0:
                loss_done[(i - 1) & 1].synchronize()
                last = float(loss_host[(i - 1) & 1])  # the previous step's result
            handle = nxt
        loss_done[(len(seq) - 1) & 1].synchronize()
        last = float(loss_host[(len(seq) - 1) & 1])
        e1.record()
        barrier()
        ms = max_over_ranks(max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3))
        n50 = rounds * distinct
        out["e2e_distinct_batches"] = {
            "value": round(total_graphs * n50 / (ms * 1e-3), 1), "unit": "graphs/s", "ms_per_step": round(ms / n50, 4), "steps": n50,
            "distinct_batches": distinct, "distinct_shapes": len({(h.num_nodes, h.edge_index.size(1)) for h in hosts50}),
            "captured_graphs": bstep.captures, "h2d_bytes_per_step": int(sum(h.h2d_bytes() for h in hosts50) / distinct),
            "d2h_bytes_per_step": 4, "last_loss": last,
            "note": "every step a batch of a different (N, E): host -> padded bucket buffers (two per bucket: the copy of batch i + 1 "
                    "overlaps step i) -> replay of the buffer's graph -> loss copied to pinned host memory every step, read by the "
                    "host one step later"}
        del hosts50, bstep, model50, opt50

    # ---- e2e_resident_dataset: the reference's InMemoryDataset (datasets/base.py:49) lives in HBM; per step the batch is
    # collated on the device by graph index (lgk_gather_rows) and its kNN graph + CSR are built inside the timed region
    # (what datasets/datamodule.py:43-48,64-69 do per sample on the CPU) -------------------------------------------------
    x_all = torch.cat([hb.x for hb in host]).to(dev)
    pos_all = torch.cat([hb.pos for hb in host]).to(dev)
    y_all = torch.cat([hb.y for hb in host]).to(dev)
    counts_all = [c for hb in host for c in hb.counts]
    starts, acc = [], 0
    for c in counts_all:
        starts.append(acc)
        acc += c
    gen = torch.Generator().manual_seed(4321 + rank)
    sets = []
    for s in range(POOL):
        gid = torch.randperm(len(counts_all), generator=gen)[:graphs]
        cnt = [counts_all[g] for g in gid.tolist()]
        node_index = torch.cat([torch.arange(starts[g], starts[g] + counts_all[g]) for g in gid.tolist()])
        batch = torch.repeat_interleave(torch.arange(graphs), torch.tensor(cnt))
        sets.append({"gid": gid.pin_memory(), "node_index": node_index.pin_memory(), "batch": batch.pin_memory(), "counts": cnt,
                     "E": transforms.knn_num_edges(cnt, spec["k"], True)})

    class _Shape:  # what train.DeviceBatch needs to size its buffers
        pass

    rbatch = []
    for st in sets:
        n = st["node_index"].numel()
        hb = _Shape()
        hb.num_graphs, hb.num_nodes = graphs, n
        hb.x = torch.empty(n, spec["feat"])
        hb.batch = torch.empty(n, dtype=torch.long)
        hb.y = torch.empty(graphs, dtype=torch.long)
        hb.edge_index = torch.empty(2, st["E"], dtype=torch.long)
        hb.edge_weight = None
        db = train.DeviceBatch(hb, dev)
        db.dst_sorted = True  # the edges come from knn_graph on the device: grouped by destination
        db.loss_weight = train.dp_loss_weight(graphs, total_graphs, world)
        st["node_index_dev"] = torch.empty(n, dtype=torch.long, device=dev)
        st["gid_dev"] = torch.empty(graphs, dtype=torch.long, device=dev)
        st["pos"] = torch.empty(n, 2, dtype=pos_all.dtype, device=dev)
        rbatch.append(db)

    def stage_resident(s):
        st, db = sets[s], rbatch[s]
        st["node_index_dev"].copy_(st["node_index"], non_blocking=True)
        st["gid_dev"].copy_(st["gid"], non_blocking=True)
        db.batch.copy_(st["batch"], non_blocking=True)
        ops.gather_rows(x_all, st["node_index_dev"], db.x)
        ops.gather_rows(pos_all, st["node_index_dev"], st["pos"])
        ops.gather_rows(y_all, st["gid_dev"], db.y)
        seg = Segments.from_batch(db.batch, graphs)
        ei = transforms.knn_graph(st["pos"], spec["k"], seg, loop=True, num_edges=st["E"])
        db.edge_index.copy_(ei)
        ei.record_stream(torch.cuda.current_stream())

    for s in range(POOL):
        stage_resident(s)
    torch.cuda.synchronize()
    for db in rbatch:
        step(db)
    h2dr = sum(st["node_index"].numel() * 8 * 2 + graphs * 8 for st in sets) / POOL
    out["e2e_resident_dataset"] = e2e_time(E2EPipe(POOL, stage_resident), lambda s: step(rbatch[s]), h2dr)
    out["e2e_resident_dataset"]["note"] = ("dataset of %d graphs resident in HBM; per step: node / graph indices host->device, "
                                           "device gather of x, pos, y, kNN graph, CSR/CSC build, fwd, bwd, optimiser" % len(counts_all))
    del x_all, pos_all, rbatch

    # ---- collective_us: the gradient exchange alone -------------------------------------------------------------------
    # fused: lgk_dp_allreduce_adam (peer-memory all-reduce + Adam + bucket release, two kernels) in a CUDA graph of 20 calls;
    # nccl: dist.all_reduce of the same bucket, eager, as the unfused step issues it
    if world > 1:
        def per_call_us(fn, graph_calls):
            for _ in range(5):
                fn()
            torch.cuda.synchronize()
            if graph_calls:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    for _ in range(graph_calls):
                        fn()
                run, calls = g.replay, 5 * graph_calls
                reps = 5
            else:
                run, calls, reps = fn, 100, 100
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                run()
            e1.record()
            torch.cuda.synchronize()
            return round(max_over_ranks(e0.elapsed_time(e1)) * 1000.0 / calls, 2)

        state = [t.clone() for t in opt.state_tensors()]
        if getattr(opt, "peer_bucket", None) is not None:
            out["collective_us"] = per_call_us(opt.step_dp, 20)
            out["collective_kind"] = "lgk_dp_allreduce_adam (NVLink peer loads, fused with Adam; in-graph)"
        out["collective_nccl_us"] = per_call_us(lambda: dist.all_reduce(opt.grad), 0)
        if "collective_us" not in out:
            out["collective_us"], out["collective_kind"] = out["collective_nccl_us"], "NCCL all-reduce (eager) + lgk_adam_step"
        out["collective_bytes"] = int(opt.grad.numel() * 4)
        for dst, src in zip(opt.state_tensors(), state):
            dst.copy_(src)
    return out


def aggregation_stress(dev, peak: float, peak_src: str, graphs: int = 128) -> dict:
    """BASELINE.json configs[3] / SURVEY.md 8d cfg 4: `graphs` superpixel graphs x 10 000 nodes, 128-d, kNN k=16
    (loop=True), GCN-normalised aggregation forward (CSR) and backward (CSC), plus mean pooling.  X = 655 MB at
    128 graphs, far beyond L2, so every launch streams from HBM.  The aggregation entries run the kernel the layers run
    at this size (the windowed, row-blocked lgk_spmm_blocked); `*_lgk_spmm` are the register-landing kernel beside it."""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import data, ops, transforms
    from lesion_gnn_b200.nn.pool import Segments
    n, k, F = 10000, 16, 128
    hb = data.synth_batch(graphs, n, F, seed=data.SEED, layout="raster")
    N = hb.num_nodes
    seg = Segments.from_batch(hb.batch.to(dev), graphs)
    ei = transforms.knn_graph(hb.pos.float().to(dev), k, seg, loop=True, num_edges=N * k)
    E = ei.size(1)
    # knn_graph emits edges grouped by destination: the sorted-input build is what the pipeline runs on them; the general
    # bucket-sort build (arbitrary edge order) is timed beside it
    t_csr = _time_cuda(lambda: lg.CSRGraph.from_edge_index(ei, N, dst_sorted=True), 3, 1)
    t_csr_general = _time_cuda(lambda: lg.CSRGraph.from_edge_index(ei, N), 3, 1)
    g = lg.CSRGraph.from_edge_index(ei, N, dst_sorted=True)
    x = hb.x.to(dev)
    dinv, _ = g.gcn_norm(False, True)
    kw = dict(pre=dinv, post=dinv, self_mode=ops.SELF_REPLACE)
    res = {"shape": {"graphs": graphs, "nodes": N, "edges": E, "feat": F, "k": k}, "peak": peak, "peak_source": peak_src, "unit": "GB/s"}
    agg_bytes = 2 * N * F * 4 + 4 * E + 4 * (N + 1) + 8 * N

    def entry(ms, nbytes):
        gbs = nbytes / (ms * 1e-3) / 1e9
        return {"us": round(ms * 1e3, 1), "bytes": nbytes, "achieved": round(gbs, 1), "frac": round(gbs / peak, 4)}

    t_quads = _time_cuda(lambda: ops.quad_build(g.rowptr, g.col), 3, 1)
    qc, qr = g.quads("csr"), g.quads("csc")
    assert ops.blocked_eligible(x, "sum", None)
    res["quad_build"] = entry(t_quads, 8 * E + 5 * N)
    # forward as GCNConv launches it on an unweighted graph: the source-side d^-1/2 was applied by the transform's epilogue
    # (nn.functional.gcn_conv), the aggregation scales its output rows and adds the self loop; `_pre_in_kernel`: both factors
    # inside the aggregation (weighted graphs, the first layer of a captured step, and the backward pass)
    kw_fwd = dict(post=dinv, self_mode=ops.SELF_REPLACE)
    res["gcn_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", quads=qc, **kw_fwd), 20), agg_bytes - 4 * N)
    res["gcn_aggregate_fwd"]["path"] = "post-scale + self loop in the aggregation, pre-scale in the transform's epilogue (GCNConv, unweighted graph)"
    res["gcn_aggregate_fwd_pre_in_kernel"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", quads=qc, **kw), 20), agg_bytes)
    res["gcn_aggregate_bwd"] = entry(_time_cuda(lambda: ops.spmm(g.colptr, g.row, x, "sum", quads=qr, **kw), 20), agg_bytes)
    plain = 2 * N * F * 4 + 4 * E + 4 * (N + 1)
    res["sum_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", quads=qc), 20), plain)
    # what SAGEConv(aggr="mean") launches: the sum variant scaled on the way out by the memoised 1 / max(deg, 1)
    inv_deg = g.inv_degree()
    res["mean_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", post=inv_deg, quads=qc), 20), plain + 4 * N)
    res["mean_aggregate_fwd_degree_in_kernel"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "mean", quads=qc), 10), plain)
    res["gcn_aggregate_fwd_lgk_spmm"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum", **kw), 10), agg_bytes)
    res["gcn_aggregate_bwd_lgk_spmm"] = entry(_time_cuda(lambda: ops.spmm(g.colptr, g.row, x, "sum", **kw), 10), agg_bytes)
    res["sum_aggregate_fwd_lgk_spmm"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "sum"), 10), plain)
    res["max_aggregate_fwd"] = entry(_time_cuda(lambda: ops.spmm(g.rowptr, g.col, x, "max", want_arg=True), 20), plain + 4 * N * F)
    H, C = 4, 32
    att = torch.randn(H * C, device=dev) * 0.1
    a_s, a_d = ops.gat_alpha(x, att, att, H, C)
    gat_bytes = 4 * (2 * N * F + 4 * N * H) + 4 * E + 4 * (N + 1)
    res["gat_softmax_aggregate_fwd"] = entry(_time_cuda(lambda: ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True, quads=qc), 20), gat_bytes)
    res["gat_softmax_aggregate_fwd_lgk_gat_fwd"] = entry(_time_cuda(lambda: ops.gat_fwd(g.rowptr, g.col, x, a_s, a_d, None, H, C, 0.2, True), 10), gat_bytes)
    pool_bytes = 4 * (N * F + graphs * F) + 4 * (graphs + 1)
    res["global_mean_pool_fwd"] = entry(_time_cuda(lambda: ops.pool_fwd(x, seg.ptr, True, True, False), 20), pool_bytes)
    res["global_mean_max_pool_fwd"] = entry(_time_cuda(lambda: ops.pool_fwd(x, seg.ptr, True, True, True), 20), pool_bytes + 8 * graphs * F)
    res["csr_csc_build"] = entry(t_csr, 28 * E + 8 * (N + 1))
    res["csr_csc_build"]["path"] = "lgk_csr_build_sorted (edges grouped by destination, as knn_graph emits them)"
    res["csr_csc_build_general"] = entry(t_csr_general, 28 * E + 8 * (N + 1))
    del x, g, ei
    torch.cuda.empty_cache()
    return res


# ------------------------------------------------------------------------------------------------------
# CPU legs: the oracle (op-for-op PyG restatement, oracle/pyg_semantics.py) on the host cores
# ------------------------------------------------------------------------------------------------------
def _oracle_stepper(spec: dict):
    from oracle import pyg_semantics as O
    torch.set_num_threads(os.cpu_count() or 1)  # explicit: torchrun exports OMP_NUM_THREADS=1
    torch.manual_seed(1234)
    model = build_model(spec, O).train()
    optim = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=spec["weight_decay"])
    graphs = min(spec["graphs"], 4096)
    batches = make_host_batches(spec, graphs, 0, 2)

    def one(i):
        hb = batches[i % len(batches)]
        return O.training_step(model, optim, hb.x, hb.edge_index, hb.batch, hb.y, hb.num_graphs, spec["loss"], 5)
    return one, graphs


def cpu_baseline(spec: dict, name: str, budget_s: float) -> dict:
    one, graphs = _oracle_stepper(spec)
    one(0)
    t0 = time.perf_counter()
    n = 0
    while True:
        one(n)
        n += 1
        dt = time.perf_counter() - t0
        if dt >= budget_s or n >= 200:
            break
    return {"value": round(graphs * n / dt, 1), "unit": "graphs/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} full steps of {name} ({graphs} graphs each, fwd+loss+bwd+Adam, fp32, dropout {spec['dropout']}) in {dt:.1f} s "
                      "after 1 warm-up", "ms_per_step": round(1e3 * dt / n, 2)}


def run_reference(args):
    """The reference's CPU path for this workload = the oracle (PyG is not installable here: SURVEY.md 8c), all
    host threads, the same model / dropout / weight decay as the CUDA arm.  Under torchrun only rank 0 works; the
    step is always one rank's batch (`graphs` graphs), whatever --gpus says: a throughput, not a scaling run."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    spec = WORKLOADS[args.workload]
    one, graphs = _oracle_stepper(spec)
    warm = max(args.warmup, 1)
    t_w = time.perf_counter()
    one(0)
    per_step = time.perf_counter() - t_w
    warm = min(warm, max(1, int(20.0 / max(per_step, 1e-3))))
    for i in range(1, warm):
        one(i)
    steps = max(1, min(args.steps, int(120.0 / max(per_step, 1e-3)), 100))
    t0 = time.perf_counter()
    for i in range(steps):
        one(i)
    dt = time.perf_counter() - t0
    v = round(graphs * steps / dt, 1)
    cores = torch.get_num_threads()
    print(json.dumps({
        "impl": "reference", "metric": "graphs/sec fwd+bwd", "value": v, "unit": "graphs/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": round(1e3 * dt / steps, 3), "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "graphs_per_gpu": graphs, "global_graphs": graphs, "model": spec["model"],
                   "feat_in": spec["feat"], "hidden": spec["hidden"], "heads": spec.get("heads"), "knn_k": spec["k"], "loss": spec["loss"],
                   "optimizer": "Adam", "dropout": spec["dropout"], "weight_decay": spec["weight_decay"], "train_mode": True,
                   "device": "cpu", "note": "one rank's batch per step on the host cores, whatever --gpus says"},
        "cpu_baseline": {"value": v, "unit": "graphs/s", "cores": cores, "kind": "port",
                         "sample": f"{steps} full steps of {graphs} graphs on {cores} host threads"},
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
