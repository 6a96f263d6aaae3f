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


def make_host_batches(spec: dict, graphs: int, rank: int, count: int, with_edges_on=None):
    """`count` distinct synthetic batches for this rank; kNN edges are built on `with_edges_on` (a CUDA device,
    lgk_knn_graph) or by the oracle on the CPU (reference arm) -- for `value` and `e2e` graph construction is outside
    the timed step, as in the reference where it happens in the DataLoader (datasets/datamodule.py:43-48); the
    `e2e_resident_dataset` leg times it."""
    from lesion_gnn_b200 import data
    out = []
    for i in range(count):
        hb = data.synth_batch(graphs, spec["nodes"], spec["feat"], seed=data.SEED + 1000 * rank + i, layout=spec["layout"],
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
