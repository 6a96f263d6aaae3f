"""Shared helpers for the parity tests: seeded synthetic lesion graphs and the error metric of BASELINE.json."""
import torch

from oracle import pyg_semantics as O

# BASELINE.json north_star: forward/backward within 1e-5 relative error in fp32, 1e-2 in bf16
TOL_F32 = 1e-5
TOL_BF16 = 1e-2


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max |a - b| / max |b|  (b = oracle)"""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.numel() == 0:
        return 0.0
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-30))


def assert_close(a, b, tol=TOL_F32, what=""):
    e = rel_err(a, b)
    assert e <= tol, f"{what}: relative error {e:.3e} > {tol:.0e}"


def row_rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max over rows of ||a_r - b_r||_2 / ||b_r||_2  (b = oracle): unlike max|a-b| / max|b| over the whole tensor this
    does not let rows of small magnitude be arbitrarily wrong.  Rows whose oracle norm is below 1e-6 of the largest
    row norm (exact zeros: empty graphs, dead ELU units) are measured against that floor; 0-d / 1-d tensors are one row."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.numel() == 0:
        return 0.0
    a2, b2 = a.reshape(1, -1) if a.dim() < 2 else a.reshape(a.size(0), -1), b.reshape(1, -1) if b.dim() < 2 else b.reshape(b.size(0), -1)
    nb = b2.norm(dim=1)
    floor = float(nb.max()) * 1e-6
    return float(((a2 - b2).norm(dim=1) / nb.clamp(min=max(floor, 1e-30))).max())


def assert_parity(got, want64, want32=None, what="", tol=TOL_F32, mult=4.0, row_tol=None):
    """The north_star bar: `got` (CUDA, fp32) within `tol` = 1e-5 of the fp64 oracle, in BOTH metrics (whole-tensor max
    and per-row L2).  Where fp32 arithmetic itself cannot hold 1e-5 (deep models, cancellation in small rows) the bound
    is `mult` x the error the ORACLE makes when it runs the same computation in fp32 on the CPU (`want32`) -- measured on
    the same inputs, never a hand-picked looser constant."""
    e, er = rel_err(got, want64), row_rel_err(got, want64)
    f = fr = 0.0
    if want32 is not None:
        f, fr = rel_err(want32, want64), row_rel_err(want32, want64)
    assert e <= max(tol, mult * f), f"{what}: relative error {e:.3e} > max({tol:.0e}, {mult} x fp32-oracle error {f:.3e})"
    rt = tol if row_tol is None else row_tol
    assert er <= max(rt, mult * fr), f"{what}: per-row error {er:.3e} > max({rt:.0e}, {mult} x fp32-oracle row error {fr:.3e})"
    return e, er, f, fr


def random_graph(n, e, seed, self_loops=True, duplicates=True):
    """random multigraph [2, E] int64 (may contain self loops and duplicate edges)"""
    g = torch.Generator().manual_seed(seed)
    if n == 0 or e == 0:
        return torch.zeros(2, 0, dtype=torch.long)
    ei = torch.randint(0, n, (2, e), generator=g)
    if not self_loops:
        ei = ei[:, ei[0] != ei[1]]
    if not duplicates:
        ei = torch.unique(ei, dim=1)
        ei = ei[:, torch.randperm(ei.size(1), generator=g)]
    return ei.contiguous()


def knn_batch(num_graphs, nodes, feat, k, seed, loop=True):
    """(HostBatch with edge_index from the oracle's kNN)"""
    from lesion_gnn_b200 import data
    hb = data.synth_batch(num_graphs, nodes, feat, seed=seed)
    hb.edge_index = O.knn_graph(hb.pos, k, hb.batch, loop=loop)
    return hb


class routed_max_pool:
    """Make the oracle's global_max_pool route its gradient through a GIVEN arg-max (the one the CUDA forward chose).

    d max / d x is discontinuous in x: when the two largest values of a (graph, channel) differ by less than fp32
    rounding, an fp32 forward (GPU or CPU alike) and the fp64 oracle can pick different nodes and the routed gradient
    lands on another row -- an ill-conditioned decision, not an arithmetic error (hard part 8).  Inside this
    context the oracle gathers x[arg] (forward value checked against its own true max at 1e-6), so the backward
    comparison tests the arithmetic; the arg-max itself is tested bit-exactly against the oracle on tie-free inputs in
    test_gpu_aggregate.py."""

    def __init__(self, arg: torch.Tensor):
        self.arg = arg.long().cpu()  # [B, F] global node index, -1 for an empty graph

    def __enter__(self):
        self._old = O.global_max_pool
        arg = self.arg

        def pool(x, batch, size=None):
            true_max = self._old(x, batch, size)
            safe = arg.clamp(min=0)
            out = torch.gather(x, 0, safe) * (arg >= 0).to(x.dtype)
            assert float((out.detach() - true_max.detach()).abs().max()) <= 1e-5 * float(true_max.detach().abs().max().clamp(min=1e-30)), \
                "the CUDA arg-max does not point at a maximal element"
            return out

        O.global_max_pool = pool
        return self

    def __exit__(self, *exc):
        O.global_max_pool = self._old
        return False


def attention_dropout_keep(seed: int, counter, slot: torch.Tensor, head: torch.Tensor, p: float) -> torch.Tensor:
    """Host restatement of liblgk's attention-dropout decision (csrc/common.cuh uniform_hash + csrc/gat.cu launch_seed):
    keep[slot, head] as bool.  slot: CSR slot of a stored edge, or 0x80000000 | i for the added self loop of node i."""
    import numpy as np
    M = np.uint64(0xFFFFFFFFFFFFFFFF)
    s = np.uint64(seed & 0xFFFFFFFFFFFFFFFF)
    with np.errstate(over="ignore"):
        if counter is not None:
            s = s + np.uint64(0xD1B54A32D192ED03) * np.uint64(counter)
        sl = slot.numpy().astype(np.uint64)
        hd = head.numpy().astype(np.uint64)
        z = s + np.uint64(0x9E3779B97F4A7C15) * ((sl << np.uint64(8)) ^ (hd + np.uint64(1)))
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    u = (z >> np.uint64(40)).astype(np.float32) * np.float32(1.0 / 16777216.0)
    return torch.from_numpy(~(u < np.float32(p)))
