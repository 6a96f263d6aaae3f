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
