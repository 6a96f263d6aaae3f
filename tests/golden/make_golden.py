"""Generates tests/golden/*.pt: small fixed inputs with the oracle's outputs and gradients (fp64).

PARITY UNPINNED at the PyG boundary: torch_geometric 2.5.1 is not installable here and the reference's tests
hold no vector for this path, so these vectors come from oracle/pyg_semantics.py after it has
been cross-checked against oracle/dense.py (tests/test_oracle.py).  The only reference-held known answers,
GaussianDistance at unit distance (/root/reference/test/test_transforms.py:20,29,38), are stored verbatim.

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pyg_semantics as O  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def graphs():
    """name -> (edge_index, num_nodes, batch)"""
    out = {}
    out["two_node"] = (torch.tensor([[0, 1], [1, 0]]), 2, torch.tensor([0, 0]))
    # star: leaves 1..5 -> hub 0, plus one stored self loop on the hub and a duplicate edge
    out["star"] = (torch.tensor([[1, 2, 3, 4, 5, 0, 3], [0, 0, 0, 0, 0, 0, 0]]), 6, torch.tensor([0] * 6))
    # path 0-1-2-3-4 (both directions) with node 5 isolated, two graphs
    s = torch.tensor([0, 1, 1, 2, 2, 3, 3, 4])
    d = torch.tensor([1, 0, 2, 1, 3, 2, 4, 3])
    out["path_isolated"] = (torch.stack([s, d]), 6, torch.tensor([0, 0, 0, 1, 1, 1]))
    gen = torch.Generator().manual_seed(1234)
    counts = [100, 7, 1, 33]
    pos = torch.rand(sum(counts), 2, generator=gen, dtype=torch.float64) * 1536
    batch = torch.repeat_interleave(torch.arange(4), torch.tensor(counts))
    out["knn_mixed"] = (O.knn_graph(pos, 6, batch, loop=True), sum(counts), batch)
    return out


def run_layer(make, x, *args):
    torch.manual_seed(0)
    layer = make().double()
    x = x.clone().requires_grad_(True)
    y = layer(x, *args)
    gen = torch.Generator().manual_seed(99)
    dy = torch.randn(y.shape, generator=gen, dtype=torch.float64)
    (y * dy).sum().backward()
    return {
        "state": {k: v.detach().clone() for k, v in layer.state_dict().items()},
        "out": y.detach(), "dy": dy, "dx": x.grad.clone(),
        "dparams": {k: p.grad.clone() for k, p in layer.named_parameters() if p.grad is not None},
    }


def main():
    gold = {"gaussian_kat": {"sigma": [1.0, 0.5, 2.0], "value": [0.2420, 0.1080, 0.1760], "tol": 1e-3}, "cases": {}}
    for name, (ei, n, batch) in graphs().items():
        gen = torch.Generator().manual_seed(hash(name) % 1000 + 5)
        gen.manual_seed(len(name) * 7 + n)
        x = torch.randn(n, 16, generator=gen, dtype=torch.float64)
        ew = torch.rand(ei.size(1), generator=gen, dtype=torch.float64) + 0.1
        B = int(batch.max()) + 1
        case = {"edge_index": ei, "num_nodes": n, "batch": batch, "x": x, "edge_weight": ew}
        rowptr, col, perm = O.to_csr(ei, n)
        colptr, row, csr2csc = O.to_csc(rowptr, col, n)
        case["csr"] = dict(rowptr=rowptr, col=col, perm=perm, colptr=colptr, row=row, csr2csc=csr2csc)
        case["gcn"] = run_layer(lambda: O.GCNConv(16, 8), x, ei)
        case["gcn_w"] = run_layer(lambda: O.GCNConv(16, 8), x, ei, ew)
        case["gat"] = run_layer(lambda: O.GATConv(16, 4, heads=2), x, ei)
        case["gin"] = run_layer(lambda: O.GINConv(O.MLP([16, 8, 8], act="ELU", norm=None)), x, ei)
        case["sage_mean"] = run_layer(lambda: O.SAGEConv(16, 8, aggr="mean"), x, ei)
        case["sage_max"] = run_layer(lambda: O.SAGEConv(16, 8, aggr="max"), x, ei)
        case["graphconv_w"] = run_layer(lambda: O.GraphConv(16, 8), x, ei, ew)
        for red in ("mean", "max", "sum"):
            xx = x.clone().requires_grad_(True)
            y = O.scatter(xx, batch, B, red)
            dy = torch.randn(y.shape, generator=gen, dtype=torch.float64)
            (y * dy).sum().backward()
            case["pool_" + red] = {"out": y.detach(), "dy": dy, "dx": xx.grad.clone()}
        gold["cases"][name] = case
    torch.save(gold, os.path.join(HERE, "layers.pt"))
    print("wrote", os.path.join(HERE, "layers.pt"), os.path.getsize(os.path.join(HERE, "layers.pt")), "bytes")


if __name__ == "__main__":
    main()
