"""torch.compile over the lgk layers (reference models/gat.py:84 `torch.compile(model, dynamic=True)`,
configs/config.py:64 `compile=True`).  CPU: Dynamo traces the models into graphs whose nodes are the opaque operators
`lgk::opaque_fwd` with symbolic shapes (fake tensors only, nothing executes).  GPU: the compiled models (backends
"eager" and "aot_eager", dynamic shapes, two different batch sizes) reproduce the eager forward and every gradient."""
import pytest
import torch


def _trace(model, *inputs):
    graphs = []

    def backend(gm, example_inputs):
        graphs.append(gm)

        def run(*args):  # nothing executes: outputs are zeros of the traced shapes
            from torch._subclasses.fake_tensor import FakeTensorMode
            with FakeTensorMode(allow_non_fake_inputs=True):
                outs = gm(*args)
            return tuple(torch.zeros([int(s) for s in o.shape], dtype=o.dtype) if isinstance(o, torch.Tensor) else o for o in outs)
        return run

    torch._dynamo.reset()
    compiled = torch.compile(model, dynamic=True, backend=backend)
    compiled(*inputs)
    return graphs


def _opaque_nodes(graphs):
    return [n for g in graphs for n in g.graph.nodes if n.op == "call_function" and "opaque_fwd" in str(n.target)]


@pytest.mark.parametrize("name", ["gat", "gin", "sage", "gcn"])
def test_dynamo_traces_the_models_as_opaque_operators(name):
    import lesion_gnn_b200 as lg
    torch.manual_seed(0)
    model = {"gat": lambda: lg.models.GAT(12, [8, 8, 8], 1, heads=2, dropout=0.35),
             "gin": lambda: lg.models.GIN(12, [8, 8, 8], 5, dropout=0.35, pooling="mean+max"),
             "sage": lambda: lg.models.SAGE(12, [8, 8], 5),
             "gcn": lambda: lg.models.GCN(12, [8, 8], 5)}[name]()
    x = torch.randn(20, 12)
    ei = torch.randint(0, 20, (2, 50))
    batch = torch.sort(torch.randint(0, 3, (20,))).values
    graphs = _trace(model, x, ei, batch)
    nodes = _opaque_nodes(graphs)
    n_layers = {"gat": 5, "gin": 7, "sage": 6, "gcn": 6}[name]  # transforms + convs (+ activations) + read-out + out_proj
    assert len(nodes) == n_layers, [str(n.target) for n in nodes]
    # symbolic row count flows through the first layer's fake implementation
    first = (nodes[0].meta.get("example_value") or nodes[0].meta.get("val"))[0]
    assert first.shape[1] == 8 and not isinstance(first.shape[0], int)


def test_copied_model_needs_prepare():
    import copy
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200 import compile as C
    m = lg.models.GCN(4, [4], 2)
    m2 = C.prepare(copy.deepcopy(m))
    assert all(id(l) in C._ENTRIES for l in m2.modules() if isinstance(l, C.OpaqueModule))


@pytest.mark.gpu
@pytest.mark.parametrize("backend", ["eager", "aot_eager"])
@pytest.mark.parametrize("name", ["gat", "gin", "drgnet"])
def test_compiled_model_equals_eager(dev, backend, name):
    import lesion_gnn_b200 as lg
    from tests.util import knn_batch
    torch.manual_seed(3)
    if name == "gat":
        model = lg.models.GAT(24, [32, 32, 32], 1, heads=4, dropout=0.0)
    elif name == "gin":
        model = lg.models.GIN(24, [32, 32], 5, dropout=0.0, pooling="mean")
    else:
        model = lg.models.DRGNet(24, 16, 2, 10, 5).eval()
    model = model.to(dev)
    torch._dynamo.reset()
    compiled = torch.compile(model, dynamic=True, backend=backend)
    for seed, graphs in ((1, 9), (2, 14)):  # two batch shapes through one compiled module
        hb = knn_batch(graphs, "ddr", 24, 6, seed=seed)
        x = hb.x.to(dev)
        ei, b = hb.edge_index.to(dev), hb.batch.to(dev)
        model.zero_grad()
        xe = x.clone().requires_grad_(True)
        want = model(xe, ei, b)
        want.square().sum().backward()
        grads = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
        model.zero_grad()
        xc = x.clone().requires_grad_(True)
        got = compiled(xc, ei, b)
        got.square().sum().backward()
        assert torch.equal(got, want), name
        # same kernels; autograd may add the partial gradients of a tensor in another order (last-bit differences)
        def same(a, b, what):
            assert float((a - b).abs().max()) <= 1e-6 * float(b.abs().max()) + 1e-30, what
        same(xc.grad, xe.grad, "dx")
        for k, p in model.named_parameters():
            if k in grads:
                same(p.grad, grads[k], k)
