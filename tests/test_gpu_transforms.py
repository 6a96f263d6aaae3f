"""The loader-level transform API on the device: the reference's own known-answer test of
GaussianDistance (reference test/test_transforms.py:8-77, every case, same tolerances) run through
lesion_gnn_b200.transforms, and get_transform(TransformConfig) pipelines (reference transforms.py:19-23,
configs/config.py:47, datasets/datamodule.py:43-48) against the oracle on a collated batch."""
import pytest
import torch

from oracle import pyg_semantics as O
from tests.util import knn_batch

pytestmark = pytest.mark.gpu


def _two_node_data(dev):
    from lesion_gnn_b200.transforms import Data
    edge_index = torch.tensor([[0, 1], [1, 0]])
    pos = torch.tensor([[0, 0], [1, 0]], dtype=torch.float)
    edge_weight = torch.tensor([1, 1], dtype=torch.float)
    edge_attr = torch.tensor([1, 1], dtype=torch.float)
    return Data(edge_index=edge_index, pos=pos, edge_attr=edge_attr, edge_weight=edge_weight).to(dev)


@pytest.mark.parametrize("sigma,expected", [(1, 0.2420), (0.5, 0.1080), (2, 0.1760)])
def test_gaussian_distance_reference_known_answers(dev, sigma, expected):
    """reference test/test_transforms.py:16-38"""
    from lesion_gnn_b200.transforms import GaussianDistance, SaveAs
    data = GaussianDistance(sigma=sigma, save_as=SaveAs.EDGE_WEIGHT_REPLACE)(_two_node_data(dev).clone())
    torch.testing.assert_close(data.edge_weight.cpu(), torch.tensor([expected, expected]), rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(data.edge_attr.cpu(), torch.tensor([1.0, 1.0]), rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("nodes", [2, 1])
def test_gaussian_distance_without_edges_warns(dev, nodes):
    """reference test/test_transforms.py:41-63"""
    from lesion_gnn_b200.transforms import Data, GaussianDistance, SaveAs
    pos = torch.tensor([[0, 0], [1, 0]][:nodes], dtype=torch.float)
    data = Data(edge_index=torch.tensor([], dtype=torch.long), pos=pos, edge_attr=torch.tensor([]),
                edge_weight=torch.tensor([])).to(dev)
    with pytest.warns(UserWarning):
        data = GaussianDistance(sigma=1, save_as=SaveAs.EDGE_WEIGHT_REPLACE)(data)
    assert data.edge_weight.numel() == 0


def test_gaussian_distance_edge_attr_modes(dev):
    """reference test/test_transforms.py:65-77"""
    from lesion_gnn_b200.transforms import GaussianDistance, SaveAs
    data = GaussianDistance(sigma=1, save_as=SaveAs.EDGE_ATTR_CAT)(_two_node_data(dev).clone())
    torch.testing.assert_close(data.edge_attr.cpu(), torch.tensor([[1, 0.2420], [1, 0.2420]]), rtol=1e-3, atol=1e-3)
    data = GaussianDistance(sigma=1, save_as=SaveAs.EDGE_ATTR_REPLACE)(_two_node_data(dev).clone())
    torch.testing.assert_close(data.edge_attr.cpu(), torch.tensor([[0.2420], [0.2420]]), rtol=1e-3, atol=1e-3)
    d = _two_node_data(dev)
    d.edge_attr = None
    d = GaussianDistance(sigma=1, save_as=SaveAs.EDGE_ATTR_CAT)(d)
    assert d.edge_attr.shape == (2, 1)


@pytest.mark.parametrize("name,kwargs", [("KNNGraph", {"k": 6, "loop": True}), ("KNNGraph", {"k": 8, "loop": False}),
                                         ("RadiusGraph", {"r": 220.0, "loop": False, "max_num_neighbors": 12})])
def test_get_transform_builds_the_batch_graph_bit_exact(dev, name, kwargs):
    """`get_transform(TransformConfig(...))(batch)`: one launch for the whole collated batch equals the oracle's
    per-graph construction (what the reference does per sample in __getitem__), bit for bit"""
    from lesion_gnn_b200 import data as D
    from lesion_gnn_b200.transforms import Data, TransformConfig, get_transform
    hb = D.synth_batch(30, "ddr", 8, seed=21)
    t = get_transform(TransformConfig(name=name, kwargs=kwargs))
    batch = Data.from_host_batch(hb, dev)
    batch.edge_attr = torch.ones(3, device=dev)
    out = t(batch)
    if name == "KNNGraph":
        want = O.knn_graph(hb.pos, kwargs["k"], hb.batch, loop=kwargs["loop"])
    else:
        want = O.radius_graph(hb.pos, kwargs["r"], hb.batch, loop=kwargs["loop"], max_num_neighbors=kwargs["max_num_neighbors"])
    assert out.edge_attr is None
    assert torch.equal(out.edge_index.cpu(), want)
    # without host-side graph sizes (one device->host read of the edge count) the result is the same
    b2 = Data.from_host_batch(hb, dev)
    b2.counts = None
    assert torch.equal(t(b2).edge_index.cpu(), want)


def test_reference_pipeline_knn_gaussian_sparse_tensor_feeds_drgnet(dev):
    """Compose([KNNGraph, GaussianDistance, ToSparseTensor]) as reference datasets/datamodule.py:43-48 builds it with
    compile=False; the resulting data.adj_t drives DRGNet (models/drgnet.py:98-99) to the oracle's logits"""
    import lesion_gnn_b200 as lg
    from lesion_gnn_b200.transforms import Compose, Data, get_transform
    hb = knn_batch(10, "ddr", 16, 6, seed=4)
    pipe = Compose([get_transform("KNNGraph", k=6, loop=True), get_transform("GaussianDistance", sigma=150.0),
                    get_transform("ToSparseTensor")])
    raw = Data.from_host_batch(hb, dev)
    raw.edge_index = None
    data = pipe(raw)
    assert data.edge_index is None and data.edge_weight is None and data.adj_t is not None
    ew = O.gaussian_distance(hb.edge_index, hb.pos, 150.0, torch.float64)
    torch.manual_seed(2)
    ref = O.DRGNet(16, 16, 2, 10, 5).double().eval()
    mod = lg.models.DRGNet(16, 16, 2, 10, 5)
    mod.load_state_dict({k: v.float() for k, v in ref.state_dict().items()})
    mod = mod.to(dev).eval()
    want = ref(hb.x.double(), hb.edge_index, hb.batch, edge_weight=ew, num_graphs=hb.num_graphs)
    got = mod(data.x, data.adj_t, data.segments())
    assert float((got.double().cpu() - want).abs().max() / want.abs().max()) < 1e-5
