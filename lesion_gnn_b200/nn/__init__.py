from .conv import GATConv, GCNConv, GINConv, GraphConv, SAGEConv, MLP, Linear, BatchNorm
from .pool import Segments, SortAggregation, global_add_pool, global_max_pool, global_mean_pool, global_mean_max_pool
from .functional import elu, aggregate, gat_attention, linear, segment_pool, sort_pool, bn_act, act_dropout
from .point import PointNetConv, fps, radius, scatter

__all__ = [
    "GATConv", "GCNConv", "GINConv", "GraphConv", "SAGEConv", "MLP", "Linear", "BatchNorm",
    "Segments", "SortAggregation", "global_add_pool", "global_max_pool", "global_mean_pool", "global_mean_max_pool",
    "elu", "aggregate", "gat_attention", "linear", "segment_pool", "sort_pool", "bn_act", "act_dropout",
    "PointNetConv", "fps", "radius", "scatter",
]
