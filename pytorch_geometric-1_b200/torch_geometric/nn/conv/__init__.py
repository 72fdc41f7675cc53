from .message_passing import MessagePassing
from .gcn_conv import GCNConv
from .gat_conv import GATConv
from .edge_conv import EdgeConv, DynamicEdgeConv
from .simple_convs import ChebConv, GraphConv, SGConv, SAGEConv, AGNNConv, GINConv, APPNP
from ._unported import (SplineConv, ARMAConv, RGCNConv, NNConv, GMMConv, PointConv, DNAConv,
                        SAGPooling)

__all__ = [
    'MessagePassing', 'GCNConv', 'GATConv', 'EdgeConv', 'DynamicEdgeConv', 'ChebConv',
    'GraphConv', 'SGConv', 'SAGEConv', 'AGNNConv', 'GINConv', 'APPNP', 'SplineConv', 'ARMAConv',
    'RGCNConv', 'NNConv', 'GMMConv', 'PointConv', 'DNAConv', 'SAGPooling',
]
