from .partition import PartitionPlan, balanced_bounds
from .halo import HaloExchange
from .gcn import DistGraph, dist_gcn_layer, dist_gcn_forward, all_reduce_grads
from .sliced import (SlicePlan, SlicedGraph, CollectiveTransport, PeerTransport,
                     sliced_gcn_forward)

__all__ = ['PartitionPlan', 'balanced_bounds', 'HaloExchange', 'DistGraph', 'dist_gcn_layer',
           'dist_gcn_forward', 'all_reduce_grads', 'SlicePlan', 'SlicedGraph', 'CollectiveTransport',
           'PeerTransport', 'sliced_gcn_forward']
