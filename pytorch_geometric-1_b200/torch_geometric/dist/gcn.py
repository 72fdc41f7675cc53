"""Destination-partitioned GCN forward: per layer, the local block
of X is transformed (X_p · W), all-gathered over NVLink, and each rank
aggregates its own rows with the same fused gather-reduce kernel as the
single-GPU path."""
import torch

from .. import ops
from ..graph import Adj
from .halo import HaloExchange
from .partition import PartitionPlan, balanced_bounds


class DistGraph(object):
    """The rank-local slice of a (loop-augmented, normalised) graph."""

    def __init__(self, edge_index, norm, num_nodes, rank, world_size, group=None, i=1):
        dev = edge_index.device
        deg = torch.bincount(edge_index[i], minlength=num_nodes).cpu()
        self.plan = PartitionPlan(balanced_bounds(deg, world_size), rank)
        ei_local, mask = self.plan.localise(edge_index, i)
        self.norm = None if norm is None else norm[mask].contiguous()
        size = [0, 0]
        size[i], size[1 - i] = self.plan.n_local, self.plan.n_padded
        self.adj = Adj(ei_local.contiguous(), i, size)
        self.halo = HaloExchange(self.plan, group)
        self.num_edges_local = int(mask.sum().item())
        self.device = dev

    def local_rows(self, x_global):
        return x_global[self.plan.lo:self.plan.hi]


def dist_gcn_layer(graph, x_local, weight, bias=None):
    """One GCNConv forward on the partitioned graph: returns the local output block."""
    xw = torch.matmul(x_local, weight)
    full = graph.halo.all_gather(xw)
    return ops.gather_reduce(full, graph.adj, graph.norm, 'sum', bias)
