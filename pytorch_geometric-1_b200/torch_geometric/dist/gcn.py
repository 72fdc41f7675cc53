"""Destination-partitioned GCN forward and backward.

Rank ``p`` owns the node range ``[lo, hi)``: its rows of the features, its rows of
every layer's output and the in-edges of those nodes.  Per layer the source
features are exchanged once (all-gather of the rank blocks over NVLink), at the
narrower of the layer's two widths — ``A (X W) = (A X) W`` — and each rank then
aggregates its own rows with the same fused gather-reduce kernel as the
single-GPU path.  The dense transform writes straight into the rank's slot of the
gather buffer, so the exchange is in place (no staging copy).

The aggregation and transform are injectable so that the host logic (partition,
renumbering, exchange, layer order) is testable on CPU with ``gloo``.
"""
import torch

from .halo import AllGatherRows, HaloExchange
from .partition import PartitionPlan, balanced_bounds


class DistGraph(object):
    """The rank-local slice of a (loop-augmented, normalised) graph."""

    def __init__(self, edge_index, norm, num_nodes, rank, world_size, group=None, i=1,
                 build_adj=True):
        deg = torch.bincount(edge_index[i], minlength=num_nodes).cpu()
        self.plan = PartitionPlan(balanced_bounds(deg, world_size), rank)
        ei_local, mask = self.plan.localise(edge_index, i)
        self.edge_index_local = ei_local.contiguous()
        self.norm = None if norm is None else norm[mask].contiguous()
        self.i = i
        size = [0, 0]
        size[i], size[1 - i] = self.plan.n_local, self.plan.n_padded
        self.size = tuple(size)
        self.adj = None
        if build_adj:
            from ..graph import Adj
            self.adj = Adj(self.edge_index_local, i, size)
        self.halo = HaloExchange(self.plan, group)
        self.num_edges_local = int(mask.sum().item())
        self.device = edge_index.device

    def local_rows(self, x_global):
        return x_global[self.plan.lo:self.plan.hi]


def _native_aggregate(graph, full, bias):
    from .. import ops
    return ops.gather_reduce(full, graph.adj, graph.norm, 'sum', bias)


def _native_transform(x, weight, bias, out=None):
    from .. import ops
    if out is not None and ops._gemm_eligible(x, weight):
        return ops.gemm_xw_raw(x, weight, bias, out=out)
    y = ops.linear(x, weight, bias)
    if out is not None:
        out[:, :y.size(1)].copy_(y)
        return out[:, :y.size(1)]
    return y


def dist_gcn_layer(graph, x_local, weight, bias=None, aggregate=_native_aggregate,
                   transform=_native_transform):
    """One GCNConv forward on the partitioned graph; returns the local output block
    ``[n_local, out]``.  Exchanges at ``min(in, out)`` channels."""
    fin, fout = weight.shape
    vec = 16 // x_local.element_size()
    if torch.is_grad_enabled() and (x_local.requires_grad or weight.requires_grad or
                                    (bias is not None and bias.requires_grad)):
        # training: the exchange is an autograd node (all-gather forward, reduce-scatter backward);
        # nothing is written in place.  Weight gradients are per-rank partial sums: all-reduce them
        # (``all_reduce_grads``) before the optimiser step, as with any data-parallel layer.
        if fin < fout:
            agg = aggregate(graph, AllGatherRows.apply(x_local, graph.halo), None)
            return transform(agg, weight, bias)
        y = transform(x_local, weight, None)
        return aggregate(graph, AllGatherRows.apply(y, graph.halo), bias)
    if fin < fout and fin % vec == 0:
        full = graph.halo.all_gather(x_local)                 # [P*block, fin]
        agg = aggregate(graph, full, None)                    # (A X) rows of this rank
        return transform(agg, weight, bias)
    slot = graph.halo.slot(fout, x_local.dtype, x_local.device)   # this rank's rows, in place
    transform(x_local, weight, None, out=slot)
    full = graph.halo.all_gather_slot(fout)                   # [P*block, fout]
    return aggregate(graph, full, bias)


def dist_gcn_forward(graph, x_local, layers, activation=torch.relu, aggregate=_native_aggregate,
                     transform=_native_transform):
    """``layers``: sequence of (weight, bias).  Activation between layers, none after the last."""
    h = x_local
    for k, (w, b) in enumerate(layers):
        h = dist_gcn_layer(graph, h, w, b, aggregate, transform)
        if k + 1 < len(layers):
            h = activation(h)
    return h


def all_reduce_grads(params, group=None):
    """Sum the per-rank partial gradients of replicated parameters (call after ``backward()``)."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    for p in params:
        if p.grad is not None:
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=group)
