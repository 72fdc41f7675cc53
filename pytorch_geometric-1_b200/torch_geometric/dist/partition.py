"""1-D destination-range partition of a graph across ranks.

Rank ``p`` owns the contiguous node range ``[bounds[p], bounds[p+1])``: the
in-edges (CSR rows) of those nodes, their feature rows and their output rows.
Range boundaries balance *edges* (prefix sum of in-degree), not nodes — under a
power-law degree distribution equal node counts would leave ranks with very
different work.  The reference has no counterpart (its only multi-GPU mode is
single-process DataParallel over lists of small graphs).

All planning is host-side integer arithmetic on CPU tensors; it is identical on
every rank, so no communication is needed to agree on it.
"""
import torch


def balanced_bounds(in_degree, world_size):
    """Node boundaries ``[0 = b_0 <= ... <= b_P = N]`` such that every range
    holds ~E/P in-edges (plus one unit per node for the output row)."""
    n = in_degree.numel()
    cost = in_degree.to(torch.int64) + 1
    csum = torch.cumsum(cost, 0)
    total = int(csum[-1]) if n > 0 else 0
    targets = torch.tensor([total * p // world_size for p in range(1, world_size)],
                           dtype=torch.int64)
    cuts = torch.searchsorted(csum, targets, right=False) + 1 if n > 0 else targets * 0
    bounds = [0] + [int(c) for c in cuts.clamp(max=n)] + [n]
    for i in range(1, len(bounds)):  # monotone
        bounds[i] = max(bounds[i], bounds[i - 1])
    return bounds


class PartitionPlan(object):
    """Everything a rank needs to aggregate its rows after the halo all-gather.

    The all-gather works on equally sized blocks, so every rank's feature block
    is padded to ``block`` rows and the gathered matrix has ``P * block`` rows;
    ``to_padded`` maps a global node id to its row in that matrix.
    """

    def __init__(self, bounds, rank):
        self.bounds = list(bounds)
        self.world_size = len(bounds) - 1
        self.rank = rank
        self.num_nodes = bounds[-1]
        self.block = max(max(bounds[p + 1] - bounds[p] for p in range(self.world_size)), 1)
        self.lo, self.hi = bounds[rank], bounds[rank + 1]

    @property
    def n_local(self):
        return self.hi - self.lo

    @property
    def n_padded(self):
        return self.world_size * self.block

    def owner(self, node):
        b = torch.tensor(self.bounds[1:], dtype=torch.int64, device=node.device)
        return torch.searchsorted(b, node, right=True)

    def to_padded(self, node):
        """global id -> row in the gathered [P*block, F] matrix."""
        own = self.owner(node)
        starts = torch.tensor(self.bounds[:-1], dtype=torch.int64, device=node.device)
        return own * self.block + (node - starts[own])

    def local_edges(self, edge_index, i=1):
        """Mask of the edges whose aggregation endpoint (row ``i``) is owned here."""
        t = edge_index[i]
        return (t >= self.lo) & (t < self.hi)

    def localise(self, edge_index, i=1):
        """edge_index of the local sub-problem: targets renumbered to
        ``[0, n_local)``, sources renumbered into the padded gathered layout.
        Returns (edge_index_local [2,E_p], mask)."""
        mask = self.local_edges(edge_index, i)
        ei = edge_index[:, mask]
        out = torch.empty_like(ei)
        out[i] = ei[i] - self.lo
        out[1 - i] = self.to_padded(ei[1 - i])
        return out, mask
