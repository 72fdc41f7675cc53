"""``TopKPooling`` (the fork's ``topk_pool_Net``,
ConvexPruning.py:301-338, enzymes_topk_pool_copy.py:26-30): per-graph top-k node
selection by a learnt projection score, then induced-subgraph filtering of the
edges.  Signatures follow PyG 1.4.3.  The selection is a per-graph sort (framework
ops on the device); the kernels of the hot path enter through the pooling calls
(``global_max_pool`` / ``global_mean_pool``) and convolutions around it."""


import torch
from torch.nn import Parameter

from ..utils import scatter_, softmax
from .inits import uniform


def topk(x, ratio, batch, min_score=None, tol=1e-7):
    """Indices of the nodes kept: graphs in order, inside a graph by descending score."""
    if min_score is not None:
        scores_max = scatter_('max', x, batch, dim=0, dim_size=int(batch.max()) + 1)[batch] - tol
        scores_min = scores_max.clamp(max=min_score)
        return (x > scores_min).nonzero().view(-1)
    num_nodes = torch.bincount(batch, minlength=int(batch.max()) + 1)
    batch_size, max_num_nodes = num_nodes.size(0), int(num_nodes.max())
    cum = torch.cat([num_nodes.new_zeros(1), num_nodes.cumsum(0)[:-1]], dim=0)
    index = torch.arange(batch.size(0), dtype=torch.long, device=x.device)
    index = (index - cum[batch]) + (batch * max_num_nodes)
    dense = x.new_full((batch_size * max_num_nodes,), torch.finfo(x.dtype).min)
    dense[index] = x
    dense = dense.view(batch_size, max_num_nodes)
    _, perm = dense.sort(dim=-1, descending=True)
    perm = perm + cum.view(-1, 1)
    perm = perm.view(-1)
    k = (ratio * num_nodes.to(torch.float)).ceil().to(torch.long)
    mask = [torch.arange(int(k[i]), dtype=torch.long, device=x.device) + i * max_num_nodes
            for i in range(batch_size)]
    return perm[torch.cat(mask, dim=0)]


def filter_adj(edge_index, edge_attr, perm, num_nodes=None):
    """Edges of the subgraph induced by ``perm``, renumbered to positions in ``perm``."""
    num_nodes = int(edge_index.max()) + 1 if num_nodes is None else num_nodes
    mask = perm.new_full((num_nodes,), -1)
    mask[perm] = torch.arange(perm.size(0), dtype=torch.long, device=perm.device)
    row, col = mask[edge_index[0]], mask[edge_index[1]]
    keep = (row >= 0) & (col >= 0)
    row, col = row[keep], col[keep]
    if edge_attr is not None:
        edge_attr = edge_attr[keep]
    return torch.stack([row, col], dim=0), edge_attr


class TopKPooling(torch.nn.Module):
    def __init__(self, in_channels, ratio=0.5, min_score=None, multiplier=1,
                 nonlinearity=torch.tanh):
        super(TopKPooling, self).__init__()
        self.in_channels, self.ratio, self.min_score = in_channels, ratio, min_score
        self.multiplier, self.nonlinearity = multiplier, nonlinearity
        self.weight = Parameter(torch.Tensor(1, in_channels))
        self.reset_parameters()

    def reset_parameters(self):
        uniform(self.in_channels, self.weight)

    def forward(self, x, edge_index, edge_attr=None, batch=None, attn=None):
        if batch is None:
            batch = edge_index.new_zeros(x.size(0))
        attn = x if attn is None else attn
        attn = attn.unsqueeze(-1) if attn.dim() == 1 else attn
        score = (attn * self.weight).sum(dim=-1)
        if self.min_score is None:
            score = self.nonlinearity(score / self.weight.norm(p=2, dim=-1))
        else:
            score = softmax(score, batch)
        perm = topk(score, self.ratio, batch, self.min_score)
        x = x[perm] * score[perm].view(-1, 1)
        x = self.multiplier * x if self.multiplier != 1 else x
        batch = batch[perm]
        edge_index, edge_attr = filter_adj(edge_index, edge_attr, perm, num_nodes=score.size(0))
        return x, edge_index, edge_attr, batch, perm, score[perm]

    def __repr__(self):
        return '{}({}, {}={}, multiplier={})'.format(
            self.__class__.__name__, self.in_channels,
            'ratio' if self.min_score is None else 'min_score',
            self.ratio if self.min_score is None else self.min_score, self.multiplier)


__all__ = ['TopKPooling', 'topk', 'filter_adj']

