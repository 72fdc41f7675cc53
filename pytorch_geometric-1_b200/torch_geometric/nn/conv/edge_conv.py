import torch

from ..inits import reset
from .message_passing import MessagePassing


class EdgeConv(MessagePassing):
    r"""Edge convolution :math:`x_i' = \max_{j \in N(i)} h_\Theta(x_i \Vert x_j - x_i)`
    (the reference's README.md:35-49 spells out exactly this class).

    ``nn`` is an arbitrary user module, so the layer takes the generic route:
    native row gathers, ``nn`` in PyTorch, then the native max-with-argmax
    reduction (first maximum wins, as on the reference's CPU path)."""

    def __init__(self, nn, aggr='max', **kwargs):
        super(EdgeConv, self).__init__(aggr=aggr, **kwargs)
        self.nn = nn
        self.reset_parameters()

    def reset_parameters(self):
        reset(self.nn)

    def forward(self, x, edge_index):
        """"""
        x = x.unsqueeze(-1) if x.dim() == 1 else x
        return self.propagate(edge_index, x=x)

    def message(self, x_i, x_j):
        return self.nn(torch.cat([x_i, x_j - x_i], dim=1))

    def __repr__(self):
        return '{}(nn={})'.format(self.__class__.__name__, self.nn)


class DynamicEdgeConv(EdgeConv):
    r"""EdgeConv on a k-NN graph rebuilt from the current features.  The k-NN
    search (``torch_cluster.knn_graph`` in the reference) is outside the hot
    path; a brute-force per-batch search is used."""

    def __init__(self, nn, k, aggr='max', **kwargs):
        super(DynamicEdgeConv, self).__init__(nn=nn, aggr=aggr, **kwargs)
        self.k = k

    def forward(self, x, batch=None):
        """"""
        from ..knn import knn_graph
        edge_index = knn_graph(x, self.k, batch, loop=False, flow=self.flow)
        return super(DynamicEdgeConv, self).forward(x, edge_index)

    def __repr__(self):
        return '{}(nn={}, k={})'.format(self.__class__.__name__, self.nn, self.k)
