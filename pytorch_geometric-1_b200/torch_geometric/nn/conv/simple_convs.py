"""Operators whose messages are ``w_e * x_j`` or a per-target softmax, i.e. that
fall straight onto the fused propagate routes (selectable
models of the fork: ConvexPruning.py:227-275).  Signatures follow PyG 1.4.3."""
import torch
import torch.nn.functional as F
from torch.nn import Parameter

from ... import ops
from ...graph import adj_for
from ...utils import add_self_loops, remove_self_loops, softmax
from ..inits import glorot, reset, uniform, zeros
from .gcn_conv import GCNConv
from .message_passing import MessagePassing


class _WeightedSum(MessagePassing):
    """``message = norm.view(-1, 1) * x_j`` routed to the fused kernel."""

    def _fused_propagate(self, edge_index, size, i, kwargs):
        cls = type(self)
        if cls.message is not _WeightedSum.message or self.node_dim != 0:
            return NotImplemented
        if self.aggr not in ('add', 'mean'):
            return NotImplemented          # weighted max: generic route (scatter_max kernel)
        x, norm = kwargs.get('x'), kwargs.get('norm')
        if isinstance(x, (tuple, list)):
            x = x[1 - i]
        if not torch.is_tensor(x) or x.dtype not in (torch.float32, torch.bfloat16):
            return NotImplemented
        adj = adj_for(edge_index, i, size)
        return ops.gather_reduce(x, adj, norm, 'mean' if self.aggr == 'mean' else 'sum')

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j if norm is not None else x_j


class ChebConv(_WeightedSum):
    def __init__(self, in_channels, out_channels, K, bias=True, **kwargs):
        super(ChebConv, self).__init__(aggr='add', **kwargs)
        assert K > 0
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = Parameter(torch.Tensor(K, in_channels, out_channels))
        if bias:
            self.bias = Parameter(torch.Tensor(out_channels))
        else:
            self.register_parameter('bias', None)
        self.reset_parameters()

    def reset_parameters(self):
        glorot(self.weight)
        zeros(self.bias)

    @staticmethod
    def norm(edge_index, num_nodes, edge_weight, dtype=None):
        from ...utils import degree  # noqa: F401
        edge_index, edge_weight = remove_self_loops(edge_index, edge_weight)
        if edge_weight is None:
            edge_weight = torch.ones((edge_index.size(1),), dtype=torch.float32,
                                     device=edge_index.device)
        row, col = edge_index
        deg = ops.segment_reduce(edge_weight, row, num_nodes, 'add')
        dis = deg.pow(-0.5)
        dis[dis == float('inf')] = 0
        return edge_index, -dis[row] * edge_weight * dis[col]

    def forward(self, x, edge_index, edge_weight=None):
        edge_index, norm = self.norm(edge_index, x.size(self.node_dim), edge_weight, x.dtype)
        Tx_0 = x
        out = ops.linear(Tx_0, self.weight[0])
        if self.weight.size(0) > 1:
            Tx_1 = self.propagate(edge_index, x=x, norm=norm)
            out = out + ops.linear(Tx_1, self.weight[1])
        for k in range(2, self.weight.size(0)):
            Tx_2 = 2 * self.propagate(edge_index, x=Tx_1, norm=norm) - Tx_0
            out = out + ops.linear(Tx_2, self.weight[k])
            Tx_0, Tx_1 = Tx_1, Tx_2
        if self.bias is not None:
            out = out + self.bias
        return out

    def __repr__(self):
        return '{}({}, {}, K={})'.format(self.__class__.__name__, self.in_channels,
                                         self.out_channels, self.weight.size(0))


class GraphConv(_WeightedSum):
    def __init__(self, in_channels, out_channels, aggr='add', bias=True, **kwargs):
        super(GraphConv, self).__init__(aggr=aggr, **kwargs)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight = Parameter(torch.Tensor(in_channels, out_channels))
        self.lin = torch.nn.Linear(in_channels, out_channels, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        uniform(self.in_channels, self.weight)
        self.lin.reset_parameters()

    def forward(self, x, edge_index, edge_weight=None, size=None):
        h = ops.linear(x, self.weight)
        return self.propagate(edge_index, size=size, x=x, h=h, norm=edge_weight)

    def _fused_propagate(self, edge_index, size, i, kwargs):
        if type(self).message is not GraphConv.message or self.node_dim != 0:
            return NotImplemented
        if self.aggr not in ('add', 'mean'):
            return NotImplemented          # aggr='max': generic route (scatter_max kernel)
        h, norm = kwargs.get('h'), kwargs.get('norm')
        if not torch.is_tensor(h) or h.dtype not in (torch.float32, torch.bfloat16):
            return NotImplemented
        adj = adj_for(edge_index, i, size)
        return ops.gather_reduce(h, adj, norm, 'mean' if self.aggr == 'mean' else 'sum')

    def message(self, h_j, norm):
        return norm.view(-1, 1) * h_j if norm is not None else h_j

    def update(self, aggr_out, x):
        return aggr_out + self.lin(x)

    def __repr__(self):
        return '{}({}, {})'.format(self.__class__.__name__, self.in_channels, self.out_channels)




class SGConv(_WeightedSum):
    def __init__(self, in_channels, out_channels, K=1, cached=False, bias=True, **kwargs):
        super(SGConv, self).__init__(aggr='add', **kwargs)
        self.in_channels, self.out_channels, self.K, self.cached = in_channels, out_channels, K, cached
        self.lin = torch.nn.Linear(in_channels, out_channels, bias=bias)
        self.reset_parameters()

    def reset_parameters(self):
        self.lin.reset_parameters()
        self.cached_result = None
        self.cached_num_edges = None

    def forward(self, x, edge_index, edge_weight=None):
        if not self.cached or self.cached_result is None:
            self.cached_num_edges = edge_index.size(1)
            edge_index, norm = GCNConv.norm(edge_index, x.size(self.node_dim), edge_weight,
                                            dtype=x.dtype)
            for _ in range(self.K):
                x = self.propagate(edge_index, x=x, norm=norm)
            self.cached_result = x
        return self.lin(self.cached_result)

    def __repr__(self):
        return '{}({}, {}, K={})'.format(self.__class__.__name__, self.in_channels,
                                         self.out_channels, self.K)


class APPNP(_WeightedSum):
    def __init__(self, K, alpha, bias=True, **kwargs):
        super(APPNP, self).__init__(aggr='add', **kwargs)
        self.K, self.alpha = K, alpha

    def forward(self, x, edge_index, edge_weight=None):
        edge_index, norm = GCNConv.norm(edge_index, x.size(self.node_dim), edge_weight,
                                        dtype=x.dtype)
        hidden = x
        for _ in range(self.K):
            x = self.propagate(edge_index, x=x, norm=norm)
            x = x * (1 - self.alpha) + self.alpha * hidden
        return x

    def __repr__(self):
        return '{}(K={}, alpha={})'.format(self.__class__.__name__, self.K, self.alpha)


class SAGEConv(MessagePassing):
    def __init__(self, in_channels, out_channels, normalize=False, concat=False, bias=True,
                 **kwargs):
        super(SAGEConv, self).__init__(aggr='mean', **kwargs)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.normalize, self.concat = normalize, concat
        in_channels = 2 * in_channels if concat else in_channels
        self.weight = Parameter(torch.Tensor(in_channels, out_channels))
        if bias:
            self.bias = Parameter(torch.Tensor(out_channels))
        else:
            self.register_parameter('bias', None)
        self.reset_parameters()

    def reset_parameters(self):
        uniform(self.weight.size(0), self.weight)
        uniform(self.weight.size(0), self.bias)

    def forward(self, x, edge_index, size=None):
        if not self.concat and torch.is_tensor(x):
            edge_index, _ = remove_self_loops(edge_index)
            edge_index, _ = add_self_loops(edge_index, num_nodes=x.size(self.node_dim))
        return self.propagate(edge_index, size=size, x=x)

    def update(self, aggr_out, x):
        if self.concat and torch.is_tensor(x):
            aggr_out = torch.cat([x, aggr_out], dim=-1)
        elif self.concat and isinstance(x, (tuple, list)):
            assert x[1] is not None
            aggr_out = torch.cat([x[1], aggr_out], dim=-1)
        aggr_out = ops.linear(aggr_out, self.weight)
        if self.bias is not None:
            aggr_out = aggr_out + self.bias
        if self.normalize:
            aggr_out = F.normalize(aggr_out, p=2, dim=-1)
        return aggr_out

    def __repr__(self):
        return '{}({}, {})'.format(self.__class__.__name__, self.in_channels, self.out_channels)


class GINConv(MessagePassing):
    def __init__(self, nn, eps=0, train_eps=False, **kwargs):
        super(GINConv, self).__init__(aggr='add', **kwargs)
        self.nn = nn
        self.initial_eps = eps
        if train_eps:
            self.eps = torch.nn.Parameter(torch.Tensor([eps]))
        else:
            self.register_buffer('eps', torch.Tensor([eps]))
        self.reset_parameters()

    def reset_parameters(self):
        reset(self.nn)
        self.eps.data.fill_(self.initial_eps)

    def forward(self, x, edge_index):
        x = x.unsqueeze(-1) if x.dim() == 1 else x
        edge_index, _ = remove_self_loops(edge_index)
        return self.nn((1 + self.eps) * x + self.propagate(edge_index, x=x))

    def __repr__(self):
        return '{}(nn={})'.format(self.__class__.__name__, self.nn)


class AGNNConv(MessagePassing):
    def __init__(self, requires_grad=True, **kwargs):
        super(AGNNConv, self).__init__(aggr='add', **kwargs)
        self.requires_grad = requires_grad
        if requires_grad:
            self.beta = Parameter(torch.Tensor(1))
        else:
            self.register_buffer('beta', torch.ones(1))
        self.reset_parameters()

    def reset_parameters(self):
        if self.requires_grad:
            self.beta.data.fill_(1)

    def forward(self, x, edge_index):
        edge_index, _ = remove_self_loops(edge_index)
        edge_index, _ = add_self_loops(edge_index, num_nodes=x.size(self.node_dim))
        x_norm = F.normalize(x, p=2, dim=-1)
        return self.propagate(edge_index, x=x, x_norm=x_norm, num_nodes=x.size(self.node_dim))

    def message(self, edge_index_i, x_j, x_norm_i, x_norm_j, num_nodes):
        beta = self.beta if self.requires_grad else self._buffers['beta']
        alpha = beta * (x_norm_i * x_norm_j).sum(dim=-1)
        alpha = softmax(alpha, edge_index_i, num_nodes)
        return x_j * alpha.view(-1, 1)

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)
