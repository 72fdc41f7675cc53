import torch
from torch.nn import Parameter

from ... import _native as nat
from ... import ops
from ...graph import adj_for, build_csr
from ...utils import add_remaining_self_loops
from ..inits import glorot, zeros
from .message_passing import MessagePassing, Updated


class GCNConv(MessagePassing):
    r"""Graph convolution :math:`X' = \hat D^{-1/2} \hat A \hat D^{-1/2} X \Theta`
    with the reference's constructor and attribute names (reference
    call sites examples/gcn.py:18-19, ConvexPruning.py:180-185, which also read
    and write ``.weight`` / ``.bias``).

    Args:
        in_channels (int), out_channels (int)
        improved (bool): :math:`\hat A = A + 2I`.
        cached (bool): compute the normalisation (and, here, the sorted graph)
            on the first call only.
        bias (bool), normalize (bool), **kwargs: as in ``MessagePassing``.
    """

    def __init__(self, in_channels, out_channels, improved=False, cached=False, bias=True,
                 normalize=True, **kwargs):
        super(GCNConv, self).__init__(aggr='add', **kwargs)

        self.in_channels = in_channels
        self.out_channels = out_channels
        self.improved = improved
        self.cached = cached
        self.normalize = normalize

        self.weight = Parameter(torch.Tensor(in_channels, out_channels))
        if bias:
            self.bias = Parameter(torch.Tensor(out_channels))
        else:
            self.register_parameter('bias', None)

        self.reset_parameters()

    def reset_parameters(self):
        glorot(self.weight)
        zeros(self.bias)
        self.cached_result = None
        self.cached_num_edges = None

    @staticmethod
    def norm(edge_index, num_nodes, edge_weight=None, improved=False, dtype=None):
        """``add_remaining_self_loops`` + degree by source row + ``D^-1/2 A D^-1/2``
        edge weights.  Integer outputs and unit-weight degrees are bit-exact with
        the reference; weighted degrees are summed in ascending edge order."""
        dev = nat.require_cuda(edge_index, edge_weight)
        fill_value = 1 if not improved else 2
        if edge_weight is not None and edge_weight.requires_grad and torch.is_grad_enabled():
            # learnable edge weights: the same formula from differentiable pieces — the native
            # scatter_add (deterministic, gather backward) and the native row gather
            # (atomic-free backward); values match the fused kernels up to fp32 rounding
            edge_index, w = add_remaining_self_loops(edge_index, edge_weight.float(), fill_value,
                                                     num_nodes)
            row, col = edge_index[0], edge_index[1]
            deg = ops.segment_reduce(w, row, int(num_nodes), 'add')
            dis = deg.pow(-0.5)
            dis = dis.masked_fill(dis == float('inf'), 0)
            norm = ops.index_select0(dis, row) * w * ops.index_select0(dis, col)
            return edge_index, norm if dtype in (None, torch.float32) else norm.to(dtype)
        unit = edge_weight is None and not improved
        if edge_weight is None and not unit:
            edge_weight = torch.ones((edge_index.size(1),), dtype=torch.float32, device=dev)
        w_in = None if unit else edge_weight.float()
        edge_index, w = add_remaining_self_loops(edge_index, w_in, fill_value, num_nodes)
        E = edge_index.size(1)
        N = int(num_nodes)
        deg = torch.empty((N,), dtype=torch.float32, device=dev)
        if unit:
            nat.call('pygb_degree', dev, nat.ptr(edge_index), E, N, 1, nat.ptr(deg))
        else:
            csr = build_csr(edge_index[0], None, N)
            nat.call('pygb_weighted_degree', dev, nat.ptr(w), nat.ptr(csr.rowptr),
                     nat.ptr(csr.perm), N, nat.ptr(deg))
        norm = torch.empty((E,), dtype=torch.float32, device=dev)
        nat.call('pygb_gcn_norm', dev, nat.ptr(edge_index), nat.ptr(w), E, N, nat.ptr(deg),
                 nat.ptr(norm))
        if dtype is not None and dtype != torch.float32:
            norm = norm.to(dtype)
        elif unit:
            # unit weights: the sorted-order copy of norm can be rebuilt from the degrees
            # (graph.Adj.permuted) instead of gathered through the permutation
            norm._pygb_unit_deg = (deg, norm._version)
        return edge_index, norm

    def _aggregate_first(self, x):
        """``A (X W) = (A X) W``: aggregate at the narrower of the two widths.  Only
        for the stock message/update on 2-D inputs whose rows are already 16-byte
        multiples (the 128-bit gather path); sums differ from the other order by
        fp32 rounding only (~1e-7 relative)."""
        cls = type(self)
        if (cls.message is not GCNConv.message or cls.update is not GCNConv.update
                or self.node_dim != 0 or not torch.is_tensor(x) or x.dim() != 2):
            return False
        if x.dtype not in (torch.float32, torch.bfloat16) or not x.is_cuda:
            return False
        vec = 16 // x.element_size()
        return self.in_channels < self.out_channels and self.in_channels % vec == 0

    def forward(self, x, edge_index, edge_weight=None):
        """"""
        agg_first = self._aggregate_first(x)
        if not agg_first:
            # X·W on the tensor cores (ops.linear); the output row is padded to 16 bytes so
            # that the aggregation stays on its 128-bit path for widths like 7 or 47.
            x = ops.linear(x, self.weight)

        if self.cached and self.cached_result is not None:
            if edge_index.size(1) != self.cached_num_edges:
                raise RuntimeError(
                    'Cached {} number of edges, but found {}. Please '
                    'disable the caching behavior of this layer by removing '
                    'the `cached=True` argument in its constructor.'.format(
                        self.cached_num_edges, edge_index.size(1)))

        if not self.cached or self.cached_result is None:
            self.cached_num_edges = edge_index.size(1)
            if self.normalize:
                edge_index, norm = self.norm(edge_index, x.size(self.node_dim), edge_weight,
                                             self.improved, x.dtype)
            else:
                norm = edge_weight
            self.cached_result = edge_index, norm

        edge_index, norm = self.cached_result

        if agg_first:
            # stock message/update (checked by _aggregate_first): the fused aggregation is
            # called directly, the bias follows the transform — no per-call state on self
            i = 0 if self.flow == 'target_to_source' else 1
            n = x.size(0)
            h = ops.gather_reduce(x, adj_for(edge_index, i, (n, n)), norm, 'sum', None)
            return ops.linear(h, self.weight, self.bias)
        return self.propagate(edge_index, x=x, norm=norm)

    def _fused_propagate(self, edge_index, size, i, kwargs):
        cls = type(self)
        if (cls.message is not GCNConv.message or cls.update is not GCNConv.update
                or self.node_dim != 0):
            return NotImplemented
        x, norm = kwargs.get('x'), kwargs.get('norm')
        if not torch.is_tensor(x) or x.dtype not in (torch.float32, torch.bfloat16):
            return NotImplemented
        adj = adj_for(edge_index, i, size)
        # update() (= + bias) runs in the kernel epilogue
        return Updated(ops.gather_reduce(x, adj, norm, 'sum', self.bias))

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j if norm is not None else x_j

    def update(self, aggr_out):
        if self.bias is not None:
            aggr_out = aggr_out + self.bias
        return aggr_out

    def __getstate__(self):
        # device caches do not travel with a pickled module (SURVEY §5 checkpointing)
        state = self.__dict__.copy()
        state['cached_result'] = None
        state['cached_num_edges'] = None
        return state

    def __repr__(self):
        return '{}({}, {})'.format(self.__class__.__name__, self.in_channels, self.out_channels)
