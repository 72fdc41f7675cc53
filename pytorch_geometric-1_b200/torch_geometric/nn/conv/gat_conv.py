import torch
import torch.nn.functional as F
from torch.nn import Parameter

from ... import ops
from ...graph import adj_for
from ...utils import add_self_loops, remove_self_loops, softmax
from ..inits import glorot, zeros
from .message_passing import MessagePassing, Updated


class GATConv(MessagePassing):
    r"""Graph attention layer with the reference's signature (SURVEY A.10;
    reference call sites ConvexPruning.py:209-214,221 and examples/ppi.py:22-28):
    ``GATConv(in, out, heads=1, concat=True, negative_slope=0.2, dropout=0, bias=True)``
    with parameters ``weight [in, heads*out]``, ``att [1, heads, 2*out]``, ``bias``.

    The attention logits split into a target and a source half, so the whole
    propagate (leaky-relu, per-target softmax, weighted sum) runs as one fused
    kernel over the destination-sorted graph; no ``[E, ...]`` tensor is built.
    With ``dropout > 0`` in training mode the attention coefficients must be
    materialised for the random mask, and the layer takes the generic route.
    """

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2,
                 dropout=0, bias=True, **kwargs):
        super(GATConv, self).__init__(aggr='add', **kwargs)

        self.in_channels = in_channels
        self.out_channels = out_channels
        self.heads = heads
        self.concat = concat
        self.negative_slope = negative_slope
        self.dropout = dropout

        self.weight = Parameter(torch.Tensor(in_channels, heads * out_channels))
        self.att = Parameter(torch.Tensor(1, heads, 2 * out_channels))

        if bias and concat:
            self.bias = Parameter(torch.Tensor(heads * out_channels))
        elif bias and not concat:
            self.bias = Parameter(torch.Tensor(out_channels))
        else:
            self.register_parameter('bias', None)

        self.reset_parameters()

    def reset_parameters(self):
        glorot(self.weight)
        glorot(self.att)
        zeros(self.bias)

    def _fusable(self, x):
        cls = type(self)
        if (cls.message is not GATConv.message or cls.update is not GATConv.update
                or self.node_dim != 0 or (self.training and self.dropout > 0)):
            return False
        xs = x if torch.is_tensor(x) else (x[0] if self.flow == 'source_to_target' else x[1])
        return (torch.is_tensor(xs) and xs.is_cuda and xs.dim() == 2
                and xs.dtype in (torch.float32, torch.bfloat16))

    def _forward_padded(self, x, edge_index, size, Cp):
        """Fused route for head widths the streaming kernels do not take as they are (the fork
        builds ``GATConv(width[i], width[i+1])`` with ``heads=1`` and arbitrary widths,
        ConvexPruning.py:204-214): every head is zero-padded to ``Cp`` channels — in the WEIGHT,
        so X·W comes out of the GEMM already in the padded layout — and the result is sliced
        back.  Exact (padded channels contribute 0 everywhere); update() is applied here."""
        H, C = self.heads, self.out_channels
        fin = self.weight.size(0)
        w = F.pad(self.weight.view(fin, H, C), (0, Cp - C)).reshape(fin, H * Cp)
        a = self.att.view(1, H, 2, C)
        att = F.pad(a, (0, Cp - C)).reshape(1, H, 2 * Cp)
        i, j = (0, 1) if self.flow == 'target_to_source' else (1, 0)
        if torch.is_tensor(x):
            xs = xd = ops.linear(x, w)
            n_src = n_dst = x.size(0)
        else:
            xs = ops.linear(x[j], w)
            xd = None if x[i] is None else ops.linear(x[i], w)
            n_src = xs.size(0)
            n_dst = xd.size(0) if xd is not None else (size[i] if size is not None else n_src)
        if size is not None:
            if size[j] is not None:
                n_src = size[j]
            if size[i] is not None:
                n_dst = size[i]
        sz = [0, 0]
        sz[i], sz[j] = n_dst, n_src
        adj = adj_for(edge_index, i, sz)
        out = ops.gat_propagate(xs, xd, att, None, adj, H, Cp, self.negative_slope)
        out = out.view(-1, H, Cp)[:, :, :C]
        out = out.reshape(-1, H * C) if self.concat else out.mean(dim=1)
        return out + self.bias if self.bias is not None else out

    def forward(self, x, edge_index, size=None):
        """"""
        if size is None and torch.is_tensor(x):
            edge_index = self._with_self_loops(edge_index, x.size(self.node_dim))

        if self._fusable(x):
            esz = 4 if (x if torch.is_tensor(x) else [t for t in x if t is not None][0]).dtype \
                == torch.float32 else 2
            Cp = ops.gat_pad_channels(self.out_channels, esz)
            if Cp != self.out_channels and Cp * esz <= 2048:
                return self._forward_padded(x, edge_index, size, Cp)

        if torch.is_tensor(x):
            x = ops.linear(x, self.weight)
        else:
            x = (None if x[0] is None else ops.linear(x[0], self.weight),
                 None if x[1] is None else ops.linear(x[1], self.weight))

        return self.propagate(edge_index, size=size, x=x)

    def _with_self_loops(self, edge_index, num_nodes):
        # remove_self_loops + add_self_loops, memoised on the edge_index identity
        # (the reference redoes it every forward).
        key = (edge_index.data_ptr(), tuple(edge_index.shape), edge_index._version, num_nodes)
        cache = self.__dict__.get('_loop_cache')
        if cache is not None and cache[0] == key:
            return cache[2]
        if edge_index.is_cuda:
            # remove_self_loops + add_self_loops as ONE native pass (PYGB_LOOPS_REMOVE_ADD): same
            # edge_index' bit for bit, one host read instead of two, ids range-checked on the way
            from ...utils.loop import _apply
            from ... import _native as nat
            ei, _, _ = _apply(edge_index, int(num_nodes), nat.LOOPS_REMOVE_ADD, None, 1.0)
        else:
            ei, _ = remove_self_loops(edge_index)
            ei, _ = add_self_loops(ei, num_nodes=num_nodes)
        self.__dict__['_loop_cache'] = (key, edge_index, ei)
        return ei

    def _fused_propagate(self, edge_index, size, i, kwargs):
        cls = type(self)
        if (cls.message is not GATConv.message or cls.update is not GATConv.update
                or self.node_dim != 0):
            return NotImplemented
        if self.training and self.dropout > 0:
            return NotImplemented
        x = kwargs.get('x')
        j = 1 - i
        if torch.is_tensor(x):
            x_src = x_dst = x
        else:
            x_src, x_dst = x[j], x[i]
        if x_src is None or x_src.dtype not in (torch.float32, torch.bfloat16):
            return NotImplemented
        adj = adj_for(edge_index, i, size)
        fuse_bias = self.bias is not None and self.concat
        out = ops.gat_propagate(x_src, x_dst, self.att, self.bias if fuse_bias else None, adj,
                                self.heads, self.out_channels, self.negative_slope)
        # update(): concat -> [N, H*C] with the bias added by the kernel; else head mean + bias
        if self.concat:
            return Updated(out.view(-1, self.heads * self.out_channels))
        out = out.view(-1, self.heads, self.out_channels).mean(dim=1)
        return Updated(out + self.bias if self.bias is not None else out)

    def message(self, edge_index_i, x_i, x_j, size_i):
        # generic route (dropout on alpha, or user subclasses)
        x_j = x_j.view(-1, self.heads, self.out_channels)
        if x_i is None:
            alpha = (x_j * self.att[:, :, self.out_channels:]).sum(dim=-1)
        else:
            x_i = x_i.view(-1, self.heads, self.out_channels)
            alpha = (torch.cat([x_i, x_j], dim=-1) * self.att).sum(dim=-1)

        alpha = F.leaky_relu(alpha, self.negative_slope)
        alpha = softmax(alpha, edge_index_i, size_i)

        alpha = F.dropout(alpha, p=self.dropout, training=self.training)

        return x_j * alpha.view(-1, self.heads, 1)

    def update(self, aggr_out):
        if self.concat is True:
            aggr_out = aggr_out.view(-1, self.heads * self.out_channels)
        else:
            aggr_out = aggr_out.mean(dim=1)

        if self.bias is not None:
            aggr_out = aggr_out + self.bias
        return aggr_out

    def __getstate__(self):
        state = self.__dict__.copy()
        state.pop('_loop_cache', None)
        return state

    def __repr__(self):
        return '{}({}, {}, heads={})'.format(self.__class__.__name__, self.in_channels,
                                             self.out_channels, self.heads)
