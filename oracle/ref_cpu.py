"""CPU ORACLE — TEST INFRASTRUCTURE ONLY.  **Parity unpinned.**

A plain-torch (CPU) restatement of the reference's message-passing hot path:
``MessagePassing.propagate`` -> gather -> ``message()`` -> ``scatter_`` with
``aggr in {add, mean, max}``, plus GCNConv's normalisation, GATConv's edge
softmax and EdgeConv's max/argmax.

Provenance.  The reference tree (/root/reference) does not vendor the code of
this path; it imports it from the pinned third-party packages
``torch-geometric==1.4.3`` (requirement.txt:7) and ``torch-scatter==2.0.4``
(requirement.txt:3), neither of which is installable here (no network, no
wheel).  This file therefore restates their *published algorithm* as written
down in SURVEY.md Appendix A (A.1-A.11), anchored on the reference's own call
sites (examples/gcn.py:18-27, ConvexPruning.py:180-214, gmm_conv.py:58,131-144,
README.md:35-49, examples/faust.py:19, examples/ppi.py:22-28).  The reference
ships no tests or golden vectors for the path, so the oracle is pinned only by
the hand-derived known-answer tests of SURVEY A.8 (tests/golden/kat.json):
"parity unpinned".

Rules.  Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import this module.  The
product package never does; it fails loudly when the CUDA library is missing.

Every function works on CPU tensors and keeps autograd, so gradients of the
CUDA path can be checked against ``torch.autograd`` through the oracle.
Quirks are numbered Q1..Q8 as in SURVEY Appendix A.
"""
import inspect

import torch
import torch.nn.functional as F

# --------------------------------------------------------------------------
# A.2  index utilities (torch_geometric/utils/{num_nodes,degree,loop}.py)
#      ref call sites: examples/faust.py:19 (degree),
#      examples/qm9_nn_conv.py:11 (remove_self_loops)
# --------------------------------------------------------------------------


def maybe_num_nodes(index, num_nodes=None):
    """A.2: ``index.max().item() + 1 if num_nodes is None else num_nodes``."""
    return int(index.max().item()) + 1 if num_nodes is None else num_nodes


def degree(index, num_nodes=None, dtype=None):
    """A.2 ``degree`` (ref site examples/faust.py:19)."""
    n = maybe_num_nodes(index, num_nodes)
    out = torch.zeros((n,), dtype=dtype, device=index.device)
    return out.scatter_add_(0, index, out.new_ones((index.size(0),)))


def contains_self_loops(edge_index):
    row, col = edge_index
    return bool((row == col).sum().item() > 0)


def remove_self_loops(edge_index, edge_attr=None):
    """A.2: mask ``row != col``; order preserved (ref examples/qm9_nn_conv.py:11)."""
    row, col = edge_index
    mask = row != col
    edge_attr = edge_attr if edge_attr is None else edge_attr[mask]
    return edge_index[:, mask], edge_attr


def add_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    """A.2: append loops 0..N-1; existing loops are NOT removed."""
    n = maybe_num_nodes(edge_index, num_nodes)
    loop = torch.arange(0, n, dtype=torch.long, device=edge_index.device)
    loop = loop.unsqueeze(0).repeat(2, 1)
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
        loop_w = edge_weight.new_full((n,), fill_value)
        edge_weight = torch.cat([edge_weight, loop_w], dim=0)
    return torch.cat([edge_index, loop], dim=1), edge_weight


def add_remaining_self_loops(edge_index, edge_weight=None, fill_value=1,
                             num_nodes=None):
    """A.2: drop existing loops, append loops 0..N-1 keeping an existing loop's
    weight (Q1: several loops on one node -> the last one in edge order)."""
    n = maybe_num_nodes(edge_index, num_nodes)
    row, col = edge_index
    mask = row != col
    inv_mask = ~mask
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
        loop_w = torch.full((n,), fill_value, dtype=edge_weight.dtype,
                            device=edge_index.device)
        remaining = edge_weight[inv_mask]
        if remaining.numel() > 0:
            # Reference: `loop_weight[row[inv_mask]] = remaining`.  With several
            # loops on one node the winner is order dependent (Q1): sequential
            # index_put_ keeps the last one, a multi-threaded one may not.  The
            # oracle pins the sequential answer: the loop with the highest edge id.
            E = edge_index.size(1)
            eid = torch.arange(E, device=row.device)[inv_mask]
            last = torch.full((n,), -1, dtype=torch.long, device=row.device)
            last = last.scatter_reduce(0, row[inv_mask], eid, 'amax', include_self=True)
            has = last >= 0
            loop_w = torch.where(has, edge_weight[last.clamp(min=0)], loop_w)
        edge_weight = torch.cat([edge_weight[mask], loop_w], dim=0)
    loop = torch.arange(0, n, dtype=row.dtype, device=row.device)
    loop = loop.unsqueeze(0).repeat(2, 1)
    edge_index = torch.cat([edge_index[:, mask], loop], dim=1)
    return edge_index, edge_weight


# --------------------------------------------------------------------------
# A.4-A.6  torch_scatter 2.0.4 scatter_{sum,mean,max}
# --------------------------------------------------------------------------


def _broadcast(index, src, dim):
    """A.4 ``broadcast``: unsqueeze then expand_as(src)."""
    if dim < 0:
        dim = src.dim() + dim
    if index.dim() == 1:
        for _ in range(0, dim):
            index = index.unsqueeze(0)
    for _ in range(index.dim(), src.dim()):
        index = index.unsqueeze(-1)
    return index.expand_as(src)


def _out_size(src, index, dim, dim_size):
    size = list(src.size())
    if dim_size is not None:
        size[dim] = dim_size
    elif index.numel() == 0:
        size[dim] = 0
    else:
        size[dim] = int(index.max()) + 1
    return size


def scatter_add(src, index, dim=0, out=None, dim_size=None):
    """A.4 ``scatter_sum``: zeros(size).scatter_add_ — sequential edge order on
    CPU (Q3)."""
    idx = _broadcast(index, src, dim)
    if out is None:
        out = torch.zeros(_out_size(src, index, dim, dim_size), dtype=src.dtype,
                          device=src.device)
    return out.scatter_add(dim, idx, src)


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    """A.5: sum / clamp(count, 1)."""
    out = scatter_add(src, index, dim, out, dim_size)
    dim_size = out.size(dim)
    index_dim = dim
    if index_dim < 0:
        index_dim = index_dim + src.dim()
    if index.dim() <= index_dim:
        index_dim = index.dim() - 1
    ones = torch.ones(index.size(), dtype=src.dtype, device=src.device)
    count = scatter_add(ones, index, index_dim, None, dim_size)
    count = count.clamp(min=1)
    count = _broadcast(count, out, dim)
    if torch.is_floating_point(out):
        return out / count
    return out // count


def scatter_max(src, index, dim=0, out=None, dim_size=None):
    """A.6 ``scatter_max`` -> (out, arg_out), CPU semantics.

    Sequential loop with strict ``>`` => first maximal edge wins (Q4); NaN and
    values <= lowest() never win; rows with no winning entry: value 0 and
    arg = E (Q5).  Vectorised restatement (dim 0 only, which is all the hot
    path uses); ``scatter_max_loop`` below is the literal loop for small cases.
    Differentiable: the value is re-gathered from ``src`` at ``arg``.
    """
    assert dim == 0 or (dim < 0 and dim + src.dim() == 0), 'oracle: dim 0 only'
    E = src.size(0)
    size = _out_size(src, index, 0, dim_size)
    n = size[0]
    flat = src.detach().reshape(E, -1)
    K = flat.size(1)
    lowest = torch.finfo(src.dtype).min if src.is_floating_point() \
        else torch.iinfo(src.dtype).min
    idx = index.reshape(E, 1).expand(E, K)
    best = torch.full((n, K), lowest, dtype=src.dtype)
    if E > 0:
        # NaN never wins under strict '>' -> treat as lowest
        cand = torch.where(torch.isnan(flat), torch.full_like(flat, lowest), flat) \
            if src.is_floating_point() else flat
        best = best.scatter_reduce(0, idx, cand, 'amax', include_self=True)
        wins = (cand == best.gather(0, idx)) & (cand > lowest)
        eid = torch.arange(E, dtype=torch.long).reshape(E, 1).expand(E, K)
        eid = torch.where(wins, eid, torch.full_like(eid, E))
        arg = torch.full((n, K), E, dtype=torch.long)
        arg = arg.scatter_reduce(0, idx, eid, 'amin', include_self=True)
    else:
        arg = torch.full((n, K), E, dtype=torch.long)
    arg = arg.reshape(size)
    empty = arg == E
    if E > 0:
        vals = src.gather(0, arg.clamp(max=E - 1))
    else:
        vals = torch.zeros(size, dtype=src.dtype)
    out_v = vals.masked_fill(empty, 0)
    return out_v, arg


def scatter_max_loop(src, index, dim_size):
    """A.6 literal restatement of scatter_cpu's loop (small cases only)."""
    E = src.size(0)
    flat = src.reshape(E, -1)
    K = flat.size(1)
    lowest = torch.finfo(src.dtype).min
    out = [[lowest] * K for _ in range(dim_size)]
    arg = [[E] * K for _ in range(dim_size)]
    fl = flat.tolist()
    il = index.tolist()
    for e in range(E):
        i = il[e]
        for k in range(K):
            if fl[e][k] > out[i][k]:
                out[i][k] = fl[e][k]
                arg[i][k] = e
    out_t = torch.tensor(out, dtype=src.dtype).reshape(dim_size, K)
    arg_t = torch.tensor(arg, dtype=torch.long).reshape(dim_size, K)
    out_t = out_t.masked_fill(arg_t == E, 0)
    shape = [dim_size] + list(src.shape[1:])
    return out_t.reshape(shape), arg_t.reshape(shape)


def scatter_min(src, index, dim=0, out=None, dim_size=None):
    v, a = scatter_max(-src, index, dim, out, dim_size)
    return -v, a


# --------------------------------------------------------------------------
# A.3  torch_geometric.utils.scatter_
# --------------------------------------------------------------------------


def scatter_(name, src, index, dim=0, dim_size=None):
    """A.3 incl. Q2: 'max' zeroes everything < -10000, 'min' everything > 10000."""
    assert name in ['add', 'mean', 'min', 'max']
    if dim != 0:
        src_t = src.movedim(dim, 0)
        return scatter_(name, src_t, index, 0, dim_size).movedim(0, dim)
    if name == 'add':
        return scatter_add(src, index, 0, None, dim_size)
    if name == 'mean':
        return scatter_mean(src, index, 0, None, dim_size)
    if name == 'max':
        out = scatter_max(src, index, 0, None, dim_size)[0]
        return out.masked_fill(out < -10000, 0)
    out = scatter_min(src, index, 0, None, dim_size)[0]
    return out.masked_fill(out > 10000, 0)


# --------------------------------------------------------------------------
# A.7  torch_geometric.utils.softmax
# --------------------------------------------------------------------------


def softmax(src, index, num_nodes=None):
    """A.7 incl. Q7 (the + 1e-16 in the denominator)."""
    n = maybe_num_nodes(index, num_nodes)
    out = src - scatter_max(src, index, 0, None, n)[0][index]
    out = out.exp()
    out = out / (scatter_add(out, index, 0, None, n)[index] + 1e-16)
    return out


# --------------------------------------------------------------------------
# A.1  MessagePassing
#      ref sites: gmm_conv.py:58,131-144; README.md:33-49
# --------------------------------------------------------------------------

_SPECIAL = ['edge_index', 'edge_index_i', 'edge_index_j', 'size', 'size_i',
            'size_j']
_SIZE_ERR = ('All tensors which should get mapped to the same source or '
             'target nodes must be of same size in dimension 0.')


class MessagePassing(torch.nn.Module):
    """A.1 restated (gather -> message -> scatter_ -> update)."""

    def __init__(self, aggr='add', flow='source_to_target', node_dim=0):
        super().__init__()
        self.aggr = aggr
        assert self.aggr in ['add', 'mean', 'max']
        self.flow = flow
        assert self.flow in ['source_to_target', 'target_to_source']
        self.node_dim = node_dim
        assert self.node_dim >= 0
        args = inspect.getfullargspec(self.message)[0][1:]
        self.__special__ = [(i, a) for i, a in enumerate(args) if a in _SPECIAL]
        self.__margs__ = [a for a in args if a not in _SPECIAL]
        self.__uargs__ = inspect.getfullargspec(self.update)[0][2:]

    def propagate(self, edge_index, size=None, **kwargs):
        dim = self.node_dim
        size = [None, None] if size is None else list(size)
        assert len(size) == 2
        i, j = (0, 1) if self.flow == 'target_to_source' else (1, 0)
        ij = {'_i': i, '_j': j}
        margs = []
        for a in self.__margs__:
            if a[-2:] in ij:
                tmp = kwargs.get(a[:-2], None)
                if tmp is None:
                    margs.append(tmp)
                    continue
                idx = ij[a[-2:]]
                if isinstance(tmp, (tuple, list)):
                    assert len(tmp) == 2
                    if tmp[1 - idx] is not None:
                        if size[1 - idx] is None:
                            size[1 - idx] = tmp[1 - idx].size(dim)
                        if size[1 - idx] != tmp[1 - idx].size(dim):
                            raise ValueError(_SIZE_ERR)
                    tmp = tmp[idx]
                if tmp is None:
                    margs.append(tmp)
                    continue
                if size[idx] is None:
                    size[idx] = tmp.size(dim)
                if size[idx] != tmp.size(dim):
                    raise ValueError(_SIZE_ERR)
                margs.append(torch.index_select(tmp, dim, edge_index[idx]))
            else:
                margs.append(kwargs.get(a, None))
        size[0] = size[1] if size[0] is None else size[0]
        size[1] = size[0] if size[1] is None else size[1]
        kwargs['edge_index'] = edge_index
        kwargs['size'] = size
        for (pos, a) in self.__special__:
            if a[-2:] in ij:
                margs.insert(pos, kwargs[a[:-2]][ij[a[-2:]]])
            else:
                margs.insert(pos, kwargs[a])
        uargs = [kwargs[a] for a in self.__uargs__]
        out = self.message(*margs)
        out = scatter_(self.aggr, out, edge_index[i], dim, dim_size=size[i])
        return self.update(out, *uargs)

    def message(self, x_j):
        return x_j

    def update(self, aggr_out):
        return aggr_out


# --------------------------------------------------------------------------
# A.9  GCNConv   (ref sites examples/gcn.py:18-19,25,27; ConvexPruning.py:180-185)
# --------------------------------------------------------------------------


def gcn_norm(edge_index, num_nodes, edge_weight=None, improved=False,
             dtype=None):
    """A.9 ``GCNConv.norm`` incl. Q8 (degree by ``row`` = source)."""
    if edge_weight is None:
        edge_weight = torch.ones((edge_index.size(1),), dtype=dtype,
                                 device=edge_index.device)
    fill = 1 if not improved else 2
    edge_index, edge_weight = add_remaining_self_loops(edge_index, edge_weight,
                                                       fill, num_nodes)
    row, col = edge_index
    deg = scatter_add(edge_weight, row, 0, None, num_nodes)
    dis = deg.pow(-0.5)
    dis = dis.masked_fill(dis == float('inf'), 0)
    return edge_index, dis[row] * edge_weight * dis[col]


def glorot_(t):
    import math
    a = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
    with torch.no_grad():
        t.uniform_(-a, a)


class GCNConv(MessagePassing):
    def __init__(self, in_channels, out_channels, improved=False, cached=False,
                 bias=True, normalize=True, **kwargs):
        super().__init__(aggr='add', **kwargs)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.improved, self.cached, self.normalize = improved, cached, normalize
        self.weight = torch.nn.Parameter(torch.empty(in_channels, out_channels))
        self.bias = torch.nn.Parameter(torch.empty(out_channels)) if bias else None
        glorot_(self.weight)
        if bias:
            torch.nn.init.zeros_(self.bias)
        self.cached_result = None
        self.cached_num_edges = None

    def forward(self, x, edge_index, edge_weight=None):
        x = torch.matmul(x, self.weight)
        if self.cached and self.cached_result is not None:
            if edge_index.size(1) != self.cached_num_edges:
                raise RuntimeError(
                    'Cached {} number of edges, but found {}. Please disable '
                    'the caching behavior of this layer by removing the '
                    '`cached=True` argument in its constructor.'.format(
                        self.cached_num_edges, edge_index.size(1)))
        if not self.cached or self.cached_result is None:
            self.cached_num_edges = edge_index.size(1)
            if self.normalize:
                edge_index, norm = gcn_norm(edge_index, x.size(self.node_dim),
                                            edge_weight, self.improved, x.dtype)
            else:
                norm = edge_weight
            self.cached_result = edge_index, norm
        edge_index, norm = self.cached_result
        return self.propagate(edge_index, x=x, norm=norm)

    def message(self, x_j, norm):
        return norm.view(-1, 1) * x_j if norm is not None else x_j

    def update(self, aggr_out):
        if self.bias is not None:
            aggr_out = aggr_out + self.bias
        return aggr_out


# --------------------------------------------------------------------------
# A.10  GATConv   (ref sites ConvexPruning.py:209-214,221; examples/ppi.py:22-28)
# --------------------------------------------------------------------------


class GATConv(MessagePassing):
    def __init__(self, in_channels, out_channels, heads=1, concat=True,
                 negative_slope=0.2, dropout=0, bias=True, **kwargs):
        super().__init__(aggr='add', **kwargs)
        self.in_channels, self.out_channels = in_channels, out_channels
        self.heads, self.concat = heads, concat
        self.negative_slope, self.dropout = negative_slope, dropout
        self.weight = torch.nn.Parameter(torch.empty(in_channels, heads * out_channels))
        self.att = torch.nn.Parameter(torch.empty(1, heads, 2 * out_channels))
        if bias and concat:
            self.bias = torch.nn.Parameter(torch.zeros(heads * out_channels))
        elif bias:
            self.bias = torch.nn.Parameter(torch.zeros(out_channels))
        else:
            self.bias = None
        glorot_(self.weight)
        glorot_(self.att)

    def forward(self, x, edge_index, size=None):
        if size is None and torch.is_tensor(x):
            edge_index, _ = remove_self_loops(edge_index)
            edge_index, _ = add_self_loops(edge_index, num_nodes=x.size(self.node_dim))
        if torch.is_tensor(x):
            x = torch.matmul(x, self.weight)
        else:
            x = (None if x[0] is None else torch.matmul(x[0], self.weight),
                 None if x[1] is None else torch.matmul(x[1], self.weight))
        return self.propagate(edge_index, size=size, x=x)

    def message(self, edge_index_i, x_i, x_j, size_i):
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


# --------------------------------------------------------------------------
# A.11  EdgeConv  (ref site README.md:35-49)
# --------------------------------------------------------------------------


class EdgeConv(MessagePassing):
    def __init__(self, nn, aggr='max', **kwargs):
        super().__init__(aggr=aggr, **kwargs)
        self.nn = nn

    def forward(self, x, edge_index):
        x = x.unsqueeze(-1) if x.dim() == 1 else x
        return self.propagate(edge_index, x=x)

    def message(self, x_i, x_j):
        return self.nn(torch.cat([x_i, x_j - x_i], dim=1))
