"""``MessagePassing`` — the hot path.

Same user contract as the reference (reference witnesses: gmm_conv.py:58,131-144
and README.md:33-49): subclasses override ``message(self, <name>_i, <name>_j,
...)`` / ``update(self, aggr_out, ...)`` and call ``self.propagate(edge_index,
size=None, **kwargs)``.

What differs is below the API.  The reference materialises an ``[E, F]`` tensor
per ``_i``/``_j`` argument, calls ``message`` and scatters with float atomics.
Here ``propagate`` first asks for a fused route:

* default ``message`` (``x_j``)            -> one segmented gather-reduce / gather-max
* ``GCNConv.message`` (``norm * x_j``)     -> the same kernel with the edge weight fused
* ``GATConv.message``                      -> fused leaky-relu/softmax/weighted-sum kernel

and only for user-defined ``message`` functions falls back to the generic route
(native gathers -> ``message`` in PyTorch -> native atomic-free ``scatter_``).
"""
import inspect

import torch

from ... import ops
from ...graph import adj_for
from ...utils.scatter import scatter_

special_args = ['edge_index', 'edge_index_i', 'edge_index_j', 'size', 'size_i', 'size_j']
__size_error_msg__ = ('All tensors which should get mapped to the same source '
                      'or target nodes must be of same size in dimension 0.')


class Updated(object):
    """Return type of a fused route that has already applied ``update()`` (bias fused
    into the kernel epilogue): ``propagate`` returns ``value`` as is.  Call-local, so
    conv modules keep no per-call state and stay re-entrant under ``parallel_apply``."""
    __slots__ = ('value',)

    def __init__(self, value):
        self.value = value


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr='add', flow='source_to_target', node_dim=0):
        super(MessagePassing, self).__init__()

        self.aggr = aggr
        assert self.aggr in ['add', 'mean', 'max']

        self.flow = flow
        assert self.flow in ['source_to_target', 'target_to_source']

        self.node_dim = node_dim
        assert self.node_dim >= 0

        names = inspect.getfullargspec(self.message)[0][1:]
        self.__special_args__ = [(pos, n) for pos, n in enumerate(names) if n in special_args]
        self.__message_args__ = [n for n in names if n not in special_args]
        self.__update_args__ = inspect.getfullargspec(self.update)[0][2:]

    # -- fused routes ------------------------------------------------------
    def _fused_propagate(self, edge_index, size, i, kwargs):
        """Return the aggregated (pre-``update``) tensor, ``Updated(final)`` when the
        route also covered ``update()``, or ``NotImplemented``."""
        cls = type(self)
        if cls.message is not MessagePassing.message or self.node_dim != 0:
            return NotImplemented
        x = kwargs.get('x', None)
        j = 1 - i
        if isinstance(x, (tuple, list)):
            x = x[j]
        if not torch.is_tensor(x) or not x.is_floating_point():
            return NotImplemented
        if x.dtype not in (torch.float32, torch.bfloat16):
            return NotImplemented
        adj = adj_for(edge_index, i, size)
        if self.aggr == 'max':
            return ops.gather_max(x, adj)
        return ops.gather_reduce(x, adj, None, 'mean' if self.aggr == 'mean' else 'sum')

    def _resolve_sizes(self, size, i, j, kwargs):
        """Sizes implied by the ``_i``/``_j`` arguments (A.1 steps 2-3), without gathering."""
        dim = self.node_dim
        ij = {'_i': i, '_j': j}
        for arg in self.__message_args__:
            if arg[-2:] not in ij:
                continue
            tmp = kwargs.get(arg[:-2], None)
            if tmp is None:
                continue
            idx = ij[arg[-2:]]
            if isinstance(tmp, (tuple, list)):
                assert len(tmp) == 2
                if tmp[1 - idx] is not None:
                    if size[1 - idx] is None:
                        size[1 - idx] = tmp[1 - idx].size(dim)
                    if size[1 - idx] != tmp[1 - idx].size(dim):
                        raise ValueError(__size_error_msg__)
                tmp = tmp[idx]
            if tmp is None:
                continue
            if size[idx] is None:
                size[idx] = tmp.size(dim)
            if size[idx] != tmp.size(dim):
                raise ValueError(__size_error_msg__)
        size[0] = size[1] if size[0] is None else size[0]
        size[1] = size[0] if size[1] is None else size[1]
        return size

    # -- the reference contract ----------------------------------------------
    def propagate(self, edge_index, size=None, **kwargs):
        r"""Gather ``*_i`` / ``*_j`` arguments by ``edge_index``, call
        :meth:`message`, aggregate with ``self.aggr`` onto the target nodes and
        call :meth:`update`.

        Args:
            edge_index (LongTensor): ``[2, E]``; row 0 = source ``j``, row 1 =
                target ``i`` for ``flow='source_to_target'``.
            size (list or tuple, optional): ``[N_source, N_target]`` for
                bipartite graphs; inferred from the node tensors when ``None``.
            **kwargs: everything ``message`` / ``update`` ask for by name.
        """
        dim = self.node_dim
        size = [None, None] if size is None else list(size)
        assert len(size) == 2

        i, j = (0, 1) if self.flow == 'target_to_source' else (1, 0)
        ij = {'_i': i, '_j': j}

        size = self._resolve_sizes(size, i, j, kwargs)
        update_args = None

        out = NotImplemented
        if size[0] is not None and size[1] is not None:
            out = self._fused_propagate(edge_index, size, i, kwargs)

        if isinstance(out, Updated):
            return out.value

        if out is NotImplemented:
            message_args = []
            for arg in self.__message_args__:
                if arg[-2:] in ij:
                    tmp = kwargs.get(arg[:-2], None)
                    idx = ij[arg[-2:]]
                    if isinstance(tmp, (tuple, list)):
                        tmp = tmp[idx]
                    if tmp is None:
                        message_args.append(None)
                    elif dim == 0:
                        message_args.append(ops.index_select0(tmp, edge_index[idx]))
                    else:
                        moved = ops.index_select0(tmp.movedim(dim, 0), edge_index[idx])
                        message_args.append(moved.movedim(0, dim))
                else:
                    message_args.append(kwargs.get(arg, None))

            kwargs['edge_index'] = edge_index
            kwargs['size'] = size
            for (pos, arg) in self.__special_args__:
                if arg[-2:] in ij:
                    message_args.insert(pos, kwargs[arg[:-2]][ij[arg[-2:]]])
                else:
                    message_args.insert(pos, kwargs[arg])
            update_args = [kwargs[arg] for arg in self.__update_args__]

            out = self.message(*message_args)
            out = scatter_(self.aggr, out, edge_index[i], dim, dim_size=size[i])
        else:
            kwargs['edge_index'] = edge_index
            kwargs['size'] = size
            update_args = [kwargs[arg] for arg in self.__update_args__]

        out = self.update(out, *update_args)
        return out

    def message(self, x_j):  # pragma: no cover
        r"""Constructs the message of edge :math:`(j, i)`; any ``propagate`` kwarg
        can be requested, ``_i`` / ``_j`` suffixes select target / source rows."""
        return x_j

    def update(self, aggr_out):  # pragma: no cover
        r"""Updates node embeddings; first argument is the aggregated output."""
        return aggr_out
