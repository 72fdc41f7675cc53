"""Self-loop utilities (reference call site examples/qm9_nn_conv.py:11).

Output order is the reference's: surviving edges in original order, then the
loops 0..N-1.  The integer results are bit-identical to the CPU reference.
"""
import torch

from .. import _native as nat
from .num_nodes import maybe_num_nodes


def _count_loops(edge_index, num_nodes=None):
    """Number of self-loops; with ``num_nodes`` the same pass also range-checks every id
    (IndexError) — one device->host read for both."""
    dev = edge_index.device
    if num_nodes is None:
        cnt = torch.empty(1, dtype=torch.long, device=dev)
        nat.call('pygb_count_self_loops', dev, nat.ptr(edge_index), edge_index.size(1),
                 nat.ptr(cnt))
        return int(cnt.item())
    cnt = torch.empty(2, dtype=torch.long, device=dev)
    nat.call('pygb_scan_edge_index', dev, nat.ptr(edge_index), edge_index.size(1),
             int(num_nodes), nat.ptr(cnt))
    loops, bad = cnt.tolist()
    if bad:
        raise IndexError('edge_index holds {} edge(s) with a node id outside [0, {})'.format(
            bad, int(num_nodes)))
    return int(loops)


def _apply(edge_index, N, mode, edge_weight, fill_value):
    dev = nat.require_cuda(edge_index, edge_weight)
    if edge_index.dtype != torch.long or edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError('edge_index must be an int64 tensor of shape [2, E]')
    edge_index = edge_index if edge_index.is_contiguous() else edge_index.contiguous()
    E = edge_index.size(1)
    n_loops = 0
    if mode != nat.LOOPS_ADD and E > 0:
        n_loops = _count_loops(edge_index, N if mode != nat.LOOPS_REMOVE else None)
    keep = E - n_loops if mode != nat.LOOPS_ADD else E
    e_out = keep + (N if mode != nat.LOOPS_REMOVE else 0)
    out = torch.empty((2, e_out), dtype=torch.long, device=dev)
    w = w_out = None
    if edge_weight is not None:
        if edge_weight.dtype != torch.float32:
            raise TypeError('edge weights must be float32 on the native path')
        w = edge_weight.contiguous()
        w_out = torch.empty((e_out,), dtype=torch.float32, device=dev)
    ws_bytes = nat.workspace('pygb_apply_loops_workspace', E, N)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    nat.call('pygb_apply_loops', dev, nat.ptr(edge_index), E, N, mode, n_loops, nat.ptr(w),
             float(fill_value), nat.ptr(out), nat.ptr(w_out), nat.ptr(ws), ws_bytes)
    if mode != nat.LOOPS_ADD and mode != nat.LOOPS_REMOVE and E > 0:
        # every id was range-checked by the scan above: the CSR build can skip its own check
        out._pygb_valid_n = int(N)
        out._pygb_valid_version = out._version
    return out, w_out, n_loops


def _needs_grad(w):
    return w is not None and w.requires_grad and torch.is_grad_enabled()


def _weights_with_grad(edge_index, w, mode, fill_value, N):
    """The float side of ``_apply`` as differentiable indexing of ``w`` (pure copies, so the
    values equal the native kernel's): used only when ``w.requires_grad`` — the native call
    still produces the integer ``edge_index'``.  Every source element is read at most once,
    so the backward (index_put with unique indices) is deterministic."""
    row, col = edge_index[0], edge_index[1]
    if mode == nat.LOOPS_ADD:
        return torch.cat([w, w.new_full((N,), fill_value)])
    mask = row != col
    kept = w[mask]
    if mode == nat.LOOPS_REMOVE:
        return kept
    loop_w = w.new_full((N,), fill_value)
    if mode == nat.LOOPS_ADD_REMAINING:
        inv = ~mask
        pos = torch.arange(row.numel(), device=row.device)[inv]
        # several existing loops on one node: the highest edge id wins (Q1, last write)
        last = torch.full((N,), -1, dtype=torch.long, device=row.device)
        last.scatter_reduce_(0, row[inv], pos, 'amax', include_self=True)
        has = last >= 0
        loop_w = torch.where(has, w[last.clamp(min=0)], loop_w)
    return torch.cat([kept, loop_w])


def contains_self_loops(edge_index):
    nat.require_cuda(edge_index)
    ei = edge_index if edge_index.is_contiguous() else edge_index.contiguous()
    return _count_loops(ei) > 0


def remove_self_loops(edge_index, edge_attr=None):
    if edge_attr is not None and (edge_attr.dim() != 1 or edge_attr.dtype != torch.float32):
        # multi-dimensional attributes: compact an edge-id vector natively, then index
        ids = torch.arange(edge_index.size(1), device=edge_index.device, dtype=torch.float32)
        if edge_index.size(1) >= (1 << 24):
            raise ValueError('remove_self_loops: edge_attr of this shape needs E < 2^24')
        out, kept, _ = _apply(edge_index, 0, nat.LOOPS_REMOVE, ids, 0.0)
        return out, edge_attr[kept.long()]
    if _needs_grad(edge_attr):
        out, _, _ = _apply(edge_index, 0, nat.LOOPS_REMOVE, None, 0.0)
        return out, _weights_with_grad(edge_index, edge_attr, nat.LOOPS_REMOVE, 0.0, 0)
    out, w, _ = _apply(edge_index, 0, nat.LOOPS_REMOVE, edge_attr, 0.0)
    return out, w


def add_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    n = maybe_num_nodes(edge_index, num_nodes)
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
    if _needs_grad(edge_weight):
        out, _, _ = _apply(edge_index, n, nat.LOOPS_ADD, None, fill_value)
        return out, _weights_with_grad(edge_index, edge_weight, nat.LOOPS_ADD, fill_value, n)
    out, w, _ = _apply(edge_index, n, nat.LOOPS_ADD, edge_weight, fill_value)
    return out, w


def add_remaining_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    n = maybe_num_nodes(edge_index, num_nodes)
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
    if _needs_grad(edge_weight):
        out, _, _ = _apply(edge_index, n, nat.LOOPS_ADD_REMAINING, None, fill_value)
        return out, _weights_with_grad(edge_index, edge_weight, nat.LOOPS_ADD_REMAINING,
                                       fill_value, n)
    out, w, _ = _apply(edge_index, n, nat.LOOPS_ADD_REMAINING, edge_weight, fill_value)
    return out, w
