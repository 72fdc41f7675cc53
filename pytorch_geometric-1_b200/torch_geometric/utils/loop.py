"""Self-loop utilities (reference call site examples/qm9_nn_conv.py:11).

Output order is the reference's: surviving edges in original order, then the
loops 0..N-1.  The integer results are bit-identical to the CPU reference.
"""
import torch

from .. import _native as nat
from .num_nodes import maybe_num_nodes


def _count_loops(edge_index):
    dev = edge_index.device
    cnt = torch.empty(1, dtype=torch.long, device=dev)
    nat.call('pygb_count_self_loops', dev, nat.ptr(edge_index), edge_index.size(1), nat.ptr(cnt))
    return int(cnt.item())


def _apply(edge_index, N, mode, edge_weight, fill_value):
    dev = nat.require_cuda(edge_index, edge_weight)
    if edge_index.dtype != torch.long or edge_index.dim() != 2 or edge_index.size(0) != 2:
        raise ValueError('edge_index must be an int64 tensor of shape [2, E]')
    edge_index = edge_index if edge_index.is_contiguous() else edge_index.contiguous()
    E = edge_index.size(1)
    n_loops = _count_loops(edge_index) if (mode != nat.LOOPS_ADD and E > 0) else 0
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
    return out, w_out, n_loops


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
    out, w, _ = _apply(edge_index, 0, nat.LOOPS_REMOVE, edge_attr, 0.0)
    return out, w


def add_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    n = maybe_num_nodes(edge_index, num_nodes)
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
    out, w, _ = _apply(edge_index, n, nat.LOOPS_ADD, edge_weight, fill_value)
    return out, w


def add_remaining_self_loops(edge_index, edge_weight=None, fill_value=1, num_nodes=None):
    n = maybe_num_nodes(edge_index, num_nodes)
    if edge_weight is not None:
        assert edge_weight.numel() == edge_index.size(1)
    out, w, _ = _apply(edge_index, n, nat.LOOPS_ADD_REMAINING, edge_weight, fill_value)
    return out, w
