import torch

from .. import _native as nat
from .num_nodes import maybe_num_nodes


def degree(index, num_nodes=None, dtype=None):
    """Number of occurrences of every node id in ``index`` (reference
    call site examples/faust.py:19).  Integer atomics only: exact and deterministic."""
    dev = nat.require_cuda(index)
    n = maybe_num_nodes(index, num_nodes)
    dtype = torch.get_default_dtype() if dtype is None else dtype
    index = index if index.is_contiguous() else index.contiguous()
    if dtype == torch.float32 or dtype == torch.long:
        out = torch.empty((n,), dtype=dtype, device=dev)
        nat.call('pygb_degree', dev, nat.ptr(index), index.numel(), n,
                 1 if dtype == torch.float32 else 0, nat.ptr(out))
        return out
    out = torch.empty((n,), dtype=torch.long, device=dev)
    nat.call('pygb_degree', dev, nat.ptr(index), index.numel(), n, 0, nat.ptr(out))
    return out.to(dtype)
