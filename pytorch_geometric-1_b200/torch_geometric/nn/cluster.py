"""``fps``, ``radius`` and ``radius_graph`` (stand in for ``torch_cluster``'s functions of the
same names: examples/pointnet++.py:9,39-41,46-48, README.md:150) — point-cloud graph
construction, the step before a propagate over the sampled neighbourhoods.

CUDA tensors only (csrc/cluster.cu); ``batch`` vectors must be sorted, as in the reference."""
import math

import torch

from .. import _native as nat


def _ptr(batch, n_examples=None):
    counts = torch.bincount(batch, minlength=n_examples or 0)
    ptr = torch.zeros(counts.numel() + 1, dtype=torch.long, device=batch.device)
    torch.cumsum(counts, 0, out=ptr[1:])
    return counts, ptr


def fps(x, batch=None, ratio=0.5, random_start=True):
    """Indices of ``ceil(ratio * n_b)`` farthest-point samples of every example, in selection
    order.  ``random_start=False`` starts at each example's first point (deterministic)."""
    n = x.size(0)
    dev = nat.require_cuda(x, batch)
    if n == 0:
        return torch.empty(0, dtype=torch.long, device=dev)
    x = x.view(n, -1)
    x = x.float()
    x = x if x.stride(-1) == 1 else x.contiguous()
    if batch is None:
        batch = torch.zeros(n, dtype=torch.long, device=dev)
    counts, ptr = _ptr(batch)
    counts_h = counts.tolist()                      # one sync: the sample counts size the output
    m = [int(math.ceil(ratio * c)) for c in counts_h]
    out_ptr = torch.tensor([0] + list(torch.tensor(m).cumsum(0).tolist()), dtype=torch.long)
    total = int(out_ptr[-1])
    if random_start:
        start = (torch.rand(len(m), device=dev) * counts.clamp(min=1).float()).long()
        start = torch.minimum(start, (counts - 1).clamp(min=0))
    else:
        start = torch.zeros(len(m), dtype=torch.long, device=dev)
    out = torch.empty(total, dtype=torch.long, device=dev)
    dmin = torch.empty(n, dtype=torch.float32, device=dev)
    nat.call('pygb_fps', dev, nat.ptr(x), x.stride(0), n, x.size(1), nat.ptr(ptr),
             nat.ptr(out_ptr.to(dev)), nat.ptr(start), len(m), nat.ptr(dmin), nat.ptr(out))
    return out


def radius(x, y, r, batch_x=None, batch_y=None, max_num_neighbors=32, _skip_same=False):
    """For every ``y`` the (first ``max_num_neighbors``, in ascending index) points of ``x`` of
    the same example within distance ``r``: ``[2, E]`` with row 0 = index into y, row 1 = index
    into x."""
    dev = nat.require_cuda(x, y, batch_x, batch_y)
    ny = y.size(0)
    if ny == 0 or x.size(0) == 0:
        return torch.empty((2, 0), dtype=torch.long, device=dev)
    x, y = x.view(x.size(0), -1).float(), y.view(ny, -1).float()
    x = x if x.stride(-1) == 1 else x.contiguous()
    y = y if y.stride(-1) == 1 else y.contiguous()
    assert x.size(1) == y.size(1)
    if batch_x is None:
        batch_x = torch.zeros(x.size(0), dtype=torch.long, device=dev)
    if batch_y is None:
        batch_y = torch.zeros(ny, dtype=torch.long, device=dev)
    nb = int(max(int(batch_x.max()), int(batch_y.max()))) + 1
    _, ptr_x = _ptr(batch_x, nb)
    k = int(max_num_neighbors)
    nbr = torch.empty((ny, k), dtype=torch.int32, device=dev)
    cnt = torch.empty(ny, dtype=torch.int32, device=dev)
    nat.call('pygb_radius', dev, nat.ptr(x), x.stride(0), nat.ptr(y), y.stride(0), x.size(1),
             nat.ptr(ptr_x), nat.ptr(batch_y.contiguous()), ny, float(r), k, int(bool(_skip_same)),
             nat.ptr(nbr), nat.ptr(cnt))
    keep = nbr >= 0
    row = torch.arange(ny, device=dev).view(-1, 1).expand(-1, k)[keep]
    col = nbr[keep].long()
    return torch.stack([row, col], dim=0)


def radius_graph(x, r, batch=None, loop=False, max_num_neighbors=32, flow='source_to_target'):
    assert flow in ['source_to_target', 'target_to_source']
    row, col = radius(x, x, r, batch, batch, max_num_neighbors if loop else max_num_neighbors + 1)
    row, col = (col, row) if flow == 'source_to_target' else (row, col)
    if not loop:
        mask = row != col
        row, col = row[mask], col[mask]
    return torch.stack([row, col], dim=0)
