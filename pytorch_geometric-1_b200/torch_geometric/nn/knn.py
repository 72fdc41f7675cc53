"""k-NN graph (stands in for ``torch_cluster.knn_graph``, README.md:149-150,
examples/pointnet++.py:39-41): the step before EdgeConv's propagate.

CUDA tensors go to the batched brute-force kernel ``pygb_knn_graph`` (csrc/knn.cu: one
thread per query, candidates streamed through shared memory, sorted k-best list per
thread; ties to the lower index).  Host tensors — dataset preparation, like the CPU build
of torch_cluster — use the framework's ``cdist`` + ``topk`` per example."""
import torch

from .. import _native as nat


def _knn_cuda(x, k, batch, loop):
    n = x.size(0)
    dev = nat.require_cuda(x, batch)
    x = x.float()
    x = x if x.stride(-1) == 1 else x.contiguous()
    counts = torch.bincount(batch)                          # (one sync: the number of examples)
    ptr = torch.zeros(counts.numel() + 1, dtype=torch.long, device=dev)
    torch.cumsum(counts, 0, out=ptr[1:])
    nbr = torch.empty((n, k), dtype=torch.int32, device=dev)
    nat.call('pygb_knn_graph', dev, nat.ptr(x), x.stride(0), n, x.size(1), nat.ptr(batch),
             nat.ptr(ptr), k, int(bool(loop)), nat.ptr(nbr))
    row = nbr.view(-1).long()
    col = torch.arange(n, device=dev).repeat_interleave(k)
    # one more sync for both checks: `batch` sorted, and every example has k candidates
    big = torch.iinfo(torch.long).max
    unsorted = (batch[1:] < batch[:-1]).any() if n > 1 else torch.zeros((), dtype=torch.bool, device=dev)
    smallest = torch.where(counts > 0, counts, torch.full_like(counts, big)).min()
    unsorted, smallest = torch.stack([unsorted.long(), smallest]).tolist()
    if unsorted:
        raise ValueError('knn_graph: `batch` must be sorted')
    if smallest < k + (0 if loop else 1):                   # some example has fewer than k candidates
        keep = row >= 0
        row, col = row[keep], col[keep]
    return row, col


def knn_graph(x, k, batch=None, loop=False, flow='source_to_target'):
    assert flow in ['source_to_target', 'target_to_source']
    n = x.size(0)
    if n == 0:
        return torch.empty((2, 0), dtype=torch.long, device=x.device)
    x = x.view(n, -1)
    if batch is None:
        batch = x.new_zeros(n, dtype=torch.long)
    if x.is_cuda and n > 0 and 0 < k <= 64 and x.size(1) <= 128:
        row, col = _knn_cuda(x, k, batch, loop)
    else:
        rows, cols = [], []
        for b in torch.unique(batch).tolist():
            ids = (batch == b).nonzero().view(-1)
            p = x[ids]
            d = torch.cdist(p, p)
            if not loop:
                d.fill_diagonal_(float('inf'))
            kk = min(k, p.size(0) - (0 if loop else 1))
            nn_idx = d.topk(kk, dim=1, largest=False)[1]
            centre = ids.view(-1, 1).expand(-1, kk).reshape(-1)
            neigh = ids[nn_idx.reshape(-1)]
            rows.append(neigh)
            cols.append(centre)
        row, col = torch.cat(rows), torch.cat(cols)
    if flow == 'source_to_target':
        return torch.stack([row, col], dim=0)
    return torch.stack([col, row], dim=0)
