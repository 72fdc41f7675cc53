"""Brute-force k-NN graph (stands in for torch_cluster.knn_graph; outside the hot path)."""
import torch


def knn_graph(x, k, batch=None, loop=False, flow='source_to_target'):
    assert flow in ['source_to_target', 'target_to_source']
    n = x.size(0)
    if batch is None:
        batch = x.new_zeros(n, dtype=torch.long)
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
