"""Seeded synthetic graphs with the shapes named in BASELINE.json.

There is no network, so the reference's datasets cannot be downloaded; every
generator here produces a graph of the same size and feature shape from a fixed
seed.  All generators run on any torch device (the large ones are meant to be
generated directly on the GPU) and use only torch ops, so the same seed gives
the same graph on the same device type.
"""

import torch

from ..data import Data


def _gen(seed, device):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return g


def _sample_by_weight(weights, n, gen):
    """n draws (with replacement) from the categorical distribution ``weights``."""
    cdf = torch.cumsum(weights.double(), 0)
    u = torch.rand(n, generator=gen, device=weights.device, dtype=torch.float64) * cdf[-1]
    return torch.searchsorted(cdf, u).clamp_(max=weights.numel() - 1)


def undirected_graph(num_nodes, num_pairs, seed, weights=None, distinct=False, device='cpu'):
    """``num_pairs`` undirected pairs without self-loops, symmetrised to
    ``2 * num_pairs`` directed edges.  ``weights`` biases the endpoint choice
    (power-law / log-normal degree); ``distinct`` removes duplicate pairs."""
    device = torch.device(device)
    gen = _gen(seed, device)
    need = num_pairs
    got = []
    have = 0
    seen = None
    while have < need:
        m = int((need - have) * 1.1) + 16
        if weights is None:
            a = torch.randint(0, num_nodes, (m,), generator=gen, device=device)
            b = torch.randint(0, num_nodes, (m,), generator=gen, device=device)
        else:
            a = _sample_by_weight(weights, m, gen)
            b = torch.randint(0, num_nodes, (m,), generator=gen, device=device)
        keep = a != b
        a, b = a[keep], b[keep]
        if distinct:
            lo, hi = torch.minimum(a, b), torch.maximum(a, b)
            code = lo * num_nodes + hi
            if seen is not None:
                code = torch.cat([seen, code])
            # stable first-occurrence de-duplication
            uniq, inv = torch.unique(code, return_inverse=True)
            first = torch.full((uniq.numel(),), code.numel(), dtype=torch.long, device=device)
            first.scatter_reduce_(0, inv, torch.arange(code.numel(), device=device), 'amin')
            code = code[torch.sort(first)[0]]
            seen = code[:need]
            have = seen.numel()
        else:
            got.append((a, b))
            have += a.numel()
    if distinct:
        a, b = seen // num_nodes, seen % num_nodes
    else:
        a = torch.cat([p[0] for p in got])[:need]
        b = torch.cat([p[1] for p in got])[:need]
    # symmetrise: (a->b) block then (b->a) block
    return torch.stack([torch.cat([a, b]), torch.cat([b, a])], dim=0)


def power_law_weights(num_nodes, alpha, max_degree, mean_degree, seed, device='cpu'):
    """Zipf-like expected degrees: w ~ u^(-1/(alpha-1)) truncated at ``max_degree``,
    rescaled to the requested mean (configuration-model endpoint weights)."""
    gen = _gen(seed, torch.device(device))
    u = torch.rand(num_nodes, generator=gen, device=device, dtype=torch.float64)
    w = (1.0 - u).clamp_(min=1e-12).pow(-1.0 / (alpha - 1.0))
    w = w.clamp_(max=float(max_degree))
    return (w * (mean_degree / w.mean())).clamp_(max=float(max_degree)).float()


def lognormal_weights(num_nodes, sigma, max_degree, mean_degree, seed, device='cpu'):
    gen = _gen(seed, torch.device(device))
    z = torch.randn(num_nodes, generator=gen, device=device, dtype=torch.float64)
    w = torch.exp(sigma * z)
    return (w * (mean_degree / w.mean())).clamp_(max=float(max_degree)).float()


def _masks(n, n_train, n_val, n_test, gen):
    perm = torch.randperm(n, generator=gen)
    out = []
    start = 0
    for k in (n_train, n_val, n_test):
        m = torch.zeros(n, dtype=torch.bool)
        m[perm[start:start + k]] = True
        out.append(m)
        start += k
    return out


def citation_like(num_nodes, num_pairs, num_features, num_classes, seed, density, scale=None,
                  splits=(140, 500, 1000)):
    """Cora-/PubMed-shaped node-classification graph (cfg 1 and cfg 2), on CPU."""
    ei = undirected_graph(num_nodes, num_pairs, seed, distinct=True)
    gen = _gen(seed + 1000, torch.device('cpu'))
    x = (torch.rand(num_nodes, num_features, generator=gen) < density).float()
    if scale is not None:
        x = x * torch.rand(num_nodes, num_features, generator=gen) * scale
    y = torch.randint(0, num_classes, (num_nodes,), generator=gen)
    tr, va, te = _masks(num_nodes, splits[0], splits[1], splits[2], gen)
    return Data(x=x, edge_index=ei, y=y, train_mask=tr, val_mask=va, test_mask=te)


def cora_like(seed=0):
    return citation_like(2708, 5278, 1433, 7, seed, 0.0127)


def citeseer_like(seed=5):
    return citation_like(3327, 4552, 3703, 6, seed, 0.0086, splits=(120, 500, 1000))


def pubmed_like(seed=1):
    return citation_like(19717, 44324, 500, 3, seed, 0.1, scale=0.1, splits=(60, 500, 1000))


def reddit_like(seed=2, device='cpu', num_nodes=232965, num_pairs=57307946, num_features=602,
                with_features=True):
    """cfg 3: log-normal degrees, mean in-degree ~492, cap 21,657."""
    w = lognormal_weights(num_nodes, 1.0, 21657, 2.0 * num_pairs / num_nodes, seed, device)
    ei = undirected_graph(num_nodes, num_pairs, seed + 1, weights=w, device=device)
    x = None
    if with_features:
        x = torch.randn(num_nodes, num_features, generator=_gen(seed + 2, torch.device(device)),
                        device=device)
    y = torch.randint(0, 41, (num_nodes,), generator=_gen(seed + 3, torch.device(device)),
                      device=device)
    return Data(x=x, edge_index=ei, y=y)


def products_like(seed=4, device='cpu', num_nodes=2449029, num_pairs=31000000, num_features=100,
                  with_features=True):
    """cfg 5: power-law (alpha 2.1, cap 17,481) in-degree, 62M directed edges."""
    w = power_law_weights(num_nodes, 2.1, 17481, 2.0 * num_pairs / num_nodes, seed, device)
    ei = undirected_graph(num_nodes, num_pairs, seed + 1, weights=w, device=device)
    x = None
    if with_features:
        x = torch.randn(num_nodes, num_features, generator=_gen(seed + 2, torch.device(device)),
                        device=device)
    y = torch.randint(0, 47, (num_nodes,), generator=_gen(seed + 3, torch.device(device)),
                      device=device)
    return Data(x=x, edge_index=ei, y=y)


def graph_collection(count, mean_nodes, mean_pairs, num_labels, num_attrs, num_classes, seed,
                     multilabel=False):
    """``count`` small graphs (TUDataset- / PPI-shaped): node count ~ Poisson-like around
    ``mean_nodes``, ``mean_pairs`` undirected pairs, x = [attrs | one-hot label]."""
    gen = _gen(seed, torch.device('cpu'))
    out = []
    for gi in range(count):
        n = max(3, int(torch.normal(float(mean_nodes), float(mean_nodes) ** 0.5 * 1.5,
                                    (1,), generator=gen).item()))
        pairs = max(2, int(n * mean_pairs / mean_nodes))
        pairs = min(pairs, n * (n - 1) // 2)
        a = torch.randint(0, n, (pairs,), generator=gen)
        b = (a + 1 + torch.randint(0, n - 1, (pairs,), generator=gen)) % n
        ei = torch.stack([torch.cat([a, b]), torch.cat([b, a])])
        cols = []
        if num_attrs:
            cols.append(torch.randn(n, num_attrs, generator=gen))
        if num_labels:
            lab = torch.randint(0, num_labels, (n,), generator=gen)
            cols.append(torch.nn.functional.one_hot(lab, num_labels).float())
        x = torch.cat(cols, dim=1)
        if multilabel:
            y = (torch.rand(n, num_classes, generator=gen) < 0.3).float()
        else:
            y = torch.randint(0, num_classes, (1,), generator=gen)
        out.append(Data(x=x, edge_index=ei, y=y))
    return out


def superpixel_collection(count, nodes, k, seed):
    """MNISTSuperpixels-shaped graphs: ``nodes`` points in [0, 28)^2 with one intensity
    feature, symmetric k-nearest-neighbour edges, ``pos`` set (for ``T.Cartesian``)."""
    gen = _gen(seed, torch.device('cpu'))
    out = []
    for gi in range(count):
        pos = torch.rand(nodes, 2, generator=gen) * 28
        d = torch.cdist(pos, pos)
        d.fill_diagonal_(float('inf'))
        nbr = d.topk(k, dim=1, largest=False)[1]
        src = torch.arange(nodes).view(-1, 1).expand(-1, k).reshape(-1)
        code = torch.unique(torch.cat([src * nodes + nbr.reshape(-1),
                                       nbr.reshape(-1) * nodes + src]))
        ei = torch.stack([code // nodes, code % nodes])
        out.append(Data(x=torch.rand(nodes, 1, generator=gen), edge_index=ei, pos=pos,
                        y=torch.randint(0, 10, (1,), generator=gen)))
    return out


def knn_clouds(num_clouds=32, points=1024, k=20, seed=3, device='cpu'):
    """cfg 4: ``num_clouds`` clouds of U(-1,1)^3 points with an exact k-NN graph
    inside each cloud; edges are (neighbour -> centre), grouped by centre."""
    device = torch.device(device)
    gen = _gen(seed, device)
    pos = torch.rand(num_clouds * points, 3, generator=gen, device=device) * 2 - 1
    rows, cols = [], []
    for c in range(num_clouds):
        p = pos[c * points:(c + 1) * points]
        d = torch.cdist(p, p)
        d.fill_diagonal_(float('inf'))
        nn_idx = d.topk(k, dim=1, largest=False)[1] + c * points
        centre = torch.arange(points, device=device).view(-1, 1).expand(-1, k) + c * points
        rows.append(nn_idx.reshape(-1))
        cols.append(centre.reshape(-1))
    ei = torch.stack([torch.cat(rows), torch.cat(cols)], dim=0)
    batch = torch.arange(num_clouds, device=device).repeat_interleave(points)
    return Data(x=pos, pos=pos, edge_index=ei, batch=batch)


def add_noise_edges(edge_index, num_nodes, frac_loops=0.01, frac_dups=0.01, seed=7):
    """Parity-test variant: append ~1% self-loops and ~1% duplicated edges."""
    gen = _gen(seed, edge_index.device)
    E = edge_index.size(1)
    nl, nd = max(1, int(E * frac_loops)), max(1, int(E * frac_dups))
    loops = torch.randint(0, num_nodes, (nl,), generator=gen, device=edge_index.device)
    dup = edge_index[:, torch.randint(0, E, (nd,), generator=gen, device=edge_index.device)]
    ei = torch.cat([edge_index, torch.stack([loops, loops]), dup], dim=1)
    perm = torch.randperm(ei.size(1), generator=gen, device=edge_index.device)
    return ei[:, perm].contiguous()
