import json
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def kat():
    with open(os.path.join(HERE, 'golden', 'kat.json')) as f:
        return json.load(f)


def rel_err(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    if a.numel() == 0:
        return 0.0
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)


def assert_close(a, b, tol=1e-5, what=''):
    assert tuple(a.shape) == tuple(b.shape), '{} shape {} vs {}'.format(what, a.shape, b.shape)
    e = rel_err(a, b)
    assert e <= tol, '{} rel err {:.3e} > {:.1e}'.format(what, e, tol)


def random_graph(n_src, n_dst, E, seed, loops=True):
    g = torch.Generator().manual_seed(seed)
    row = torch.randint(0, n_src, (E,), generator=g)
    col = torch.randint(0, n_dst, (E,), generator=g)
    if not loops and n_src == n_dst:
        col = torch.where(row == col, (col + 1) % n_dst, col)
    return torch.stack([row, col])


def skewed_graph(n, E, seed, hubs=3):
    """Some rows empty, a few hub rows with thousands of edges."""
    g = torch.Generator().manual_seed(seed)
    row = torch.randint(0, n, (E,), generator=g)
    col = torch.randint(0, max(n // 2, 1), (E,), generator=g)  # upper half never a target
    hub = torch.randint(0, hubs, (E,), generator=g)
    col = torch.where(torch.rand(E, generator=g) < 0.3, hub, col)
    return torch.stack([row, col])
