import json
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))


def kat():
    with open(os.path.join(HERE, 'golden', 'kat.json')) as f:
        return json.load(f)


def rel_err(a, b):
    """Max-norm error relative to the largest reference entry (reported, not the gate)."""
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    if a.numel() == 0:
        return 0.0
    return (a - b).abs().max().item() / max(b.abs().max().item(), 1e-30)


def elem_err(a, b, tol=1e-5, floor=None):
    """Worst element of ``|a-b| / (tol*|b| + floor*max|b|)`` — <= 1 passes.

    The gate of every floating-point parity assertion: each element is held to ``tol``
    relative to ITS OWN reference value, plus an absolute floor of ``floor`` (default
    ``0.1*tol``) times the largest reference entry.  The floor is there because two
    summation orders of the same addends differ by ~eps*sum|t_i|, which near a cancelling
    sum is not small relative to |sum t_i|; it is 10x below what the max-norm criterion of
    round 1 allowed everywhere."""
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    if a.numel() == 0:
        return 0.0
    floor = 0.1 * tol if floor is None else floor
    bound = tol * b.abs() + floor * max(b.abs().max().item(), 1e-30)
    bad = ~torch.isfinite(a) & torch.isfinite(b)
    if bad.any():
        return float('inf')
    return ((a - b).abs() / bound).max().item()


def assert_close(a, b, tol=1e-5, what='', floor=None):
    assert tuple(a.shape) == tuple(b.shape), '{} shape {} vs {}'.format(what, a.shape, b.shape)
    e = elem_err(a, b, tol, floor)
    assert e <= 1.0, ('{} element-wise error {:.3f}x the bound (tol {:.1e}*|ref| + {:.1e}*max|ref|); '
                      'max-norm rel err {:.3e}'.format(what, e, tol,
                                                       0.1 * tol if floor is None else floor,
                                                       rel_err(a, b)))


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
