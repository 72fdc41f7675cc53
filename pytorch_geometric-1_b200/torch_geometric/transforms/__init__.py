"""``NormalizeFeatures`` and ``Compose`` — the only transforms on the fork's GCN
path (examples/gcn.py:11, ConvexPruning.py:459)."""
import torch


class Compose(object):
    def __init__(self, transforms):
        self.transforms = transforms

    def __call__(self, data):
        for t in self.transforms:
            data = t(data)
        return data


class NormalizeFeatures(object):
    r"""Row-normalises node features to sum to one (rows summing to 0 stay 0)."""

    def __call__(self, data):
        data.x = data.x / data.x.sum(1, keepdim=True).clamp(min=1)
        return data

    def __repr__(self):
        return '{}()'.format(self.__class__.__name__)


class Cartesian(object):
    r"""Saves the relative Cartesian coordinates of linked nodes in ``edge_attr``
    (``pos[col] - pos[row]``, scaled into [0, 1] when ``norm``); used by the fork for
    MNISTSuperpixels (ConvexPruning.py:513)."""

    def __init__(self, norm=True, max_value=None, cat=True):
        self.norm, self.max, self.cat = norm, max_value, cat

    def __call__(self, data):
        (row, col), pos, pseudo = data.edge_index, data.pos, data.edge_attr
        cart = pos[col] - pos[row]
        cart = cart.view(-1, 1) if cart.dim() == 1 else cart
        if self.norm and cart.numel() > 0:
            max_value = cart.abs().max() if self.max is None else self.max
            cart = cart / (2 * max_value) + 0.5
        if pseudo is not None and self.cat:
            pseudo = pseudo.view(-1, 1) if pseudo.dim() == 1 else pseudo
            data.edge_attr = torch.cat([pseudo, cart.type_as(pseudo)], dim=-1)
        else:
            data.edge_attr = cart
        return data

    def __repr__(self):
        return '{}(norm={}, max_value={})'.format(self.__class__.__name__, self.norm, self.max)


__all__ = ['Compose', 'NormalizeFeatures', 'Cartesian']
