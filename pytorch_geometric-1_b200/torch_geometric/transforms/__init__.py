"""``NormalizeFeatures`` and ``Compose`` — the only transforms on the fork's GCN
path (examples/gcn.py:11, ConvexPruning.py:459)."""


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


__all__ = ['Compose', 'NormalizeFeatures']
