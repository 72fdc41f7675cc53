"""Dataset classes under the reference's names (ConvexPruning.py:19,458-517,
examples/gcn.py:5,11).  The files cannot be downloaded (no network), so each
class builds the seeded synthetic graph of the same shape."""
from ..data import InMemoryDataset
from . import synthetic


class Planetoid(InMemoryDataset):
    def __init__(self, root, name, transform=None, pre_transform=None):
        super(Planetoid, self).__init__(root, transform, pre_transform)
        self.name = name
        make = {'cora': synthetic.cora_like, 'citeseer': synthetic.citeseer_like,
                'pubmed': synthetic.pubmed_like}[name.lower()]
        self._graphs = [make()]

    def __repr__(self):
        return '{}()'.format(self.name)


class Reddit(InMemoryDataset):
    def __init__(self, root, transform=None, pre_transform=None, device='cpu'):
        super(Reddit, self).__init__(root, transform, pre_transform)
        self._graphs = [synthetic.reddit_like(device=device)]


def _unavailable(name):
    class _DS(InMemoryDataset):
        def __init__(self, *args, **kwargs):
            raise NotImplementedError(
                '{} is not available offline and has no synthetic stand-in in this build'.format(
                    name))
    _DS.__name__ = _DS.__qualname__ = name
    return _DS


TUDataset = _unavailable('TUDataset')
MNISTSuperpixels = _unavailable('MNISTSuperpixels')
PPI = _unavailable('PPI')
Amazon = _unavailable('Amazon')
CoraFull = _unavailable('CoraFull')
ModelNet = _unavailable('ModelNet')
FAUST = _unavailable('FAUST')
QM9 = _unavailable('QM9')
Entities = _unavailable('Entities')

__all__ = ['Planetoid', 'Reddit', 'TUDataset', 'MNISTSuperpixels', 'PPI', 'Amazon', 'CoraFull',
           'ModelNet', 'FAUST', 'QM9', 'Entities', 'synthetic']
