"""Dataset classes under the reference's names (ConvexPruning.py:19,458-517,
examples/ConvexPruning.py:19,266-300, examples/gcn.py:5,11).

``Planetoid`` reads the real raw files (``<root>/raw/ind.<name>.*``) when they are
there.  Nothing can be downloaded (no network), so when the files are missing every
class falls back to a seeded synthetic graph of the same shape and SAYS
SO: a ``SyntheticDatasetWarning`` is raised once per construction and the dataset's
``__repr__`` / ``.synthetic`` flag name it a stand-in, so that accuracies printed by
the fork's scripts cannot be mistaken for results on the real data.  Set
``PYGB_REQUIRE_REAL_DATA=1`` to turn the fallback into an error.
"""
import os
import warnings

import torch

from ..data import Data, InMemoryDataset
from . import synthetic
from .planetoid_io import read_planetoid_raw


class SyntheticDatasetWarning(UserWarning):
    pass


def _fallback(cls_name, what, root):
    msg = ('{}: no raw files under {!r} and no network — using a SEEDED SYNTHETIC stand-in '
           '({}). Metrics computed on it say nothing about the real dataset.'.format(
               cls_name, root, what))
    if os.environ.get('PYGB_REQUIRE_REAL_DATA', '0') not in ('', '0'):
        raise FileNotFoundError(msg)
    warnings.warn(msg, SyntheticDatasetWarning, stacklevel=3)


class _Named(InMemoryDataset):
    synthetic = False
    name = ''

    def __repr__(self):
        tag = ' [synthetic stand-in]' if self.synthetic else ''
        return '{}({}){}'.format(self.name or self.__class__.__name__, len(self), tag)


class Planetoid(_Named):
    """Cora / CiteSeer / PubMed (examples/gcn.py:11, ConvexPruning.py:459)."""

    def __init__(self, root, name, transform=None, pre_transform=None):
        super(Planetoid, self).__init__(root, transform, pre_transform)
        self.name = name
        raw = os.path.join(os.path.expanduser(str(root)), 'raw')
        if os.path.exists(os.path.join(raw, 'ind.{}.x'.format(name.lower()))):
            d = read_planetoid_raw(raw, name.lower())
            g = Data(**d)
        else:
            make = {'cora': synthetic.cora_like, 'citeseer': synthetic.citeseer_like,
                    'pubmed': synthetic.pubmed_like}[name.lower()]
            _fallback('Planetoid', '{}-shaped citation graph'.format(name), root)
            self.synthetic = True
            g = make()
        if pre_transform is not None:
            g = pre_transform(g)
        self._graphs = [g]


class Reddit(_Named):
    name = 'Reddit'

    def __init__(self, root, transform=None, pre_transform=None, device='cpu'):
        super(Reddit, self).__init__(root, transform, pre_transform)
        _fallback('Reddit', 'Reddit-shaped log-normal graph, 233K nodes / 114M edges', root)
        self.synthetic = True
        self._graphs = [synthetic.reddit_like(device=device)]


class CoraFull(_Named):
    name = 'CoraFull'

    def __init__(self, root, transform=None, pre_transform=None):
        super(CoraFull, self).__init__(root, transform, pre_transform)
        _fallback('CoraFull', '19,793 nodes / 126,842 edges / 8,710 features / 70 classes', root)
        self.synthetic = True
        self._graphs = [synthetic.citation_like(19793, 63421, 8710, 70, 6, 0.0065,
                                                splits=(1400, 500, 1000))]


class Amazon(_Named):
    SHAPES = {'photo': (7650, 119081, 745, 8), 'computers': (13752, 245861, 767, 10)}

    def __init__(self, root, name, transform=None, pre_transform=None):
        super(Amazon, self).__init__(root, transform, pre_transform)
        self.name = 'Amazon' + name
        n, pairs, f, c = self.SHAPES[name.lower()]
        _fallback('Amazon', '{}: {} nodes / {} edges / {} features'.format(name, n, 2 * pairs, f),
                  root)
        self.synthetic = True
        self._graphs = [synthetic.citation_like(n, pairs, f, c, 7, 0.35, splits=(160, 240, 1000))]


class TUDataset(_Named):
    """ENZYMES / MUTAG graph-classification sets (examples/ConvexPruning.py:267,
    ConvexPruning.py:486)."""
    SHAPES = {  # graphs, mean nodes, mean undirected edges, node labels, node attrs, classes
        'ENZYMES': (600, 32.6, 62.1, 3, 18, 6),
        'MUTAG': (188, 17.9, 19.8, 7, 0, 2),
        'PROTEINS': (1113, 39.1, 72.8, 3, 1, 2),
    }

    def __init__(self, root, name, transform=None, pre_transform=None, pre_filter=None,
                 use_node_attr=False):
        super(TUDataset, self).__init__(root, transform, pre_transform)
        self.name = name
        if name not in self.SHAPES:
            raise NotImplementedError('TUDataset({!r}) has no synthetic stand-in'.format(name))
        count, nodes, edges, labels, attrs, classes = self.SHAPES[name]
        _fallback('TUDataset', '{}: {} graphs, ~{} nodes each'.format(name, count, nodes), root)
        self.synthetic = True
        self._graphs = synthetic.graph_collection(count, nodes, edges, labels,
                                                  attrs if use_node_attr else 0, classes, seed=8)


class MNISTSuperpixels(_Named):
    name = 'MNISTSuperpixels'

    def __init__(self, root, train=True, transform=None, pre_transform=None, pre_filter=None,
                 num_graphs=2000):
        super(MNISTSuperpixels, self).__init__(root, transform, pre_transform)
        _fallback('MNISTSuperpixels', '{} graphs of 75 superpixels (the real set has {})'.format(
            num_graphs, 60000 if train else 10000), root)
        self.synthetic = True
        self._graphs = synthetic.superpixel_collection(num_graphs, 75, 8, seed=9 if train else 10)


class PPI(_Named):
    name = 'PPI'

    def __init__(self, root, split='train', transform=None, pre_transform=None, pre_filter=None):
        assert split in ('train', 'val', 'test')
        super(PPI, self).__init__(root, transform, pre_transform)
        count = {'train': 20, 'val': 2, 'test': 2}[split]
        _fallback('PPI', '{} graphs of ~2,245 nodes, 50 features, 121 labels'.format(count), root)
        self.synthetic = True
        seed = {'train': 11, 'val': 12, 'test': 13}[split]
        self._graphs = synthetic.graph_collection(count, 2245.0, 30500.0, 0, 50, 121, seed=seed,
                                                  multilabel=True)

    @property
    def num_classes(self):
        return self._graphs[0].y.size(-1)


def _unavailable(name):
    class _DS(InMemoryDataset):
        def __init__(self, *args, **kwargs):
            raise NotImplementedError(
                '{} is not available offline and has no synthetic stand-in in this build'.format(
                    name))
    _DS.__name__ = _DS.__qualname__ = name
    return _DS


ModelNet = _unavailable('ModelNet')
FAUST = _unavailable('FAUST')
QM9 = _unavailable('QM9')
Entities = _unavailable('Entities')

__all__ = ['Planetoid', 'Reddit', 'TUDataset', 'MNISTSuperpixels', 'PPI', 'Amazon', 'CoraFull',
           'ModelNet', 'FAUST', 'QM9', 'Entities', 'synthetic', 'SyntheticDatasetWarning']
