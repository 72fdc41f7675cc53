"""Names the fork's scripts import next to the hot-path operators
(examples/gcn.py:7, ConvexPruning.py:17,22).  They are outside the scope of this
build and exist so that ``from torch_geometric.nn import
...`` resolves; constructing one raises."""
import torch


def _unported(name):
    class _Op(torch.nn.Module):
        def __init__(self, *args, **kwargs):
            raise NotImplementedError(
                '{} is outside the B200 hot-path build (MessagePassing.propagate, GCNConv, '
                'GATConv, EdgeConv); see SURVEY.md §8f.'.format(name))

    _Op.__name__ = _Op.__qualname__ = name
    return _Op


SplineConv = _unported('SplineConv')
ARMAConv = _unported('ARMAConv')
RGCNConv = _unported('RGCNConv')
NNConv = _unported('NNConv')
GMMConv = _unported('GMMConv')
PointConv = _unported('PointConv')
DNAConv = _unported('DNAConv')
SAGPooling = _unported('SAGPooling')
