"""B200-native message passing behind the ``torch_geometric`` API of
ggzhang0071/pytorch_geometric-1 (hot path only: MessagePassing.propagate,
GCNConv, GATConv, EdgeConv and the utils they call).  CUDA (sm_100a) only."""
from .debug import is_debug_enabled, debug, set_debug
from . import _native  # noqa: F401
from . import graph, ops  # noqa: F401
from . import utils, data, transforms, nn, datasets  # noqa: F401

__version__ = '1.4.3+b200'

__all__ = ['is_debug_enabled', 'debug', 'set_debug', '__version__']
