from .. import ops
from .num_nodes import maybe_num_nodes


def softmax(src, index, num_nodes=None):
    """Softmax of ``src`` over the entries that share an ``index`` value
    (subtract the group max, exp, divide by group sum + 1e-16).
    One fused kernel instead of scatter_max/exp/scatter_add/div."""
    num_nodes = maybe_num_nodes(index, num_nodes)
    return ops.edge_softmax(src, index, num_nodes)
