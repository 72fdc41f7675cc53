from .degree import degree
from .loop import (add_remaining_self_loops, add_self_loops, contains_self_loops,
                   remove_self_loops)
from .num_nodes import maybe_num_nodes
from .scatter import scatter_
from .softmax import softmax

__all__ = [
    'degree', 'softmax', 'scatter_', 'contains_self_loops', 'remove_self_loops',
    'add_self_loops', 'add_remaining_self_loops', 'maybe_num_nodes',
]
