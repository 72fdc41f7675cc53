from .conv import *  # noqa
from .conv import __all__ as _conv_all
from .glob import global_add_pool, global_mean_pool, global_max_pool
from .data_parallel import DataParallel
from .knn import knn_graph
from .cluster import fps, radius, radius_graph
from .pool import TopKPooling

__all__ = list(_conv_all) + ['global_add_pool', 'global_mean_pool', 'global_max_pool',
                             'DataParallel', 'knn_graph', 'TopKPooling', 'fps', 'radius',
                             'radius_graph']
