"""Global pooling = the same segmented reductions keyed on ``batch``."""
from ..utils import scatter_


def _size(batch, size):
    return int(batch.max().item()) + 1 if size is None else size


def global_add_pool(x, batch, size=None):
    return scatter_('add', x, batch, dim=0, dim_size=_size(batch, size))


def global_mean_pool(x, batch, size=None):
    return scatter_('mean', x, batch, dim=0, dim_size=_size(batch, size))


def global_max_pool(x, batch, size=None):
    return scatter_('max', x, batch, dim=0, dim_size=_size(batch, size))
