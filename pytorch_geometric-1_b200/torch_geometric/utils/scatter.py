
from .. import ops


def scatter_(name, src, index, dim=0, dim_size=None):
    """Aggregate ``src`` rows into ``dim_size`` buckets given by ``index``
    (called from ``MessagePassing.propagate``).

    name: 'add' | 'mean' | 'min' | 'max'.  Keeps the reference's quirk Q2:
    'max' zeroes results below -10000 and 'min' results above 10000 (this is
    also how empty buckets become 0).  Atomic-free: ``index`` is sorted once
    (memoised) and every bucket is reduced by one warp in a fixed order.
    """
    assert name in ['add', 'mean', 'min', 'max']
    if dim_size is None:
        dim_size = int(index.max().item()) + 1 if index.numel() > 0 else 0
    if dim < 0:
        dim += src.dim()
    if index.dim() != 1:
        raise ValueError('scatter_: index must be one-dimensional on the native path')
    if dim != 0:
        out = scatter_(name, src.movedim(dim, 0).contiguous(), index, 0, dim_size)
        return out.movedim(0, dim)
    zero_below = -10000.0 if name in ('max', 'min') else ops.NEG_INF
    return ops.segment_reduce(src, index, dim_size, name, zero_below)
