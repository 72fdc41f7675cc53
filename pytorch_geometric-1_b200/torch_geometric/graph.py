"""Device-side graph structures for the B200 path.

``CSR``  — a stable sort of E entries by an int64 key (``rowptr``, ``perm`` and
optionally ``col``), produced by ``pygb_csr_build``.  Stability makes the order
inside a row the ascending original edge id, i.e. the order in which the CPU
reference (torch_scatter's sequential loop) visits them.

``Adj``  — the two CSRs of one ``edge_index``: by aggregation index (forward)
and, lazily, by the opposite row (transposed walk for the backward).

Both are memoised on the identity of the index tensor (storage pointer, shape,
version counter), so ``GATConv`` — which the reference rebuilds every forward —
and the generic ``scatter_`` path sort once per graph.
"""
import collections
import ctypes
import threading

import torch

from . import _native as nat


class CSR(object):
    __slots__ = ('rowptr', 'col', 'perm', 'n_rows', 'n_cols', 'E', '_keep', '_items')

    def __init__(self, rowptr, col, perm, n_rows, n_cols, E, keep=None):
        self.rowptr, self.col, self.perm = rowptr, col, perm
        self.n_rows, self.n_cols, self.E = n_rows, n_cols, E
        self._keep = keep
        self._items = {}

    def items(self, T=None):
        """Work decomposition of the streaming kernels: ``(Ts, n_sub, nz_rows,
        sub_info)`` (see include/pygb.h, pygb_seg_tables); built once, cached.
        ``T=None`` lets the library choose Ts from E and the SM count."""
        hit = self._items.get(T)
        if hit is not None:
            return hit
        dev = self.rowptr.device
        if T is None:
            t, n = ctypes.c_int32(0), ctypes.c_int64(0)
            with torch.cuda.device(dev):
                rc = nat.load().pygb_seg_plan(self.E, ctypes.byref(t), ctypes.byref(n))
            if rc != 0:
                raise nat.NativeError(nat.load().pygb_last_error().decode())
            t, n = int(t.value), int(n.value)
        else:
            t, n = int(T), (self.E + int(T) - 1) // int(T)
        nz_rows = torch.empty(self.n_rows + 2, dtype=torch.int32, device=dev)
        sub_info = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        if self.n_rows > 0:
            ws_bytes = nat.workspace('pygb_seg_tables_workspace', self.n_rows)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            nat.call('pygb_seg_tables', dev, nat.ptr(self.rowptr), self.n_rows, self.E, t, n,
                     nat.ptr(nz_rows), nat.ptr(sub_info), nat.ptr(ws), ws_bytes)
        plan = (t, n, nz_rows, sub_info)
        self._items[T] = plan
        if T is None:
            self._items[t] = plan
        return plan

    def flagged(self, which, idx=None):
        """``col``, ``perm`` or a caller-supplied int32 index ``idx`` (cached under the name
        ``which``) with bit 31 set on the last slot of every row — the index stream of the
        streaming kernels; built once, cached."""
        key = ('flag', which)
        hit = self._items.get(key)
        if hit is not None:
            return hit
        if idx is None:
            idx = self.col if which == 'col' else self.perm
        out = torch.empty_like(idx)
        if self.E > 0:
            nat.call('pygb_seg_flag_idx', idx.device, nat.ptr(self.rowptr), self.n_rows,
                     nat.ptr(idx), self.E, nat.ptr(out))
        self._items[key] = out
        return out

    def flagged_hot(self, n_hot):
        """``flagged('col')`` with bit 30 set on the slots of the ``n_hot`` most frequently
        gathered source rows (pygb_seg_mark_hot): the index stream of
        pygb_seg_gather_reduce_stream_hinted, whose gathers ask L2 to keep those rows.  Built once
        per (graph, n_hot); ``None`` when every source is gathered equally often or ids need
        more than 30 bits."""
        key = ('hot', int(n_hot))
        if key in self._items:
            return self._items[key]
        out = None
        n_src = int(self.col.max().item()) + 1 if self.E > 0 else 0
        if 0 < n_hot < n_src <= (1 << 30):
            freq = torch.bincount(self.col, minlength=n_src)
            thr = int(torch.topk(freq, int(n_hot), sorted=True).values[-1].item())
            if thr > int(freq.min().item()):
                out = self.flagged('col').clone()
                freq32 = freq.to(torch.int32)
                nat.call('pygb_seg_mark_hot', out.device, nat.ptr(out), self.E, nat.ptr(freq32), n_src,
                         thr)
        self._items[key] = out
        return out

    @property
    def device(self):
        return self.rowptr.device


def build_csr(key, other, n_rows, n_cols=None, validated=False):
    """Sort ``key`` (int64 [E]) into a CSR with ``n_rows`` rows; ``other`` (int64
    [E], optional) becomes ``col``.  Raises ``IndexError`` for out-of-range
    indices (one device->host read of a status word per build) — unless the caller
    vouches that the ids were range-checked already (``validated``: an ``edge_index'``
    that came out of the self-loop pass, which checks every id behind the read it
    needs anyway), so the GCN first call pays one host synchronisation, not three."""
    dev = nat.require_cuda(key, other)
    if key.dtype != torch.long or (other is not None and other.dtype != torch.long):
        raise TypeError('index tensors must be int64')
    E = key.numel()
    n_rows = int(n_rows)
    n_cols = int(n_cols) if n_cols is not None else n_rows
    key = key if key.is_contiguous() else key.contiguous()
    if other is not None and not other.is_contiguous():
        other = other.contiguous()
    rowptr = torch.empty(n_rows + 1, dtype=torch.int32, device=dev)
    perm = torch.empty(E, dtype=torch.int32, device=dev)
    col = torch.empty(E, dtype=torch.int32, device=dev) if other is not None else None
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws_bytes = nat.workspace('pygb_csr_build_workspace', E, n_rows)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    nat.call('pygb_csr_build', dev, nat.ptr(key), nat.ptr(other), E, n_rows, n_cols,
             nat.ptr(rowptr), nat.ptr(perm), nat.ptr(col), nat.ptr(status), nat.ptr(ws), ws_bytes)
    st = 0 if validated else int(status.item())
    if st == 1:
        raise IndexError('index out of range: expected values in [0, {})'.format(n_rows))
    if st == 2:
        raise IndexError('index out of range: expected values in [0, {})'.format(n_cols))
    return CSR(rowptr, col, perm, n_rows, n_cols, E)


# --------------------------------------------------------------------------
# memoisation keyed on tensor identity
# --------------------------------------------------------------------------
class _LRU(object):
    def __init__(self, capacity):
        self.capacity = capacity
        self.data = collections.OrderedDict()
        self.lock = threading.Lock()

    def get(self, key):
        with self.lock:
            v = self.data.get(key)
            if v is not None:
                self.data.move_to_end(key)
            return v

    def put(self, key, value):
        with self.lock:
            self.data[key] = value
            self.data.move_to_end(key)
            while len(self.data) > self.capacity:
                self.data.popitem(last=False)

    def clear(self):
        with self.lock:
            self.data.clear()


def _ident(t):
    """Cache key of an index tensor: storage pointer, shape, strides, autograd version counter,
    device.  The version counter sees every in-place torch op, but NOT writes through ``.data``
    or from foreign kernels: code that rewrites an ``edge_index`` buffer that way must call
    ``torch_geometric.graph.clear_caches()``."""
    return (t.data_ptr(), tuple(t.shape), tuple(t.stride()), t._version, t.device.index)


_csr_cache = _LRU(8)
_adj_cache = _LRU(4)


def clear_caches():
    _csr_cache.clear()
    _adj_cache.clear()


def csr_for(index, n_rows):
    """Memoised ``build_csr(index, None, n_rows)``."""
    key = ('csr', _ident(index), int(n_rows))
    hit = _csr_cache.get(key)
    if hit is not None:
        return hit
    csr = build_csr(index, None, n_rows)
    csr._keep = index  # pins the storage so the pointer cannot be recycled
    _csr_cache.put(key, csr)
    return csr


class Adj(object):
    """Adjacency of one ``edge_index`` for aggregation over row ``i``.

    ``dst``: CSR keyed on ``edge_index[i]`` with ``col = edge_index[j]``.
    ``src``: CSR keyed on ``edge_index[j]`` with ``col = edge_index[i]`` (lazy).
    """

    def __init__(self, edge_index, i, size):
        self.edge_index = edge_index
        self.i, self.j = i, 1 - i
        self.size = (int(size[0]), int(size[1]))
        self.E = edge_index.size(1)
        # ids already range-checked against this node count by utils.loop._apply?
        vn = getattr(edge_index, '_pygb_valid_n', None)
        self._validated = vn is not None and vn <= min(self.size) and \
            getattr(edge_index, '_pygb_valid_version', -1) == edge_index._version
        self.dst = build_csr(edge_index[self.i], edge_index[self.j], self.size[self.i],
                             self.size[self.j], self._validated)
        self._src = None
        self._wcache = {}
        self._idx = {}

    @property
    def n_dst(self):
        return self.size[self.i]

    @property
    def n_src(self):
        return self.size[self.j]

    @property
    def src(self):
        if self._src is None:
            ei = self.edge_index
            # the dst build (or the self-loop pass) has range-checked both rows
            self._src = build_csr(ei[self.j], ei[self.i], self.size[self.j], self.size[self.i],
                                  True)
        return self._src

    @property
    def iota(self):
        """0..E-1 (int32): the identity index over the slots of a sorted graph."""
        hit = self._idx.get('iota')
        if hit is None:
            hit = torch.arange(self.E, dtype=torch.int32, device=self.edge_index.device)
            self._idx['iota'] = hit
        return hit

    @property
    def src_to_dst_slot(self):
        """For slot p of the source-sorted graph, the slot of the same edge in the
        destination-sorted graph (int32 [E]); built once."""
        hit = self._idx.get('s2d')
        if hit is None:
            inv = torch.empty(self.E, dtype=torch.int32, device=self.edge_index.device)
            inv[self.dst.perm.long()] = self.iota
            hit = inv[self.src.perm.long()].contiguous()
            self._idx['s2d'] = hit
        return hit

    def permuted(self, w, which):
        """``w[perm]`` for the 'dst' or 'src' order, memoised per weight tensor."""
        key = (which, w.data_ptr(), w._version)
        hit = self._wcache.get(key)
        if hit is not None:
            return hit[0]
        csr = self.dst if which == 'dst' else self.src
        out = torch.empty_like(w)
        tag = getattr(w, '_pygb_unit_deg', None)
        deg = tag[0] if tag is not None and tag[1] == w._version else None
        if (deg is not None and w.dtype == torch.float32 and w.numel() == csr.E
                and deg.numel() == csr.n_rows == csr.n_cols and csr.col is not None):
            # GCNConv.norm with unit weights: D^-1/2 A D^-1/2 straight in slot order
            nat.call('pygb_gcn_norm_csr', w.device, nat.ptr(csr.rowptr), nat.ptr(csr.col),
                     nat.ptr(deg), csr.n_rows, nat.ptr(out))
        else:
            nat.call('pygb_permute_f32', w.device, nat.ptr(w), nat.ptr(csr.perm), w.numel(),
                     nat.ptr(out))
        if len(self._wcache) > 4:
            self._wcache.clear()
        self._wcache[key] = (out, w)
        return out


def adj_for(edge_index, i, size):
    key = ('adj', _ident(edge_index), int(i), int(size[0]), int(size[1]))
    hit = _adj_cache.get(key)
    if hit is not None:
        return hit
    adj = Adj(edge_index, i, size)
    _adj_cache.put(key, adj)
    return adj
