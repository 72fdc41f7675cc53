"""Feature-sliced multi-GPU GCN forward.

The destination-range partition of ``gcn.py`` has to replicate every source
feature row on every rank before it can aggregate: ``(P-1)/P * N * F`` bytes
received per GPU and layer, which grows with P and bounds the step by NVLink.
This module keeps the *graph* whole on every rank (the sorted graph of the 62 M
edge configuration is 0.8 GB) and slices the *features* instead:

* **column layout** — rank ``p`` holds columns ``[p*Fc, (p+1)*Fc)`` of all N
  rows.  Aggregation ``A @ X`` is independent per column, so every rank runs the
  same fused gather-reduce kernel over the whole graph at width ``Fc = F/P`` with
  no exchange at all;
* **row layout** — rank ``p`` holds all columns of the rows ``[p*blk, (p+1)*blk)``:
  what the dense transform ``X @ W`` needs.

Going from one layout to the other is a transpose across ranks that moves
``(P-1)/P**2 * N * F`` bytes per GPU — P times less than the all-gather.  Two
transports implement it:

* ``CollectiveTransport`` — ``all_to_all_single`` (NCCL, or gloo in the CPU tests);
* ``PeerTransport`` — no collective on the data path: the aggregation kernel
  stores every finished row slice straight into the HBM of the rank that owns
  the row (``pygb_seg_gather_reduce_stream_routed``) and ``pygb_cols_scatter``
  stores column blocks into the peers' column buffers, all over NVLink through
  CUDA-IPC mapped memory; ranks meet at a one-element stream-ordered all-reduce
  after each such kernel (no flags, no spinning).

The reference has no counterpart (single device).  Aggregation and
transform are injectable so the host logic runs on CPU under gloo.
"""
import ctypes

import torch
import torch.distributed as dist


def _ceil_div(a, b):
    return (a + b - 1) // b


class SlicePlan(object):
    """Row blocks and column slices of one rank."""

    def __init__(self, num_nodes, world_size, rank):
        self.num_nodes, self.world_size, self.rank = int(num_nodes), int(world_size), int(rank)
        self.blk = max(_ceil_div(self.num_nodes, self.world_size), 2)
        self.n_padded = self.blk * self.world_size
        self.lo = min(self.rank * self.blk, self.num_nodes)
        self.hi = min(self.lo + self.blk, self.num_nodes)

    @property
    def n_local(self):
        return self.hi - self.lo

    def widths(self, F, itemsize=4):
        """(Fp, Fc): F padded so that every rank's slice is a 32-byte multiple — rows
        that start inside a 32-byte sector slow the gather by 20-25 %
        (profiles/r01_k1_shapes.md)."""
        unit = max(32 // itemsize, 1) * self.world_size
        Fp = _ceil_div(max(F, 1), unit) * unit
        return Fp, Fp // self.world_size

    def col_slice(self, F, itemsize=4, rank=None):
        Fp, Fc = self.widths(F, itemsize)
        r = self.rank if rank is None else rank
        return slice(r * Fc, (r + 1) * Fc)

    def exchange_bytes(self, F, itemsize=4):
        """Bytes one rank sends in one layout change of an [N, F] matrix."""
        Fp, Fc = self.widths(F, itemsize)
        return (self.world_size - 1) * self.blk * Fc * itemsize


def _pad_cols(t, width):
    if t.size(1) == width:
        return t
    out = t.new_zeros((t.size(0), width))
    out[:, :t.size(1)] = t
    return out


def _pad_rows(t, rows):
    if t.size(0) == rows:
        return t
    out = t.new_zeros((rows,) + tuple(t.shape[1:]))
    out[:t.size(0)] = t
    return out


class CollectiveTransport(object):
    """Layout changes through ``all_to_all_single`` (any backend)."""

    fused = False

    def __init__(self, plan, group=None):
        self.plan, self.group = plan, group
        self.bytes_sent_total = 0
        self.collectives = 0
        self.event_log = None

    def _a2a(self, send):
        p = self.plan
        recv = torch.empty_like(send)
        if p.world_size == 1:
            recv.copy_(send)
            return recv
        nbytes = send.numel() // p.world_size * (p.world_size - 1) * send.element_size()
        if self.event_log is not None and send.is_cuda:
            e0 = torch.cuda.Event(enable_timing=True)
            e1 = torch.cuda.Event(enable_timing=True)
            e0.record()
            dist.all_to_all_single(recv, send, group=self.group)
            e1.record()
            self.event_log.append((e0, e1, nbytes))
        else:
            dist.all_to_all_single(recv, send, group=self.group)
        self.bytes_sent_total += nbytes
        self.collectives += 1
        return recv

    def cols_to_rows(self, cols, F):
        """cols [>=N, Fc] (this rank's columns of all rows) -> rows [n_local, Fp]."""
        p = self.plan
        Fp, Fc = p.widths(F, cols.element_size())
        send = _pad_rows(cols[:, :Fc], p.n_padded).contiguous().view(p.world_size, p.blk, Fc)
        recv = self._a2a(send)                               # [q] = my rows, columns of rank q
        rows = recv.permute(1, 0, 2).reshape(p.blk, Fp)
        return rows[:p.n_local]

    def rows_to_cols(self, rows, F, relu=False):
        """rows [n_local, >=F] -> cols [n_padded, Fc] (rows beyond N are padding)."""
        p = self.plan
        Fp, Fc = p.widths(F, rows.element_size())
        r = _pad_rows(_pad_cols(rows[:, :min(rows.size(1), Fp)], Fp), p.blk)
        if relu:
            r = torch.relu(r)
        send = r.view(p.blk, p.world_size, Fc).permute(1, 0, 2).contiguous()
        return self._a2a(send).view(p.n_padded, Fc)

    def aggregate_to_rows(self, graph, cols, F, bias, aggregate):
        p = self.plan
        Fp, Fc = p.widths(F, cols.element_size())
        b = None
        if bias is not None:
            b = _pad_cols(bias.view(1, -1), Fp).view(-1)[p.col_slice(F, cols.element_size())]
        z = aggregate(graph, cols, b)                        # [N, Fc]
        return self.cols_to_rows(z, F)


class _DevMem(object):
    """A raw device range exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {'shape': (int(nbytes),), 'typestr': '|u1',
                                         'data': (int(ptr), False), 'version': 2}


class PeerTransport(object):
    """Layout changes by plain stores into peer HBM (CUDA IPC over NVLink)."""

    fused = True

    def __init__(self, plan, device, group=None):
        from .. import _native as nat
        self.nat = nat
        self.plan, self.group, self.device = plan, group, torch.device(device)
        self.bytes_sent_total = 0
        self.collectives = 0            # fences (one-element all-reduces), not data
        self._cap = None                # (Fp_max elements) the arena was sized for
        self._base = None               # this rank's arena
        self._peer = []                 # arena base address of every rank, as seen from here
        self._opened = []
        self._flip = 0
        self._tabs = {}
        self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
        nat.load()

    # ---- arena ---------------------------------------------------------------------
    def _layout(self, fp_max):
        p = self.plan
        r_bytes = (p.blk * fp_max * 4 + 255) // 256 * 256
        c_bytes = (p.n_padded * (fp_max // p.world_size) * 4 + 255) // 256 * 256
        return r_bytes, c_bytes

    def ensure(self, fp_max):
        """Collective: (re)allocate the arena for row widths up to ``fp_max`` fp32 columns."""
        if self._cap is not None and fp_max <= self._cap:
            return
        nat, p = self.nat, self.plan
        self.close()
        r_bytes, c_bytes = self._layout(fp_max)
        total = 2 * r_bytes + c_bytes
        with torch.cuda.device(self.device):
            base = ctypes.c_void_p()
            nat.call_host('pygb_peer_alloc', ctypes.c_size_t(total), ctypes.byref(base))
            handle = (ctypes.c_ubyte * 64)()
            nat.call_host('pygb_peer_export', base, handle, ctypes.c_size_t(64))
            mine = torch.tensor(list(handle), dtype=torch.uint8).to(self.device)
            everyone = torch.empty((p.world_size, 64), dtype=torch.uint8, device=self.device)
            if p.world_size > 1:
                dist.all_gather_into_tensor(everyone, mine, group=self.group)
            else:
                everyone[0] = mine
            everyone = everyone.cpu()
            self._peer, self._opened = [], []
            for q in range(p.world_size):
                if q == p.rank:
                    self._peer.append(int(base.value))
                    continue
                h = (ctypes.c_ubyte * 64)(*everyone[q].tolist())
                ptr = ctypes.c_void_p()
                nat.call_host('pygb_peer_open', h, ctypes.c_size_t(64), ctypes.byref(ptr))
                self._peer.append(int(ptr.value))
                self._opened.append(int(ptr.value))
        self._base, self._cap = int(base.value), fp_max
        self._r_off = (0, r_bytes)
        self._c_off = 2 * r_bytes
        self._mem = torch.as_tensor(_DevMem(self._base, total), device=self.device)
        self._tabs = {}
        self.fence()

    def close(self):
        if self._base is None:
            return
        torch.cuda.synchronize(self.device)
        if self.plan.world_size > 1:
            dist.barrier(group=self.group)      # nobody still stores into an arena being freed
        with torch.cuda.device(self.device):
            for ptr in self._opened:
                self.nat.call_host('pygb_peer_close', ctypes.c_void_p(ptr))
            self._mem = None
            self.nat.call_host('pygb_peer_free', ctypes.c_void_p(self._base))
        self._base, self._cap, self._peer, self._opened = None, None, [], []

    def _view(self, off, rows, width):
        n = rows * width * 4
        return self._mem[off:off + n].view(torch.float32).view(rows, width)

    def _table(self, key, addrs):
        t = self._tabs.get(key)
        if t is None:
            t = torch.tensor(addrs, dtype=torch.int64).to(self.device)
            self._tabs[key] = t
        return t

    def fence(self):
        """All ranks' earlier kernels (and their peer stores) are complete before any
        rank's later kernels start: a one-element all-reduce on the current stream."""
        if self.plan.world_size > 1:
            dist.all_reduce(self._flag, group=self.group)
        self.collectives += 1

    # ---- layout changes ---------------------------------------------------------------
    def _check(self, t):
        if t.dtype != torch.float32 or not t.is_cuda:
            raise TypeError('PeerTransport moves fp32 CUDA tensors')

    def aggregate_to_rows(self, graph, cols, F, bias, aggregate=None):
        from .. import ops
        self._check(cols)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        b = None
        if bias is not None:
            b = _pad_cols(bias.float().view(1, -1), Fp).view(-1)[p.col_slice(F)].contiguous()
        self._flip ^= 1
        off = self._r_off[self._flip]
        tab = self._table(('r', self._flip, Fc),
                          [a + off + p.rank * Fc * 4 for a in self._peer])
        ops.seg_gather_reduce_routed_raw(cols[:, :Fc], graph.adj.dst, graph.w_csr, b, tab, p.blk,
                                         Fp)
        self.bytes_sent_total += (p.world_size - 1) * p.blk * Fc * 4
        self.fence()
        return self._view(off, p.blk, Fp)[:p.n_local]

    def rows_to_cols(self, rows, F, relu=False):
        from .. import ops
        self._check(rows)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        if rows.size(1) < Fp or rows.stride(1) != 1 or rows.stride(0) % 4 or \
                rows.data_ptr() % 16:
            rows = _pad_cols(rows[:, :min(rows.size(1), Fp)], Fp)
        tab = self._table(('c',), [a + self._c_off for a in self._peer])
        if rows.size(0) > 0:
            ops.cols_scatter_raw(rows, Fc, p.world_size, tab, p.lo, Fc, relu)
        self.bytes_sent_total += (p.world_size - 1) * p.n_local * Fc * 4
        self.fence()
        return self._view(self._c_off, p.n_padded, Fc)

    def cols_to_rows(self, cols, F):
        """Pure redistribution (no aggregation): strided device copies into the peers."""
        self._check(cols)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        self._flip ^= 1
        off = self._r_off[self._flip]
        for q in range(p.world_size):
            n = min(p.blk, max(p.num_nodes - q * p.blk, 0))
            if n == 0:
                continue
            dst = torch.as_tensor(_DevMem(self._peer[q] + off, p.blk * Fp * 4), device=self.device)
            dst = dst.view(torch.float32).view(p.blk, Fp)
            dst[:n, p.rank * Fc:(p.rank + 1) * Fc].copy_(cols[q * p.blk:q * p.blk + n, :Fc])
        self.bytes_sent_total += (p.world_size - 1) * p.blk * Fc * 4
        self.fence()
        return self._view(off, p.blk, Fp)[:p.n_local]


class SlicedGraph(object):
    """The whole (loop-augmented, normalised) graph on every rank plus this rank's
    row block / column slice bookkeeping."""

    def __init__(self, edge_index, norm, num_nodes, rank, world_size, group=None, i=1,
                 transport='peer', build_adj=True):
        self.plan = SlicePlan(num_nodes, world_size, rank)
        self.edge_index, self.norm, self.i = edge_index, norm, i
        self.device = edge_index.device
        self.adj = self.w_csr = None
        if build_adj:
            from ..graph import Adj
            self.adj = Adj(edge_index, i, (num_nodes, num_nodes))
            self.w_csr = None if norm is None else self.adj.permuted(norm.float(), 'dst')
        if transport == 'peer':
            self.transport = PeerTransport(self.plan, self.device, group)
        elif transport == 'collective':
            self.transport = CollectiveTransport(self.plan, group)
        else:
            raise ValueError("transport must be 'peer' or 'collective'")
        self._bufs = {}

    def local_rows(self, x_global):
        return x_global[self.plan.lo:self.plan.hi]

    def slice_cols(self, x_global):
        """This rank's column slice ``[n_padded, Fc]`` of a full ``[N, F]`` matrix
        (the input layout of an aggregate-first first layer)."""
        p = self.plan
        F = x_global.size(1)
        Fp, Fc = p.widths(F, x_global.element_size())
        out = x_global.new_zeros((p.n_padded, Fc))
        sl = p.col_slice(F, x_global.element_size())
        hi = min(sl.stop, F)
        if hi > sl.start:
            out[:p.num_nodes, :hi - sl.start] = x_global[:, sl.start:hi]
        return out

    def buffer(self, key, rows, width, dtype):
        buf = self._bufs.get(key)
        if buf is None or buf.shape != (rows, width) or buf.dtype != dtype:
            buf = torch.zeros((rows, width), dtype=dtype, device=self.device)
            self._bufs[key] = buf
        return buf


def _native_aggregate(graph, cols, bias):
    from .. import ops
    return ops.gather_reduce(cols, graph.adj, graph.norm, 'sum', bias)


def _native_transform(x, weight, bias, out, relu=False):
    from .. import ops
    n = weight.size(1)
    if x.size(0) == 0:
        return out
    if ops._gemm_eligible(x, weight):
        ops.gemm_xw_raw(x, weight, bias, out=out, relu=relu)     # bias and ReLU in the epilogue
    else:
        y = torch.matmul(x, weight)
        y = y + bias if bias is not None else y
        out[:, :n].copy_(torch.relu(y) if relu else y)
    return out


def sliced_gcn_forward(graph, x, layers, relu=True, x_layout='cols', aggregate=_native_aggregate,
                       transform=_native_transform):
    """GCN forward with ``layers = [(weight [in,out], bias|None), ...]``; ReLU between
    layers when ``relu``.  ``x``: this rank's column slice ``[n_padded, Fc]``
    (``x_layout='cols'``, see ``SlicedGraph.slice_cols``) or its rows ``[n_local, in]``
    (``'rows'``).  Returns this rank's rows ``[n_local, out]`` of the last layer."""
    p, tr = graph.plan, graph.transport
    state, kind = x, x_layout
    esz = x.element_size()

    def padded_weight(k, w, rows):
        if w.size(0) == rows:
            return w
        hit = graph._bufs.get(('w', k))
        if hit is None or hit[0] is not w or hit[1] != w._version or hit[2].size(0) != rows:
            hit = (w, w._version, _pad_rows(w.detach(), rows).contiguous())
            graph._bufs[('w', k)] = hit
        return hit[2]

    for k, (w, b) in enumerate(layers):
        fin, fout = w.shape
        fp_in, _ = p.widths(fin, esz)
        fp_out, _ = p.widths(fout, esz)
        act = bool(relu) and k + 1 < len(layers)
        out = graph.buffer(('y', k), p.n_local, fp_out, w.dtype)
        if fin <= fout:                                   # (A X) W: aggregate at width fin
            if kind == 'rows':
                state = tr.rows_to_cols(state, fin, False)
            rows = tr.aggregate_to_rows(graph, state, fin, None, aggregate)   # [n_local, fp_in]
            state = transform(rows, padded_weight(k, w, fp_in), b, out, act)  # bias, ReLU fused
        else:                                             # A (X W): aggregate at width fout
            if kind == 'cols':
                state = tr.cols_to_rows(state, fin)
            t = transform(state, padded_weight(k, w, state.size(1)), None, out, False)
            cols = tr.rows_to_cols(t, fout, False)
            state = tr.aggregate_to_rows(graph, cols, fout, b, aggregate)     # bias fused
            if act:
                state = torch.relu_(state)
        kind = 'rows'
    if kind == 'cols':                                    # no layers
        return state
    return state[:, :layers[-1][0].size(1)]
