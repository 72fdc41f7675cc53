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

With ``row_groups = G > 1`` the layout is two-dimensional: the P ranks form G row
groups of P/G ranks; a rank aggregates only the destination rows of its row group,
at width ``F / (P/G)``.  Wider slices keep the gather kernel near the HBM roofline
(its cost per edge has a fixed part: index, weight, issue slots), at the price of
sending every column block to G ranks.  On 8 GPUs 2 x 4 beats 1 x 8.

The reference has no counterpart (single device).  Aggregation and
transform are injectable so the host logic runs on CPU under gloo.
"""
import ctypes

import torch
import torch.distributed as dist


def _ceil_div(a, b):
    return (a + b - 1) // b


class SlicePlan(object):
    """Row blocks, row groups and column slices of one rank.

    rank = rg * Gc + cg: row group ``rg`` (of ``Gr``), column slice ``cg`` (of ``Gc``).
    Row block ``q`` (``blk`` rows) is owned by rank ``q``; row group ``rg`` is the union
    of the blocks of its ranks."""

    def __init__(self, num_nodes, world_size, rank, row_groups=1, blk_align=1):
        self.num_nodes, self.world_size, self.rank = int(num_nodes), int(world_size), int(rank)
        self.Gr = int(row_groups)
        if self.Gr < 1 or self.world_size % self.Gr:
            raise ValueError('row_groups must divide the world size')
        self.Gc = self.world_size // self.Gr
        self.rg, self.cg = divmod(self.rank, self.Gc)
        # blk_align: row blocks that split into equal row chunks (the chunked pipeline below)
        self.blk = _ceil_div(max(_ceil_div(self.num_nodes, self.world_size), 2),
                             int(blk_align)) * int(blk_align)
        self.n_padded = self.blk * self.world_size
        self.lo = min(self.rank * self.blk, self.num_nodes)
        self.hi = min(self.lo + self.blk, self.num_nodes)
        self.g_lo = min(self.rg * self.Gc * self.blk, self.num_nodes)
        self.g_hi = min(self.g_lo + self.Gc * self.blk, self.num_nodes)

    @property
    def n_local(self):
        return self.hi - self.lo

    @property
    def n_group(self):
        return self.g_hi - self.g_lo

    def widths(self, F, itemsize=4):
        """(Fp, Fc): F padded so that every column slice is a 32-byte multiple — rows
        that start inside a 32-byte sector slow the gather by 20-25 %
        (profiles/r01_k1_shapes.md)."""
        u32, u128 = max(32 // itemsize, 1), max(128 // itemsize, 1)
        per = _ceil_div(max(F, 1), self.Gc)
        Fc = _ceil_div(per, u32) * u32
        # whole 128-byte lines per slice when that costs <= 15 % more columns: the gather
        # is faster (2.1 vs 2.4 ms for 64 vs 56 columns) and the row buffers can be kept
        # slice-major (contiguous peer stores, PeerTransport.aggregate_to_rows)
        wide = _ceil_div(per, u128) * u128
        if wide * 100 <= Fc * 115:
            Fc = wide
        return Fc * self.Gc, Fc

    def col_slice(self, F, itemsize=4, cg=None):
        Fp, Fc = self.widths(F, itemsize)
        c = self.cg if cg is None else cg
        return slice(c * Fc, (c + 1) * Fc)

    def exchange_bytes(self, F, itemsize=4):
        """Bytes one rank sends when an [N, F] matrix goes from the row layout to the
        column layout (every column block to the Gr ranks that aggregate it)."""
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


def _row_major(rows, n):
    """``[n, width]`` view/copy of a row-layout state that may be slice-major ``[S, blk, Fc]``."""
    if rows.dim() == 3:
        return rows.permute(1, 0, 2).reshape(rows.size(1), -1)[:n]
    return rows


def _bias_slice(plan, bias, F, itemsize):
    if bias is None:
        return None
    Fp, _ = plan.widths(F, itemsize)
    return _pad_cols(bias.view(1, -1), Fp).view(-1)[plan.col_slice(F, itemsize)].contiguous()


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

    def prepare(self, fp_max):
        pass

    def group_cols_to_rows(self, z, F):
        """z [n_group, Fc]: this rank's column slice of the rows of its row group ->
        rows [n_local, Fp].  Only the ranks of the same row group exchange data."""
        p = self.plan
        Fp, Fc = p.widths(F, z.element_size())
        send = z.new_zeros((p.world_size, p.blk, Fc))
        mine = _pad_rows(z[:, :Fc], p.Gc * p.blk).view(p.Gc, p.blk, Fc)
        send[p.rg * p.Gc:(p.rg + 1) * p.Gc] = mine
        recv = self._a2a(send)[p.rg * p.Gc:(p.rg + 1) * p.Gc]    # [cg'] = my rows, slice cg'
        return recv.permute(1, 0, 2).reshape(p.blk, Fp)[:p.n_local]

    def cols_to_rows(self, cols, F):
        """cols [>=N, Fc] (all rows) -> rows [n_local, Fp]: pure redistribution."""
        p = self.plan
        return self.group_cols_to_rows(cols[p.g_lo:p.g_hi], F)

    def rows_to_cols(self, rows, F, relu=False):
        """rows [n_local, >=F] -> cols [n_padded, Fc] (rows beyond N are padding)."""
        p = self.plan
        Fp, Fc = p.widths(F, rows.element_size())
        r = _pad_rows(_pad_cols(rows[:, :min(rows.size(1), Fp)], Fp), p.blk)
        if relu:
            r = torch.relu(r)
        blocks = r.view(p.blk, p.Gc, Fc).permute(1, 0, 2)          # [cg', blk, Fc]
        send = blocks.repeat(p.Gr, 1, 1).contiguous()              # rank (rg', cg') gets block cg'
        return self._a2a(send).view(p.n_padded, Fc)

    def aggregate_to_rows(self, graph, cols, F, bias, aggregate):
        b = _bias_slice(self.plan, bias, F, cols.element_size())
        return self.group_cols_to_rows(aggregate(graph, cols, b), F)       # z: [n_group, Fc]

    def transform_to_cols(self, graph, rows, w, bias, F, relu, transform, out):
        return self.rows_to_cols(transform(rows, w, bias, out, relu), F, False)


class _DevMem(object):
    """A raw device range exposed to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {'shape': (int(nbytes),), 'typestr': '|u1',
                                         'data': (int(ptr), False), 'version': 2}


class PeerTransport(object):
    """Layout changes by plain stores into peer HBM (CUDA IPC over NVLink)."""

    fused = True
    _FLAG_BYTES = 1024                  # 64 ranks x 8-byte epoch flags at the head of the arena
    FENCE_TIMEOUT_NS = 20 * 10 ** 9

    def __init__(self, plan, device, group=None, fence_mode='nccl'):
        from .. import _native as nat
        self.fence_mode = fence_mode
        self.nat = nat
        self.plan, self.group, self.device = plan, group, torch.device(device)
        if plan.Gr > 4:
            raise ValueError('PeerTransport supports at most 4 row groups')
        self.bytes_sent_total = 0
        self.collectives = 0            # fences (one-element all-reduces), not data
        self.event_log = None
        self._cap = None                # (Fp_max elements) the arena was sized for
        self._base = None               # this rank's arena
        self._peer = []                 # arena base address of every rank, as seen from here
        self._opened = []
        self._flip = 0
        self.rotate = True
        self.slice_major = True
        self._tabs = {}
        self._flag = torch.zeros(1, dtype=torch.float32, device=self.device)
        self._side = None               # high-priority stream of the chunked pipeline
        self.gemm_ctas = 0              # SMs of the overlapped routed transform (0 = from the sizes)
        nat.load()

    # ---- arena ---------------------------------------------------------------------
    def _layout(self, fp_max):
        p = self.plan
        r_bytes = (p.blk * fp_max * 4 + 255) // 256 * 256
        c_bytes = (p.n_padded * (fp_max // p.Gc) * 4 + 255) // 256 * 256
        return r_bytes, c_bytes

    def prepare(self, fp_max):
        if fp_max > 0:
            self.ensure(fp_max)

    def ensure(self, fp_max):
        """Collective: (re)allocate the arena for row widths up to ``fp_max`` fp32 columns.
        Invalidates every view handed out before: call it (``prepare``) ahead of a forward."""
        if self._cap is not None and fp_max <= self._cap:
            return
        nat, p = self.nat, self.plan
        self.close()
        r_bytes, c_bytes = self._layout(fp_max)
        total = self._FLAG_BYTES + 2 * r_bytes + c_bytes
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
        self._r_off = (self._FLAG_BYTES, self._FLAG_BYTES + r_bytes)
        self._c_off = self._FLAG_BYTES + 2 * r_bytes
        self._flag_tab = torch.tensor(self._peer, dtype=torch.int64).to(self.device)
        self._epoch = 0
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
        rank's later kernels start.  ``fence_mode='flags'``: a one-block kernel that
        exchanges epoch flags through the peer-mapped arenas (pygb_peer_barrier);
        ``'nccl'``: a one-element all-reduce on the current stream."""
        p = self.plan
        if p.world_size > 1:
            if self.fence_mode == 'flags' and self._base is not None:
                self._epoch += 1
                self.nat.call('pygb_peer_barrier', self.device, self.nat.ptr(self._flag_tab),
                              p.rank, p.world_size, self._epoch, self.FENCE_TIMEOUT_NS)
            else:
                dist.all_reduce(self._flag, group=self.group)
        self.collectives += 1

    def _check(self, t):
        if t.dtype != torch.float32 or not t.is_cuda:
            raise TypeError('PeerTransport moves fp32 CUDA tensors')

    def _cols_table(self):
        """Column block j of this rank's rows goes to the ranks (rg', j) of every row group."""
        p = self.plan
        return self._table(('c',), [self._peer[r * p.Gc + j] + self._c_off
                                    for j in range(p.Gc) for r in range(p.Gr)])

    def _sent_cols(self, Fc):
        p = self.plan
        self.bytes_sent_total += (p.world_size - 1) * p.n_local * Fc * 4

    # ---- layout changes ---------------------------------------------------------------
    def aggregate_to_rows(self, graph, cols, F, bias, aggregate=None):
        """The aggregation kernel stores every finished row slice into the row buffer of
        the rank that owns the row (peers of the same row group)."""
        from .. import ops
        self._check(cols)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        b = _bias_slice(p, None if bias is None else bias.float(), F, 4)
        self._flip ^= 1
        off = self._r_off[self._flip]
        # Row buffer layout.  Row-major [blk, Fp]: every rank of the group stores 4*Fc-byte
        # pieces at a 4*Fp-byte stride into every peer.  Slice-major [Gc, blk, Fc] (slices of
        # whole 128-byte lines, which the GEMM reads through a 3-D TMA map): every rank's
        # stores into a peer are one contiguous stream — ~10 % faster at 64 columns x 4 ranks.
        sliced = self.slice_major and Fc % 32 == 0 and p.Gc > 1
        if sliced:
            tab = self._table(('rs', self._flip, Fc),
                              [self._peer[p.rg * p.Gc + j] + off + p.cg * p.blk * Fc * 4
                               for j in range(p.Gc)])
        else:
            tab = self._table(('r', self._flip, Fc),
                              [self._peer[p.rg * p.Gc + j] + off + p.cg * Fc * 4
                               for j in range(p.Gc)])
        # start at the row block of the next rank of the group: at any time every rank
        # receives from one peer instead of all ranks storing into the same one
        start = ((p.cg + 1) % p.Gc) * p.blk if self.rotate else 0
        ops.seg_gather_reduce_routed_raw(cols[:, :Fc], graph.adj.dst, graph.w_csr, b, tab, p.blk,
                                         Fc if sliced else Fp,
                                         start_row=start if start < p.n_group else 0)
        self.bytes_sent_total += (p.Gc - 1) * p.blk * Fc * 4
        self.fence()
        if sliced:
            return self._view(off, p.blk, Fp).view(p.Gc, p.blk, Fc)      # [slice, row, column]
        return self._view(off, p.blk, Fp)[:p.n_local]

    def transform_to_cols(self, graph, rows, w, bias, F, relu, transform=None, out=None):
        """``relu?(rows @ w + bias)`` with the epilogue of the GEMM storing the column
        blocks straight into the column buffers of their ranks."""
        from .. import ops
        self._check(rows)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        if p.n_local > 0 and not ops._gemm_eligible(rows, w):
            return self.rows_to_cols(ops.linear(_row_major(rows, p.n_local), w, bias), F, relu)
        if p.n_local > 0:
            ops.gemm_xw_routed_raw(rows, w, bias, Fc, self._cols_table(), p.lo, Fc, relu, p.Gr,
                                   rows=p.n_local)
        self._sent_cols(Fc)
        self.fence()
        return self._view(self._c_off, p.n_padded, Fc)

    def in_arena(self, t):
        return (self._base is not None and self._mem is not None
                and self._base <= t.data_ptr() < self._base + self._mem.numel())

    def aggregate_transform_pipelined(self, graph, cols, F_agg, w, bias, F_out, relu):
        """``relu?((A cols) @ w + bias)`` -> column layout, as a pipeline over row chunks.

        The aggregation of row chunk c stores straight into the row buffers of the owners (as
        in ``aggregate_to_rows``); after a fence the owner's routed transform of chunk c runs
        on a second, high-priority stream, with a capped grid, while the aggregation of chunk
        c+1 runs on the SMs left to it.  The transform's epilogue is NVLink-bound (every rank
        receives its column slice of all N rows: 548 MB at P = 8) and the aggregation is
        HBM-bound, so the two overlap instead of adding up.  Results are
        bit-identical to the unchunked path: every row is summed in the same slot order."""
        from .. import ops
        self._check(cols)
        p = self.plan
        C = graph.chunks
        blk_c = p.blk // C
        Fpa, Fca = p.widths(F_agg, 4)
        Fpo, Fco = p.widths(F_out, 4)
        self.ensure(max(Fpa, Fpo))
        self._flip ^= 1
        off = self._r_off[self._flip]
        sliced = self.slice_major and Fca % 32 == 0 and p.Gc > 1
        ldo = Fca if sliced else Fpa
        main = torch.cuda.current_stream(self.device)
        if self._side is None:
            self._side = torch.cuda.Stream(self.device, priority=-1)
        side = self._side
        cols_tab = self._cols_table()
        rows_all = self._view(off, p.blk, Fpa)
        if sliced:
            rows_all = rows_all.view(p.Gc, p.blk, Fca)
        rot = ((p.cg + 1) % p.Gc) * blk_c if self.rotate else 0
        # SMs for the overlapped transform: enough that its tensor work (3xTF32, ~4.5 TFLOP/s
        # per SM) keeps pace with what NVLink takes (~700 GB/s of the epilogue's peer stores)
        gemm_ctas = self.gemm_ctas
        if gemm_ctas <= 0:
            flops = 6.0 * p.n_local * w.size(0) * w.size(1)
            t_link = max((p.world_size - 1) * p.n_local * Fco * 4 / 700e9, 1e-6)
            gemm_ctas = int(min(max(flops / (4.5e12 * t_link) + 1, 24), 96))
        for c in range(C):
            adj_c, w_c = graph.chunk(c)
            if sliced:
                base = off + (p.cg * p.blk + c * blk_c) * Fca * 4
            else:
                base = off + c * blk_c * Fpa * 4 + p.cg * Fca * 4
            tab = self._table(('rc', self._flip, Fca, c, C),
                              [self._peer[p.rg * p.Gc + j] + base for j in range(p.Gc)])
            ops.seg_gather_reduce_routed_raw(cols[:, :Fca], adj_c.dst, w_c, None, tab, blk_c, ldo,
                                             start_row=rot,
                                             max_ctas=-gemm_ctas if c > 0 else 0)
            self.fence()
            ev = torch.cuda.Event()
            ev.record(main)
            n_c = min(max(p.n_local - c * blk_c, 0), blk_c)
            with torch.cuda.stream(side):
                side.wait_event(ev)
                if n_c > 0:
                    a = rows_all[:, c * blk_c:c * blk_c + n_c] if sliced else \
                        rows_all[c * blk_c:c * blk_c + n_c]
                    ops.gemm_xw_routed_raw(a, w, bias, Fco, cols_tab, p.lo + c * blk_c, Fco, relu,
                                           p.Gr, rows=n_c,
                                           max_ctas=gemm_ctas if c + 1 < C else 0)
        main.wait_stream(side)
        self.bytes_sent_total += (p.Gc - 1) * p.blk * Fca * 4
        self._sent_cols(Fco)
        self.fence()
        return self._view(self._c_off, p.n_padded, Fco)

    def own_rows_transform_pipelined(self, graph, x, F0, w, bias, F_out, relu):
        """First layer from a replicated input, pipelined: the aggregation of row chunk c+1 of
        the rank's OWN rows (local output, no fence) runs while the routed transform of chunk c
        sends its column blocks over NVLink from a second, high-priority stream with a capped
        grid.  Unlike the column-sliced pipeline there is nothing to wait for between chunks —
        the rows are local — and at full width the aggregation is HBM-bound, not issue-bound."""
        from .. import ops
        p = self.plan
        C = graph.chunks
        Fpo, Fco = p.widths(F_out, 4)
        self.ensure(Fpo)
        main = torch.cuda.current_stream(self.device)
        if self._side is None:
            self._side = torch.cuda.Stream(self.device, priority=-1)
        side = self._side
        cols_tab = self._cols_table()
        vec = 4
        ld = (F0 + vec - 1) // vec * vec
        rows = graph.buffer(('own_rows', F0), max(p.n_local, 1), ld, torch.float32)
        gemm_ctas = self.gemm_ctas
        if gemm_ctas <= 0:
            flops = 6.0 * p.n_local * w.size(0) * w.size(1)
            t_link = max((p.world_size - 1) * p.n_local * Fco * 4 / 700e9, 1e-6)
            gemm_ctas = int(min(max(flops / (4.5e12 * t_link) + 1, 24), 96))
        for c in range(C):
            adj_c, w_c, r0, n_c = graph.own_chunk(c)
            if n_c == 0:
                continue
            tab = self._table(('own', F0, c, C), [rows.data_ptr() + r0 * ld * 4])
            ops.seg_gather_reduce_routed_raw(x[:, :F0], adj_c.dst, w_c, None, tab, max(n_c, 2), ld,
                                             max_ctas=-gemm_ctas if c > 0 else 0)
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(side):
                side.wait_event(ev)
                ops.gemm_xw_routed_raw(rows[r0:r0 + n_c, :F0], w, bias, Fco, cols_tab, p.lo + r0,
                                       Fco, relu, p.Gr, rows=n_c,
                                       max_ctas=gemm_ctas if c + 1 < C else 0)
        main.wait_stream(side)
        self._sent_cols(Fco)
        self.fence()
        return self._view(self._c_off, p.n_padded, Fco)

    def rows_to_cols(self, rows, F, relu=False):
        from .. import ops
        self._check(rows)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        if rows.size(1) < Fp or rows.stride(1) != 1 or rows.stride(0) % 4 or \
                rows.data_ptr() % 16:
            rows = _pad_cols(rows[:, :min(rows.size(1), Fp)], Fp)
        if rows.size(0) > 0:
            ops.cols_scatter_raw(rows, Fc, p.Gc, self._cols_table(), p.lo, Fc, relu, p.Gr)
        self._sent_cols(Fc)
        self.fence()
        return self._view(self._c_off, p.n_padded, Fc)

    def cols_to_rows(self, cols, F):
        """Pure redistribution (no aggregation): strided device copies into the peers
        of the same row group."""
        self._check(cols)
        p = self.plan
        Fp, Fc = p.widths(F, 4)
        self.ensure(Fp)
        self._flip ^= 1
        off = self._r_off[self._flip]
        for j in range(p.Gc):
            q = p.rg * p.Gc + j
            n = min(p.blk, max(p.num_nodes - q * p.blk, 0))
            if n == 0:
                continue
            dst = torch.as_tensor(_DevMem(self._peer[q] + off, p.blk * Fp * 4), device=self.device)
            dst = dst.view(torch.float32).view(p.blk, Fp)
            dst[:n, p.cg * Fc:(p.cg + 1) * Fc].copy_(cols[q * p.blk:q * p.blk + n, :Fc])
        self.bytes_sent_total += (p.Gc - 1) * p.blk * Fc * 4
        self.fence()
        return self._view(off, p.blk, Fp)[:p.n_local]


class SlicedGraph(object):
    """The (loop-augmented, normalised) in-edges of this rank's row group — the whole
    graph when there is one row group — plus the row block / column slice bookkeeping."""

    def __init__(self, edge_index, norm, num_nodes, rank, world_size, group=None, i=1,
                 transport='peer', build_adj=True, row_groups=1, fence_mode='nccl'):
        self.plan = p = SlicePlan(num_nodes, world_size, rank, row_groups,
                                  blk_align=8 if transport == 'peer' else 1)
        self.i = i
        self.chunks = 1
        self._chunk_graphs = {}
        self.device = edge_index.device
        # in-edges of this rank's OWN row block (destinations renumbered from 0): the first
        # layer of a forward whose input is replicated (`x_layout='full'`) aggregates them at
        # full width with no exchange at all
        self.own_adj = self.own_w = self.own_norm = None
        if build_adj:
            from ..graph import Adj
            t = edge_index[i]
            m_own = (t >= p.lo) & (t < p.hi)
            ei_o = edge_index[:, m_own].clone()
            ei_o[i] -= p.lo
            size = [0, 0]
            size[i], size[1 - i] = p.n_local, num_nodes
            self.own_adj = Adj(ei_o, i, size)
            self.own_norm = None if norm is None else norm[m_own].float().contiguous()
            self.own_w = None if norm is None else self.own_adj.permuted(self.own_norm, 'dst')
            self.num_edges_own = int(ei_o.size(1))
            del t, m_own
        if p.Gr > 1:                                   # in-edges of the row group, renumbered
            t = edge_index[i]
            mask = (t >= p.g_lo) & (t < p.g_hi)
            ei = edge_index[:, mask].clone()
            ei[i] -= p.g_lo
            edge_index, norm = ei, (None if norm is None else norm[mask].contiguous())
        self.edge_index, self.norm = edge_index, norm
        self.num_edges_local = edge_index.size(1)
        self.adj = self.w_csr = None
        if build_adj:
            from ..graph import Adj
            size = [0, 0]
            size[i], size[1 - i] = p.n_group, num_nodes
            self.adj = Adj(edge_index, i, size)
            self.w_csr = None if norm is None else self.adj.permuted(norm.float(), 'dst')
        if transport == 'peer':
            self.transport = PeerTransport(p, self.device, group, fence_mode)
        elif transport == 'collective':
            self.transport = CollectiveTransport(p, group)
        else:
            raise ValueError("transport must be 'peer' or 'collective'")
        self._bufs = {}

    def set_chunks(self, chunks):
        """Row chunks of the pipelined aggregate -> transform step (peer transport; 1 = off).
        Must divide the row block (a multiple of 8)."""
        chunks = max(int(chunks), 1)
        if chunks > 1 and (self.plan.blk % chunks or not getattr(self.transport, 'fused', False)):
            raise ValueError('chunks must divide the row block ({}) and need the peer '
                             'transport'.format(self.plan.blk))
        if chunks != self.chunks:
            self._chunk_graphs = {}
        self.chunks = chunks
        for c in range(chunks if chunks > 1 else 0):
            self.chunk(c)                           # sort the sub-graphs now, not in a timed step
            if self.own_adj is not None:
                self.own_chunk(c)

    def chunk(self, c):
        """(Adj, w_csr) of row chunk ``c``: the in-edges of rows ``[c*blk_c, (c+1)*blk_c)`` of
        EVERY row block of the group, destinations renumbered ``block*blk_c + offset`` — the
        stable sort keeps each row's edge order, so sums equal the whole graph's bit for bit."""
        hit = self._chunk_graphs.get(c)
        if hit is not None:
            return hit
        from ..graph import Adj
        p, i = self.plan, self.i
        blk_c = p.blk // self.chunks
        t = self.edge_index[i]
        r = t % p.blk
        mask = (r // blk_c) == c
        ei = self.edge_index[:, mask].clone()
        ei[i] = (t[mask] // p.blk) * blk_c + (r[mask] % blk_c)
        size = [0, 0]
        size[i], size[1 - i] = p.Gc * blk_c, p.num_nodes
        adj = Adj(ei, i, size)
        w = None if self.norm is None else adj.permuted(self.norm[mask].float().contiguous(), 'dst')
        adj.dst.items(None)
        adj.dst.flagged('col')
        self._chunk_graphs[c] = (adj, w)
        return adj, w

    def own_chunk(self, c):
        """(Adj, w_csr, first row, rows) of chunk ``c`` of the rank's own rows: the in-edges of
        the contiguous row range, destinations renumbered from 0 (stable: same per-row order)."""
        key = ('own', c, self.chunks)
        hit = self._chunk_graphs.get(key)
        if hit is not None:
            return hit
        from ..graph import Adj
        p, i = self.plan, self.i
        per = -(-p.n_local // self.chunks)
        per = -(-per // 8) * 8
        r0 = min(c * per, p.n_local)
        n_c = min(per, p.n_local - r0)
        if n_c <= 0:
            hit = (None, None, r0, 0)
        else:
            ei_o = self.own_adj.edge_index
            t = ei_o[i]
            mask = (t >= r0) & (t < r0 + n_c)
            ei = ei_o[:, mask].clone()
            ei[i] -= r0
            size = [0, 0]
            size[i], size[1 - i] = max(n_c, 2), p.num_nodes
            adj = Adj(ei, i, size)
            w = None if self.own_norm is None else \
                adj.permuted(self.own_norm[mask].contiguous(), 'dst')
            adj.dst.items(None)
            adj.dst.flagged('col')
            hit = (adj, w, r0, n_c)
        self._chunk_graphs[key] = hit
        return hit

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

    def padded(self, key, t, rows, cols):
        """``t`` zero-padded to ``[rows, cols]`` (or ``[cols]``), cached per tensor version."""
        if t is None:
            return None
        if tuple(t.shape) == ((rows, cols) if t.dim() == 2 else (cols,)):
            return t
        hit = self._bufs.get(key)
        if hit is None or hit[0] is not t or hit[1] != t._version or \
                hit[2].shape[-1] != cols or (t.dim() == 2 and hit[2].size(0) != rows):
            d = t.detach()
            out = d.new_zeros((rows, cols) if t.dim() == 2 else (cols,))
            if t.dim() == 2:
                out[:t.size(0), :t.size(1)] = d
            else:
                out[:t.numel()] = d
            hit = (t, t._version, out)
            self._bufs[key] = hit
        return hit[2]


def _native_aggregate(graph, cols, bias):
    from .. import ops
    return ops.gather_reduce(cols, graph.adj, graph.norm, 'sum', bias)


def _native_transform(x, weight, bias, out, relu=False):
    from .. import ops
    n = weight.size(1)
    if out.size(0) == 0:
        return out
    if ops._gemm_eligible(x, weight):                           # bias and ReLU in the epilogue
        ops.gemm_xw_raw(x, weight, bias, out=out, relu=relu, rows=out.size(0))
    else:
        y = torch.matmul(_row_major(x, out.size(0)), weight)
        y = y + bias if bias is not None else y
        out[:, :n].copy_(torch.relu(y) if relu else y)
    return out


def sliced_gcn_forward(graph, x, layers, relu=True, x_layout='cols', aggregate=_native_aggregate,
                       transform=_native_transform):
    """GCN forward with ``layers = [(weight [in,out], bias|None), ...]``; ReLU between
    layers when ``relu``.  ``x``: this rank's column slice ``[n_padded, Fc]``
    (``x_layout='cols'``, see ``SlicedGraph.slice_cols``), its rows ``[n_local, in]``
    (``'rows'``), or the whole ``[N, in]`` matrix replicated on every rank (``'full'``: a static
    input such as the dataset features — the first aggregation then runs over the rank's own
    row block at full width, with nothing to exchange before or after it).  Returns this
    rank's rows ``[n_local, out]`` of the last layer.

    Every layer aggregates at ``min(in, out)`` channels (``A (X W) = (A X) W``).  The
    forward is a chain of two primitives — aggregate (column layout -> row layout) and
    transform (row layout -> row layout, or -> column layout when an aggregation follows:
    then the layout change is fused into the GEMM epilogue)."""
    p, tr = graph.plan, graph.transport
    esz = x.element_size()
    ops_seq = []                                           # ('agg', F, bias, act) | ('gemm', k, w, b, act)
    for k, (w, b) in enumerate(layers):
        act = bool(relu) and k + 1 < len(layers)
        if w.size(0) <= w.size(1):
            ops_seq += [('agg', w.size(0), None, False), ('gemm', k, w, b, act)]
        else:
            ops_seq += [('gemm', k, w, None, False), ('agg', w.size(1), b, act)]
    # size the peer arena once, before any view of it exists (a later, wider layer must not
    # re-allocate it under a live row buffer)
    tr.prepare(max([p.widths(f, esz)[0] for w, _ in layers for f in w.shape] or [0]))
    state, kind = x, x_layout
    skip = False
    if x_layout == 'full':
        first = ops_seq[0] if ops_seq else None
        if (first is not None and first[0] == 'agg' and aggregate is _native_aggregate
                and graph.own_adj is not None and p.n_local > 0):
            from .. import ops
            _, F0, b0, act0 = first
            nxt_ = ops_seq[1] if len(ops_seq) > 1 else None
            nxt2_ = ops_seq[2] if len(ops_seq) > 2 else None
            if (getattr(graph, 'chunks', 1) > 1 and getattr(tr, 'fused', False) and b0 is None
                    and not act0 and transform is _native_transform and F0 % 4 == 0
                    and x.dtype == torch.float32 and nxt_ is not None and nxt_[0] == 'gemm'
                    and nxt2_ is not None and nxt2_[0] == 'agg'):
                _, k_, w_, b_, act_ = nxt_
                fpo_, _ = p.widths(w_.size(1), esz)
                wp_ = graph.padded(('w', k_), w_, F0, fpo_)
                bp_ = graph.padded(('b', k_), b_, 0, fpo_)
                state = tr.own_rows_transform_pipelined(graph, x, F0, wp_, bp_, w_.size(1), act_)
                kind, skip = 'cols', 2
            else:
                csr = graph.own_adj.dst
                state = ops.seg_gather_reduce_raw(x[:, :F0], csr, csr.col, 'sum', graph.own_w, None,
                                                  b0)
                if act0:
                    state = torch.relu_(state)
                kind, skip = 'rows', 1
        elif first is not None and first[0] == 'agg':
            state, kind = graph.slice_cols(x), 'cols'
        else:
            state, kind = x[p.lo:p.hi], 'rows'
    for n_op, op in enumerate(ops_seq):
        if skip:
            skip = int(skip) - 1
            continue
        nxt = ops_seq[n_op + 1] if n_op + 1 < len(ops_seq) else None
        nxt2 = ops_seq[n_op + 2] if n_op + 2 < len(ops_seq) else None
        if (op[0] == 'agg' and getattr(graph, 'chunks', 1) > 1 and getattr(tr, 'fused', False)
                and aggregate is _native_aggregate and transform is _native_transform
                and kind == 'cols' and op[2] is None and not op[3]
                and nxt is not None and nxt[0] == 'gemm' and nxt2 is not None and nxt2[0] == 'agg'
                and state.dtype == torch.float32 and not tr.in_arena(state)):
            # aggregate -> transform -> (next aggregation): pipelined over row chunks
            _, F, _, _ = op
            _, k, w, b, act = nxt
            fout = w.size(1)
            fpa, _ = p.widths(F, esz)
            fpo, _ = p.widths(fout, esz)
            wp = graph.padded(('w', k), w, fpa, fpo)
            bp = graph.padded(('b', k), b, 0, fpo)
            state = tr.aggregate_transform_pipelined(graph, state, F, wp, bp, fout, act)
            kind = 'cols'
            skip = True
            continue
        if op[0] == 'agg':
            _, F, b, act = op
            if kind == 'rows':
                state = tr.rows_to_cols(_row_major(state, p.n_local), F, False)
            state = tr.aggregate_to_rows(graph, state, F, b, aggregate)      # bias fused
            if act:
                state = torch.relu_(state)
            kind = 'rows'
        else:
            _, k, w, b, act = op
            fin, fout = w.shape
            if kind == 'cols':
                state = tr.cols_to_rows(state, fin)
            fp_out, _ = p.widths(fout, esz)
            k_in = state.size(0) * state.size(2) if state.dim() == 3 else state.size(1)
            wp = graph.padded(('w', k), w, k_in, fp_out)            # zero rows / columns: padding
            bp = graph.padded(('b', k), b, 0, fp_out)
            out = graph.buffer(('y', k), p.n_local, fp_out, w.dtype)
            if n_op + 1 < len(ops_seq) and ops_seq[n_op + 1][0] == 'agg':
                state = tr.transform_to_cols(graph, state, wp, bp, fout, act, transform, out)
                kind = 'cols'
            else:
                if state.dim() == 3 and transform is not _native_transform:
                    state = _row_major(state, p.n_local)
                state = transform(state, wp, bp, out, act)
                kind = 'rows'
    if not layers:
        return state
    fout = layers[-1][0].size(1)
    if kind == 'cols':
        state = tr.cols_to_rows(state, fout)
    return _row_major(state, p.n_local)[:, :fout]
