"""Tensor-level operators over the C ABI, with autograd.

These are the pieces ``MessagePassing.propagate``, ``utils.scatter_`` and
``utils.softmax`` are assembled from.  Everything here launches kernels of
``libpygb.so`` on the tensor's current CUDA stream; there is no CPU path.
"""
import collections
import os

import torch

from . import _native as nat
from .graph import adj_for, csr_for  # noqa: F401  (re-exported)

NEG_INF = float('-inf')
# K1 routing: the streaming kernel (seg_stream.cu) when the rows are 16-byte aligned and
# padded, else the one-warp-per-row kernel.  STREAM_T forces the item size (tests).
STREAM = True
STREAM_T = None
# K1: megabytes of L2 the most frequently gathered source rows may claim (evict_last hints on
# their gathers, pygb_seg_mark_hot); 0 disables the hints
# (products graph, tools/k1_hot_probe.py: F=256 9.3-9.4 -> 8.9-9.05 ms, F=100 -3..5 %, F=64 -7 %
# with 16-32 MB; larger budgets lose again)
K1_HOT_MB = int(os.environ.get('PYGB_K1_HOT_MB', '24'))
LINE_ALIGN = os.environ.get('PYGB_LINE_ALIGN', '1') not in ('', '0')


def _rows2d(x):
    """View ``x`` as [n, F] with unit inner stride; returns (x2d, F)."""
    n = x.size(0)
    if x.numel() == 0:
        f = 1
        for d in x.shape[1:]:
            f *= d
        return x.new_empty((n, f))
    x2 = x.reshape(n, -1)
    if x2.numel() > 0 and x2.stride(1) != 1:
        x2 = x2.contiguous()
    if x2.size(1) == 1 and x2.stride(0) < 1:
        x2 = x2.contiguous()
    return x2


def _unflat(out2, n, tail):
    """[n, F] (possibly a padded-row view) -> [n, *tail]."""
    tail = tuple(tail)
    if len(tail) == 1:
        return out2
    if len(tail) == 0:
        return out2.reshape(n)
    return out2.reshape((n,) + tail)


def _ld(x2):
    return x2.stride(0) if x2.size(0) > 1 else max(x2.size(1), 1)


def _f32(t):
    if t is None:
        return None
    t = t if t.dtype == torch.float32 else t.float()
    return t if t.is_contiguous() else t.contiguous()


# --------------------------------------------------------------------------
# raw (no autograd) launches
# --------------------------------------------------------------------------
_seen_wide = collections.OrderedDict()     # identities of large gather sources seen once
_line_cache = collections.OrderedDict()    # identity -> copy with a 128-byte multiple row stride


def _line_aligned_rows(x2):
    """A large, constant gather source whose rows are not whole 128-byte lines (cfg5 layer 1:
    F = 100 fp32 = 400-byte rows that start mid-sector, 0.89 of the HBM peak against 1.1 at
    F = 256 — profiles/r01_k1_shapes.md) is gathered from a copy with the row stride padded to
    128 bytes.  Like the sorted graph the copy is made once per tensor identity — and only for
    a tensor that is seen a second time unchanged, so a source that is rewritten every step
    (end-to-end uploads) is never copied."""
    esz = x2.element_size()
    row_b = x2.size(1) * esz
    if (row_b % 128 == 0 or row_b < 256 or x2.numel() * esz < (64 << 20) or x2.requires_grad
            or x2.stride(1) != 1 or (x2.stride(0) * esz) % 128 == 0):
        return x2
    key = (x2.data_ptr(), tuple(x2.shape), tuple(x2.stride()), x2._version, x2.device.index,
           x2.dtype)
    hit = _line_cache.get(key)
    if hit is not None:
        _line_cache.move_to_end(key)
        return hit[0]
    if key not in _seen_wide:
        _seen_wide[key] = True
        while len(_seen_wide) > 16:
            _seen_wide.popitem(last=False)
        return x2
    ld = (row_b + 127) // 128 * 128 // esz
    xp = torch.empty((x2.size(0), ld), dtype=x2.dtype, device=x2.device)
    xp[:, :x2.size(1)].copy_(x2)
    xp[:, x2.size(1):].zero_()
    out = xp[:, :x2.size(1)]
    _line_cache[key] = (out, x2)
    while len(_line_cache) > 2:
        _line_cache.popitem(last=False)
    return out

def seg_gather_reduce_raw(x, csr, idx, reduce='sum', w=None, wperm=None, bias=None,
                          heads=None, idx_key=None):
    """out[i] = reduce_p w_p * x[idx[p]] over the rows of ``csr``.  ``idx`` is ``csr.col``,
    ``csr.perm`` or — with a cache key ``idx_key`` — any int32 ``[E]`` index in slot order."""
    dev = nat.require_cuda(x, idx, w, bias)
    x2 = _rows2d(x)
    F = x2.size(1)
    if heads is None and wperm is None and STREAM and LINE_ALIGN:
        x2 = _line_aligned_rows(x2)
    # rows padded to a 16-byte multiple keep the kernel on its 128-bit path
    vec = 16 // x2.element_size()
    ldo = (F + vec - 1) // vec * vec if F > 1 else F
    out_full = torch.empty((csr.n_rows, ldo), dtype=x2.dtype, device=dev)
    out = out_full[:, :F] if ldo != F else out_full
    if csr.n_rows == 0 or F == 0:
        return out
    if csr.E == 0:
        out.zero_()
        if bias is not None:
            out += bias.to(out.dtype)
        return out
    dt = nat.dtype_code(x2)
    w = _f32(w)
    bias = _f32(bias)
    ldx = _ld(x2)
    if (heads is None and wperm is None and STREAM and F > 1
            and (x2.data_ptr() % 16 or ldx % vec or ldx < ldo)):
        # rows that are not 16-byte multiples (e.g. a contiguous [N, 47] gradient): one padded
        # copy (2 passes over N*F) buys the streaming kernel (vs ~3x slower per-row kernel)
        xp = torch.empty((x2.size(0), ldo), dtype=x2.dtype, device=dev)
        xp[:, :F].copy_(x2)
        if ldo > F:
            xp[:, F:].zero_()
        x2, ldx = xp, ldo
    if (heads is None and wperm is None and STREAM and F > 1 and x2.data_ptr() % 16 == 0
            and ldx % vec == 0 and ldx >= ldo):
        # K1 v2: streaming kernel over equal-slot work items
        Ts, n_sub, nz_rows, sub_info = csr.items(STREAM_T)
        if idx is csr.col:
            idxf = csr.flagged('col')
        elif idx is csr.perm:
            idxf = csr.flagged('perm')
        elif idx_key is not None:
            idxf = csr.flagged(idx_key, idx)
        else:
            raise ValueError('seg_gather_reduce_raw: idx must be csr.col or csr.perm '
                             '(or come with an idx_key)')
        entry = 'pygb_seg_gather_reduce_stream'
        if idx is csr.col and K1_HOT_MB > 0 and x2.size(0) * ldx * x2.element_size() > (K1_HOT_MB << 21):
            # source rows far beyond L2: ask it to keep the most frequently gathered ones
            hot = csr.flagged_hot((K1_HOT_MB << 20) // (ldx * x2.element_size()))
            if hot is not None:
                idxf, entry = hot, 'pygb_seg_gather_reduce_stream_hinted'
        ws_bytes = nat.workspace('pygb_seg_stream_workspace', dt, n_sub, F)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        nat.call(entry, dev, nat.MEAN if reduce == 'mean' else nat.SUM,
                 dt, nat.ptr(x2), ldx, F, nat.ptr(csr.rowptr), nat.ptr(idxf), nat.ptr(w),
                 nat.ptr(bias), csr.n_rows, csr.E, nat.ptr(nz_rows), nat.ptr(sub_info), Ts, n_sub,
                 nat.ptr(out), ldo, nat.ptr(ws), ws_bytes)
    elif heads is None:
        nat.call('pygb_seg_gather_reduce', dev, nat.MEAN if reduce == 'mean' else nat.SUM, dt,
                 nat.ptr(x2), ldx, F, nat.ptr(csr.rowptr), nat.ptr(idx), nat.ptr(w),
                 nat.ptr(wperm), nat.ptr(bias), csr.n_rows, nat.ptr(out), ldo)
    else:
        H, C = heads
        assert H * C == F and reduce == 'sum'
        nat.call('pygb_seg_gather_reduce_heads', dev, dt, nat.ptr(x2), _ld(x2), H, C,
                 nat.ptr(csr.rowptr), nat.ptr(idx), nat.ptr(w), nat.ptr(wperm), nat.ptr(bias),
                 csr.n_rows, nat.ptr(out), ldo)
    return out


def seg_gather_reduce_routed_raw(x, csr, w_csr, bias, out_tab, blk_rows, ldo, reduce='sum',
                                 start_row=0, max_ctas=0):
    """The streaming fused gather-reduce with a routed output: row ``r`` of the
    result is stored at ``out_tab[r // blk_rows] + (r % blk_rows) * ldo`` (element
    units of ``x.dtype``) — ``out_tab`` is an int64 CUDA tensor of base addresses,
    local or peer-GPU (pygb_seg_gather_reduce_stream_routed; feature-sliced
    multi-GPU aggregation, dist/sliced.py).  ``start_row``: the launch begins its work
    at (the work item holding) that destination row and wraps around — ranks that
    store into each other start at different row blocks.  Nothing is returned: the
    caller owns the destinations and orders the stores against their consumers."""
    dev = nat.require_cuda(x, w_csr, bias, out_tab)
    x2 = _rows2d(x)
    if LINE_ALIGN:
        x2 = _line_aligned_rows(x2)
    F = x2.size(1)
    vec = 16 // x2.element_size()
    if not (F > 1 and F % vec == 0 and x2.data_ptr() % 16 == 0 and _ld(x2) % vec == 0
            and ldo % vec == 0 and out_tab.dtype == torch.int64 and out_tab.is_contiguous()
            and out_tab.numel() * blk_rows >= csr.n_rows):
        raise ValueError('seg_gather_reduce_routed_raw: rows must be 16-byte multiples and the '
                         'pointer table must cover every destination row')
    if csr.n_rows == 0 or csr.E == 0:
        raise ValueError('seg_gather_reduce_routed_raw: empty graph')
    dt = nat.dtype_code(x2)
    Ts, n_sub, nz_rows, sub_info = csr.items(STREAM_T)
    idxf = csr.flagged('col')
    sub_rot = 0
    if start_row:
        key = ('slot_of_row', int(start_row))
        slot = csr._items.get(key)
        if slot is None:
            slot = int(csr.rowptr[min(int(start_row), csr.n_rows)].item())   # once per CSR
            csr._items[key] = slot
        sub_rot = min(slot // Ts, n_sub)
    entry = 'pygb_seg_gather_reduce_stream_routed'
    row_bytes = _ld(x2) * x2.element_size()
    if K1_HOT_MB > 0 and x2.size(0) * row_bytes > (K1_HOT_MB << 21):
        hot = csr.flagged_hot((K1_HOT_MB << 20) // row_bytes)     # see seg_gather_reduce_raw
        if hot is not None:
            idxf, entry = hot, 'pygb_seg_gather_reduce_stream_routed_hinted'
    ws_bytes = nat.workspace('pygb_seg_stream_workspace', dt, n_sub, F)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    nat.call(entry, dev,
             nat.MEAN if reduce == 'mean' else nat.SUM, dt, nat.ptr(x2), _ld(x2), F,
             nat.ptr(csr.rowptr), nat.ptr(idxf), nat.ptr(_f32(w_csr)), nat.ptr(_f32(bias)),
             csr.n_rows, csr.E, nat.ptr(nz_rows), nat.ptr(sub_info), Ts, n_sub, nat.ptr(out_tab),
             int(blk_rows), int(sub_rot), int(ldo), int(max_ctas), nat.ptr(ws), ws_bytes)


def cols_scatter_raw(src, Fc, n_parts, dst_tab, dst_row0, ldd, relu=False, n_dup=1):
    """Row layout -> column layout (pygb_cols_scatter): column block ``q`` of
    ``src [n, n_parts*Fc]`` is stored at ``dst_tab[q*n_dup + d] + (dst_row0 + r) * ldd``
    for every ``d < n_dup``."""
    dev = nat.require_cuda(src, dst_tab)
    assert src.dim() == 2 and src.stride(1) == 1 and src.size(1) >= Fc * n_parts
    assert dst_tab.dtype == torch.int64 and dst_tab.numel() >= n_parts * n_dup
    nat.call('pygb_cols_scatter', dev, nat.dtype_code(src), nat.ptr(src), _ld(src), src.size(0),
             int(Fc), int(n_parts), int(n_dup), nat.ptr(dst_tab), int(dst_row0), int(ldd),
             int(bool(relu)))


def gemm_xw_routed_raw(x, w, bias, Fc, dst_tab, dst_row0, ldd, relu=False, n_dup=1, passes=None,
                       rows=None, max_ctas=0):
    """``relu?(x @ w + bias)`` with the result's column blocks of ``Fc`` columns stored
    through ``dst_tab`` (pygb_gemm_xw_routed): the row -> column layout change of the
    feature-sliced multi-GPU forward fused into the GEMM epilogue.  ``x`` as in
    ``gemm_xw_raw``."""
    dev = nat.require_cuda(x, w, bias, dst_tab)
    if not _gemm_eligible(x, w):
        raise ValueError('gemm_xw_routed_raw: operands not eligible for the tcgen05 kernel')
    a_ptr, lda, a_slices, a_plane, M, K = _gemm_a_args(x, rows)
    N = w.size(1)
    assert w.size(0) == K
    assert N % Fc == 0 and dst_tab.dtype == torch.int64 and dst_tab.numel() >= N // Fc * n_dup
    ws_bytes = nat.workspace('pygb_gemm_weight_workspace', K, N)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    nat.call('pygb_gemm_xw_routed', dev, a_ptr, lda, a_slices, a_plane, M, K, nat.ptr(w),
             w.stride(0), N, nat.ptr(_f32(bias)), int(bool(relu)), nat.ptr(dst_tab), int(Fc),
             int(n_dup), int(dst_row0), int(ldd), int(passes or GEMM_PASSES), int(max_ctas),
             nat.ptr(ws), ws_bytes)


def gather_rows_raw(x, index):
    dev = nat.require_cuda(x, index)
    x2 = _rows2d(x)
    F = x2.size(1)
    E = index.numel()
    out = torch.empty((E, F), dtype=x2.dtype, device=dev)
    if E > 0 and F > 0:
        index = index if index.is_contiguous() else index.contiguous()
        nat.call('pygb_gather_rows', dev, nat.dtype_code(x2), nat.ptr(x2), _ld(x2), F,
                 nat.ptr(index), E, nat.ptr(out), F)
    return out


def edge_dot_raw(g, x, csr):
    """``out[e] = <g[dst_e], x[src_e]>`` (fp32 ``[E]``, original edge order) over the
    destination-sorted ``csr`` (pygb_edge_dot)."""
    dev = nat.require_cuda(g, x)
    g2, x2 = _rows2d(g), _rows2d(x)
    if g2.dtype != x2.dtype:
        g2, x2 = g2.float(), x2.float()
    assert g2.size(1) == x2.size(1) and g2.size(0) == csr.n_rows
    out = torch.empty((csr.E,), dtype=torch.float32, device=dev)
    if csr.E > 0:
        nat.call('pygb_edge_dot', dev, nat.dtype_code(g2), nat.ptr(g2), _ld(g2), nat.ptr(x2),
                 _ld(x2), g2.size(1), nat.ptr(csr.rowptr), nat.ptr(csr.col), nat.ptr(csr.perm),
                 csr.n_rows, csr.E, nat.ptr(out))
    return out


def seg_max_raw(src, csr, idx, eid, e_fill, zero_below=NEG_INF, mask_arg=False, want_arg=True):
    dev = nat.require_cuda(src, idx, eid)
    s2 = _rows2d(src)
    F = s2.size(1)
    out = torch.empty((csr.n_rows, F), dtype=s2.dtype, device=dev)
    arg = torch.empty((csr.n_rows, F), dtype=torch.long, device=dev) if want_arg else None
    if csr.n_rows > 0 and F > 0:
        nat.call('pygb_seg_max', dev, nat.dtype_code(s2), nat.ptr(s2), _ld(s2), F,
                 nat.ptr(csr.rowptr), nat.ptr(idx), nat.ptr(eid), csr.n_rows, int(e_fill),
                 float(zero_below), int(bool(mask_arg)), nat.ptr(out), F, nat.ptr(arg))
    return out, arg


def _row_counts(csr, dtype):
    cnt = (csr.rowptr[1:] - csr.rowptr[:-1]).clamp(min=1)
    return cnt.to(dtype).unsqueeze(1)


# --------------------------------------------------------------------------
# fused propagate for messages of the form  w_e * x_j   (GCNConv and the default)
# --------------------------------------------------------------------------
class _GatherReduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias, adj, reduce):
        shape = x.shape
        w_csr = adj.permuted(_f32(w), 'dst') if w is not None else None
        out = seg_gather_reduce_raw(x, adj.dst, adj.dst.col, reduce, w_csr, None, bias)
        ctx.adj, ctx.reduce = adj, reduce
        ctx.has_bias = bias is not None
        ctx.save_for_backward(x if (w is not None and w.requires_grad) else None, w)
        ctx.xshape = shape
        return _unflat(out, adj.n_dst, shape[1:])

    @staticmethod
    def backward(ctx, g):
        adj, reduce = ctx.adj, ctx.reduce
        x, w = ctx.saved_tensors
        g2 = _rows2d(g)
        gx = gw = gb = None
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g2.sum(0)
        gs = g2 / _row_counts(adj.dst, g2.dtype) if reduce == 'mean' else g2
        if ctx.needs_input_grad[0]:
            w_t = adj.permuted(_f32(w), 'src') if w is not None else None
            gx = seg_gather_reduce_raw(gs, adj.src, adj.src.col, 'sum', w_t, None, None)
            gx = _unflat(gx, adj.n_src, ctx.xshape[1:])
        if w is not None and ctx.needs_input_grad[1]:
            # d/dw_e = <g[dst_e], x[src_e]>: one walk over the sorted graph, no [E, F] tensor
            gw = edge_dot_raw(gs, x, adj.dst).to(w.dtype)
        return gx, gw, gb, None, None


def gather_reduce(x, adj, w=None, reduce='sum', bias=None):
    """``scatter_(reduce, w.view(-1,1) * x[edge_index[j]], edge_index[i])`` fused
    (SURVEY A.1 step 2+6, A.9 message/update)."""
    return _GatherReduce.apply(x, w, bias, adj, reduce)


# --------------------------------------------------------------------------
# gathers for generic message() functions
# --------------------------------------------------------------------------
class _GatherRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, index):
        ctx.n = x.size(0)
        ctx.xshape = x.shape
        ctx.save_for_backward(index)
        out = gather_rows_raw(x, index)
        return out.view((index.numel(),) + tuple(x.shape[1:]))

    @staticmethod
    def backward(ctx, g):
        (index,) = ctx.saved_tensors
        csr = csr_for(index, ctx.n)
        gx = seg_gather_reduce_raw(g, csr, csr.perm, 'sum')
        return _unflat(gx, ctx.xshape[0], ctx.xshape[1:]), None


def index_select0(x, index):
    """``torch.index_select(x, 0, index)`` (A.1 step 2) with an atomic-free backward."""
    return _GatherRows.apply(x, index)


# --------------------------------------------------------------------------
# scatter of a materialised message tensor (generic path, A.3-A.6)
# --------------------------------------------------------------------------
class _SegReduce(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, index, dim_size, name, zero_below):
        csr = csr_for(index, dim_size)
        E = index.numel()
        ctx.csr, ctx.name, ctx.E = csr, name, E
        ctx.sshape = src.shape
        tail = tuple(src.shape[1:])
        if name in ('add', 'mean'):
            out = seg_gather_reduce_raw(src, csr, csr.perm, 'mean' if name == 'mean' else 'sum')
            ctx.save_for_backward(index)
        else:
            s = src if name == 'max' else -src
            out, arg = seg_max_raw(s, csr, csr.perm, csr.perm, E, zero_below, True)
            if name == 'min':
                out = -out
            ctx.save_for_backward(arg)
        return _unflat(out, dim_size, tail)

    @staticmethod
    def backward(ctx, g):
        name, csr = ctx.name, ctx.csr
        g2 = _rows2d(g)
        if name in ('add', 'mean'):
            (index,) = ctx.saved_tensors
            if name == 'mean':
                g2 = g2 / _row_counts(csr, g2.dtype)
            gs = gather_rows_raw(g2, index)
        else:
            (arg,) = ctx.saved_tensors
            gf = _f32(g2)
            if name == 'min':
                pass  # d(-max(-x))/dx routes +g to the arg position
            F = gf.size(1)
            gs = torch.zeros((ctx.E, F), dtype=torch.float32, device=g.device)
            if ctx.E > 0 and F > 0:
                nat.call('pygb_seg_max_backward', g.device, nat.ptr(gf), _ld(gf), nat.ptr(arg),
                         csr.n_rows, F, ctx.E, nat.ptr(gs), F)
            gs = gs.to(g.dtype)
        return gs.view(ctx.sshape), None, None, None, None


def segment_reduce(src, index, dim_size, name, zero_below=NEG_INF):
    return _SegReduce.apply(src, index, int(dim_size), name, zero_below)


def scatter_max(src, index, dim_size):
    """torch_scatter.scatter_max semantics (A.6): (values, argmax), first maximum
    wins, empty rows -> 0 / E.  Not differentiable through this entry point."""
    csr = csr_for(index, dim_size)
    out, arg = seg_max_raw(src, csr, csr.perm, csr.perm, index.numel())
    tail = tuple(src.shape[1:])
    return out.view((dim_size,) + tail), arg.view((dim_size,) + tail)


# --------------------------------------------------------------------------
# fused gather-max for the default message (x_j) with aggr='max'
# --------------------------------------------------------------------------
class _GatherMax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, adj, zero_below):
        csr = adj.dst
        out, arg = seg_max_raw(x, csr, csr.col, csr.perm, adj.E, zero_below, True)
        ctx.adj, ctx.xshape = adj, x.shape
        ctx.save_for_backward(arg)
        return out.view((adj.n_dst,) + tuple(x.shape[1:]))

    @staticmethod
    def backward(ctx, g):
        (arg,) = ctx.saved_tensors
        adj = ctx.adj
        gf = _f32(_rows2d(g))
        F = gf.size(1)
        ge = torch.zeros((adj.E, F), dtype=torch.float32, device=g.device)
        if adj.E > 0 and F > 0:
            nat.call('pygb_seg_max_backward', g.device, nat.ptr(gf), _ld(gf), nat.ptr(arg),
                     adj.n_dst, F, adj.E, nat.ptr(ge), F)
        csr = csr_for(adj.edge_index[adj.j], adj.n_src)
        gx = seg_gather_reduce_raw(ge, csr, csr.perm, 'sum').to(g.dtype)
        return _unflat(gx, ctx.xshape[0], ctx.xshape[1:]), None, None


def gather_max(x, adj, zero_below=-10000.0):
    return _GatherMax.apply(x, adj, zero_below)


# --------------------------------------------------------------------------
# utils.softmax (A.7)
# --------------------------------------------------------------------------
class _EdgeSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, index, num_nodes):
        dev = nat.require_cuda(src, index)
        csr = csr_for(index, num_nodes)
        s2 = _f32(_rows2d(src))
        H = s2.size(1)
        out = torch.empty_like(s2)
        if s2.numel() > 0:
            nat.call('pygb_edge_softmax', dev, nat.ptr(s2), H, nat.ptr(csr.rowptr),
                     nat.ptr(csr.perm), csr.n_rows, nat.ptr(out))
        ctx.csr = csr
        ctx.save_for_backward(out)
        ctx.sdtype = src.dtype
        return out.to(src.dtype).view(src.shape)

    @staticmethod
    def backward(ctx, g):
        (y,) = ctx.saved_tensors
        csr = ctx.csr
        g2 = _f32(_rows2d(g))
        gs = torch.empty_like(y)
        if y.numel() > 0:
            nat.call('pygb_edge_softmax_backward', g.device, nat.ptr(y), nat.ptr(g2), y.size(1),
                     nat.ptr(csr.rowptr), nat.ptr(csr.perm), csr.n_rows, nat.ptr(gs))
        return gs.to(ctx.sdtype).view(g.shape), None, None


def edge_softmax(src, index, num_nodes):
    return _EdgeSoftmax.apply(src, index, int(num_nodes))


# --------------------------------------------------------------------------
# fused GAT propagate (A.10)
# --------------------------------------------------------------------------
def _gat_scores(xw2, att2, H, C, want_dst=True):
    dev = xw2.device
    n = xw2.size(0)
    a_src = torch.empty((n, H), dtype=torch.float32, device=dev)
    a_dst = torch.empty((n, H), dtype=torch.float32, device=dev) if want_dst else None
    if n > 0:
        nat.call('pygb_gat_scores', dev, nat.dtype_code(xw2), nat.ptr(xw2), _ld(xw2),
                 nat.ptr(att2), n, H, C, nat.ptr(a_src), nat.ptr(a_dst))
    return a_src, a_dst


class _GATFused(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_src, x_dst, att, bias, adj, H, C, slope, same):
        dev = nat.require_cuda(x_src, x_dst, att, bias)
        xs = _rows2d(x_src)
        att2 = _f32(att.reshape(H, 2 * C))
        a_src, a_dst = _gat_scores(xs, att2, H, C, want_dst=same)
        if not same and x_dst is not None:
            _, a_dst = _gat_scores(_rows2d(x_dst), att2, H, C, want_dst=True)
        n_dst = adj.n_dst
        F = H * C
        out = torch.empty((n_dst, F), dtype=xs.dtype, device=dev)
        row_max = torch.empty((n_dst, H), dtype=torch.float32, device=dev)
        row_sum = torch.empty((n_dst, H), dtype=torch.float32, device=dev)
        bias2 = _f32(bias)
        vec = 16 // xs.element_size()
        if n_dst > 0 and STREAM and C % vec == 0 and F // vec <= 128 and \
                xs.data_ptr() % 16 == 0 and _ld(xs) % vec == 0 and adj.E > 0:
            csr = adj.dst
            Ts, n_sub, nz_rows, sub_info = csr.items(STREAM_T)
            idxf = csr.flagged('col')
            ws_bytes = nat.workspace('pygb_gat_stream_workspace', n_sub, H, C)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            nat.call('pygb_gat_forward_stream', dev, nat.dtype_code(xs), nat.ptr(xs), _ld(xs),
                     nat.ptr(a_src), nat.ptr(a_dst), H, C, float(slope), nat.ptr(csr.rowptr),
                     nat.ptr(idxf), n_dst, csr.E, nat.ptr(nz_rows), nat.ptr(sub_info), Ts, n_sub,
                     nat.ptr(bias2), nat.ptr(out), F, nat.ptr(row_max), nat.ptr(row_sum),
                     nat.ptr(ws), ws_bytes)
        elif n_dst > 0:
            nat.call('pygb_gat_forward', dev, nat.dtype_code(xs), nat.ptr(xs), _ld(xs),
                     nat.ptr(a_src), nat.ptr(a_dst), H, C, float(slope), nat.ptr(adj.dst.rowptr),
                     nat.ptr(adj.dst.col), n_dst, nat.ptr(bias2), nat.ptr(out), F,
                     nat.ptr(row_max), nat.ptr(row_sum))
        ctx.adj, ctx.H, ctx.C, ctx.slope, ctx.same = adj, H, C, slope, same
        ctx.att_shape = att.shape
        ctx.save_for_backward(x_src, None if same else x_dst, att2, bias2, a_src, a_dst, out,
                              row_max, row_sum)
        return out

    @staticmethod
    def backward(ctx, g):
        x_src, x_dst, att2, bias2, a_src, a_dst, out, row_max, row_sum = ctx.saved_tensors
        adj, H, C, slope = ctx.adj, ctx.H, ctx.C, ctx.slope
        dev = g.device
        F = H * C
        gf = _f32(_rows2d(g))
        xs = _f32(_rows2d(x_src))
        outf = _f32(out)
        E = adj.E
        dst, src = adj.dst, adj.src
        lh = C // 4
        d_a_dst = None
        if (STREAM and adj.n_dst > 0 and E > 0 and C % 4 == 0 and lh & (lh - 1) == 0 and lh <= 32
                and F // 4 <= 128 and _ld(xs) % 4 == 0 and _ld(gf) % 4 == 0
                and xs.data_ptr() % 16 == 0 and gf.data_ptr() % 16 == 0):
            # streaming backward: alpha / ds are recomputed from per-node records in both
            # walks, no [E, H] tensor exists (csrc/gat_bwd_stream.cu)
            rec = torch.empty((adj.n_dst, H, 4), dtype=torch.float32, device=dev)
            nat.call('pygb_gat_backward_prep', dev, nat.ptr(a_dst), nat.ptr(row_max),
                     nat.ptr(row_sum), nat.ptr(outf), _ld(outf), nat.ptr(bias2), nat.ptr(gf),
                     _ld(gf), adj.n_dst, H, C, nat.ptr(rec))

            def walk(mode, csr, row_feat, gat_feat, d_feat, d_a):
                Ts, n_sub, nz_rows, sub_info = csr.items(STREAM_T)
                ws_bytes = nat.workspace('pygb_gat_backward_stream_workspace', mode, n_sub, H, C)
                ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
                nat.call('pygb_gat_backward_stream', dev, mode, nat.ptr(row_feat), _ld(row_feat),
                         nat.ptr(gat_feat), _ld(gat_feat), nat.ptr(a_src), nat.ptr(rec), H, C,
                         float(slope), nat.ptr(csr.rowptr), nat.ptr(csr.flagged('col')),
                         csr.n_rows, csr.E, nat.ptr(nz_rows), nat.ptr(sub_info), Ts, n_sub,
                         nat.ptr(d_feat), F, nat.ptr(d_a), nat.ptr(ws), ws_bytes)

            if a_dst is not None:
                d_a_dst = torch.empty((adj.n_dst, H), dtype=torch.float32, device=dev)
                walk(0, dst, gf, xs, None, d_a_dst)
            d_xs = torch.empty((adj.n_src, F), dtype=torch.float32, device=dev)
            d_a_src = torch.empty((adj.n_src, H), dtype=torch.float32, device=dev)
            walk(1, src, xs, gf, d_xs, d_a_src)
        else:
            # alpha and ds per (edge, head) in the slot order of the destination-sorted graph:
            # coalesced writes, and the sums over a destination row read consecutive slots
            alpha = torch.empty((E, H), dtype=torch.float32, device=dev)
            ds = torch.empty((E, H), dtype=torch.float32, device=dev)
            if adj.n_dst > 0 and E > 0:
                nat.call('pygb_gat_edge_grad', dev, nat.ptr(xs), _ld(xs), nat.ptr(a_src),
                         nat.ptr(a_dst), H, C, float(slope), nat.ptr(dst.rowptr), nat.ptr(dst.col),
                         None, adj.n_dst, nat.ptr(outf), F, nat.ptr(bias2), nat.ptr(gf),
                         _ld(gf), nat.ptr(row_max), nat.ptr(row_sum), nat.ptr(alpha), nat.ptr(ds))
            s2d = adj.src_to_dst_slot                 # slot of the source-sorted graph -> dst slot
            # d_xw[j] = sum_p alpha_p * g[dst_p]   (transposed walk, per-head weights)
            d_xs = seg_gather_reduce_raw(gf, src, src.col, 'sum', alpha, s2d, None, (H, C))
            d_a_src = seg_gather_reduce_raw(ds, src, s2d, 'sum', idx_key='s2d')   # [n_src, H]
            if a_dst is not None:
                d_a_dst = seg_gather_reduce_raw(ds, dst, adj.iota, 'sum', idx_key='iota')
        att_d, att_s = att2[:, :C], att2[:, C:]
        d_xs = d_xs.view(-1, H, C) + d_a_src.unsqueeze(-1) * att_s.unsqueeze(0)
        d_att_s = torch.einsum('nh,nhc->hc', d_a_src, xs.view(-1, H, C))
        d_xd = None
        d_att_d = torch.zeros_like(d_att_s)
        if a_dst is not None:
            xd = xs if ctx.same else _f32(_rows2d(x_dst))
            d_att_d = torch.einsum('nh,nhc->hc', d_a_dst, xd.view(-1, H, C))
            dd = d_a_dst.unsqueeze(-1) * att_d.unsqueeze(0)
            if ctx.same:
                d_xs = d_xs + dd
            else:
                d_xd = dd.reshape(-1, F).to(x_dst.dtype)
        d_att = torch.cat([d_att_d, d_att_s], dim=1).view(ctx.att_shape)
        d_bias = gf.sum(0) if bias2 is not None else None
        d_xs = d_xs.reshape(-1, F).to(x_src.dtype)
        if ctx.same:
            # x was passed twice (as x_src and x_dst): give all of it to x_src
            return d_xs, None, d_att, d_bias, None, None, None, None, None
        return d_xs, d_xd, d_att, d_bias, None, None, None, None, None


def gat_pad_channels(C, itemsize=4):
    """Channels per head the streaming GAT kernels want: a power of two in [4, 128] (forward and
    both backward walks stream), else the next multiple of the 16-byte vector (forward streams,
    backward uses the two-pass kernels).  Zero padding is exact: padded channels add 0 to every
    logit and produce 0 outputs."""
    vec = 16 // itemsize
    if C <= 128:
        cp = max(vec, 4)
        while cp < C:
            cp *= 2
        return cp
    return (C + vec - 1) // vec * vec


def gat_propagate(x_src, x_dst, att, bias, adj, H, C, slope):
    """Fused GATConv message+softmax+aggregate (A.10).  ``x_dst`` is ``x_src``
    itself (ordinary graphs), another tensor (bipartite) or None (x_i = None).

    Heads are independent, so wide layers (H*C beyond the 2 KB a lane group stages per row,
    e.g. examples/ppi.py:22 GATConv(.., 256, heads=4)) run as one launch per group of heads on
    column views of the same matrices — the streaming kernels instead of the two-pass ones."""
    same = x_dst is x_src
    per = max(1, 512 // max(C, 1)) if x_src.dtype == torch.float32 else max(1, 1024 // max(C, 1))
    if H <= per or C * x_src.element_size() > 2048 or x_src.dim() != 2 or not STREAM:
        return _GATFused.apply(x_src, None if same else x_dst, att, bias, adj, H, C, slope, same)
    outs = []
    att3 = att.reshape(1, H, 2 * C)
    for h0 in range(0, H, per):
        h1 = min(h0 + per, H)
        xs = x_src[:, h0 * C:h1 * C]
        xd = xs if same else (None if x_dst is None else x_dst[:, h0 * C:h1 * C])
        b = None if bias is None else bias[h0 * C:h1 * C]
        outs.append(_GATFused.apply(xs, None if same else xd, att3[:, h0:h1], b, adj, h1 - h0, C,
                                    slope, same))
    return torch.cat(outs, dim=1)


# --------------------------------------------------------------------------
# K5: dense transform X @ W (+ bias) on the tensor cores
# --------------------------------------------------------------------------
# '3' = 3xTF32 split (fp32-faithful, default), '1' = plain TF32, '0' = library GEMM.
GEMM_PASSES = int(os.environ.get('PYGB_GEMM', '3'))
# PYGB_STRICT_GEMM=1: `linear` raises instead of handing anything to the library GEMM
# (tests use it to prove that a configuration runs on the tcgen05 kernels).
STRICT_GEMM = os.environ.get('PYGB_STRICT_GEMM', '0') not in ('', '0')


def _pad4(n, itemsize):
    vec = 16 // itemsize
    return (n + vec - 1) // vec * vec if n > 1 else n


def _gemm_eligible(x, w):
    """fp32 (tf32 kernel) or bf16 (f16-kind kernel) CUDA operands of one dtype with 16-byte
    aligned rows; an fp32 ``x`` may also be slice-major: ``[S, M, Kc]`` planes of ``Kc``
    (multiple of 32) columns, standing for ``[M, S*Kc]`` (internal callers only)."""
    vec = 16 // x.element_size()
    ok = (GEMM_PASSES in (1, 3) and x.is_cuda and w.dim() == 2 and x.dtype == w.dtype
          and x.dtype in (torch.float32, torch.bfloat16) and x.size(-2) > 0
          and x.size(-1) > 0 and w.size(1) > 0 and x.stride(-1) == 1 and x.stride(-2) % vec == 0
          and x.stride(-2) >= x.size(-1) and x.data_ptr() % 16 == 0 and w.stride(1) == 1)
    if x.dim() == 3:
        return (ok and x.dtype == torch.float32 and x.size(2) % 32 == 0 and x.stride(0) % 4 == 0
                and x.stride(0) >= x.stride(1) * x.size(1))
    return ok and x.dim() == 2


_pad_cache = collections.OrderedDict()      # identity of a constant input -> its padded copy


def _aligned_rows(x, cache=True):
    """``x`` [M, K] as a tensor whose rows start on 16-byte boundaries (what TMA needs):
    ``x`` itself, or the ``[:, :K]`` view of a zero-padded copy ``[M, Kp]``.  K = 602 or 1433
    fp32 rows (BASELINE cfg3 / cfg1) are not 16-byte multiples.  For an input that does not
    require grad (dataset features) the copy is made once and memoised on the tensor's
    identity, like the sorted graph; otherwise it is a differentiable ``F.pad``."""
    vec = 16 // x.element_size()
    if x.stride(1) == 1 and x.stride(0) % vec == 0 and x.stride(0) >= x.size(1) \
            and x.data_ptr() % 16 == 0:
        return x
    K = x.size(1)
    Kp = (K + vec - 1) // vec * vec
    if x.requires_grad and torch.is_grad_enabled():
        return torch.nn.functional.pad(x, (0, Kp - K))[:, :K]
    key = (x.data_ptr(), tuple(x.shape), tuple(x.stride()), x._version, x.device.index, x.dtype)
    if cache:
        hit = _pad_cache.get(key)
        if hit is not None:
            _pad_cache.move_to_end(key)
            return hit[0]
    xp = x.new_zeros((x.size(0), Kp))
    xp[:, :K].copy_(x)
    out = xp[:, :K]
    if cache:
        _pad_cache[key] = (out, x)           # keeps x alive: the pointer cannot be recycled
        while len(_pad_cache) > 4:
            _pad_cache.popitem(last=False)
    return out


def _gemm_a_args(x, rows=None):
    """(ptr, lda, a_slices, a_plane, M, K) of a row-major or slice-major A operand."""
    if x.dim() == 3:
        M = x.size(1) if rows is None else int(rows)
        return nat.ptr(x), x.stride(1), x.size(0), x.stride(0), M, x.size(0) * x.size(2)
    M = x.size(0) if rows is None else int(rows)
    return nat.ptr(x), x.stride(0), 1, 0, M, x.size(1)


def gemm_xw_raw(x, w, bias=None, passes=None, out=None, relu=False, rows=None):
    """``x @ w (+ bias)`` by the tcgen05 kernel; the result's rows are padded to 16 bytes
    (a view ``[:, :N]`` is returned) so that a following aggregation stays vectorised.
    ``out``: optional ``[M, >=N]`` destination of x's dtype with unit inner stride.
    ``x``: ``[M, K]`` fp32 / bf16, or fp32 slice-major ``[S, >=M, K/S]`` (``rows`` = M)."""
    dev = nat.require_cuda(x, w, bias)
    a_ptr, lda, a_slices, a_plane, M, K = _gemm_a_args(x, rows)
    N = w.size(1)
    assert w.size(0) == K and w.dtype == x.dtype
    esz = x.element_size()
    if out is None:
        ldc = _pad4(N, esz)
        out_full = torch.empty((M, ldc), dtype=x.dtype, device=dev)
    else:
        assert out.dtype == x.dtype and out.size(0) == M and out.size(1) >= N
        assert out.stride(1) == 1
        out_full, ldc = out, (out.stride(0) if M > 1 else max(out.size(1), 1))
    ws_bytes = nat.workspace('pygb_gemm_weight_workspace', K, N)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    b = _f32(bias)
    if x.dtype == torch.bfloat16:
        assert x.dim() == 2
        nat.call('pygb_gemm_xw_bf16', dev, a_ptr, lda, M, K, nat.ptr(w), w.stride(0), N,
                 nat.ptr(b), int(bool(relu)), nat.ptr(out_full), ldc, nat.ptr(ws), ws_bytes)
    else:
        nat.call('pygb_gemm_xw', dev, a_ptr, lda, a_slices, a_plane, M, K, nat.ptr(w),
                 w.stride(0), N, nat.ptr(b), int(bool(relu)), nat.ptr(out_full), ldc,
                 int(passes or GEMM_PASSES), nat.ptr(ws), ws_bytes)
    return out_full[:, :N] if out_full.size(1) != N else out_full


def _xtg_eligible(x, g):
    return (GEMM_PASSES in (1, 3) and x.is_cuda and x.dim() == 2 and g.dim() == 2
            and x.dtype == torch.float32 and g.dtype == torch.float32
            and x.size(0) == g.size(0) and x.size(0) >= 1024
            and 0 < x.size(1) and 0 < g.size(1)
            and x.stride(1) == 1 and g.stride(1) == 1 and x.stride(0) % 4 == 0
            and g.stride(0) % 4 == 0 and x.stride(0) >= x.size(1) and g.stride(0) >= g.size(1)
            and x.data_ptr() % 16 == 0 and g.data_ptr() % 16 == 0)


def gemm_xtg_raw(x, g, passes=None):
    """``x.t() @ g`` (the weight gradient) by the split-K tcgen05 kernel: both operands are
    read row-major as they are (MN-major tensor-core operands).  The kernel keeps a
    ``[<=256, <=256]`` partial in TMEM; wider outputs are tiled over column blocks of x / g
    (views with the same row stride — nothing is copied)."""
    dev = nat.require_cuda(x, g)
    R, Kx, Ng = x.size(0), x.size(1), g.size(1)
    out = torch.empty((Kx, Ng), dtype=torch.float32, device=dev)
    for k0 in range(0, Kx, 256):
        k1 = min(k0 + 256, Kx)
        for n0 in range(0, Ng, 256):
            n1 = min(n0 + 256, Ng)
            xs, gs, o = x[:, k0:k1], g[:, n0:n1], out[k0:k1, n0:n1]
            ws_bytes = nat.workspace('pygb_gemm_xtg_workspace', R, k1 - k0, n1 - n0)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            nat.call('pygb_gemm_xtg', dev, nat.ptr(xs), x.stride(0), nat.ptr(gs), g.stride(0), R,
                     k1 - k0, n1 - n0, nat.ptr(o), Ng, int(passes or GEMM_PASSES), nat.ptr(ws),
                     ws_bytes)
    return out


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, bias):
        ctx.save_for_backward(x, w)
        ctx.has_bias = bias is not None
        return gemm_xw_raw(x, w, bias)

    @staticmethod
    def backward(ctx, g):
        x, w = ctx.saved_tensors
        gx = gw = gb = None
        ga = None
        if g.dim() == 2 and g.dtype == x.dtype:
            ga = _aligned_rows(g if g.stride(-1) == 1 else g.contiguous(), cache=False)
        if ctx.needs_input_grad[0]:
            wt = w.t().contiguous()
            if ga is not None and _gemm_eligible(ga, wt):    # dX = g W^T on the same kernel
                gx = gemm_xw_raw(ga, wt)
                gx = gx if gx.is_contiguous() else gx.contiguous()
            else:
                gx = torch.matmul(g, w.t())
        if ctx.needs_input_grad[1]:
            if ga is not None and _xtg_eligible(x, ga):      # dW = x^T g: split-K on tcgen05
                gw = gemm_xtg_raw(x, ga)
            else:
                gw = torch.matmul(x.t(), g)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            gb = g.sum(0)
        return gx, gw, gb


def linear(x, w, bias=None):
    """``x @ w (+ bias)`` (the X·W of GCNConv / GATConv).  2-D fp32 / bf16 CUDA inputs run on
    the tcgen05 kernels (rows that are not 16-byte multiples through one padded copy, see
    ``_aligned_rows``); the output row is padded to 16 bytes.  Anything else (other dtypes,
    ``[N, T, F]`` inputs) is ``torch.matmul`` — or an error under ``PYGB_STRICT_GEMM=1``."""
    if (x.dim() == 2 and w.dim() == 2 and x.is_cuda and x.dtype == w.dtype
            and x.dtype in (torch.float32, torch.bfloat16) and x.numel() > 0 and w.numel() > 0
            and GEMM_PASSES in (1, 3)):
        xa = _aligned_rows(x)
        wa = w if w.stride(1) == 1 else w.contiguous()
        if _gemm_eligible(xa, wa):
            return _Linear.apply(xa, wa, bias)
    if STRICT_GEMM:
        raise RuntimeError('ops.linear: {} x {} ({}, {}) would fall back to the library GEMM'.format(
            tuple(x.shape), tuple(w.shape), x.dtype, x.device))
    pad = 0
    if x.dim() == 2 and x.is_floating_point():
        pad = _pad4(w.size(1), x.element_size()) - w.size(1)
    if pad:
        out = torch.matmul(x, torch.nn.functional.pad(w, (0, pad)))[:, :w.size(1)]
    else:
        out = torch.matmul(x, w)
    return out + bias if bias is not None else out
