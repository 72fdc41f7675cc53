"""GPU: ``pygb_seg_gather_reduce_stream_static`` — the streaming gather-reduce with its fix-up
scratch in a file-scope ``__device__`` array (no workspace argument, no device allocation;
include/pygb.h).  No counterpart in the reference.

Called through the C ABI and compared with ``pygb_seg_gather_reduce_stream`` (caller-provided
workspace) — same kernels, so the comparison is bit-exact — on calls in a row that share the
array: a large problem, then a smaller one that finds the first one's leftovers, then the first
again.  Named ``test_zz_*`` so that pytest collects it last."""
import pytest
import torch

from util import skewed_graph

pytestmark = pytest.mark.gpu

STATIC_WS_BYTES = 16 << 20        # PYGB_SEG_STREAM_STATIC_WS_BYTES (include/pygb.h)


def _case(dev, n, E, F, dtype, reduce, weighted, with_bias, seed):
    """Every tensor either entry point touches, allocated up front."""
    from torch_geometric import _native as nat
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    g = torch.Generator().manual_seed(seed)
    ei = skewed_graph(n, E, seed)
    csr = Adj(ei.to(dev), 1, (n, n)).dst
    vec = 16 // torch.empty((), dtype=dtype).element_size()
    ld = (F + vec - 1) // vec * vec
    x = torch.zeros((n, ld), dtype=dtype, device=dev)
    x[:, :F] = torch.randn(n, F, generator=g).to(dtype).to(dev)
    w = (torch.rand(E, generator=g) + 0.5).to(dev) if weighted else None
    bias = torch.randn(F, generator=g).to(dev) if with_bias else None
    Ts, n_sub, nz_rows, sub_info = csr.items(ops.STREAM_T)
    idxf = csr.flagged('col')
    dt = nat.dtype_code(x)
    ws_bytes = nat.workspace('pygb_seg_stream_workspace', dt, n_sub, F)
    assert ws_bytes <= STATIC_WS_BYTES, 'test case does not fit the built-in array'
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    head = (nat.MEAN if reduce == 'mean' else nat.SUM, dt, nat.ptr(x), ld, F, nat.ptr(csr.rowptr),
            nat.ptr(idxf), nat.ptr(w), nat.ptr(bias), csr.n_rows, csr.E, nat.ptr(nz_rows),
            nat.ptr(sub_info), Ts, n_sub)
    keep = (x, w, bias, csr, idxf, nz_rows, sub_info, ws)

    def with_workspace():
        out = torch.full((n, ld), float('nan'), dtype=dtype, device=dev)
        nat.call('pygb_seg_gather_reduce_stream', dev, *head, nat.ptr(out), ld, nat.ptr(ws), ws_bytes)
        return out[:, :F]

    def static(out):
        nat.call('pygb_seg_gather_reduce_stream_static', dev, *head, nat.ptr(out), ld)
        return out[:, :F]

    return with_workspace, static, (n, ld), keep


def test_static_scratch_entry_point_matches_workspace_entry_point(dev):
    big = _case(dev, 6000, 250000, 100, torch.float32, 'sum', True, True, 11)
    small = _case(dev, 900, 20000, 47, torch.float32, 'mean', False, False, 12)
    half = _case(dev, 3000, 90000, 64, torch.bfloat16, 'sum', True, False, 13)
    want = [c[0]() for c in (big, small, half)]
    outs = [torch.full(c[2], float('nan'), dtype=w.dtype, device=dev)
            for c, w in zip((big, small, half, big), want + [want[0]])]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    # four calls in a row on one stream, sharing the array, nothing allocated in between
    got = [c[1](o) for c, o in zip((big, small, half, big), outs)]
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == before
    for k, (a, b) in enumerate(zip(got, want + [want[0]])):
        assert not torch.isnan(a.float()).any(), 'call {}: rows left unwritten'.format(k)
        assert torch.equal(a, b), 'call {} differs from the workspace entry point'.format(k)


def test_static_scratch_entry_point_rejects_oversized_problems(dev):
    from torch_geometric import _native as nat
    # 4 M slots in sub-items of 32 at F = 256: 268 MB of partial rows
    E, Ts, F = 1 << 22, 32, 256
    n_sub = E // Ts
    assert nat.workspace('pygb_seg_stream_workspace', nat.F32, n_sub, F) > STATIC_WS_BYTES
    dummy = torch.zeros(64, dtype=torch.int32, device=dev)
    p = nat.ptr(dummy)
    with pytest.raises(nat.NativeError, match='built-in array'):
        nat.call('pygb_seg_gather_reduce_stream_static', dev, nat.SUM, nat.F32, p, F, F, p, p,
                 None, None, 8, E, p, p, Ts, n_sub, p, F)
