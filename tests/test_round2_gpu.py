"""GPU: round-2 additions — the SDDMM edge-weight gradient (pygb_edge_dot), learnable edge
weights through ``GCNConv.norm`` and the self-loop utilities, re-entrant conv
modules (no per-call state on ``self``), index validation."""
import threading

import pytest
import torch

from oracle import ref_cpu as O
from util import assert_close, random_graph, skewed_graph

pytestmark = pytest.mark.gpu


def _feat(n, F, seed, dtype=torch.float32):
    return torch.randn(n, F, generator=torch.Generator().manual_seed(seed)).to(dtype)


def _sync(conv, ref):
    ref.load_state_dict({k: v.detach().cpu().float() for k, v in conv.state_dict().items()})


@pytest.mark.parametrize('F', [1, 3, 4, 7, 16, 47, 64, 100, 256, 300])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_edge_dot_matches_gathered_product(dev, F, dtype):
    """out[e] = <g[dst_e], x[src_e]> against the [E, F] gathers it replaces (fp64 on the CPU),
    on a graph with hub rows and empty rows; every edge written exactly once."""
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    ns, nd, E = 500, 400, 9000
    ei = skewed_graph(max(ns, nd), E, F + 3)
    ei[0] %= ns
    ei[1] %= nd
    g, x = _feat(nd, F, 1, dtype), _feat(ns, F, 2, dtype)
    adj = Adj(ei.to(dev), 1, (ns, nd))
    out = torch.full((E,), float('nan'), device=dev)
    got = ops.edge_dot_raw(g.to(dev), x.to(dev), adj.dst)
    assert got.shape == out.shape and got.dtype == torch.float32
    ref = (g.double()[ei[1]] * x.double()[ei[0]]).sum(1)
    # inputs are exact in both dtypes (bf16 values are fp32 values); accumulation is fp32
    assert_close(got, ref, 1e-5, 'edge_dot F={}'.format(F), floor=2e-6)
    assert torch.equal(got, ops.edge_dot_raw(g.to(dev), x.to(dev), adj.dst))


def test_edge_dot_empty_and_single_row(dev):
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    ei = torch.zeros((2, 0), dtype=torch.long, device=dev)
    adj = Adj(ei, 1, (5, 4))
    assert ops.edge_dot_raw(torch.randn(4, 8, device=dev), torch.randn(5, 8, device=dev),
                            adj.dst).shape == (0,)
    ei = torch.stack([torch.arange(700) % 5, torch.full((700,), 2)]).to(dev)   # one hub row
    adj = Adj(ei, 1, (5, 4))
    g, x = torch.randn(4, 12, device=dev), torch.randn(5, 12, device=dev)
    got = ops.edge_dot_raw(g, x, adj.dst)
    assert_close(got, (g[ei[1]] * x[ei[0]]).sum(1), 1e-5)


@pytest.mark.parametrize('improved', [False, True])
def test_gcn_norm_learnable_edge_weight(dev, improved):
    """with ``edge_weight.requires_grad`` the normalisation must be
    differentiable like the reference's (A.9) — values, d/dw through norm AND through the
    aggregation, on a graph with existing and duplicated self-loops (Q1)."""
    from torch_geometric.nn import GCNConv
    n, E = 200, 2500
    ei = random_graph(n, n, E, 5)
    ei[1, :40] = ei[0, :40]                       # existing loops
    ei[:, 40:50] = ei[:, 30:40]                   # duplicated loops: the later one wins
    x = _feat(n, 24, 1)
    w = torch.rand(E, generator=torch.Generator().manual_seed(2)) + 0.25
    conv = GCNConv(24, 10, improved=improved).to(dev)
    ref = O.GCNConv(24, 10, improved=improved)
    _sync(conv, ref)
    wa = w.clone().requires_grad_(True)
    wb = w.to(dev).requires_grad_(True)
    xa, xb = x.clone().requires_grad_(True), x.to(dev).requires_grad_(True)
    ra, rb = ref(xa, ei, wa), conv(xb, ei.to(dev), wb)
    assert_close(rb, ra, 1e-5, 'forward')
    gout = _feat(n, 10, 3)
    ra.backward(gout)
    rb.backward(gout.to(dev))
    assert wb.grad is not None, 'edge_weight gradient silently dropped'
    assert_close(wb.grad, wa.grad, 1e-4, 'd edge_weight')
    assert_close(xb.grad, xa.grad, 1e-4, 'dx')
    assert_close(conv.weight.grad, ref.weight.grad, 1e-4, 'dW')
    # the no-grad path (fused kernels) computes the same norm
    ei1, n1 = GCNConv.norm(ei.to(dev), n, w.to(dev), improved)
    ei2, n2 = GCNConv.norm(ei.to(dev), n, wb, improved)
    assert torch.equal(ei1, ei2)
    assert_close(n2, n1, 1e-6, 'norm: autograd path vs fused path')


def test_loop_utils_keep_the_weight_graph(dev):
    from torch_geometric import utils as U
    ei = torch.tensor([[0, 1, 1, 2, 2, 3, 1], [1, 1, 2, 2, 0, 3, 1]])
    w = torch.tensor([.1, .2, .3, .4, .5, .6, .7])
    for fn, ofn, kw in ((U.remove_self_loops, O.remove_self_loops, {}),
                        (U.add_self_loops, O.add_self_loops, {'num_nodes': 5}),
                        (U.add_remaining_self_loops, O.add_remaining_self_loops,
                         {'fill_value': 2, 'num_nodes': 5})):
        wa = w.clone().requires_grad_(True)
        wb = w.to(dev).requires_grad_(True)
        ea, oa = ofn(ei, wa, **kw)
        eb, ob = fn(ei.to(dev), wb, **kw)
        assert torch.equal(eb.cpu(), ea) and torch.equal(ob.detach().cpu(), oa.detach())
        coef = torch.arange(1, oa.numel() + 1, dtype=torch.float32)
        (oa * coef).sum().backward()
        (ob * coef.to(dev)).sum().backward()
        assert torch.equal(wb.grad.cpu(), wa.grad), fn.__name__


def test_conv_modules_are_reentrant(dev):
    """no ``_bias_fused`` / ``_defer_bias`` state on the module — an
    exception inside one call must not change the next, and concurrent calls from worker
    threads (the reference runs convs under ``parallel_apply``) must agree with serial ones."""
    from torch_geometric.nn import GCNConv, GATConv
    n, E = 400, 5000
    ei = skewed_graph(n, E, 9).to(dev)
    x = _feat(n, 16, 1).to(dev)
    gcn = GCNConv(16, 32).to(dev)          # aggregate-first route (in < out)
    gcn2 = GCNConv(16, 8).to(dev)          # transform-first route, bias in the kernel epilogue
    gat = GATConv(16, 8, heads=4).to(dev)
    with torch.no_grad():
        for m in (gcn, gcn2, gat):
            m.bias.uniform_(-1, 1)
        want = [m(x, ei) for m in (gcn, gcn2, gat)]
        for m in (gcn, gcn2, gat):         # a failing call in between leaves nothing behind
            with pytest.raises(Exception):
                m(x, ei[:, :0].float())
            assert '_bias_fused' not in m.__dict__ and '_defer_bias' not in m.__dict__
        got = [m(x, ei) for m in (gcn, gcn2, gat)]
        for a, b in zip(got, want):
            assert torch.equal(a, b)

        # generic route right after a fused call: the bias must still be added once
        class Plain(GCNConv):
            def message(self, x_j, norm):
                return norm.view(-1, 1) * x_j
        p = Plain(16, 8).to(dev)
        p.load_state_dict(gcn2.state_dict())
        assert_close(p(x, ei), want[1], 1e-5, 'generic route after fused route')

        results = {}

        def work(k):
            torch.cuda.set_device(dev)
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                results[k] = [m(x, ei) for m in (gcn, gcn2, gat)]
            s.synchronize()

        ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        for k in range(4):
            for a, b in zip(results[k], want):
                assert torch.equal(a, b), 'thread {} differs'.format(k)


def test_gcn_norm_rejects_out_of_range_indices(dev):
    """ids >= num_nodes or < 0 must raise instead of reading out of bounds."""
    from torch_geometric.nn import GCNConv
    ei = torch.tensor([[0, 1, 7], [1, 0, 2]], device=dev)
    with pytest.raises(IndexError):
        GCNConv.norm(ei, 5)
    with pytest.raises(IndexError):
        GCNConv.norm(torch.tensor([[0, -1], [1, 0]], device=dev), 5)
    with pytest.raises(IndexError):
        GCNConv.norm(ei, 5, torch.ones(3, device=dev))


def test_linear_3d_input_falls_back_to_matmul_semantics(dev):
    """an ordinary [N, T, F] input is ``torch.matmul(x, w)``, not the
    internal slice-major layout."""
    from torch_geometric import ops
    x = torch.randn(6, 5, 32, device=dev)
    w = torch.randn(32, 8, device=dev)
    assert_close(ops.linear(x, w), torch.matmul(x.double().cpu(), w.double().cpu()), 1e-5)


@pytest.mark.parametrize('H,C,concat,bwd_stream', [
    (1, 7, True, True), (1, 100, True, True), (1, 33, False, True), (6, 121, False, True),
    (4, 256, True, False), (1, 300, True, False), (2, 130, True, False)])
def test_gat_fork_shapes_run_on_streaming_kernels(dev, H, C, concat, bwd_stream):
    """the fork builds ``GATConv(w_i, w_{i+1})`` with ``heads=1`` and
    arbitrary widths (ConvexPruning.py:204-214); examples/ppi.py:22 uses 4 x 256.  Heads are
    zero-padded to a streamable width (in the weight) and wide layers run per head group — the
    native event log proves which entry points ran; values against the oracle."""
    from torch_geometric import _native as nat
    from torch_geometric.nn import GATConv
    n, E, fin = 700, 9000, 24
    ei = skewed_graph(n, E, H * C)
    x = _feat(n, fin, 1)
    conv = GATConv(fin, C, heads=H, concat=concat).to(dev)
    ref = O.GATConv(fin, C, heads=H, concat=concat)
    with torch.no_grad():
        conv.bias.uniform_(-1, 1)
    _sync(conv, ref)
    a = x.clone().requires_grad_(True)
    b = x.to(dev).requires_grad_(True)
    nat.event_log = []
    try:
        rb = conv(b, ei.to(dev))
        fwd_names = [e[0] for e in nat.event_log]
        nat.event_log = []
        g = torch.randn(rb.shape, generator=torch.Generator().manual_seed(3))
        rb.backward(g.to(dev))
        bwd_names = [e[0] for e in nat.event_log]
    finally:
        nat.event_log = None
    assert 'pygb_gat_forward_stream' in fwd_names and 'pygb_gat_forward' not in fwd_names, fwd_names
    if bwd_stream:
        assert 'pygb_gat_backward_stream' in bwd_names and 'pygb_gat_edge_grad' not in bwd_names
    ra = ref(a, ei)
    assert_close(rb, ra, 2e-5, 'fwd', floor=5e-6)
    ra.backward(g)
    assert_close(b.grad, a.grad, 2e-4, 'dx')
    assert_close(conv.weight.grad, ref.weight.grad, 2e-4, 'dW')
    assert_close(conv.att.grad, ref.att.grad, 2e-4, 'datt')
    assert_close(conv.bias.grad, ref.bias.grad, 2e-4, 'dbias')


def test_k1_hot_row_hints_are_bit_identical(dev):
    """pygb_seg_mark_hot sets bit 30 exactly on the slots of the most frequently gathered source
    rows and leaves the row-end flags alone; pygb_seg_gather_reduce_stream_hinted (L2 evict_last
    on those gathers) returns the bits of pygb_seg_gather_reduce_stream."""
    from torch_geometric import _native as nat
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    n, E, F = 3000, 60000, 64
    ei = skewed_graph(n, E, 21)
    ei = torch.stack([ei[1], ei[0]])            # hub SOURCES: a few rows are gathered thousands of times
    csr = Adj(ei.to(dev), 1, (n, n)).dst
    base = csr.flagged('col')
    hot = csr.flagged_hot(40)
    assert hot is not None and hot is csr.flagged_hot(40)
    col = csr.col.long()
    freq = torch.bincount(col, minlength=n)
    thr = torch.topk(freq, 40).values[-1]
    want = freq[col] >= thr
    assert torch.equal(((hot >> 30) & 1).bool(), want) and 0 < int(want.sum()) < E
    assert torch.equal(hot & ~(1 << 30), base)
    assert csr.flagged_hot(n + 5) is None       # everything "hot": no hint
    x = _feat(n, F, 3).to(dev)
    w = torch.rand(E, device=dev) + 0.5
    Ts, n_sub, nz_rows, sub_info = csr.items(ops.STREAM_T)
    ws_bytes = nat.workspace('pygb_seg_stream_workspace', nat.F32, n_sub, F)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    outs = []
    for entry, idxf in (('pygb_seg_gather_reduce_stream', base),
                        ('pygb_seg_gather_reduce_stream_hinted', hot)):
        out = torch.full((n, F), float('nan'), device=dev)
        nat.call(entry, dev, nat.SUM, nat.F32, nat.ptr(x), F, F, nat.ptr(csr.rowptr), nat.ptr(idxf),
                 nat.ptr(w), None, n, E, nat.ptr(nz_rows), nat.ptr(sub_info), Ts, n_sub, nat.ptr(out),
                 F, nat.ptr(ws), ws_bytes)
        outs.append(out)
    assert torch.equal(outs[0], outs[1]) and not torch.isnan(outs[0]).any()
