"""GPU, NCCL: the two multi-GPU layouts of the GCN forward against the oracle on
the whole graph — the destination partition (native kernels + halo all-gather)
and the feature-sliced layout (dist/sliced.py) with both transports (all-to-all
and plain stores into CUDA-IPC peer memory).  Two-rank tests are skipped when
fewer than two GPUs are visible; the routed kernels and a one-rank PeerTransport
are also checked on a single GPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _join_all(procs, q=None, limit=150):
    """Join with one shared deadline; a rank that hangs (e.g. in a collective its peer never
    reached because it failed) is killed instead of being left on the GPU."""
    import time
    deadline = time.time() + limit
    for p in procs:
        p.join(max(deadline - time.time(), 1))
    hung = [p for p in procs if p.exitcode is None]
    for p in hung:
        p.kill()
        p.join(10)
    codes = [p.exitcode for p in procs]
    if hung or any(c != 0 for c in codes):
        msgs = []
        if q is not None:
            try:
                while True:
                    msgs.append(q.get_nowait())
            except Exception:
                pass
        raise AssertionError('rank exit codes {} ({} hung): {}'.format(codes, len(hung), msgs))


def _worker(rank, world, port, q):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, 'pytorch_geometric-1_b200'), os.path.join(root, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        from oracle import ref_cpu as O
        from util import skewed_graph
        from torch_geometric.dist import DistGraph, dist_gcn_forward
        from torch_geometric.nn import GCNConv
        torch.manual_seed(0)
        n = 3001
        ei = skewed_graph(n, 60000, 11)
        x = torch.randn(n, 100)
        dims = [(100, 256), (256, 256), (256, 47)]
        layers = [(torch.randn(a, b) / a ** 0.5, torch.randn(b)) for a, b in dims]
        ei2, norm = GCNConv.norm(ei.to(dev), n, None, False, torch.float32)
        g = DistGraph(ei2, norm, n, rank, world)
        out_local = dist_gcn_forward(g, g.local_rows(x.to(dev)).contiguous(),
                                     [(w.to(dev), b.to(dev)) for w, b in layers])
        parts = [torch.zeros(g.plan.block, 47, device=dev) for _ in range(world)]
        pad = torch.zeros(g.plan.block, 47, device=dev)
        pad[:g.plan.n_local] = out_local
        dist.all_gather(parts, pad)
        if rank == 0:
            got = torch.cat([parts[p][:g.plan.bounds[p + 1] - g.plan.bounds[p]]
                             for p in range(world)]).cpu()
            e2, nrm = O.gcn_norm(ei, n, dtype=torch.float32)
            assert torch.equal(e2, ei2.cpu())
            h = x
            for k, (w, b) in enumerate(layers):
                h = O.scatter_('add', nrm.view(-1, 1) * (h @ w)[e2[0]], e2[1], 0, n) + b
                if k + 1 < len(layers):
                    h = torch.relu(h)
            q.put(('ok', (got - h).abs().max().item() / h.abs().max().item()))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def test_dist_gcn_two_gpus_nccl():
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    _join_all(procs, q)
    status, err = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 2e-5, err


# ---- feature-sliced layout ---------------------------------------------------------
def _sliced_worker(rank, world, port, q, transport, x_layout, row_groups, fence='flags', chunks=1):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, 'pytorch_geometric-1_b200'), os.path.join(root, 'tests')):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        from oracle import ref_cpu as O
        from util import skewed_graph
        from torch_geometric.dist import SlicedGraph, sliced_gcn_forward
        from torch_geometric.nn import GCNConv
        torch.manual_seed(0)
        n = 3001
        ei = skewed_graph(n, 60000, 11)
        x = torch.randn(n, 100)
        dims = [(100, 256), (256, 256), (256, 47)]
        layers = [(torch.randn(a, b) / a ** 0.5, torch.randn(b)) for a, b in dims]
        ei2, norm = GCNConv.norm(ei.to(dev), n, None, False, torch.float32)
        g = SlicedGraph(ei2, norm, n, rank, world, transport=transport, row_groups=row_groups,
                        fence_mode=fence)
        pl = g.plan
        if x_layout == 'cols':
            xin = g.slice_cols(x.to(dev))
        elif x_layout == 'full':
            xin = x.to(dev)                   # replicated static input: first layer over own rows
        else:
            xin = g.local_rows(x.to(dev)).contiguous()
        lay = [(w.to(dev), b.to(dev)) for w, b in layers]
        for _ in range(3):                    # buffer reuse across steps (double-buffered rows)
            out_local = sliced_gcn_forward(g, xin, lay, x_layout=x_layout)
        if chunks > 1:
            # the chunked pipeline (aggregation of row chunk c+1 concurrent with the routed
            # transform of chunk c on a second stream): same bits as the unchunked forward
            plain = out_local.clone()
            g.set_chunks(chunks)
            from torch_geometric import _native as nat
            for _ in range(3):
                nat.event_log = []
                out_local = sliced_gcn_forward(g, xin, lay, x_layout=x_layout)
                names = [e[0] for e in nat.event_log]
                nat.event_log = None
            torch.cuda.synchronize()
            # same addends per row in the same order; a row cut by a work-item boundary is
            # combined from pieces, and the cut points move with the chunk's slot numbering,
            # so the last bits may differ (fp32 association) — compared at 1e-6 of the maximum
            rel = float((out_local - plain).abs().max() / plain.abs().max())
            ok = torch.tensor([1.0 if rel <= 1e-6 else 0.0], device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)       # fail on every rank together
            assert float(ok) == 1.0, 'chunked pipeline changed the result: {:.3e}'.format(rel)
            assert names.count('pygb_gemm_xw_routed') == chunks + 1, names       # both pipelines
        parts = [torch.zeros(pl.blk, 47, device=dev) for _ in range(world)]
        pad = torch.zeros(pl.blk, 47, device=dev)
        pad[:pl.n_local] = out_local
        dist.all_gather(parts, pad)
        if transport == 'peer':
            g.transport.close()
        if rank == 0:
            got = torch.cat(parts)[:n].cpu()
            e2, nrm = O.gcn_norm(ei, n, dtype=torch.float32)
            h = x
            for k, (w, b) in enumerate(layers):
                h = O.scatter_('add', nrm.view(-1, 1) * (h @ w)[e2[0]], e2[1], 0, n) + b
                if k + 1 < len(layers):
                    h = torch.relu(h)
            q.put(('ok', (got - h).abs().max().item() / h.abs().max().item()))
    except Exception as e:  # pragma: no cover
        q.put(('error', repr(e)))
        raise
    finally:
        dist.destroy_process_group()


def _run_sliced(world, transport, x_layout, row_groups=1, fence='flags', chunks=1):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sliced_worker,
                         args=(r, world, port, q, transport, x_layout, row_groups, fence, chunks))
             for r in range(world)]
    for p in procs:
        p.start()
    _join_all(procs, q)
    status, err = q.get(timeout=10)
    assert status == 'ok', err
    assert err <= 2e-5, err


@pytest.mark.parametrize('transport,x_layout,row_groups', [
    ('collective', 'cols', 1), ('peer', 'cols', 1), ('peer', 'rows', 1), ('peer', 'cols', 2),
    ('peer', 'full', 1), ('peer', 'full', 2), ('collective', 'full', 1)])
def test_sliced_gcn_two_gpus(transport, x_layout, row_groups):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    _run_sliced(2, transport, x_layout, row_groups)


def test_sliced_gcn_two_gpus_nccl_fence():
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    _run_sliced(2, 'peer', 'cols', 1, fence='nccl')


@pytest.mark.parametrize('transport', ['collective', 'peer'])
def test_sliced_gcn_four_gpus_two_row_groups(transport):
    if torch.cuda.device_count() < 4:
        pytest.skip('needs four GPUs')
    _run_sliced(4, transport, 'cols', 2)


@pytest.mark.parametrize('transport', ['collective', 'peer'])
@pytest.mark.parametrize('x_layout', ['cols', 'full'])
def test_sliced_gcn_one_gpu(transport, x_layout):
    """world_size 1: the same code path (routed stores into the rank's own arena)."""
    _run_sliced(1, transport, x_layout)


@pytest.mark.parametrize('x_layout,chunks', [('cols', 2), ('cols', 8), ('full', 2), ('full', 8)])
def test_sliced_gcn_one_gpu_chunked_pipeline(x_layout, chunks):
    """world_size 1, peer transport: the chunked aggregate -> transform pipelines (two streams,
    capped grids) on the rank's own arena; same result as the unchunked forward."""
    _run_sliced(1, 'peer', x_layout, fence='nccl', chunks=chunks)


@pytest.mark.parametrize('x_layout,row_groups,chunks', [('cols', 1, 2), ('cols', 1, 4), ('cols', 2, 4),
                                                         ('full', 1, 4), ('full', 2, 2)])
def test_sliced_gcn_two_gpus_chunked_pipeline(x_layout, row_groups, chunks):
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    _run_sliced(2, 'peer', x_layout, row_groups, fence='nccl', chunks=chunks)


@pytest.mark.parametrize('x_layout,chunks', [('cols', 1), ('cols', 4), ('full', 1), ('full', 4)])
def test_sliced_gcn_eight_gpus_two_by_four(x_layout, chunks):
    """The layout the driver's 8-GPU SCALE run uses: 2 row groups x 4 column slices, peer stores
    ('full': the first aggregation over the rank's own rows from the replicated input)."""
    if torch.cuda.device_count() < 8:
        pytest.skip('needs eight GPUs')
    _run_sliced(8, 'peer', x_layout, 2, fence='nccl', chunks=chunks)


def test_routed_kernels_with_capped_grids(dev):
    """max_ctas (persistent-grid caps used by the chunked pipeline) never changes a result."""
    from util import skewed_graph
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    n, E, F, blk = 1200, 40000, 32, 300
    ei = skewed_graph(n, E, 5).to(dev)
    g = torch.Generator().manual_seed(2)
    x = torch.randn(n, F, generator=g).to(dev)
    w = torch.rand(E, generator=g).to(dev)
    adj = Adj(ei, 1, (n, n))
    outs = []
    for cap in (0, 3, 40, -20, -147, -500):
        bufs = [torch.full((blk, F), -7.0, device=dev) for _ in range(4)]
        tab = torch.tensor([b.data_ptr() for b in bufs], dtype=torch.int64).to(dev)
        ops.seg_gather_reduce_routed_raw(x, adj.dst, adj.permuted(w, 'dst'), None, tab, blk, F,
                                         max_ctas=cap)
        outs.append(torch.cat(bufs))
    assert all(torch.equal(o, outs[0]) for o in outs[1:])
    M, K, N, Fc = 3000, 128, 256, 64
    a = torch.randn(M, K, generator=g).to(dev)
    wt = (torch.randn(K, N, generator=g) / K ** 0.5).to(dev)
    res = []
    for cap in (0, 1, 7, 32):
        o = [torch.zeros(M, Fc, device=dev) for _ in range(N // Fc)]
        tab = torch.tensor([t.data_ptr() for t in o], dtype=torch.int64).to(dev)
        ops.gemm_xw_routed_raw(a, wt, None, Fc, tab, 0, Fc, relu=True, max_ctas=cap)
        res.append(torch.cat(o, dim=1))
    assert all(torch.equal(r, res[0]) for r in res[1:])
    assert torch.equal(res[0], torch.relu(ops.gemm_xw_raw(a, wt)))


def test_routed_gather_reduce_and_cols_scatter(dev):
    """The routed K1 and the column scatter on one GPU: the pointer table sends row
    blocks / column blocks to separate local buffers."""
    from util import skewed_graph, assert_close
    from torch_geometric import ops
    from torch_geometric.graph import Adj
    n, E, F, blk = 1000, 30000, 24, 300
    ei = skewed_graph(n, E, 3).to(dev)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, F, generator=g).to(dev)
    w = torch.rand(E, generator=g).to(dev)
    bias = torch.randn(F, generator=g).to(dev)
    adj = Adj(ei, 1, (n, n))
    ref = ops.gather_reduce(x, adj, w, 'sum', bias)
    nblk = (n + blk - 1) // blk
    ldo = 40                                            # rows live inside a wider matrix
    bufs = [torch.full((blk, ldo), -7.0, device=dev) for _ in range(nblk)]
    tab = torch.tensor([b.data_ptr() + 8 * 4 for b in bufs], dtype=torch.int64).to(dev)
    old = ops.STREAM_T
    try:
        for T in (32, None):
            ops.STREAM_T = T
            ops.seg_gather_reduce_routed_raw(x, adj.dst, adj.permuted(w, 'dst'), bias, tab, blk, ldo)
            got = torch.cat([b[:, 8:8 + F] for b in bufs])[:n]
            assert_close(got, ref, 1e-6, 'routed rows')
            for start in (1, 300, 450, n - 1):           # rotated work order: same bits
                ops.seg_gather_reduce_routed_raw(x, adj.dst, adj.permuted(w, 'dst'), bias, tab,
                                                 blk, ldo, start_row=start)
                assert torch.equal(torch.cat([b[:, 8:8 + F] for b in bufs])[:n], got)
            assert all(float(b[:, :8].min()) == -7.0 and float(b[:, 8 + F:].max()) == -7.0
                       for b in bufs), 'stores outside the slice'
    finally:
        ops.STREAM_T = old
    # column scatter: [n, 3*8] -> 3 column blocks x 2 copies of [n0 + n, 8], with / without relu
    src = torch.randn(n, 28, generator=g).to(dev)
    for relu in (False, True):
        outs = [torch.full((n + 5, 12), -7.0, device=dev) for _ in range(6)]
        tab = torch.tensor([o.data_ptr() for o in outs], dtype=torch.int64).to(dev)
        ops.cols_scatter_raw(src, 8, 3, tab, 5, 12, relu, n_dup=2)
        for k_, o in enumerate(outs):
            q_ = k_ // 2
            want = src[:, q_ * 8:(q_ + 1) * 8]
            assert torch.equal(o[5:, :8], torch.relu(want) if relu else want)
            assert float(o[:5].max()) == -7.0 and float(o[:, 8:].max()) == -7.0


@pytest.mark.parametrize('M,K,N,Fc,dup', [(1000, 104, 256, 32, 1), (777, 256, 64, 8, 2),
                                           (130, 64, 48, 24, 1), (5, 16, 16, 16, 4)])
def test_gemm_routed_epilogue(dev, M, K, N, Fc, dup):
    """pygb_gemm_xw_routed == pygb_gemm_xw followed by the column scatter, bit for bit."""
    from torch_geometric import ops
    g = torch.Generator().manual_seed(M + N)
    x = torch.randn(M, K, generator=g).to(dev)
    w = (torch.randn(K, N, generator=g) / K ** 0.5).to(dev)
    b = torch.randn(N, generator=g).to(dev)
    for relu in (False, True):
        plain = ops.gemm_xw_raw(x, w, b, relu=relu)
        nblk, row0, ldd = N // Fc, 3, Fc + 4
        outs = [torch.full((M + row0 + 2, ldd), -7.0, device=dev) for _ in range(nblk * dup)]
        tab = torch.tensor([o.data_ptr() for o in outs], dtype=torch.int64).to(dev)
        ops.gemm_xw_routed_raw(x, w, b, Fc, tab, row0, ldd, relu=relu, n_dup=dup)
        for k_, o in enumerate(outs):
            q_ = k_ // dup
            assert torch.equal(o[row0:row0 + M, :Fc], plain[:, q_ * Fc:(q_ + 1) * Fc])
            assert float(o[:row0].max()) == -7.0 and float(o[row0 + M:].max()) == -7.0
            assert float(o[:, Fc:].max()) == -7.0
