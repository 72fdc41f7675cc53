"""SURVEY §8(f) rows 2–3 on the GPU: the convolutions that fall onto the fused
propagate routes (ChebConv, GraphConv, SGConv, APPNP, SAGEConv, GINConv, AGNNConv)
and global pooling, against their unfused CPU restatement (oracle/ref_next.py):
forward within 1e-5, gradients within 1e-4, max-pool bit-exact."""
import pytest
import torch

from oracle import ref_next as R
from util import assert_close, random_graph, skewed_graph

pytestmark = pytest.mark.gpu


def _graph(n, E, seed):
    ei = random_graph(n, n, E, seed)                 # self-loops and duplicates included
    ei[:, :5] = ei[:, 5:10]
    return ei


def _copy_params(dst, src):
    sd = {k: v.detach().cpu().clone() for k, v in src.state_dict().items()}
    missing = dst.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing


def _check(gpu_mod, ref_mod, inputs_gpu, inputs_ref, fwd_tol=1e-5, grad_tol=1e-4):
    xg = inputs_gpu[0].clone().requires_grad_(True)
    xr = inputs_ref[0].clone().requires_grad_(True)
    out = gpu_mod(xg, *inputs_gpu[1:])
    ref = ref_mod(xr, *inputs_ref[1:])
    assert_close(out, ref, fwd_tol, 'forward')
    g = torch.randn(ref.shape, generator=torch.Generator().manual_seed(9))
    out.backward(g.to(out.device))
    ref.backward(g)
    assert_close(xg.grad, xr.grad, grad_tol, 'grad x')
    ref_params = dict(ref_mod.named_parameters())
    for name, p in gpu_mod.named_parameters():
        if name in ref_params and ref_params[name].grad is not None:
            assert_close(p.grad, ref_params[name].grad, grad_tol, 'grad ' + name)


@pytest.mark.parametrize('K', [1, 2, 4])
@pytest.mark.parametrize('weighted', [False, True])
def test_cheb_conv(dev, K, weighted):
    from torch_geometric.nn import ChebConv
    n, E, fin, fout = 300, 3000, 20, 12
    ei = _graph(n, E, K)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    w = torch.rand(E, generator=torch.Generator().manual_seed(2)) + 0.1 if weighted else None
    torch.manual_seed(0)
    m = ChebConv(fin, fout, K).to(dev)
    r = R.ChebConv(fin, fout, K)
    _copy_params(r, m)
    extra_g = (ei.to(dev),) + ((w.to(dev),) if weighted else ())
    extra_r = (ei,) + ((w,) if weighted else ())
    _check(m, r, (x.to(dev),) + extra_g, (x,) + extra_r)


@pytest.mark.parametrize('aggr', ['add', 'mean', 'max'])
def test_graph_conv(dev, aggr):
    from torch_geometric.nn import GraphConv
    n, E, fin, fout = 250, 2500, 16, 24
    ei = _graph(n, E, 3)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(0)
    m = GraphConv(fin, fout, aggr=aggr).to(dev)
    r = R.GraphConv(fin, fout, aggr=aggr)
    _copy_params(r, m)
    _check(m, r, (x.to(dev), ei.to(dev)), (x, ei))
    if aggr == 'add':
        w = torch.rand(E, generator=torch.Generator().manual_seed(2))
        _check(m, r, (x.to(dev), ei.to(dev), w.to(dev)), (x, ei, w))


@pytest.mark.parametrize('K', [1, 3])
def test_sg_conv_and_appnp(dev, K):
    from torch_geometric.nn import SGConv, APPNP
    n, E, fin, fout = 400, 5000, 32, 7
    ei = skewed_graph(n, E, 4)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(0)
    m = SGConv(fin, fout, K=K).to(dev)
    r = R.SGConv(fin, fout, K=K)
    _copy_params(r, m)
    _check(m, r, (x.to(dev), ei.to(dev)), (x, ei))
    _check(APPNP(K, 0.1).to(dev), R.APPNP(K, 0.1), (x.to(dev), ei.to(dev)), (x, ei))


def test_sg_conv_cached(dev):
    from torch_geometric.nn import SGConv
    n, E = 200, 1500
    ei = _graph(n, E, 6).to(dev)
    x = torch.randn(n, 8, generator=torch.Generator().manual_seed(1)).to(dev)
    m = SGConv(8, 4, K=2, cached=True).to(dev)
    a = m(x, ei)
    b = m(x * 0 + 1, ei)                                # cached propagation is reused
    assert torch.equal(a, b)


@pytest.mark.parametrize('normalize', [False, True])
def test_sage_conv(dev, normalize):
    from torch_geometric.nn import SAGEConv
    n, E, fin, fout = 300, 2000, 12, 20
    ei = _graph(n, E, 7)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(0)
    m = SAGEConv(fin, fout, normalize=normalize).to(dev)
    r = R.SAGEConv(fin, fout, normalize=normalize)
    _copy_params(r, m)
    _check(m, r, (x.to(dev), ei.to(dev)), (x, ei))


def test_gin_conv(dev):
    from torch_geometric.nn import GINConv
    n, E, fin = 300, 2500, 10
    ei = _graph(n, E, 8)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(fin, 16), torch.nn.ReLU(), torch.nn.Linear(16, 6))
    mlp_r = torch.nn.Sequential(torch.nn.Linear(fin, 16), torch.nn.ReLU(), torch.nn.Linear(16, 6))
    m = GINConv(mlp, eps=0.3)                           # re-initialises mlp (reset_parameters)
    mlp_r.load_state_dict(m.nn.state_dict())
    m = m.to(dev)
    r = R.GINConv(mlp_r, eps=0.3)
    _check(m, r, (x.to(dev), ei.to(dev)), (x, ei))


def test_agnn_conv(dev):
    from torch_geometric.nn import AGNNConv
    n, E, fin = 300, 4000, 16
    ei = skewed_graph(n, E, 9)
    x = torch.randn(n, fin, generator=torch.Generator().manual_seed(1))
    m = AGNNConv(requires_grad=True).to(dev)
    m.beta.data.fill_(1.7)
    r = R.AGNNConv(beta=1.7)
    _check(m, r, (x.to(dev), ei.to(dev)), (x, ei), fwd_tol=2e-5)


def test_global_pools(dev):
    from torch_geometric.nn import global_add_pool, global_mean_pool, global_max_pool
    g = torch.Generator().manual_seed(3)
    sizes = [1, 40, 7, 300, 2, 90]
    batch = torch.cat([torch.full((s,), i, dtype=torch.long) for i, s in enumerate(sizes)])
    x = (torch.randint(-4, 5, (batch.numel(), 33), generator=g).float() / 4)     # ties for max
    for fn, rf, exact in ((global_add_pool, R.global_add_pool, False),
                          (global_mean_pool, R.global_mean_pool, False),
                          (global_max_pool, R.global_max_pool, True)):
        xg = x.to(dev).requires_grad_(True)
        xr = x.clone().requires_grad_(True)
        out, ref = fn(xg, batch.to(dev)), rf(xr, batch)
        if exact:
            assert torch.equal(out.cpu(), ref)
        else:
            assert_close(out, ref, 1e-6, fn.__name__)
        w = torch.randn(ref.shape, generator=g)
        (out * w.to(dev)).sum().backward()
        (ref * w).sum().backward()
        assert_close(xg.grad, xr.grad, 1e-6, fn.__name__ + ' grad')   # max: first maximum gets it
    # explicit size with trailing empty graphs
    out = global_add_pool(x.to(dev), batch.to(dev), size=9)
    assert out.shape == (9, 33) and float(out[6:].abs().max()) == 0.0


@pytest.mark.parametrize('ratio', [0.8, 0.5, 0.34])
def test_topk_pooling(dev, ratio):
    """TopKPooling as the fork's topk_pool_Net calls it (ConvexPruning.py:306-327)."""
    from torch_geometric.nn import TopKPooling, global_max_pool, global_mean_pool
    g = torch.Generator().manual_seed(int(ratio * 100))
    sizes = [13, 1, 40, 7, 25]
    batch = torch.cat([torch.full((s,), i, dtype=torch.long) for i, s in enumerate(sizes)])
    n = batch.numel()
    x = torch.randn(n, 12, generator=g)
    cum = [0]
    for s_ in sizes:
        cum.append(cum[-1] + s_)
    ei = torch.cat([random_graph(s_, s_, 4 * s_, 10 + i) + cum[i] for i, s_ in enumerate(sizes)], 1)
    torch.manual_seed(0)
    pool = TopKPooling(12, ratio=ratio).to(dev)
    xg = x.to(dev).requires_grad_(True)
    xo, eio, _, bo, perm, score = pool(xg, ei.to(dev), None, batch.to(dev))
    xr = x.clone().requires_grad_(True)
    w = pool.weight.detach().cpu().view(-1).clone().requires_grad_(True)
    rx, rei, rb, rperm, rscore = R.topk_pooling(xr, ei, batch, w, ratio)
    assert torch.equal(perm.cpu(), rperm) and torch.equal(bo.cpu(), rb)
    assert torch.equal(eio.cpu(), rei)
    assert_close(xo, rx, 1e-6, 'pooled x')
    assert_close(score, rscore, 1e-6, 'score')
    # readout of the fork's model: cat(global max pool, global mean pool), gradients
    ro = torch.cat([global_max_pool(xo, bo), global_mean_pool(xo, bo)], dim=1)
    rr = torch.cat([R.global_max_pool(rx, rb), R.global_mean_pool(rx, rb)], dim=1)
    assert_close(ro, rr, 1e-6, 'readout')
    gw = torch.randn(rr.shape, generator=g)
    (ro * gw.to(dev)).sum().backward()
    (rr * gw).sum().backward()
    assert_close(xg.grad, xr.grad, 1e-5, 'grad x')
    assert_close(pool.weight.grad.view(-1), w.grad, 1e-5, 'grad weight')


@pytest.mark.parametrize('D,k,loop', [(3, 20, False), (3, 6, True), (16, 8, False), (64, 20, False)])
def test_knn_graph_kernel(dev, D, k, loop):
    """pygb_knn_graph (csrc/knn.cu) against an fp64 brute force: edges grouped by centre,
    neighbours in ascending distance, the k-th distance exact to fp32 rounding, examples of
    different sizes (one smaller than k)."""
    from torch_geometric.nn import knn_graph
    gen = torch.Generator().manual_seed(D + k)
    sizes = [300, 5, 1024, 130, 77]
    n = sum(sizes)
    x = torch.rand(n, D, generator=gen) * 2 - 1
    batch = torch.cat([torch.full((s,), i) for i, s in enumerate(sizes)]).long()
    ei = knn_graph(x.to(dev), k, batch.to(dev), loop=loop).cpu()
    assert ei.dtype == torch.long and torch.equal(batch[ei[0]], batch[ei[1]])
    assert bool((ei[1][1:] >= ei[1][:-1]).all()), 'edges not grouped by centre'
    xd = x.double()
    start = 0
    e = 0
    for s in sizes:
        d = torch.cdist(xd[start:start + s], xd[start:start + s]) ** 2
        if not loop:
            d.fill_diagonal_(float('inf'))
        kk = min(k, s - (0 if loop else 1))
        ref_d = d.topk(kk, dim=1, largest=False)[0]
        for i in range(s):
            assert bool((ei[1][e:e + kk] == start + i).all())
            got = d[i, ei[0][e:e + kk] - start]
            assert bool((got[1:] >= got[:-1] - 1e-6).all()), 'neighbours not in ascending distance'
            assert torch.allclose(got, ref_d[i], rtol=1e-5, atol=1e-7)
            if not loop:
                assert bool((ei[0][e:e + kk] != start + i).all())
            e += kk
        start += s
    assert e == ei.size(1)
    # flow and the default batch
    assert torch.equal(knn_graph(x.to(dev), k, batch.to(dev), loop=loop, flow='target_to_source').cpu(),
                       ei.flip(0))
    one = knn_graph(x[:300].to(dev), 4).cpu()
    assert one.shape == (2, 1200)


def test_dynamic_edge_conv_uses_knn_kernel(dev):
    from torch_geometric.nn import DynamicEdgeConv
    from torch_geometric import _native as nat
    torch.manual_seed(0)
    mlp = torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.ReLU(), torch.nn.Linear(16, 16))
    conv = DynamicEdgeConv(mlp, k=6).to(dev)
    pos = torch.rand(512, 3, device=dev)
    batch = torch.arange(4, device=dev).repeat_interleave(128)
    nat.event_log = []
    out = conv(pos, batch)
    names = [e[0] for e in nat.event_log]
    nat.event_log = None
    assert 'pygb_knn_graph' in names and out.shape == (512, 16)


# ---- f4: fps / radius / radius_graph (examples/pointnet++.py:39-41,46-48) -----------------
def _grid_cloud(n, seed):
    """Coordinates on a 1/8 grid: every squared distance is exact in fp32 (and ties are
    frequent), so the selection order can be compared index for index."""
    g = torch.Generator().manual_seed(seed)
    return ((torch.rand(n, 3, generator=g) * 2 - 1) * 8).round() / 8


@pytest.mark.parametrize('ratio', [0.5, 0.25, 1.0])
def test_fps_matches_loop_reference_including_ties(dev, ratio):
    from torch_geometric.nn import fps
    sizes = [130, 1, 77, 300]
    x = _grid_cloud(sum(sizes), 3)
    batch = torch.cat([torch.full((s,), b) for b, s in enumerate(sizes)]).long()
    got = fps(x.to(dev), batch.to(dev), ratio=ratio, random_start=False)
    ref = R.fps(x, batch, ratio)
    assert got.dtype == torch.long and torch.equal(got.cpu(), ref)
    assert torch.equal(got, fps(x.to(dev), batch.to(dev), ratio=ratio, random_start=False))
    # random start: a valid farthest-point sequence from whatever start was drawn
    rnd = fps(x.to(dev), batch.to(dev), ratio=ratio, random_start=True).cpu()
    assert rnd.numel() == ref.numel()
    assert torch.equal(batch[rnd], batch[ref])
    starts, o = [], 0
    for s in sizes:
        m = -(-int(round(ratio * s * 1e6)) // 10 ** 6)
        starts.append(int(rnd[o]) - int((batch < len(starts)).sum()))
        o += m
    assert torch.equal(rnd, R.fps(x, batch, ratio, start=starts))
    assert fps(torch.zeros(0, 3, device=dev)).numel() == 0


def test_fps_on_cfg4_sized_clouds(dev):
    """32 x 1,024 points, ratio 0.5 (examples/pointnet++.py:39): the samples are distinct, per
    cloud, and each one maximises the distance to the earlier ones (checked in fp64)."""
    from torch_geometric.nn import fps
    g = torch.Generator().manual_seed(1)
    x = torch.rand(32 * 1024, 3, generator=g) * 2 - 1
    batch = torch.arange(32).repeat_interleave(1024)
    idx = fps(x.to(dev), batch.to(dev), ratio=0.5, random_start=False).cpu()
    assert idx.numel() == 32 * 512 and idx.unique().numel() == idx.numel()
    assert torch.equal(batch[idx], torch.arange(32).repeat_interleave(512))
    xd = x.double()
    for c in (0, 17, 31):
        sel = idx[c * 512:(c + 1) * 512]
        pts = xd[c * 1024:(c + 1) * 1024]
        assert int(sel[0]) == c * 1024
        dmin = torch.full((1024,), float('inf'), dtype=torch.float64)
        for t in range(40):
            dmin = torch.minimum(dmin, ((pts - xd[sel[t]]) ** 2).sum(1))
            nxt = int(sel[t + 1]) - c * 1024
            assert dmin[nxt] >= dmin.max() * (1 - 1e-6)


@pytest.mark.parametrize('max_nn', [4, 32])
def test_radius_and_radius_graph_match_loop_reference(dev, max_nn):
    from torch_geometric.nn import radius, radius_graph
    sizes = [90, 40, 1, 120]
    x = _grid_cloud(sum(sizes), 5)
    bx = torch.cat([torch.full((s,), b) for b, s in enumerate(sizes)]).long()
    sel = torch.arange(0, x.size(0), 3)
    y, by = x[sel] + 0.125, bx[sel]
    got = radius(x.to(dev), y.to(dev), 0.5, bx.to(dev), by.to(dev), max_num_neighbors=max_nn)
    ref = R.radius(x, y, 0.5, bx, by, max_nn)
    assert torch.equal(got.cpu(), ref)
    for loop in (False, True):
        for flow in ('source_to_target', 'target_to_source'):
            gg = radius_graph(x.to(dev), 0.5, bx.to(dev), loop=loop, max_num_neighbors=max_nn,
                              flow=flow)
            rr = R.radius_graph(x, 0.5, bx, loop=loop, max_num_neighbors=max_nn, flow=flow)
            assert torch.equal(gg.cpu(), rr), (loop, flow)
    assert radius(x.to(dev), torch.zeros(0, 3, device=dev), 0.5).shape == (2, 0)


def test_knn_graph_kernel_large_k_and_ties(dev):
    """k-best list in shared memory (round 2): k up to 64, tie rule = lower index."""
    from torch_geometric.nn import knn_graph
    x = _grid_cloud(400, 9)
    batch = torch.cat([torch.zeros(150), torch.ones(250)]).long()
    for k in (1, 20, 64):
        got = knn_graph(x.to(dev), k, batch.to(dev)).cpu()
        ref = R.knn_graph(x, k, batch)
        assert torch.equal(got, ref), k
