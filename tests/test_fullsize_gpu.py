"""GPU: size-independent properties at BASELINE configs[4]'s full size (2,449,029
nodes, 62 M directed edges, E' = 64.4 M after self-loops) — sizes the CPU oracle
cannot finish.  Integer results are checked bit-exactly against independent
torch integer ops (bincount / sort / scatter_reduce), floating-point results
through linearity, a checksum of column sums, agreement of the two kernel
generations, run-to-run bit equality and an fp64 recomputation of sampled rows."""
import pytest
import torch

from util import assert_close

pytestmark = pytest.mark.gpu

N, PAIRS = 2449029, 31000000


@pytest.fixture(scope='module')
def big(dev):
    if torch.cuda.get_device_properties(dev).total_memory < 60 * 2 ** 30:
        pytest.skip('needs a large-memory GPU')
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    from torch_geometric.graph import adj_for, clear_caches
    data = synthetic.products_like(seed=4, device=dev, num_nodes=N, num_pairs=PAIRS, num_features=4)
    ei = data.edge_index
    ei2, norm = GCNConv.norm(ei, N, None, False, torch.float32)
    adj = adj_for(ei2, 1, (N, N))
    yield {'ei': ei, 'ei2': ei2, 'norm': norm, 'adj': adj}
    clear_caches()
    torch.cuda.empty_cache()


def test_self_loops_exact_and_idempotent(dev, big):
    from torch_geometric.utils import add_remaining_self_loops, contains_self_loops
    ei, ei2 = big['ei'], big['ei2']
    E = ei.size(1)
    assert not contains_self_loops(ei)
    assert ei2.size(1) == E + N
    assert torch.equal(ei2[:, :E], ei)
    loops = torch.arange(N, device=dev)
    assert torch.equal(ei2[0, E:], loops) and torch.equal(ei2[1, E:], loops)
    again, _ = add_remaining_self_loops(ei2, None, 1, N)
    assert torch.equal(again, ei2), 'add_remaining_self_loops is not idempotent'


def test_csr_is_the_stable_sort(dev, big):
    ei2, csr = big['ei2'], big['adj'].dst
    E2 = ei2.size(1)
    rowptr, col, perm = csr.rowptr.long(), csr.col.long(), csr.perm.long()
    assert rowptr[0].item() == 0 and rowptr[-1].item() == E2
    assert torch.equal(rowptr[1:] - rowptr[:-1], torch.bincount(ei2[1], minlength=N))
    assert torch.equal(torch.sort(perm)[0], torch.arange(E2, device=dev)), 'perm is not a permutation'
    assert torch.equal(col, ei2[0][perm])
    dst = ei2[1][perm]
    assert bool((dst[1:] >= dst[:-1]).all()), 'slots not sorted by destination'
    same = dst[1:] == dst[:-1]
    assert bool((perm[1:] > perm[:-1])[same].all()), 'sort is not stable inside a row'


def test_degree_and_norm_exact(dev, big):
    from torch_geometric.utils import degree
    ei2, norm = big['ei2'], big['norm']
    deg = degree(ei2[0], N)
    assert torch.equal(deg, torch.bincount(ei2[0], minlength=N).float())   # integers < 2^24: exact
    dis = deg.double().pow(-0.5)
    ref = dis[ei2[0]] * dis[ei2[1]]
    assert_close(norm, ref, 1e-6, 'gcn_norm')


def _x(F, seed, dev, dtype=torch.float32):
    g = torch.Generator(device=dev).manual_seed(seed)
    return torch.rand(N, F, generator=g, device=dev, dtype=torch.float32).to(dtype)


@pytest.mark.parametrize('F', [48, 100, 256])
def test_gather_reduce_checksum_v1_and_determinism(dev, big, F):
    from torch_geometric import ops
    ei2, norm, adj = big['ei2'], big['norm'], big['adj']
    x = _x(F, F, dev)
    old = ops.STREAM
    try:
        ops.STREAM = True
        out = ops.gather_reduce(x, adj, norm, 'sum')
        again = ops.gather_reduce(x, adj, norm, 'sum')
        ops.STREAM = False
        v1 = ops.gather_reduce(x, adj, norm, 'sum')
    finally:
        ops.STREAM = old
    assert torch.equal(out, again), 'not bit-identical from run to run'
    assert_close(out, v1, 1e-5, 'streaming vs one-warp-per-row')
    # checksum of checksums: sum_i out[i,:] == (sum of the weights leaving each source) @ x
    wsrc = torch.zeros(N, dtype=torch.float64, device=dev).index_add_(0, ei2[0], norm.double())
    ref = wsrc @ x.double()
    assert_close(out.double().sum(0), ref, 1e-6, 'column checksum')
    # fp64 recomputation of a sample of rows straight from the edge list (hubs included)
    indeg = torch.bincount(ei2[1], minlength=N)
    rows = torch.cat([torch.topk(indeg, 8)[1],
                      torch.randint(0, N, (3000,), device=dev, generator=torch.Generator(device=dev).manual_seed(1))])
    rows = torch.unique(rows)
    pick = torch.zeros(N, dtype=torch.bool, device=dev)
    pick[rows] = True
    e = pick[ei2[1]].nonzero().view(-1)
    pos = torch.full((N,), -1, dtype=torch.long, device=dev)
    pos[rows] = torch.arange(rows.numel(), device=dev)
    ref = torch.zeros(rows.numel(), F, dtype=torch.float64, device=dev)
    ref.index_add_(0, pos[ei2[1][e]], norm[e].double().view(-1, 1) * x[ei2[0][e]].double())
    assert_close(out[rows], ref, 1e-5, 'sampled rows vs fp64')
    del out, again, v1, x


def test_gather_reduce_linearity_and_unit_rows(dev, big):
    from torch_geometric import ops
    norm, adj, ei2 = big['norm'], big['adj'], big['ei2']
    F = 64
    x, y = _x(F, 11, dev), _x(F, 12, dev) - 0.5
    ax = ops.gather_reduce(x, adj, norm, 'sum')
    ay = ops.gather_reduce(y, adj, norm, 'sum')
    az = ops.gather_reduce(2.0 * x - 3.0 * y, adj, norm, 'sum')
    assert_close(az, 2.0 * ax - 3.0 * ay, 1e-5, 'linearity')
    del ax, ay, az, y
    # A·1 with unit weights is the in-degree: integer-valued fp32, exact
    ones = torch.ones(N, 4, device=dev)
    cnt = ops.gather_reduce(ones, adj, None, 'sum')
    assert torch.equal(cnt[:, 0], torch.bincount(ei2[1], minlength=N).float())
    mean = ops.gather_reduce(ones, adj, None, 'mean')
    assert torch.equal(mean, ones)                     # every row has its self-loop


def test_bf16_features_within_tolerance(dev, big):
    from torch_geometric import ops
    norm, adj = big['norm'], big['adj']
    x = _x(256, 21, dev)
    ref = ops.gather_reduce(x.bfloat16().float(), adj, norm, 'sum')
    got = ops.gather_reduce(x.bfloat16(), adj, norm, 'sum')
    assert got.dtype == torch.bfloat16
    assert_close(got.float(), ref, 1e-2, 'bf16')


def test_max_argmax_first_wins_exact(dev, big):
    """scatter_max over all E' slots with heavy ties: value and argmax bit-exact
    against torch's integer scatter_reduce (first maximum in edge order wins)."""
    from torch_geometric.ops import scatter_max
    ei2 = big['ei2']
    E2, F = ei2.size(1), 8
    g = torch.Generator(device=dev).manual_seed(5)
    src = torch.randint(-3, 4, (E2, F), generator=g, device=dev).float()
    out, arg = scatter_max(src, ei2[1], N)
    idx = ei2[1].view(-1, 1).expand(-1, F)
    ref = torch.full((N, F), -1e30, device=dev).scatter_reduce_(0, idx, src, 'amax')
    assert torch.equal(out, ref)                       # every row is non-empty (self-loops)
    eid = torch.arange(E2, device=dev).view(-1, 1).expand(-1, F)
    cand = torch.where(src == ref[ei2[1]], eid, torch.full_like(eid, E2))
    first = torch.full((N, F), E2, dtype=torch.long, device=dev).scatter_reduce_(0, idx, cand, 'amin')
    assert torch.equal(arg, first)


def test_edge_softmax_rows_sum_to_one(dev, big):
    from torch_geometric.utils import softmax
    ei2 = big['ei2']
    E2, H = ei2.size(1), 4
    g = torch.Generator(device=dev).manual_seed(6)
    src = torch.randn(E2, H, generator=g, device=dev) * 4
    a = softmax(src, ei2[1], N)
    assert bool((a >= 0).all()) and bool((a <= 1).all())
    tot = torch.zeros(N, H, dtype=torch.float64, device=dev).index_add_(0, ei2[1], a.double())
    assert_close(tot, torch.ones_like(tot), 1e-5, 'softmax row sums')
    # shift invariance per destination
    shift = torch.randn(N, H, generator=g, device=dev)
    b = softmax(src + shift[ei2[1]], ei2[1], N)
    assert float((a - b).abs().max()) < 1e-5


def test_gat_fused_matches_unfused_full_size(dev, big):
    """Fused streaming GAT == (gat_alpha weights) x (weighted K1 sum) at full size."""
    from torch_geometric import ops
    adj, ei2 = big['adj'], big['ei2']
    H, C = 2, 16
    g = torch.Generator(device=dev).manual_seed(7)
    xw = torch.randn(N, H * C, generator=g, device=dev)
    att = torch.randn(1, H, 2 * C, generator=g, device=dev)
    out = ops.gat_propagate(xw, xw, att, None, adj, H, C, 0.2)
    again = ops.gat_propagate(xw, xw, att, None, adj, H, C, 0.2)
    assert torch.equal(out, again)
    x3 = xw.view(N, H, C)
    a_src = (x3 * att[:, :, C:]).sum(-1)
    a_dst = (x3 * att[:, :, :C]).sum(-1)
    logit = torch.nn.functional.leaky_relu(a_dst[ei2[1]] + a_src[ei2[0]], 0.2)
    from torch_geometric.utils import softmax
    alpha = softmax(logit, ei2[1], N)
    rows = torch.randint(0, N, (2000,), device=dev, generator=torch.Generator(device=dev).manual_seed(2))
    rows = torch.unique(torch.cat([rows, torch.topk(torch.bincount(ei2[1], minlength=N), 4)[1]]))
    pick = torch.zeros(N, dtype=torch.bool, device=dev)
    pick[rows] = True
    e = pick[ei2[1]].nonzero().view(-1)
    pos = torch.full((N,), -1, dtype=torch.long, device=dev)
    pos[rows] = torch.arange(rows.numel(), device=dev)
    ref = torch.zeros(rows.numel(), H, C, dtype=torch.float64, device=dev)
    ref.index_add_(0, pos[ei2[1][e]], alpha[e].double().unsqueeze(-1) * x3[ei2[0][e]].double())
    assert_close(out[rows].view(-1, H, C), ref, 2e-5, 'fused GAT vs unfused')


def test_gat_backward_stream_full_size(dev, big):
    """Streaming GAT backward (two walks, alpha / ds recomputed) at full size: bit-identical
    from run to run, equal to the per-edge kernels it replaces, and consistent with the
    forward through a directional derivative <dL/dx, v> ~ (L(x + hv) - L(x - hv)) / 2h for a
    linear loss L = <out, g> evaluated in the forward kernel."""
    from torch_geometric import ops
    adj = big['adj']
    H, C = 4, 8
    gen = torch.Generator(device=dev).manual_seed(9)
    xw0 = torch.randn(N, H * C, generator=gen, device=dev)
    att0 = torch.randn(1, H, 2 * C, generator=gen, device=dev) * 0.5
    g = torch.randn(N, H * C, generator=gen, device=dev)
    old_stream = ops.STREAM

    def grads(stream_bwd):
        xw = xw0.clone().requires_grad_(True)
        att = att0.clone().requires_grad_(True)
        ops.STREAM = True
        out = ops.gat_propagate(xw, xw, att, None, adj, H, C, 0.2)
        ops.STREAM = stream_bwd
        out.backward(g)
        return xw.grad, att.grad

    try:
        dx, datt = grads(True)
        dx2, datt2 = grads(True)
        assert torch.equal(dx, dx2) and torch.equal(datt, datt2), 'not deterministic'
        dx_old, datt_old = grads(False)
        assert_close(dx, dx_old, 2e-5, 'dx: streaming vs per-edge kernels')
        assert_close(datt, datt_old, 2e-4, 'datt: streaming vs per-edge kernels')
    finally:
        ops.STREAM = old_stream
    v = torch.randn(N, H * C, generator=gen, device=dev)
    h = 1e-2
    with torch.no_grad():
        xp, xm = xw0 + h * v, xw0 - h * v
        lp = (ops.gat_propagate(xp, xp, att0, None, adj, H, C, 0.2).double() * g).sum()
        lm = (ops.gat_propagate(xm, xm, att0, None, adj, H, C, 0.2).double() * g).sum()
    fd = float((lp - lm) / (2 * h))
    an = float((dx.double() * v).sum())
    assert abs(fd - an) <= 2e-3 * max(abs(an), 1.0) + 5e-2 * float(N) ** 0.5 * h, (fd, an)
