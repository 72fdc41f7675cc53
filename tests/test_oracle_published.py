"""CPU: the oracle against the PUBLISHED algorithms it restates, computed a second, independent
way (dense linear algebra / literal per-node loops in fp64) — a pin that does not go through
``oracle/ref_cpu.py``'s own gather / scatter helpers:

* GCN (Kipf & Welling, ICLR 2017, eq. 2):  H' = D^-1/2 (A + I) D^-1/2 X W + b, via ``scipy.sparse``;
  on a directed multigraph the matrix form of the pinned code (degree by source row);
* GAT (Velickovic et al., ICLR 2018, eq. 1-6): e_ij = LeakyReLU(a^T [W h_i || W h_j]), softmax over
  N(i) with a self-loop, concatenated / averaged heads;
* EdgeConv (Wang et al., DGCNN, eq. 8):  x_i' = max_{j in N(i)} h(x_i, x_j - x_i);
* the scatter reductions against ATen's own ``index_add_`` / ``scatter_reduce`` (what
  torch-scatter 2.0.4's CPU path reduces to).

"Parity unpinned"  is about the reference's CODE, which is not in the tree;
these tests tie the oracle to the papers' formulas, which are."""
import numpy as np
import pytest
import torch

from oracle import ref_cpu as O

F64 = torch.float64


def _sym_graph(n, pairs, seed, dup=0):
    g = torch.Generator().manual_seed(seed)
    a = torch.randint(0, n, (pairs,), generator=g)
    b = torch.randint(0, n, (pairs,), generator=g)
    keep = a != b
    a, b = a[keep], b[keep]
    a, b = torch.cat([a, a[:dup]]), torch.cat([b, b[:dup]])       # duplicate pairs count twice
    ei = torch.stack([torch.cat([a, b]), torch.cat([b, a])])
    return ei[:, torch.randperm(ei.size(1), generator=g)]


@pytest.mark.parametrize('improved', [False, True])
def test_gcn_oracle_is_kipf_welling_eq2(improved):
    import scipy.sparse as sp
    n, fin, fout = 60, 7, 5
    ei = _sym_graph(n, 150, 1, dup=20)
    g = torch.Generator().manual_seed(2)
    x = torch.randn(n, fin, generator=g, dtype=F64)
    conv = O.GCNConv(fin, fout, improved=improved).double()
    with torch.no_grad():
        conv.bias.uniform_(-1, 1)
    got = conv(x, ei).detach().numpy()
    A = sp.coo_matrix((np.ones(ei.size(1)), (ei[0].numpy(), ei[1].numpy())), shape=(n, n)).tocsr()
    A_hat = A + (2.0 if improved else 1.0) * sp.identity(n)
    d = np.asarray(A_hat.sum(axis=1)).ravel()
    D = sp.diags(d ** -0.5)
    want = (D @ A_hat @ D) @ x.numpy() @ conv.weight.detach().numpy() + conv.bias.detach().numpy()
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)


def test_gcn_oracle_matrix_form_on_a_directed_weighted_multigraph():
    """out = (diag(s) A_hat diag(s))^T X W + b with s = rowsum(A_hat)^-1/2: degree by SOURCE row,
    messages source -> target, duplicate edges add up, isolated rows give 0."""
    n, fin, fout, E = 40, 6, 4, 300
    g = torch.Generator().manual_seed(3)
    src = torch.randint(0, n - 3, (E,), generator=g)       # the last 3 nodes never send
    dst = torch.randint(0, n, (E,), generator=g)
    keep = src != dst
    ei = torch.stack([src[keep], dst[keep]])
    w = torch.rand(ei.size(1), generator=g, dtype=F64) + 0.1
    x = torch.randn(n, fin, generator=g, dtype=F64)
    conv = O.GCNConv(fin, fout).double()
    got = conv(x, ei, w).detach()
    A_hat = torch.zeros(n, n, dtype=F64)
    A_hat.index_put_((ei[0], ei[1]), w, accumulate=True)
    A_hat += torch.eye(n, dtype=F64)
    s = A_hat.sum(1).pow(-0.5)
    M = s.view(-1, 1) * A_hat * s.view(1, -1)
    want = M.t() @ x @ conv.weight.detach() + conv.bias.detach()
    assert torch.allclose(got, want, rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize('heads,concat', [(1, True), (3, True), (3, False)])
def test_gat_oracle_is_velickovic_eq1_to_6(heads, concat):
    n, fin, C, slope = 14, 5, 4, 0.2
    g = torch.Generator().manual_seed(4)
    E = 45
    src = torch.randint(0, n, (E,), generator=g)
    dst = torch.randint(0, n, (E,), generator=g)
    ei = torch.stack([src, dst])                          # includes some self-loops: replaced by one each
    x = torch.randn(n, fin, generator=g, dtype=F64)
    conv = O.GATConv(fin, C, heads=heads, concat=concat, negative_slope=slope).double()
    with torch.no_grad():
        conv.bias.uniform_(-1, 1)
    got = conv(x, ei).detach().numpy()
    W = conv.weight.detach().numpy().reshape(fin, heads, C)
    a = conv.att.detach().numpy().reshape(heads, 2 * C)
    h = np.einsum('nf,fhc->nhc', x.numpy(), W)            # W h_i per head
    out = np.zeros((n, heads, C))
    for i in range(n):
        # neighbourhood of i: one entry per edge j -> i with j != i (a multigraph keeps duplicates),
        # plus i itself exactly once
        nbrs = [int(s) for s, d in zip(src.tolist(), dst.tolist()) if d == i and s != i] + [i]
        for k in range(heads):
            e = np.array([np.dot(a[k], np.concatenate([h[i, k], h[j, k]])) for j in nbrs])
            e = np.where(e > 0, e, slope * e)
            alpha = np.exp(e - e.max())
            alpha /= alpha.sum()
            out[i, k] = sum(al * h[j, k] for al, j in zip(alpha, nbrs))
    want = out.reshape(n, heads * C) if concat else out.mean(axis=1)
    want = want + conv.bias.detach().numpy()
    np.testing.assert_allclose(got, want, rtol=1e-11, atol=1e-12)


def test_edge_conv_oracle_is_dgcnn_eq8():
    n, k = 30, 4
    g = torch.Generator().manual_seed(5)
    x = torch.randn(n, 3, generator=g, dtype=F64)
    d = (x[:, None] - x[None]).pow(2).sum(-1)
    d.fill_diagonal_(float('inf'))
    nbr = d.topk(k, largest=False).indices               # k nearest neighbours of every point
    ei = torch.stack([nbr.reshape(-1), torch.arange(n).repeat_interleave(k)])   # j -> i
    mlp = torch.nn.Sequential(torch.nn.Linear(6, 8), torch.nn.ReLU(), torch.nn.Linear(8, 5)).double()
    got = O.EdgeConv(mlp)(x, ei).detach()
    want = torch.stack([
        torch.stack([mlp(torch.cat([x[i], x[j] - x[i]])) for j in nbr[i].tolist()]).max(0).values
        for i in range(n)]).detach()
    assert torch.allclose(got, want, rtol=1e-12, atol=1e-12)


def test_scatter_oracle_equals_aten_reductions():
    g = torch.Generator().manual_seed(6)
    n, E, F = 25, 400, 6
    index = torch.randint(0, n - 4, (E,), generator=g)     # rows n-4.. stay empty
    src = torch.randn(E, F, generator=g, dtype=F64)
    ix = index.view(-1, 1).expand(-1, F)
    add = torch.zeros(n, F, dtype=F64).index_add_(0, index, src)
    assert torch.allclose(O.scatter_add(src, index, 0, None, n), add, rtol=1e-13, atol=1e-13)
    mean = torch.zeros(n, F, dtype=F64).scatter_reduce(0, ix, src, 'mean', include_self=False)
    assert torch.allclose(O.scatter_mean(src, index, 0, None, n), mean, rtol=1e-13, atol=1e-13)
    amax = torch.zeros(n, F, dtype=F64).scatter_reduce(0, ix, src, 'amax', include_self=False)
    val, arg = O.scatter_max(src, index, 0, None, n)
    assert torch.equal(val[:n - 4], amax[:n - 4])
    assert torch.equal(src.gather(0, arg[:n - 4]), val[:n - 4])           # argmax points at the maximum
    assert torch.equal(index[arg[:n - 4]], torch.arange(n - 4).view(-1, 1).expand(-1, F))
    assert torch.equal(O.scatter_('max', src, index, 0, n)[n - 4:], torch.zeros(4, F, dtype=F64))
    # softmax over groups == per-group torch.softmax
    logits = torch.randn(E, 3, generator=g, dtype=F64)
    sm = O.softmax(logits, index, n)
    for r in range(n - 4):
        m = index == r
        if m.any():
            assert torch.allclose(sm[m], torch.softmax(logits[m], dim=0), rtol=1e-12, atol=1e-14)


# ---- the widened rows (oracle/ref_next.py) -------------------------------------------------
def _dense(ei, n, w=None):
    """A[src, dst] with duplicate edges added up."""
    A = torch.zeros(n, n, dtype=F64)
    A.index_put_((ei[0], ei[1]), torch.ones(ei.size(1), dtype=F64) if w is None else w, accumulate=True)
    return A


def _digraph(n, E, seed, loops=False):
    g = torch.Generator().manual_seed(seed)
    src = torch.randint(0, n, (E,), generator=g)
    dst = torch.randint(0, n, (E,), generator=g)
    if not loops:
        keep = src != dst
        src, dst = src[keep], dst[keep]
    return torch.stack([src, dst])


def test_sgc_appnp_cheb_oracles_are_their_matrix_recurrences():
    from oracle import ref_next as R
    n, fin, fout = 30, 5, 4
    ei = _digraph(n, 200, 7)
    g = torch.Generator().manual_seed(8)
    x = torch.randn(n, fin, generator=g, dtype=F64)
    A_hat = _dense(ei, n) + torch.eye(n, dtype=F64)
    s = A_hat.sum(1).pow(-0.5)
    S = (s.view(-1, 1) * A_hat * s.view(1, -1)).t()          # propagation matrix: out = S x
    # SGC (Wu et al. 2019): S^K X Theta
    sg = R.SGConv(fin, fout, K=3).double()
    want = torch.linalg.matrix_power(S, 3) @ x @ sg.lin.weight.detach().t() + sg.lin.bias.detach()
    assert torch.allclose(sg(x, ei).detach(), want, rtol=1e-12, atol=1e-12)
    # APPNP (Klicpera et al. 2019): Z^(k+1) = (1 - alpha) S Z^(k) + alpha H
    z = x
    for _ in range(4):
        z = 0.85 * (S @ z) + 0.15 * x
    assert torch.allclose(R.APPNP(4, 0.15)(x, ei), z, rtol=1e-12, atol=1e-12)
    # ChebNet (Defferrard et al. 2016) with lambda_max = 2: L~ = -D^-1/2 A D^-1/2
    A = _dense(ei, n)
    d = A.sum(1).pow(-0.5)
    d[torch.isinf(d)] = 0
    L = -(d.view(-1, 1) * A * d.view(1, -1)).t()
    ch = R.ChebConv(fin, fout, K=4).double()
    with torch.no_grad():
        ch.weight.normal_(generator=g)
        ch.bias.uniform_(-1, 1, generator=g)
    T = [x, L @ x]
    for _ in range(2, 4):
        T.append(2 * L @ T[-1] - T[-2])
    want = sum(t @ ch.weight[k].detach() for k, t in enumerate(T)) + ch.bias.detach()
    assert torch.allclose(ch(x, ei).detach(), want, rtol=1e-12, atol=1e-12)


def test_sage_gin_graphconv_agnn_oracles_are_their_formulas():
    from oracle import ref_next as R
    n, fin, fout = 24, 5, 3
    ei = _digraph(n, 150, 9, loops=True)
    nl = ei[:, ei[0] != ei[1]]
    g = torch.Generator().manual_seed(10)
    x = torch.randn(n, fin, generator=g, dtype=F64)
    A = _dense(nl, n)                                         # without the input's self-loops
    # GraphSAGE-mean (Hamilton et al. 2017), self included: mean over N(i) + {i}, then W
    sage = R.SAGEConv(fin, fout).double()
    with torch.no_grad():
        sage.weight.normal_(generator=g)
        sage.bias.uniform_(-1, 1, generator=g)
    Ah = A + torch.eye(n, dtype=F64)
    want = (Ah.t() @ x) / Ah.sum(0).view(-1, 1) @ sage.weight.detach() + sage.bias.detach()
    assert torch.allclose(sage(x, ei).detach(), want, rtol=1e-12, atol=1e-12)
    # GIN (Xu et al. 2019): h((1 + eps) x_i + sum_{j in N(i)} x_j)
    mlp = torch.nn.Sequential(torch.nn.Linear(fin, 6), torch.nn.ReLU(), torch.nn.Linear(6, fout)).double()
    want = mlp(1.3 * x + A.t() @ x).detach()
    assert torch.allclose(R.GINConv(mlp, eps=0.3)(x, ei).detach(), want, rtol=1e-12, atol=1e-12)
    # GraphConv (Morris et al. 2019): Theta_1 x_i + sum_j e_ji Theta_2 x_j  (input loops kept)
    w = torch.rand(ei.size(1), generator=g, dtype=F64)
    gc = R.GraphConv(fin, fout).double()
    with torch.no_grad():
        gc.weight.normal_(generator=g)
    want = _dense(ei, n, w).t() @ (x @ gc.weight.detach()) + gc.lin(x).detach()
    assert torch.allclose(gc(x, ei, w).detach(), want, rtol=1e-12, atol=1e-12)
    # AGNN (Thekumparampil et al. 2018): softmax_j(beta cos(x_i, x_j)) over N(i) + {i}
    beta = 1.75          # exactly representable: the module holds beta in fp32 before .double()
    xn = x / x.norm(dim=1, keepdim=True)
    want = torch.zeros(n, fin, dtype=F64)
    for i in range(n):
        nb = [int(s) for s, d in nl.t().tolist() if d == i] + [i]
        a = torch.softmax(torch.stack([beta * (xn[i] * xn[j]).sum() for j in nb]), 0)
        want[i] = sum(al * x[j] for al, j in zip(a, nb))
    assert torch.allclose(R.AGNNConv(beta).double()(x, ei).detach(), want, rtol=1e-12, atol=1e-12)


def test_knn_radius_fps_oracles_against_a_kd_tree_and_the_greedy_definition():
    from scipy.spatial import cKDTree
    from oracle import ref_next as R
    g = torch.Generator().manual_seed(11)
    sizes = [40, 25, 33]
    x = torch.rand(sum(sizes), 3, generator=g, dtype=F64)
    batch = torch.cat([torch.full((m,), b) for b, m in enumerate(sizes)])
    k, r = 5, 0.3
    knn = R.knn_graph(x, k, batch)
    rad = R.radius(x, x, r, batch, batch, max_num_neighbors=10 ** 6)
    samp = R.fps(x, batch, ratio=0.25)
    off = 0
    picked = 0
    for b, m in enumerate(sizes):
        pts = x[off:off + m].numpy()
        tree = cKDTree(pts)
        _, nb = tree.query(pts, k=k + 1)                      # the point itself comes first
        got = knn[0][knn[1] >= off][:m * k].view(m, k) - off if b == 0 else \
            knn[0][(knn[1] >= off) & (knn[1] < off + m)].view(m, k) - off
        assert np.array_equal(got.numpy(), nb[:, 1:])         # same neighbours, same distance order
        balls = tree.query_ball_point(pts, r)
        for q in range(m):
            mine = rad[1][rad[0] == off + q] - off
            assert mine.tolist() == sorted(balls[q])          # every point within r, ascending index
        # farthest point sampling: ceil(0.25 m) picks, each the arg max of the distance to the picks
        cnt = -(-m // 4)
        sel = (samp[picked:picked + cnt] - off).tolist()
        assert sel[0] == 0 and len(set(sel)) == cnt
        for t in range(1, cnt):
            dmin = ((pts[:, None] - pts[sel[:t]][None]) ** 2).sum(-1).min(1)
            assert sel[t] == int(dmin.argmax())
        picked += cnt
        off += m
    assert picked == samp.numel()
