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
