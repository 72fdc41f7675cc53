"""GPU: K4 — integer graph build, bit-exact against the oracle."""
import pytest
import torch

from oracle import ref_cpu as O
from util import kat, random_graph, skewed_graph

pytestmark = pytest.mark.gpu
T = torch.tensor
K = kat()


def test_kats(dev):
    from torch_geometric import utils as U
    from torch_geometric.nn import GCNConv
    k = K['remove_self_loops']
    assert U.remove_self_loops(T(k['edge_index']).to(dev))[0].cpu().tolist() == k['out']
    k = K['add_self_loops']
    ei, w = U.add_self_loops(T(k['edge_index']).to(dev), T(k['w']).to(dev))
    assert ei.cpu().tolist() == k['out'] and w.cpu().tolist() == k['w_out']
    k = K['add_remaining_self_loops']
    ei, w = U.add_remaining_self_loops(T(k['edge_index']).to(dev), T(k['w']).to(dev))
    assert ei.cpu().tolist() == k['out'] and w.cpu().tolist() == k['w_out']
    k = K['degree']
    d = U.degree(T(k['index']).to(dev), dtype=torch.long)
    assert d.dtype == torch.long and d.cpu().tolist() == k['out']
    k = K['gcn_norm']
    ei, norm = GCNConv.norm(T(k['edge_index']).to(dev), k['num_nodes'], dtype=torch.float32)
    assert ei.cpu().tolist() == k['out']
    assert torch.allclose(norm.cpu(), T(k['norm']), rtol=1e-6, atol=0)
    k = K['softmax']
    assert torch.allclose(U.softmax(T(k['src']).to(dev), T(k['index']).to(dev)).cpu(), T(k['out']))
    k = K['scatter_max']
    from torch_geometric import ops
    out, arg = ops.scatter_max(T(k['src']).to(dev), T(k['index']).to(dev), k['dim_size'])
    assert out.cpu().tolist() == k['out'] and arg.cpu().tolist() == k['arg']
    assert U.scatter_('max', T(k['src']).to(dev), T(k['index']).to(dev), 0,
                      k['dim_size']).cpu().tolist() == k['scatter_']
    k = K['scatter_mean']
    assert U.scatter_('mean', T(k['src']).to(dev), T(k['index']).to(dev), 0,
                      k['dim_size']).cpu().tolist() == k['out']


@pytest.mark.parametrize('n,E,seed', [(1, 1, 0), (7, 0, 1), (50, 400, 2), (300, 5000, 3),
                                      (70000, 300000, 4), (5000, 200000, 5)])
def test_csr_build_is_a_stable_sort(dev, n, E, seed):
    from torch_geometric.graph import build_csr
    ei = random_graph(n, n, E, seed) if E else torch.zeros(2, 0, dtype=torch.long)
    if E > 1000:
        ei = skewed_graph(n, E, seed)
    csr = build_csr(ei[1].to(dev), ei[0].to(dev), n)
    key = ei[1]
    order = torch.sort(key, stable=True)[1] if E else torch.zeros(0, dtype=torch.long)
    counts = torch.bincount(key, minlength=n) if E else torch.zeros(n, dtype=torch.long)
    rowptr = torch.cat([torch.zeros(1, dtype=torch.long), counts.cumsum(0)])
    assert torch.equal(csr.rowptr.cpu().long(), rowptr)
    assert torch.equal(csr.perm.cpu().long(), order)
    assert torch.equal(csr.col.cpu().long(), ei[0][order])


def test_csr_build_rejects_out_of_range(dev):
    from torch_geometric.graph import build_csr
    with pytest.raises(IndexError):
        build_csr(T([0, 5, 1]).to(dev), None, 3)
    with pytest.raises(IndexError):
        build_csr(T([0, 1, 1]).to(dev), T([0, 9, 0]).to(dev), 3, 3)
    with pytest.raises(IndexError):
        build_csr(T([0, -1, 1]).to(dev), None, 3)


@pytest.mark.parametrize('n,E,seed', [(5, 0, 0), (40, 300, 1), (3000, 40000, 2),
                                      (20000, 150000, 3)])
@pytest.mark.parametrize('weighted', [False, True])
def test_loop_utils_bit_exact(dev, n, E, seed, weighted):
    from torch_geometric import utils as U
    ei = random_graph(n, n, E, seed) if E else torch.zeros(2, 0, dtype=torch.long)
    if E:  # make ~5% self loops, some nodes with several
        g = torch.Generator().manual_seed(seed)
        m = torch.rand(E, generator=g) < 0.05
        ei[1] = torch.where(m, ei[0], ei[1])
    w = torch.rand(E, generator=torch.Generator().manual_seed(seed + 9)) if weighted else None
    wd = None if w is None else w.to(dev)
    for fn in ('remove_self_loops', 'add_self_loops', 'add_remaining_self_loops'):
        if fn == 'remove_self_loops':
            ref = O.remove_self_loops(ei, w)
            got = U.remove_self_loops(ei.to(dev), wd)
        else:
            ref = getattr(O, fn)(ei, w, 2.0, n)
            got = getattr(U, fn)(ei.to(dev), wd, 2.0, n)
        assert torch.equal(got[0].cpu(), ref[0]), fn
        if weighted:
            assert torch.equal(got[1].cpu(), ref[1]), fn
        else:
            assert got[1] is None
    assert U.contains_self_loops(ei.to(dev)) == (O.contains_self_loops(ei) if E else False)


def test_remove_self_loops_with_matrix_attr(dev):
    from torch_geometric import utils as U
    ei = T([[0, 1, 2, 2], [0, 2, 2, 1]])
    attr = torch.arange(8.).view(4, 2)
    got = U.remove_self_loops(ei.to(dev), attr.to(dev))
    ref = O.remove_self_loops(ei, attr)
    assert torch.equal(got[0].cpu(), ref[0]) and torch.equal(got[1].cpu(), ref[1])


@pytest.mark.parametrize('improved', [False, True])
@pytest.mark.parametrize('weighted', [False, True])
def test_gcn_norm_matches_oracle(dev, improved, weighted):
    from torch_geometric.nn import GCNConv
    n, E = 2000, 30000
    ei = skewed_graph(n, E, 5)
    ei[1, :200] = ei[0, :200]  # pre-existing self loops
    w = torch.rand(E, generator=torch.Generator().manual_seed(1)) + 0.1 if weighted else None
    ref_ei, ref_norm = O.gcn_norm(ei, n, w, improved, torch.float32)
    got_ei, got_norm = GCNConv.norm(ei.to(dev), n, None if w is None else w.to(dev), improved,
                                    torch.float32)
    assert torch.equal(got_ei.cpu(), ref_ei)
    assert torch.allclose(got_norm.cpu(), ref_norm, rtol=2e-6, atol=0)


def test_degree_exact_and_dtypes(dev):
    from torch_geometric.utils import degree
    idx = skewed_graph(1000, 50000, 1)[1]
    for dt in (torch.long, torch.float32, torch.int32, torch.float64):
        got = degree(idx.to(dev), 1000, dtype=dt)
        ref = O.degree(idx, 1000, dtype=dt)
        assert got.dtype == dt and torch.equal(got.cpu(), ref)
    assert degree(idx.to(dev)).numel() == int(idx.max()) + 1


def test_weighted_degree_bit_exact(dev):
    """Sequential ascending-edge-order fp32 sum == CPU scatter_add_ order."""
    from torch_geometric import _native as nat
    from torch_geometric.graph import build_csr
    n, E = 500, 60000
    ei = skewed_graph(n, E, 2)
    w = torch.rand(E, generator=torch.Generator().manual_seed(3))
    ref = O.scatter_add(w, ei[0], 0, None, n)
    csr = build_csr(ei[0].to(dev), None, n)
    deg = torch.empty(n, device=dev)
    wd = w.to(dev)
    nat.call('pygb_weighted_degree', dev, nat.ptr(wd), nat.ptr(csr.rowptr), nat.ptr(csr.perm), n,
             nat.ptr(deg))
    assert torch.equal(deg.cpu(), ref)


def test_gcn_norm_in_slot_order_is_bit_identical(dev):
    """Adj.permuted rebuilds the unit-weight GCN norm from the degrees in slot order
    (pygb_gcn_norm_csr) — the same bits as gathering norm through the permutation, for both
    sorted graphs; a norm that was modified in place falls back to the gather."""
    from torch_geometric.nn import GCNConv
    from torch_geometric.graph import Adj
    from torch_geometric import _native as nat
    ei = skewed_graph(900, 12000, 21).to(dev)
    ei2, norm = GCNConv.norm(ei, 900)
    assert hasattr(norm, '_pygb_unit_deg')
    adj = Adj(ei2, 1, (900, 900))
    nat.event_log = []
    got_d, got_s = adj.permuted(norm, 'dst'), adj.permuted(norm, 'src')
    names = [e[0] for e in nat.event_log]
    nat.event_log = None
    assert names.count('pygb_gcn_norm_csr') == 2 and 'pygb_permute_f32' not in names
    assert torch.equal(got_d, norm[adj.dst.perm.long()])
    assert torch.equal(got_s, norm[adj.src.perm.long()])
    norm.mul_(2.0)                                   # stale tag: version changed
    adj2 = Adj(ei2, 1, (900, 900))
    assert torch.equal(adj2.permuted(norm, 'dst'), norm[adj2.dst.perm.long()])
    # weighted / improved norms carry no tag
    _, nw = GCNConv.norm(ei, 900, torch.rand(ei.size(1), device=dev))
    assert not hasattr(nw, '_pygb_unit_deg')
    _, ni = GCNConv.norm(ei, 900, None, True)
    assert not hasattr(ni, '_pygb_unit_deg')
