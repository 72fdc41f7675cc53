"""GPU: K1/K3/K2 and the nn modules against the oracle, through the C ABI."""
import copy

import pytest
import torch

from oracle import ref_cpu as O
from util import assert_close, random_graph, skewed_graph

pytestmark = pytest.mark.gpu
T = torch.tensor

WIDTHS = [1, 3, 4, 7, 16, 47, 64, 100, 128, 256, 300, 512, 700]


def _feat(n, F, seed, dtype=torch.float32):
    return torch.randn(n, F, generator=torch.Generator().manual_seed(seed)).to(dtype)


@pytest.mark.parametrize('F', WIDTHS)
@pytest.mark.parametrize('name', ['add', 'mean'])
def test_scatter_sum_mean(dev, F, name):
    from torch_geometric.utils import scatter_
    n, E = 700, 9000
    idx = skewed_graph(n, E, F)[1]
    src = _feat(E, F, 1)
    ref = O.scatter_(name, src, idx, 0, n)
    got = scatter_(name, src.to(dev), idx.to(dev), 0, n)
    assert_close(got, ref, 1e-5, name)
    again = scatter_(name, src.to(dev), idx.to(dev), 0, n)
    assert torch.equal(got, again), 'not deterministic'


@pytest.mark.parametrize('F', WIDTHS)
def test_scatter_max_bit_exact_with_ties(dev, F):
    from torch_geometric import ops
    from torch_geometric.utils import scatter_
    n, E = 700, 9000
    idx = skewed_graph(n, E, F + 1)[1]
    src = (_feat(E, F, 2) * 2).round() / 4  # quantised: ties are frequent
    src[::97] = -20000.0
    src[5] = float('nan')
    rv, ra = O.scatter_max(src, idx, 0, None, n)
    gv, ga = ops.scatter_max(src.to(dev), idx.to(dev), n)
    assert torch.equal(gv.cpu(), rv) and torch.equal(ga.cpu(), ra)
    assert torch.equal(scatter_('max', src.to(dev), idx.to(dev), 0, n).cpu(),
                       O.scatter_('max', src, idx, 0, n))
    assert torch.equal(scatter_('min', src.to(dev), idx.to(dev), 0, n).cpu().nan_to_num(),
                       O.scatter_('min', src, idx, 0, n).nan_to_num())


def test_scatter_edge_cases(dev):
    from torch_geometric.utils import scatter_
    # empty input
    out = scatter_('add', torch.zeros(0, 5).to(dev), torch.zeros(0, dtype=torch.long).to(dev), 0, 4)
    assert out.shape == (4, 5) and out.abs().sum().item() == 0
    out = scatter_('max', torch.zeros(0, 5).to(dev), torch.zeros(0, dtype=torch.long).to(dev), 0, 4)
    assert out.shape == (4, 5) and out.abs().sum().item() == 0
    # 1-D src, inferred dim_size, 3-D src, dim=1
    src = _feat(50, 1, 0).view(-1)
    idx = torch.randint(0, 9, (50,), generator=torch.Generator().manual_seed(0))
    assert_close(scatter_('add', src.to(dev), idx.to(dev)), O.scatter_('add', src, idx))
    src3 = _feat(50, 12, 1).view(50, 3, 4)
    assert_close(scatter_('mean', src3.to(dev), idx.to(dev), 0, 9), O.scatter_('mean', src3, idx, 0, 9))
    src_t = _feat(6, 50, 2)
    assert_close(scatter_('add', src_t.to(dev), idx.to(dev), 1, 9), O.scatter_('add', src_t, idx, 1, 9))
    assert_close(scatter_('max', src_t.to(dev), idx.to(dev), -1, 9), O.scatter_('max', src_t, idx, 1, 9))
    # one giant row
    idx1 = torch.zeros(40000, dtype=torch.long)
    big = _feat(40000, 8, 3)
    assert_close(scatter_('add', big.to(dev), idx1.to(dev), 0, 2), O.scatter_('add', big, idx1, 0, 2), 2e-5)
    assert torch.equal(scatter_('max', big.to(dev), idx1.to(dev), 0, 2).cpu(), O.scatter_('max', big, idx1, 0, 2))


def test_scatter_backward(dev):
    from torch_geometric.utils import scatter_
    n, E, F = 300, 4000, 24
    idx = skewed_graph(n, E, 3)[1]
    for name in ['add', 'mean', 'max', 'min']:
        src = (_feat(E, F, 4) * 2).round() / 4
        a = src.clone().requires_grad_(True)
        b = src.to(dev).requires_grad_(True)
        g = _feat(n, F, 5)
        O.scatter_(name, a, idx, 0, n).backward(g)
        scatter_(name, b, idx.to(dev), 0, n).backward(g.to(dev))
        assert_close(b.grad, a.grad, 1e-5, name)


def test_index_select_and_backward(dev):
    from torch_geometric import ops
    x = _feat(200, 37, 0)
    idx = torch.randint(0, 200, (5000,), generator=torch.Generator().manual_seed(1))
    a = x.clone().requires_grad_(True)
    b = x.to(dev).requires_grad_(True)
    ra = torch.index_select(a, 0, idx)
    rb = ops.index_select0(b, idx.to(dev))
    assert torch.equal(rb.cpu(), ra.detach())
    g = _feat(5000, 37, 2)
    ra.backward(g)
    rb.backward(g.to(dev))
    assert_close(b.grad, a.grad, 1e-5)


@pytest.mark.parametrize('H', [1, 8, 40])
def test_softmax(dev, H):
    from torch_geometric.utils import softmax
    n, E = 400, 6000
    idx = skewed_graph(n, E, H)[1]
    src = _feat(E, H, 7) * 3
    if H == 1:
        src = src.view(-1)
    a = src.clone().requires_grad_(True)
    b = src.to(dev).requires_grad_(True)
    ra = O.softmax(a, idx, n)
    rb = softmax(b, idx.to(dev), n)
    assert_close(rb, ra, 1e-5)
    g = torch.randn(ra.shape, generator=torch.Generator().manual_seed(8))
    ra.backward(g)
    rb.backward(g.to(dev))
    assert_close(b.grad, a.grad, 2e-5)


# ---------------------------------------------------------------- modules


def _sync(conv, ref):
    ref.load_state_dict({k: v.detach().cpu() for k, v in conv.state_dict().items()})


@pytest.mark.parametrize('fin,fout', [(33, 16), (64, 47), (100, 256), (50, 7)])
@pytest.mark.parametrize('improved', [False, True])
def test_gcn_conv_forward_backward(dev, fin, fout, improved):
    from torch_geometric.nn import GCNConv
    n, E = 1500, 20000
    ei = skewed_graph(n, E, fin)
    x = _feat(n, fin, 1)
    conv = GCNConv(fin, fout, improved=improved, cached=True).to(dev)
    ref = O.GCNConv(fin, fout, improved=improved, cached=True)
    with torch.no_grad():
        conv.bias.uniform_(-1, 1)
    _sync(conv, ref)
    a = x.clone().requires_grad_(True)
    b = x.to(dev).requires_grad_(True)
    ra, rb = ref(a, ei), conv(b, ei.to(dev))
    assert_close(rb, ra, 1e-5, 'fwd')
    assert torch.equal(conv.cached_result[0].cpu(), ref.cached_result[0])
    assert torch.allclose(conv.cached_result[1].cpu(), ref.cached_result[1], rtol=2e-6, atol=0)
    g = _feat(n, fout, 2)
    ra.backward(g)
    rb.backward(g.to(dev))
    assert_close(b.grad, a.grad, 2e-5, 'dx')
    assert_close(conv.weight.grad, ref.weight.grad, 2e-5, 'dW')
    assert_close(conv.bias.grad, ref.bias.grad, 2e-5, 'db')
    # second call hits the cache; a different edge count must raise like the reference
    assert_close(conv(b, ei.to(dev)), ra, 1e-5)
    with pytest.raises(RuntimeError, match='Cached'):
        conv(b, ei[:, :-1].to(dev))


def test_gcn_conv_edge_weight_grad_and_uncached(dev):
    from torch_geometric.nn import GCNConv
    n, E = 300, 3000
    ei = random_graph(n, n, E, 0)
    x = _feat(n, 20, 1)
    w = torch.rand(E, generator=torch.Generator().manual_seed(2)) + 0.5
    conv = GCNConv(20, 12, normalize=False).to(dev)
    ref = O.GCNConv(20, 12, normalize=False)
    _sync(conv, ref)
    wa = w.clone().requires_grad_(True)
    wb = w.to(dev).requires_grad_(True)
    ra, rb = ref(x, ei, wa), conv(x.to(dev), ei.to(dev), wb)
    assert_close(rb, ra, 1e-5)
    ra.sum().backward()
    rb.sum().backward()
    assert_close(wb.grad, wa.grad, 2e-5)


def test_gcn_bf16(dev):
    from torch_geometric.nn import GCNConv
    n, E = 800, 10000
    ei = skewed_graph(n, E, 3)
    x = _feat(n, 64, 1)
    conv = GCNConv(64, 128).to(dev)
    ref = O.GCNConv(64, 128)
    _sync(conv, ref)
    out = conv.to(torch.bfloat16)(x.to(dev).to(torch.bfloat16), ei.to(dev))
    assert out.dtype == torch.bfloat16
    # inputs, weights and the result are rounded to bf16 (2^-8 relative each): an element that
    # is a cancelling sum inherits that error relative to its largest addend -> absolute floor
    # of one bf16 ulp of the largest entry
    assert_close(out.float(), ref(x, ei), 1e-2, floor=4e-3)


@pytest.mark.parametrize('H,C,concat', [(8, 8, True), (1, 16, True), (4, 7, False), (6, 121, False),
                                        (2, 64, True), (8, 8, False), (4, 128, True), (3, 12, True)])
def test_gat_conv_forward_backward(dev, H, C, concat):
    from torch_geometric.nn import GATConv
    n, E, fin = 900, 12000, 40
    ei = skewed_graph(n, E, H * C)
    ei[1, :50] = ei[0, :50]
    x = _feat(n, fin, 1)
    conv = GATConv(fin, C, heads=H, concat=concat).to(dev)
    ref = O.GATConv(fin, C, heads=H, concat=concat)
    with torch.no_grad():
        conv.bias.uniform_(-1, 1)
    _sync(conv, ref)
    a = x.clone().requires_grad_(True)
    b = x.to(dev).requires_grad_(True)
    ra, rb = ref(a, ei), conv(b, ei.to(dev))
    assert_close(rb, ra, 2e-5, 'fwd')
    g = torch.randn(ra.shape, generator=torch.Generator().manual_seed(3))
    ra.backward(g)
    rb.backward(g.to(dev))
    assert_close(b.grad, a.grad, 1e-4, 'dx')
    assert_close(conv.weight.grad, ref.weight.grad, 1e-4, 'dW')
    assert_close(conv.att.grad, ref.att.grad, 1e-4, 'datt')
    assert_close(conv.bias.grad, ref.bias.grad, 1e-4, 'db')


def test_gat_dropout_takes_generic_route(dev):
    from torch_geometric.nn import GATConv
    n, E = 200, 2000
    ei = random_graph(n, n, E, 0)
    x = _feat(n, 16, 1)
    conv = GATConv(16, 8, heads=2, dropout=0.5).to(dev)
    ref = O.GATConv(16, 8, heads=2, dropout=0.5)
    _sync(conv, ref)
    conv.eval(), ref.eval()
    assert_close(conv(x.to(dev), ei.to(dev)), ref(x, ei), 2e-5)
    conv.train()
    out = conv(x.to(dev).requires_grad_(True), ei.to(dev))  # random mask: shape/grad only
    out.sum().backward()
    assert out.shape == (n, 16) and conv.weight.grad is not None


def test_gat_bipartite(dev):
    from torch_geometric.nn import GATConv
    ns, nd, E = 300, 120, 4000
    ei = random_graph(ns, nd, E, 0)
    xs, xd = _feat(ns, 24, 1), _feat(nd, 24, 2)
    conv = GATConv(24, 8, heads=3).to(dev)
    ref = O.GATConv(24, 8, heads=3)
    _sync(conv, ref)
    for pair in [(xs, xd), (xs, None)]:
        pd = tuple(None if t is None else t.to(dev) for t in pair)
        assert_close(conv(pd, ei.to(dev), size=(ns, nd)), ref(pair, ei, size=(ns, nd)), 2e-5)


@pytest.mark.parametrize('tie_stress', [False, True])
def test_edge_conv_bit_exact_argmax(dev, tie_stress):
    from torch_geometric.nn import EdgeConv
    from torch_geometric.datasets import synthetic
    data = synthetic.knn_clouds(4, 256, 20, seed=3)
    mlp = torch.nn.Sequential(torch.nn.Linear(6, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64))
    if tie_stress:
        class Quant(torch.nn.Module):
            def forward(self, v):
                return (v * 4).round() / 4
        mlp = torch.nn.Sequential(mlp, Quant())
    ref = O.EdgeConv(copy.deepcopy(mlp))
    conv = EdgeConv(mlp)
    conv.nn.load_state_dict(ref.nn.state_dict())
    conv = conv.to(dev)
    a = data.x.clone().requires_grad_(True)
    b = data.x.to(dev).requires_grad_(True)
    ra, rb = ref(a, data.edge_index), conv(b, data.edge_index.to(dev))
    assert_close(rb, ra, 1e-5)
    # argmax parity on identical messages: feed the CPU messages to the GPU reduction
    from torch_geometric import ops
    msg = ref.message(a[data.edge_index[1]], a[data.edge_index[0]]).detach()
    rv, rarg = O.scatter_max(msg, data.edge_index[1], 0, None, a.size(0))
    gv, garg = ops.scatter_max(msg.to(dev), data.edge_index[1].to(dev), a.size(0))
    assert torch.equal(gv.cpu(), rv) and torch.equal(garg.cpu(), rarg)
    ra.sum().backward()
    rb.sum().backward()
    assert_close(b.grad, a.grad, 1e-4)


def test_default_message_fused_routes(dev):
    from torch_geometric.nn import MessagePassing
    n, E = 400, 5000
    ei = skewed_graph(n, E, 1)
    x = (_feat(n, 19, 1) * 2).round() / 4
    for aggr in ['add', 'mean', 'max']:
        for flow in ['source_to_target', 'target_to_source']:
            a = x.clone().requires_grad_(True)
            b = x.to(dev).requires_grad_(True)
            ra = O.MessagePassing(aggr=aggr, flow=flow).propagate(ei, x=a)
            rb = MessagePassing(aggr=aggr, flow=flow).propagate(ei.to(dev), x=b)
            if aggr == 'max':
                assert torch.equal(rb.cpu(), ra.detach())
            else:
                assert_close(rb, ra, 1e-5)
            ra.sum().backward()
            rb.sum().backward()
            assert_close(b.grad, a.grad, 1e-5, aggr + flow)


def test_user_defined_message_generic_route(dev):
    from torch_geometric.nn import MessagePassing

    def make(base):
        class Conv(base):
            def __init__(self):
                super(Conv, self).__init__(aggr='mean')

            def message(self, x_i, x_j, edge_index_i, size_i, scale):
                assert edge_index_i.numel() == x_j.size(0)
                return scale * (x_j - x_i)

            def update(self, aggr_out, x):
                return aggr_out + x
        return Conv()

    n, E = 300, 4000
    ei = random_graph(n, n, E, 0)
    x = _feat(n, 10, 1)
    ra = make(O.MessagePassing).propagate(ei, x=x, scale=0.5)
    rb = make(MessagePassing).propagate(ei.to(dev), x=x.to(dev), scale=0.5)
    assert_close(rb, ra, 1e-5)
    with pytest.raises(ValueError, match='same size'):
        make(MessagePassing).propagate(ei.to(dev), x=(x.to(dev), x[:10].to(dev)), size=(n, n),
                                       scale=1.0)


def test_two_layer_gcn_training_step_matches_oracle(dev):
    """cfg 1 shape: examples/gcn.py's Net on the Cora-shaped synthetic graph."""
    import torch.nn.functional as F
    from torch_geometric.datasets import Planetoid
    import torch_geometric.transforms as Tr
    from torch_geometric.nn import GCNConv
    ds = Planetoid('/tmp/unused', 'Cora', Tr.NormalizeFeatures())
    data = ds[0]
    c1, c2 = GCNConv(ds.num_features, 16, cached=True), GCNConv(16, ds.num_classes, cached=True)
    r1, r2 = O.GCNConv(ds.num_features, 16, cached=True), O.GCNConv(16, ds.num_classes, cached=True)
    _sync(c1, r1), _sync(c2, r2)
    c1, c2 = c1.to(dev), c2.to(dev)
    dd = data.to(dev)

    def fwd(l1, l2, d):
        h = F.relu(l1(d.x, d.edge_index))
        return F.log_softmax(l2(h, d.edge_index), dim=1)
    out, ref = fwd(c1, c2, dd), fwd(r1, r2, ds[0])
    assert_close(out, ref, 1e-5)
    F.nll_loss(out[dd.train_mask], dd.y[dd.train_mask]).backward()
    d0 = ds[0]
    F.nll_loss(ref[d0.train_mask], d0.y[d0.train_mask]).backward()
    assert_close(c1.weight.grad, r1.weight.grad, 1e-4)
    assert_close(c2.weight.grad, r2.weight.grad, 1e-4)
