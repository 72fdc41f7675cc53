"""GPU: parity at BASELINE.json's OWN shapes (VERDICT r01 "What's weak" #3).

cfg2 exactly (PubMed-shaped 19,717 nodes / 88,648 edges, 500 -> 512 -> 3 with x*sigmoid(x), the
fork's ``Net`` of examples/ConvexPruning.py:103-125), cfg4 exactly (32 x 1,024 points, k = 20,
MLP 6 -> 64 -> 64, README.md:35-49), cfg3 at 1/16 scale against the oracle with H = 8, C = 8,
K = 602 (ConvexPruning.py:209-221 / examples/ppi.py:22 shapes) plus a full-size fp64
recomputation of sampled rows on the Reddit-shaped graph itself, cfg5 at 1/16 scale end to
end against the oracle, and the SURVEY 8d "1 % self-loops + 1 % duplicate edges" variant of
every config.  The oracle is oracle/ref_cpu.py ("parity unpinned": a restatement of PyG 1.4.3 /
torch-scatter 2.0.4 — see DESIGN.md §0)."""
import copy

import pytest
import torch
import torch.nn.functional as F

from oracle import ref_cpu as O
from util import assert_close

pytestmark = pytest.mark.gpu


def _sync(conv, ref):
    ref.load_state_dict({k: v.detach().cpu().float() for k, v in conv.state_dict().items()})


class _PruningNet(torch.nn.Module):
    """examples/ConvexPruning.py:103-125 ``Net(datasetroot, width)`` with the conv class
    injected (product or oracle)."""

    def __init__(self, conv_cls, fin, width, ncls):
        super().__init__()
        dims = [fin] + list(width) + [ncls]
        self.layers = torch.nn.ModuleList(
            [conv_cls(dims[i], dims[i + 1], cached=True) for i in range(len(dims) - 1)])

    def forward(self, x, edge_index):
        for layer in self.layers[:-1]:
            x = layer(x, edge_index)
            x = x * torch.sigmoid(x)
        return F.log_softmax(self.layers[-1](x, edge_index), dim=1)


def _noisy(data, seed=7):
    from torch_geometric.datasets import synthetic
    return synthetic.add_noise_edges(data.edge_index, data.x.size(0), 0.01, 0.01, seed)


@pytest.mark.parametrize('noise', [False, True])
def test_cfg2_pruning_gcn_pubmed_shape_forward_backward(dev, noise):
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    from torch_geometric.transforms import NormalizeFeatures
    data = NormalizeFeatures()(synthetic.pubmed_like())
    assert data.x.shape == (19717, 500) and data.edge_index.size(1) == 88648
    ei = _noisy(data) if noise else data.edge_index
    torch.manual_seed(0)
    net = _PruningNet(GCNConv, 500, [512], 3).to(dev)
    ref = _PruningNet(O.GCNConv, 500, [512], 3)
    _sync(net, ref)
    out = net(data.x.to(dev), ei.to(dev))
    out_ref = ref(data.x, ei)
    # the cached edge_index' (self-loop handling) is bit-exact, the norm to fp32 rounding
    for a, b in zip(net.layers, ref.layers):
        assert torch.equal(a.cached_result[0].cpu(), b.cached_result[0])
        assert torch.allclose(a.cached_result[1].cpu(), b.cached_result[1], rtol=2e-6, atol=0)
    assert_close(out, out_ref, 1e-5, 'cfg2 log_softmax output')
    loss = F.nll_loss(out[data.train_mask.to(dev)], data.y[data.train_mask].to(dev))
    loss_ref = F.nll_loss(out_ref[data.train_mask], data.y[data.train_mask])
    assert abs(loss.item() - loss_ref.item()) <= 1e-5 * abs(loss_ref.item())
    loss.backward()
    loss_ref.backward()
    for (name, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
        # 60 training rows feed the gradient: sums of a few hundred terms, 1e-4 class
        assert_close(p.grad, q.grad, 1e-4, 'cfg2 grad ' + name)


@pytest.mark.parametrize('tie_stress', [False, True])
def test_cfg4_edge_conv_full_shape_bit_exact_argmax(dev, tie_stress):
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import EdgeConv
    data = synthetic.knn_clouds(32, 1024, 20, seed=3)
    ei = data.edge_index
    assert data.x.shape == (32768, 3) and ei.size(1) == 655360
    assert torch.equal(torch.bincount(ei[1], minlength=32768), torch.full((32768,), 20))
    torch.manual_seed(1)
    mlp = torch.nn.Sequential(torch.nn.Linear(6, 64), torch.nn.ReLU(), torch.nn.Linear(64, 64))
    if tie_stress:
        class Quant(torch.nn.Module):
            def forward(self, v):
                return (v * 4).round() / 4
        mlp = torch.nn.Sequential(mlp, Quant())
    ref = O.EdgeConv(copy.deepcopy(mlp))
    conv = EdgeConv(mlp)
    conv.nn.load_state_dict(ref.nn.state_dict())     # both constructors re-initialise `nn`
    conv = conv.to(dev)
    a = data.x.clone().requires_grad_(True)
    b = data.x.to(dev).requires_grad_(True)
    ra, rb = ref(a, ei), conv(b, ei.to(dev))
    if not tie_stress:          # (quantisation turns 1-ulp MLP differences into 0.25 steps)
        assert_close(rb, ra, 1e-5, 'cfg4 output')
    # bit-exact max / argmax: the SAME message tensor through both reductions
    msg = ref.message(a[ei[1]], a[ei[0]]).detach()
    rv, rarg = O.scatter_max(msg, ei[1], 0, None, 32768)
    gv, garg = ops.scatter_max(msg.to(dev), ei[1].to(dev), 32768)
    assert torch.equal(gv.cpu(), rv), 'max values differ'
    assert torch.equal(garg.cpu(), rarg), 'argmax differs'
    if not tie_stress:
        ra.sum().backward()
        rb.sum().backward()
        assert_close(b.grad, a.grad, 1e-4, 'cfg4 dx')


def test_cfg4_with_noise_edges_bit_exact(dev):
    """k-NN graph + 1 % self-loops + 1 % duplicates: rows of unequal length, ties by duplicates."""
    from torch_geometric import ops
    from torch_geometric.datasets import synthetic
    data = synthetic.knn_clouds(32, 1024, 20, seed=3)
    ei = synthetic.add_noise_edges(data.edge_index, 32768, 0.01, 0.01, 5)
    g = torch.Generator().manual_seed(2)
    msg = (torch.randn(ei.size(1), 64, generator=g) * 2).round() / 2
    rv, rarg = O.scatter_max(msg, ei[1], 0, None, 32768)
    gv, garg = ops.scatter_max(msg.to(dev), ei[1].to(dev), 32768)
    assert torch.equal(gv.cpu(), rv) and torch.equal(garg.cpu(), rarg)


@pytest.mark.parametrize('scale,backward,slope', [(16, False, 0.2), (64, True, 1.0), (64, True, 0.2)])
def test_cfg3_gat_reddit_shape_scaled_vs_oracle(dev, scale, backward, slope):
    """Reddit-shaped log-normal graph at 1/scale (same mean in-degree ~492), x [N, 602],
    GATConv(602, 8, heads=8): K = 602 exercises the unaligned-row tcgen05 path, H = 8 / C = 8
    the streaming fused kernels.  1/16 forward (7.2 M edges through the CPU oracle), 1/64 with
    the backward (and 1 % self-loops + 1 % duplicate edges).

    Backward at ``negative_slope=0.2``: leaky-relu has a kink at 0, and with 29 M (edge, head)
    logits a handful lie within the forward rounding error of it, so their derivative (1 vs 0.2)
    differs between ANY two fp32 implementations; an edge with a large attention weight then
    moves one row of d xw by ~1e-3 of the maximum (measured: 5 outlier rows, both backward
    kernel generations agree with each other to 1e-7 and differ from the oracle in the same
    rows — tools/debug_r2b.py).  So the strict element-wise comparison runs with slope 1.0 (no
    kink, everything else identical) and the slope-0.2 run must agree in all but 0.5 % of the
    rows of d x."""
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GATConv
    from util import elem_err
    n, pairs = 232965 // scale, 57307946 // scale
    data = synthetic.reddit_like(seed=2, num_nodes=n, num_pairs=pairs)
    ei = synthetic.add_noise_edges(data.edge_index, n, 0.01, 0.01, 3) if backward else data.edge_index
    torch.manual_seed(2)
    conv = GATConv(602, 8, heads=8, negative_slope=slope).to(dev)
    ref = O.GATConv(602, 8, heads=8, negative_slope=slope)
    with torch.no_grad():
        conv.bias.uniform_(-0.5, 0.5)
    _sync(conv, ref)
    a = data.x.clone().requires_grad_(backward)
    b = data.x.to(dev).requires_grad_(backward)
    if backward:
        ra, rb = ref(a, ei), conv(b, ei.to(dev))
    else:
        with torch.no_grad():
            ra, rb = ref(a, ei), conv(b, ei.to(dev))
    # X.W at K = 602 (3xTF32, split accumulator) feeds an exponential: 2e-5 relative with a
    # floor of 5e-6 of the largest entry
    assert_close(rb, ra, 2e-5, 'cfg3 forward 1/{}'.format(scale), floor=5e-6)
    if not backward:
        return
    g = torch.randn(ra.shape, generator=torch.Generator().manual_seed(3))
    ra.backward(g)
    rb.backward(g.to(dev))
    pairs_ = [(b.grad, a.grad, 'dx'), (conv.weight.grad, ref.weight.grad, 'dW'),
              (conv.att.grad, ref.att.grad, 'datt'), (conv.bias.grad, ref.bias.grad, 'dbias')]
    if slope == 1.0:
        for got, want, what in pairs_:
            assert_close(got, want, 2e-4, 'cfg3 ' + what)
    else:
        assert_close(conv.bias.grad, ref.bias.grad, 2e-4, 'cfg3 dbias')
        got, want = b.grad.detach().cpu().double(), a.grad.double()
        bound = 2e-4 * want.abs() + 2e-5 * want.abs().max()
        bad_rows = int(((got - want).abs() > bound).any(dim=1).sum())
        # measured: 5 of 3,640 rows (each one a node with an incident kink edge)
        assert bad_rows <= max(n // 200, 1), 'cfg3 dx: {} of {} rows outside the bound'.format(bad_rows, n)
        # the parameter gradients are sums over all rows: the few kink rows move them by ~1e-3
        for got, want, what in pairs_[1:3]:
            assert_close(got, want, 2e-3, 'cfg3 ' + what + ' (kink-limited)', floor=2e-3)


def test_cfg3_gat_full_size_sampled_rows_fp64(dev):
    """The Reddit-shaped graph itself (232,965 nodes, 114.6 M edges), GATConv(602, 8, heads=8)
    forward on the fused kernels; ~500 sampled destination rows plus the 8 highest-degree rows
    are recomputed in fp64 straight from the edge list (A.10: leaky-relu logits, max-subtracted
    softmax with +1e-16, weighted sum, bias)."""
    if torch.cuda.get_device_properties(dev).total_memory < 60 * 2 ** 30:
        pytest.skip('needs a large-memory GPU')
    from torch_geometric.datasets import synthetic
    from torch_geometric.graph import clear_caches
    from torch_geometric.nn import GATConv
    data = synthetic.reddit_like(seed=2, device=dev)
    n = data.x.size(0)
    assert n == 232965 and data.edge_index.size(1) == 114615892
    H, C = 8, 8
    torch.manual_seed(4)
    conv = GATConv(602, C, heads=H).to(dev)
    with torch.no_grad():
        conv.bias.uniform_(-0.5, 0.5)
        out = conv(data.x, data.edge_index)
        again = conv(data.x, data.edge_index)
    assert torch.equal(out, again), 'not bit-identical from run to run'
    ei = conv._with_self_loops(data.edge_index, n)          # remove + add self loops (A.10)
    assert ei.size(1) == data.edge_index.size(1) + n
    indeg = torch.bincount(ei[1], minlength=n)
    gen = torch.Generator(device=dev).manual_seed(1)
    rows = torch.unique(torch.cat([torch.topk(indeg, 8)[1],
                                   torch.randint(0, n, (500,), device=dev, generator=gen)]))
    pick = torch.zeros(n, dtype=torch.bool, device=dev)
    pick[rows] = True
    e = pick[ei[1]].nonzero().view(-1)
    pos = torch.full((n,), -1, dtype=torch.long, device=dev)
    pos[rows] = torch.arange(rows.numel(), device=dev)
    src, dst = ei[0][e], pos[ei[1][e]]
    with torch.no_grad():
        xw = (data.x.double() @ conv.weight.double()).view(n, H, C)
        att = conv.att.double()
        a_src = (xw * att[:, :, C:]).sum(-1)
        a_dst = (xw * att[:, :, :C]).sum(-1)
        logit = F.leaky_relu(a_dst[rows][dst] + a_src[src], 0.2)
        R = rows.numel()
        idx = dst.view(-1, 1).expand(-1, H)
        mx = torch.full((R, H), -1e300, dtype=torch.float64, device=dev).scatter_reduce_(
            0, idx, logit, 'amax')
        ex = (logit - mx[dst]).exp()
        den = torch.zeros(R, H, dtype=torch.float64, device=dev).index_add_(0, dst, ex) + 1e-16
        alpha = ex / den[dst]
        ref = torch.zeros(R, H, C, dtype=torch.float64, device=dev)
        ref.index_add_(0, dst, alpha.unsqueeze(-1) * xw[src])
        ref = ref.view(R, H * C) + conv.bias.double()
    assert_close(out[rows], ref, 2e-5, 'cfg3 full size, sampled rows vs fp64')
    del data, out, again, xw
    clear_caches()
    torch.cuda.empty_cache()


@pytest.mark.parametrize('noise', [False, True])
def test_cfg5_products_small_end_to_end_vs_oracle(dev, noise):
    """cfg5's model (3-layer GCNConv 100 -> 256 -> 256 -> 47, relu, log_softmax) on the 1/16-scale
    products-shaped graph (153 K nodes, 3.9 M edges) against the oracle, forward + backward."""
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    n, pairs = 2449029 // 16, 31000000 // 16
    data = synthetic.products_like(seed=4, num_nodes=n, num_pairs=pairs, num_features=100)
    ei = synthetic.add_noise_edges(data.edge_index, n, 0.01, 0.01, 9) if noise else data.edge_index
    torch.manual_seed(0)
    convs = [GCNConv(100, 256, cached=True), GCNConv(256, 256, cached=True),
             GCNConv(256, 47, cached=True)]
    refs = [O.GCNConv(100, 256, cached=True), O.GCNConv(256, 256, cached=True),
            O.GCNConv(256, 47, cached=True)]
    for c, r in zip(convs, refs):
        c.to(dev)
        _sync(c, r)

    def fwd(layers, x, e):
        h = x
        for k, c in enumerate(layers):
            h = c(h, e)
            if k < 2:
                h = F.relu(h)
        return F.log_softmax(h, dim=1)

    out = fwd(convs, data.x.to(dev), ei.to(dev))
    # oracle forward keeping every layer's input and pre-activation output
    acts, pre = [data.x], []
    h = data.x
    for k, r in enumerate(refs):
        h = r(h, ei)
        h.retain_grad()
        pre.append(h)
        h = F.relu(h) if k < 2 else h
        if k < 2:
            acts.append(h.detach())
    out_ref = F.log_softmax(h, dim=1)
    assert torch.equal(convs[0].cached_result[0].cpu(), refs[0].cached_result[0])
    assert_close(out, out_ref, 1e-5, 'cfg5 (1/16) output')
    y = data.y[:n]
    F.nll_loss(out_ref, y).backward()
    # Backward, layer by layer: the same input and the same upstream gradient (the oracle's)
    # through both implementations.  A chained end-to-end comparison is ill-posed here: the
    # model has ReLUs, 39 M pre-activations per layer, and ~1e-5 of them lie within the
    # forward rounding error of zero, so their masks differ between ANY two fp32
    # implementations (measured: d pre-activation differs by 0.3 of its max in a few hundred
    # elements, dW0 by 4e-4, while dW0 recomputed in fp64 from the kernels' own upstream
    # gradient agrees to 6e-6 — tools/debug_r2.py).  cfg2 (x*sigmoid(x), smooth) is compared
    # end to end in test_cfg2_*.
    eid = ei.to(dev)
    for k, (c, r) in enumerate(zip(convs, refs)):
        xin = acts[k].to(dev).requires_grad_(k > 0)
        c.zero_grad()
        ho = c(xin, eid)
        assert_close(ho, pre[k], 1e-5, 'cfg5 layer {} forward from the oracle input'.format(k))
        ho.backward(pre[k].grad.to(dev))
        assert_close(c.weight.grad, r.weight.grad, 1e-4, 'cfg5 dW{}'.format(k))
        assert_close(c.bias.grad, r.bias.grad, 1e-4, 'cfg5 db{}'.format(k))
        if k > 0:
            # oracle d(input of layer k) = d(relu output of layer k-1): pre[k-1].grad / mask
            mask = (pre[k - 1] > 0).detach()
            assert_close(xin.grad.cpu() * mask, pre[k - 1].grad, 1e-4, 'cfg5 dx into layer {}'.format(k))


def test_cfg1_cora_shape_with_noise_edges(dev):
    """cfg1 (examples/gcn.py model) on the Cora-shaped graph with 1 % loops + 1 % duplicates;
    K = 1433 goes through the padded-row tcgen05 path."""
    from torch_geometric.datasets import synthetic
    from torch_geometric.nn import GCNConv
    from torch_geometric.transforms import NormalizeFeatures
    data = NormalizeFeatures()(synthetic.cora_like())
    ei = _noisy(data, 11)
    torch.manual_seed(0)
    c1, c2 = GCNConv(1433, 16, cached=True).to(dev), GCNConv(16, 7, cached=True).to(dev)
    r1, r2 = O.GCNConv(1433, 16, cached=True), O.GCNConv(16, 7, cached=True)
    _sync(c1, r1)
    _sync(c2, r2)
    eid = ei.to(dev)
    out = F.log_softmax(c2(F.relu(c1(data.x.to(dev), eid)), eid), dim=1)
    h1 = r1(data.x, ei)
    h1.retain_grad()
    a1 = F.relu(h1)
    h2 = r2(a1, ei)
    h2.retain_grad()
    out_ref = F.log_softmax(h2, dim=1)
    assert torch.equal(c1.cached_result[0].cpu(), r1.cached_result[0])
    assert_close(out, out_ref, 1e-5, 'cfg1 output')
    F.nll_loss(out_ref[data.train_mask], data.y[data.train_mask]).backward()
    # per layer, from the oracle's activations and upstream gradients (ReLU masks of two fp32
    # forwards differ in elements that are zero up to rounding — see the cfg5 test)
    c1.zero_grad()
    c2.zero_grad()
    g1 = c1(data.x.to(dev), eid)
    g1.backward(h1.grad.to(dev))
    assert_close(c1.weight.grad, r1.weight.grad, 1e-4, 'cfg1 dW1')
    assert_close(c1.bias.grad, r1.bias.grad, 1e-4, 'cfg1 db1')
    xin = a1.detach().to(dev).requires_grad_(True)
    g2 = c2(xin, eid)
    g2.backward(h2.grad.to(dev))
    assert_close(c2.weight.grad, r2.weight.grad, 1e-4, 'cfg1 dW2')
    assert_close(xin.grad.cpu() * (h1 > 0), h1.grad, 1e-4, 'cfg1 dx into layer 2')
