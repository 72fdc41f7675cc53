"""GPU: a forward pass captured into a CUDA graph (torch_geometric.cuda_graph) replays
to the same bits as the eager launches, also on new input values."""
import pytest
import torch
import torch.nn.functional as F

from util import random_graph

pytestmark = pytest.mark.gpu


def test_gcn_forward_graph_replay(dev):
    from torch_geometric.cuda_graph import CapturedForward
    from torch_geometric.nn import GCNConv
    n, E = 3000, 20000
    ei = random_graph(n, n, E, 1).to(dev)
    g = torch.Generator().manual_seed(0)
    x1 = torch.randn(n, 500, generator=g).to(dev)
    x2 = torch.randn(n, 500, generator=g).to(dev)
    torch.manual_seed(0)
    c1, c2 = GCNConv(500, 64, cached=True).to(dev), GCNConv(64, 3, cached=True).to(dev)

    def net(x):
        h = c1(x, ei)
        h = h * torch.sigmoid(h)
        return F.log_softmax(c2(h, ei), dim=1)

    with torch.no_grad():
        ref1, ref2 = net(x1).clone(), net(x2).clone()
    cap = CapturedForward(net, x1)
    assert torch.equal(cap(x1), ref1)
    assert torch.equal(cap(x2), ref2)
    assert torch.equal(cap(x1), ref1)


def test_gat_and_edgeconv_graph_replay(dev):
    from torch_geometric.cuda_graph import CapturedForward
    from torch_geometric.nn import GATConv, EdgeConv
    n, E = 2000, 30000
    ei = random_graph(n, n, E, 2, loops=False).to(dev)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, 32, generator=g).to(dev)
    torch.manual_seed(0)
    gat = GATConv(32, 8, heads=4).to(dev)
    mlp = torch.nn.Sequential(torch.nn.Linear(64, 32), torch.nn.ReLU(), torch.nn.Linear(32, 16))
    ec = EdgeConv(mlp, aggr='max').to(dev)

    def net(x):
        return ec(gat(x, ei), ei)

    with torch.no_grad():
        ref = net(x).clone()
        ref_b = net(x * 0.5).clone()
    cap = CapturedForward(net, x)
    assert torch.equal(cap(x), ref)
    assert torch.equal(cap(x * 0.5), ref_b)
