"""CPU: host-side components next to the hot path against
their loop-level restatement in oracle/ref_next.py — ``Batch.from_data_list``
(block-diagonal stacking, the step before propagate for mini-batches of graphs),
``DataLoader`` / ``DataListLoader``, ``NormalizeFeatures`` and ``knn_graph``.
These are index arithmetic on the host / framework ops; no CUDA kernel is involved."""
import torch

from oracle import ref_next as R
from util import random_graph


def _graphs(seed, count=5):
    from torch_geometric.data import Data
    g = torch.Generator().manual_seed(seed)
    out, raw = [], []
    for i in range(count):
        n = int(torch.randint(1, 30, (1,), generator=g))
        e = int(torch.randint(0, 80, (1,), generator=g))
        ei = random_graph(n, n, e, seed + i) if e else torch.zeros(2, 0, dtype=torch.long)
        x = torch.randn(n, 6, generator=g)
        y = torch.randint(0, 3, (1,), generator=g)
        out.append(Data(x=x, edge_index=ei, y=y))
        raw.append({'x': x, 'edge_index': ei, 'y': y})
    return out, raw


def test_batch_from_data_list_is_block_diagonal():
    from torch_geometric.data import Batch
    data_list, raw = _graphs(0)
    b = Batch.from_data_list(data_list)
    ref = R.batch_from_list(raw)
    assert torch.equal(b.x, ref['x'])
    assert torch.equal(b.edge_index, ref['edge_index'])
    assert torch.equal(b.batch, ref['batch'])
    assert torch.equal(b.y, ref['y'])
    assert b.num_graphs == len(data_list) and b.num_nodes == ref['x'].size(0)
    # no edge crosses two graphs
    assert torch.equal(b.batch[b.edge_index[0]], b.batch[b.edge_index[1]])


def test_loaders_yield_batches_and_lists():
    from torch_geometric.data import DataLoader, DataListLoader, Batch
    data_list, raw = _graphs(1, count=7)
    got = list(DataLoader(data_list, batch_size=3, shuffle=False))
    assert [b.num_graphs for b in got] == [3, 3, 1]
    assert torch.equal(got[1].edge_index, R.batch_from_list(raw[3:6])['edge_index'])
    lists = list(DataListLoader(data_list, batch_size=4, shuffle=False))
    assert [len(chunk) for chunk in lists] == [4, 3] and lists[0][0] is data_list[0]
    assert isinstance(got[0], Batch)


def test_normalize_features():
    from torch_geometric.data import Data
    from torch_geometric.transforms import NormalizeFeatures, Compose
    x = torch.tensor([[1., 3., 0.], [0., 0., 0.], [0.25, 0.25, 0.]])
    d = Compose([NormalizeFeatures()])(Data(x=x.clone()))
    want = torch.tensor([[0.25, 0.75, 0.], [0., 0., 0.], [0.25, 0.25, 0.]])   # divisor clamped at 1
    assert torch.equal(d.x, want)


def test_knn_graph_matches_brute_force():
    from torch_geometric.nn import knn_graph
    g = torch.Generator().manual_seed(4)
    x = torch.rand(120, 3, generator=g)
    batch = torch.cat([torch.zeros(50), torch.ones(70)]).long()
    for k, loop in ((4, False), (6, True)):
        ei = knn_graph(x, k, batch, loop=loop)
        ref = R.knn_graph(x, k, batch, loop=loop)
        assert ei.shape == ref.shape
        # same neighbour sets per centre (distances are distinct with probability one)
        key = lambda e: set(zip(e[1].tolist(), e[0].tolist()))
        assert key(ei) == key(ref)
        assert torch.equal(batch[ei[0]], batch[ei[1]])
    flipped = knn_graph(x, 4, batch, flow='target_to_source')
    assert torch.equal(flipped.flip(0), knn_graph(x, 4, batch))
    assert knn_graph(torch.zeros(0, 3), 4).shape == (2, 0)
