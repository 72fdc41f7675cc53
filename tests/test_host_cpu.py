"""CPU: host-side components next to the hot path against
their loop-level restatement in oracle/ref_next.py — ``Batch.from_data_list``
(block-diagonal stacking, the step before propagate for mini-batches of graphs),
``DataLoader`` / ``DataListLoader``, ``NormalizeFeatures`` and ``knn_graph``.
These are index arithmetic on the host / framework ops; no CUDA kernel is involved."""
import pytest
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


def test_planetoid_reads_real_raw_files_and_warns_on_synthetic(tmp_path, monkeypatch):
    """``Planetoid(root, name)`` must read ``<root>/raw/ind.<name>.*`` when the
    files exist, and must never hand out synthetic data silently."""
    import pickle
    import warnings
    import numpy as np
    import scipy.sparse as sp
    from collections import defaultdict
    from torch_geometric.datasets import Planetoid, SyntheticDatasetWarning

    raw = tmp_path / 'Cora' / 'raw'
    raw.mkdir(parents=True)
    rng = np.random.RandomState(0)
    n_all, n_test, f, c, n_train = 12, 4, 5, 3, 6
    allx = sp.csr_matrix(rng.rand(n_all, f).astype(np.float32))
    tx = sp.csr_matrix(rng.rand(n_test, f).astype(np.float32))
    onehot = lambda k: np.eye(c, dtype=np.int32)[rng.randint(0, c, k)]
    ally, ty = onehot(n_all), onehot(n_test)
    graph = defaultdict(list)
    for a, b in [(0, 1), (1, 2), (2, 2), (3, 15), (14, 13), (0, 1), (12, 5)]:   # loop + duplicate
        graph[a].append(b)
    files = {'x': allx[:n_train], 'tx': tx, 'allx': allx, 'y': ally[:n_train], 'ty': ty,
             'ally': ally, 'graph': graph}
    for k, v in files.items():
        with open(raw / ('ind.cora.' + k), 'wb') as fh:
            pickle.dump(v, fh)
    test_index = [14, 12, 15, 13]                       # unsorted, as in the real files
    (raw / 'ind.cora.test.index').write_text('\n'.join(map(str, test_index)))

    with warnings.catch_warnings():
        warnings.simplefilter('error')                  # the real path must not warn
        ds = Planetoid(str(tmp_path / 'Cora'), 'Cora')
    assert not ds.synthetic and 'synthetic' not in repr(ds)
    d = ds[0]
    assert d.x.shape == (16, f) and d.y.shape == (16,)
    # rows of the test block are permuted so that node test_index[k] carries tx[k]
    dense_tx = np.asarray(tx.todense())
    for k, node in enumerate(test_index):
        assert np.allclose(d.x[node].numpy(), dense_tx[k])
        assert int(d.y[node]) == int(ty[k].argmax())
    assert d.train_mask.nonzero().view(-1).tolist() == list(range(n_train))
    assert sorted(d.test_mask.nonzero().view(-1).tolist()) == sorted(test_index)
    ei = d.edge_index
    pairs = set(zip(ei[0].tolist(), ei[1].tolist()))
    assert pairs == {(0, 1), (1, 0), (1, 2), (2, 1), (3, 15), (15, 3), (13, 14), (14, 13),
                     (5, 12), (12, 5)}                  # symmetric, loop-free, duplicate-free
    assert ei.size(1) == len(pairs)
    code = ei[0] * 16 + ei[1]
    assert torch.equal(code, code.sort()[0])            # sorted by (row, col)

    with pytest.warns(SyntheticDatasetWarning, match='SYNTHETIC stand-in'):
        syn = Planetoid(str(tmp_path / 'nowhere'), 'Cora')
    assert syn.synthetic and 'synthetic stand-in' in repr(syn)
    assert syn.num_features == 1433 and syn.num_classes == 7
    monkeypatch.setenv('PYGB_REQUIRE_REAL_DATA', '1')
    with pytest.raises(FileNotFoundError):
        Planetoid(str(tmp_path / 'nowhere'), 'Cora')


def test_synthetic_stand_ins_have_the_published_shapes():
    import warnings
    import torch_geometric.transforms as T
    from torch_geometric.data import DataLoader
    from torch_geometric.datasets import TUDataset, MNISTSuperpixels, Amazon, PPI
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        enz = TUDataset('/nowhere', 'ENZYMES', use_node_attr=True)
        mut = TUDataset('/nowhere', 'MUTAG')
        mn = MNISTSuperpixels('/nowhere', transform=T.Cartesian(), num_graphs=8)
        am = Amazon('/nowhere', 'Photo')
        ppi = PPI('/nowhere', split='val')
    assert (len(enz), enz.num_features, enz.num_classes) == (600, 21, 6)
    assert (len(mut), mut.num_features, mut.num_classes) == (188, 7, 2)
    g = mn[0]
    assert g.x.shape == (75, 1) and g.pos.shape == (75, 2) and g.edge_attr.size(1) == 2
    assert float(g.edge_attr.min()) >= 0 and float(g.edge_attr.max()) <= 1
    assert am[0].x.shape == (7650, 745) and am[0].edge_index.size(1) == 238162
    assert (len(ppi), ppi.num_features, ppi.num_classes) == (2, 50, 121)
    b = next(iter(DataLoader(enz, batch_size=32, shuffle=False)))
    assert b.num_graphs == 32 and b.y.shape == (32,)
    for ds in (enz, mut):
        for g in (ds[0], ds[len(ds) - 1]):
            ei = g.edge_index
            assert int(ei.max()) < g.num_nodes and not bool((ei[0] == ei[1]).any())


def test_fork_script_launcher_stubs_only_what_is_missing(tmp_path):
    import importlib
    import sys
    sys.path.insert(0, __import__('os').path.join(__import__('os').path.dirname(
        __import__('os').path.dirname(__import__('os').path.abspath(__file__))), 'tools'))
    rfs = importlib.import_module('run_fork_script')
    before = set(sys.modules)
    added = rfs.install_stubs()
    try:
        assert 'numpy' not in added and 'scipy' not in added
        for name in added:
            assert isinstance(sys.modules[name], rfs._Inert)
        if 'pyts.image' in added:
            from pyts.image import RecurrencePlot
            with pytest.raises(ImportError):
                RecurrencePlot()
    finally:
        for name in set(sys.modules) - before:
            if isinstance(sys.modules[name], rfs._Inert):
                del sys.modules[name]
