"""Reader for the raw Planetoid files (``ind.<name>.{x,tx,allx,y,ty,ally,graph,test.index}``,
the format of Kipf & Welling's GCN release that ``Planetoid`` downloads).  Host-side IO,
off the hot path: numpy / scipy only.

x / tx / allx are pickled scipy CSR matrices, y / ty / ally one-hot numpy arrays, graph a
``{node: [neighbours]}`` dict, test.index one integer per line.  Node order, masks and the
symmetrised, duplicate-free, loop-free ``edge_index`` follow the published split protocol
(train = the first ``len(y)`` nodes, val = the next 500, test = the listed indices)."""
import os
import pickle

import numpy as np
import torch


def _load(folder, prefix, name):
    path = os.path.join(folder, 'ind.{}.{}'.format(prefix, name))
    if name == 'test.index':
        with open(path) as f:
            return np.array([int(line) for line in f.read().split()], dtype=np.int64)
    with open(path, 'rb') as f:
        obj = pickle.load(f, encoding='latin1')
    if name == 'graph':
        return obj
    obj = obj.todense() if hasattr(obj, 'todense') else obj
    return np.asarray(obj, dtype=np.float32)


def _edges_from_dict(graph, num_nodes):
    rows, cols = [], []
    for key, nbrs in graph.items():
        rows.append(np.full(len(nbrs), key, dtype=np.int64))
        cols.append(np.asarray(nbrs, dtype=np.int64))
    row, col = np.concatenate(rows), np.concatenate(cols)
    keep = row != col
    row, col = row[keep], col[keep]
    # symmetrise, then sort by (row, col) and drop duplicates
    code = np.unique(np.concatenate([row * num_nodes + col, col * num_nodes + row]))
    return torch.from_numpy(np.stack([code // num_nodes, code % num_nodes]))


def read_planetoid_raw(folder, prefix):
    x, tx, allx, y, ty, ally = [_load(folder, prefix, n) for n in
                                ('x', 'tx', 'allx', 'y', 'ty', 'ally')]
    graph = _load(folder, prefix, 'graph')
    test_index = _load(folder, prefix, 'test.index')
    sorted_test = np.sort(test_index)
    if prefix == 'citeseer':
        # some test nodes are isolated and missing from tx: give them zero rows
        lo = test_index.min()
        span = int(test_index.max() - lo) + 1
        tx_ext = np.zeros((span, tx.shape[1]), dtype=np.float32)
        ty_ext = np.zeros((span, ty.shape[1]), dtype=np.float32)
        tx_ext[sorted_test - lo] = tx
        ty_ext[sorted_test - lo] = ty
        tx, ty = tx_ext, ty_ext
    feats = np.concatenate([allx, tx], axis=0)
    labels = np.concatenate([ally, ty], axis=0).argmax(axis=1)
    feats[test_index] = feats[sorted_test]
    labels[test_index] = labels[sorted_test]
    n = labels.shape[0]

    def mask(idx):
        m = torch.zeros(n, dtype=torch.bool)
        m[torch.as_tensor(idx)] = True
        return m

    n_train = y.shape[0]
    return {'x': torch.from_numpy(feats), 'y': torch.from_numpy(labels).long(),
            'edge_index': _edges_from_dict(graph, n),
            'train_mask': mask(np.arange(n_train)),
            'val_mask': mask(np.arange(n_train, min(n_train + 500, n))),
            'test_mask': mask(test_index)}
