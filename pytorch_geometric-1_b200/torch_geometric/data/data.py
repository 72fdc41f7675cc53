"""Minimal ``Data`` / ``Batch`` containers: just enough of the
reference's attribute protocol for the fork's scripts (examples/gcn.py:12,32,46:
``dataset[0]``, ``data.to(device)``, ``data('train_mask', ...)``)."""
import torch


class Data(object):
    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, **kwargs):
        self.x, self.edge_index, self.edge_attr, self.y, self.pos = x, edge_index, edge_attr, y, pos
        for k, v in kwargs.items():
            setattr(self, k, v)

    @property
    def keys(self):
        return [k for k in self.__dict__.keys() if self.__dict__[k] is not None and k[:2] != '__']

    def __getitem__(self, key):
        return getattr(self, key, None)

    def __setitem__(self, key, value):
        setattr(self, key, value)

    def __contains__(self, key):
        return key in self.keys

    def __iter__(self):
        for k in sorted(self.keys):
            yield k, self[k]

    def __call__(self, *keys):
        for k in (sorted(self.keys) if not keys else keys):
            if k in self:
                yield k, self[k]

    def __len__(self):
        return len(self.keys)

    def __cat_dim__(self, key, value):
        return -1 if key in ('edge_index', 'face') else 0

    def __inc__(self, key, value):
        return self.num_nodes if key in ('edge_index', 'face') else 0

    @property
    def num_nodes(self):
        if getattr(self, '__num_nodes__', None) is not None:
            return self.__num_nodes__
        for k in ('x', 'pos', 'batch'):
            v = self[k]
            if v is not None:
                return v.size(0)
        if self.edge_index is not None and self.edge_index.numel() > 0:
            return int(self.edge_index.max()) + 1
        return None

    @num_nodes.setter
    def num_nodes(self, n):
        self.__num_nodes__ = n

    @property
    def num_edges(self):
        return None if self.edge_index is None else self.edge_index.size(1)

    @property
    def num_features(self):
        if self.x is None:
            return 0
        return 1 if self.x.dim() == 1 else self.x.size(1)

    @property
    def num_node_features(self):
        return self.num_features

    def apply(self, func, *keys):
        for k, v in self(*keys):
            if torch.is_tensor(v):
                self[k] = func(v)
        return self

    def contiguous(self, *keys):
        return self.apply(lambda t: t.contiguous(), *keys)

    def to(self, device, *keys, **kwargs):
        return self.apply(lambda t: t.to(device, **kwargs), *keys)

    def clone(self):
        out = self.__class__()
        for k, v in self.__dict__.items():
            out.__dict__[k] = v.clone() if torch.is_tensor(v) else v
        return out

    def __repr__(self):
        info = ['{}={}'.format(k, list(v.size()) if torch.is_tensor(v) else v) for k, v in self]
        return '{}({})'.format(self.__class__.__name__, ', '.join(info))


class Batch(Data):
    def __init__(self, batch=None, **kwargs):
        super(Batch, self).__init__(**kwargs)
        self.batch = batch

    @staticmethod
    def from_data_list(data_list, follow_batch=()):
        """Block-diagonal stacking: node tensors are concatenated, ``edge_index`` is
        shifted by the running node count, ``batch`` maps nodes to their graph."""
        keys = set()
        for d in data_list:
            keys |= set(d.keys)
        keys.discard('batch')
        out = Batch()
        cols = {k: [] for k in keys}
        batch_vec = []
        cum = 0
        for gi, d in enumerate(data_list):
            n = d.num_nodes
            for k in keys:
                v = d[k]
                if torch.is_tensor(v) and v.dtype != torch.bool and d.__inc__(k, v) != 0:
                    v = v + cum
                cols[k].append(v)
            batch_vec.append(torch.full((n,), gi, dtype=torch.long))
            cum += n
        for k in keys:
            v0 = cols[k][0]
            if torch.is_tensor(v0):
                out[k] = torch.cat(cols[k], dim=data_list[0].__cat_dim__(k, v0))
            else:
                out[k] = torch.tensor(cols[k]) if isinstance(v0, (int, float)) else cols[k]
        dev = None
        for k in keys:
            if torch.is_tensor(out[k]):
                dev = out[k].device
                break
        out.batch = torch.cat(batch_vec, dim=0).to(dev)
        return out.contiguous()

    @property
    def num_graphs(self):
        return int(self.batch.max()) + 1
