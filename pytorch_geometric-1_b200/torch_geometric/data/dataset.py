import copy

import torch


class InMemoryDataset(torch.utils.data.Dataset):
    """A list of ``Data`` objects held in memory, with the attribute protocol the
    fork's scripts use (``num_features``, ``num_classes``, ``dataset[0]``,
    ``shuffle()``, ``transform``)."""

    def __init__(self, root=None, transform=None, pre_transform=None, pre_filter=None):
        self.root, self.transform = root, transform
        self._graphs = []

    @property
    def num_features(self):
        return self._graphs[0].num_features

    @property
    def num_node_features(self):
        return self.num_features

    @property
    def num_classes(self):
        ys = torch.cat([g.y.view(-1) for g in self._graphs])
        return int(ys.max().item()) + 1 if ys.dim() == 1 or ys.size(-1) == 1 else ys.size(-1)

    def __len__(self):
        return len(self._graphs)

    def __getitem__(self, idx):
        if isinstance(idx, int):
            g = self._graphs[idx].clone()
            return g if self.transform is None else self.transform(g)
        out = copy.copy(self)
        if isinstance(idx, slice):
            out._graphs = self._graphs[idx]
        else:
            idx = idx.tolist() if torch.is_tensor(idx) else list(idx)
            if idx and isinstance(idx[0], bool):
                idx = [i for i, b in enumerate(idx) if b]
            out._graphs = [self._graphs[i] for i in idx]
        return out

    def shuffle(self, return_perm=False):
        perm = torch.randperm(len(self))
        ds = self[perm]
        return (ds, perm) if return_perm else ds

    def __repr__(self):
        return '{}({})'.format(self.__class__.__name__, len(self))


Dataset = InMemoryDataset
