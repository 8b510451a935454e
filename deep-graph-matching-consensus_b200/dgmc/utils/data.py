"""Pair datasets (reference: dgmc/utils/data.py:9-133) — host-side plumbing that defines the input
contract of DGMC.forward: `edge_index_s/_t` are offset by the own graph's node
count when pairs are batched, and ValidPairDataset emits graph-local target columns in `y`.
PyTorch Geometric is not a dependency here: `PairData` is a small attribute container with the
`__inc__` rule the reference relies on, and `collate_pairs` is the `follow_batch=['x_s', 'x_t']`
collation the example scripts use (examples/pascal.py:42-43)."""
import random
import re
from itertools import chain

import torch


class PairData(object):
    """Attribute bag for one (source, target) pair; repr follows PyG's old `Data` style."""
    def __init__(self, **fields):
        self._keys = []
        for key, value in fields.items():
            setattr(self, key, value)
            if value is not None and key != 'num_nodes':
                self._keys.append(key)

    @property
    def keys(self):
        return sorted(self._keys)

    def __len__(self):
        return len(self._keys)

    def __iter__(self):
        for key in self.keys:
            yield key, getattr(self, key)

    def __inc__(self, key, value):
        if re.search('index_s', key):
            return self.x_s.size(0)
        if re.search('index_t', key):
            return self.x_t.size(0)
        return 0

    def to(self, device):
        for key in self._keys:
            value = getattr(self, key)
            if torch.is_tensor(value):
                moved = value.to(device)
                hint = getattr(value, '__dgmc_layout__', None)
                if hint is not None:     # (number of graphs, largest graph) of a batch vector: see collate_pairs
                    moved.__dgmc_layout__ = hint
                setattr(self, key, moved)
        return self

    def __repr__(self):
        body = ', '.join('{}={}'.format(k, list(v.size()) if torch.is_tensor(v) else v)
                         for k, v in self)
        return '{}({})'.format(type(self).__name__, body)


def collate_pairs(pairs):
    """Batch PairData objects: concatenates node tensors, offsets edge indices with `__inc__`,
    and adds `x_s_batch` / `x_t_batch` assignment vectors."""
    out, off_s, off_t = {}, 0, 0
    cols = {k: [] for k in pairs[0].keys}
    cols['x_s_batch'], cols['x_t_batch'] = [], []
    for g, pair in enumerate(pairs):
        for key in pair.keys:
            value = getattr(pair, key)
            if 'index_s' in key:
                value = value + off_s
            elif 'index_t' in key:
                value = value + off_t
            cols[key].append(value)
        cols['x_s_batch'].append(torch.full((pair.x_s.size(0), ), g, dtype=torch.long))
        cols['x_t_batch'].append(torch.full((pair.x_t.size(0), ), g, dtype=torch.long))
        off_s += pair.x_s.size(0)
        off_t += pair.x_t.size(0)
    for key, parts in cols.items():
        out[key] = torch.cat(parts, dim=-1 if 'index' in key else 0)
    # The sizes are known HERE, on the host: attach (number of graphs, largest graph) to the batch
    # vectors so that DGMC.forward can lay out the padded batch without asking the device for them
    # (the reference's to_dense_batch synchronises twice per side and call).
    out['x_s_batch'].__dgmc_layout__ = (len(pairs), max(p.x_s.size(0) for p in pairs))
    out['x_t_batch'].__dgmc_layout__ = (len(pairs), max(p.x_t.size(0) for p in pairs))
    return PairData(**out)


def _pair(data_s, data_t, **extra):
    return PairData(x_s=data_s.x, edge_index_s=data_s.edge_index, edge_attr_s=data_s.edge_attr,
                    x_t=data_t.x, edge_index_t=data_t.edge_index, edge_attr_t=data_t.edge_attr,
                    num_nodes=None, **extra)


class PairDataset(torch.utils.data.Dataset):
    """All (or, with sample=True, one random target per source) pairs of two datasets."""
    def __init__(self, dataset_s, dataset_t, sample=False):
        self.dataset_s, self.dataset_t, self.sample = dataset_s, dataset_t, sample

    def __len__(self):
        n_s, n_t = len(self.dataset_s), len(self.dataset_t)
        return n_s if self.sample else n_s * n_t

    def __getitem__(self, idx):
        n_t = len(self.dataset_t)
        if self.sample:
            return _pair(self.dataset_s[idx], self.dataset_t[random.randint(0, n_t - 1)])
        return _pair(self.dataset_s[idx // n_t], self.dataset_t[idx % n_t])

    def __repr__(self):
        return '{}({}, {}, sample={})'.format(type(self).__name__, self.dataset_s, self.dataset_t,
                                              self.sample)


class ValidPairDataset(torch.utils.data.Dataset):
    """Pairs in which every node class of the source graph also occurs in the target graph; `y`
    maps each source node to the graph-local index of the target node of the same class."""
    def __init__(self, dataset_s, dataset_t, sample=False):
        self.dataset_s, self.dataset_t, self.sample = dataset_s, dataset_t, sample
        self.pairs, self.cumdeg = self.__compute_pairs__()

    def __compute_pairs__(self):
        n_cls = 1 + max(int(d.y.max()) for d in chain(self.dataset_s, self.dataset_t))

        def presence(dataset):
            table = torch.zeros(len(dataset), n_cls, dtype=torch.bool)
            for i, d in enumerate(dataset):
                table[i, d.y] = True
            return table

        has_s, has_t = presence(self.dataset_s), presence(self.dataset_t)
        shared = (has_s.unsqueeze(1) & has_t.unsqueeze(0)).sum(dim=-1)
        pairs = (shared == has_s.sum(dim=-1, keepdim=True)).nonzero()
        per_source = torch.bincount(pairs[:, 0], minlength=len(self.dataset_s))
        return pairs.tolist(), [0] + per_source.cumsum(0).tolist()

    def __len__(self):
        return len(self.dataset_s) if self.sample else len(self.pairs)

    def __getitem__(self, idx):
        if self.sample:
            data_s = self.dataset_s[idx]
            pick = random.randint(self.cumdeg[idx], self.cumdeg[idx + 1] - 1)
            data_t = self.dataset_t[self.pairs[pick][1]]
        else:
            data_s = self.dataset_s[self.pairs[idx][0]]
            data_t = self.dataset_t[self.pairs[idx][1]]
        lookup = data_s.y.new_full((int(data_t.y.max()) + 1, ), -1)
        lookup[data_t.y] = torch.arange(data_t.y.numel())
        return _pair(data_s, data_t, y=lookup[data_s.y])

    def __repr__(self):
        return '{}({}, {}, sample={})'.format(type(self).__name__, self.dataset_s, self.dataset_t,
                                              self.sample)
