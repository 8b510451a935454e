"""Stand-ins for torch_geometric.data.{Data, Batch} (reference: data.py:6, test_dgmc.py:2,
test_data.py:2).  Only what the reference's tests touch: keyword storage, `num_nodes`,
`num_node_features`, key iteration / len, the old-style repr pinned by test_data.py:13-17,
and `Batch.from_data_list` with `edge_index` offsetting and a `batch` vector."""
import torch


class Data(object):
    def __init__(self, x=None, edge_index=None, edge_attr=None, y=None, pos=None, **kwargs):
        self.x, self.edge_index, self.edge_attr, self.y, self.pos = x, edge_index, edge_attr, y, pos
        for key, item in kwargs.items():
            setattr(self, key, item)

    @property
    def keys(self):
        return sorted(k for k, v in self.__dict__.items()
                      if v is not None and not k.startswith('__'))

    def __len__(self):
        return len(self.keys)

    def __iter__(self):
        for k in self.keys:
            yield k, getattr(self, k)

    def __inc__(self, key, value):
        return self.num_nodes if 'index' in key else 0

    @property
    def num_nodes(self):
        if getattr(self, '__num_nodes__', None) is not None:
            return self.__num_nodes__
        for k in ('x', 'pos'):
            if getattr(self, k, None) is not None:
                return getattr(self, k).size(0)
        return None

    @num_nodes.setter
    def num_nodes(self, n):
        self.__dict__['__num_nodes__'] = n

    @property
    def num_node_features(self):
        return 0 if self.x is None else (1 if self.x.dim() == 1 else self.x.size(1))

    def to(self, device):
        for k, v in self:
            if torch.is_tensor(v):
                setattr(self, k, v.to(device))
        return self

    def __repr__(self):
        info = ['{}={}'.format(k, list(v.size()) if torch.is_tensor(v) else v) for k, v in self]
        return '{}({})'.format(self.__class__.__name__, ', '.join(info))


class Batch(Data):
    @staticmethod
    def from_data_list(data_list, follow_batch=()):
        out = Batch()
        keys = data_list[0].keys
        cols = {k: [] for k in keys}
        batch = []
        follow = {k: [] for k in follow_batch}
        inc = {k: 0 for k in keys}
        for g, d in enumerate(data_list):
            for k in keys:
                v = getattr(d, k)
                if torch.is_tensor(v) and v.dtype == torch.long and 'index' in k:
                    cols[k].append(v + inc[k])
                else:
                    cols[k].append(v)
                inc[k] += d.__inc__(k, v)
                if k in follow:
                    follow[k].append(torch.full((v.size(0), ), g, dtype=torch.long))
            batch.append(torch.full((d.num_nodes, ), g, dtype=torch.long))
        for k in keys:
            dim = -1 if 'index' in k else 0
            setattr(out, k, torch.cat(cols[k], dim=dim))
        for k in follow:
            setattr(out, k + '_batch', torch.cat(follow[k]))
        out.batch = torch.cat(batch)
        return out
