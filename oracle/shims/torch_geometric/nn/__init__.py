"""Stand-ins for torch_geometric.nn.{MessagePassing, SplineConv, GINConv}.

* MessagePassing (rel.py:4,7-37): `propagate(edge_index, x=...)` gathers `x_j` from
  `edge_index[0]` and reduces at `edge_index[1]` when `flow == 'source_to_target'`, the
  reverse otherwise; `aggr='mean'` divides by max(count, 1).  `flow` is read at call time
  because RelConv flips it between two propagate calls (rel.py:27-30).
* SplineConv (spline.py:4,21): degree-1 open B-spline continuous kernel
  (Fey et al., SplineCNN), mean aggregation, root weight and bias.
* GINConv (gin.py:3,23): nn((1 + eps) * x + sum_j x_j), trainable eps initialised to 0.
"""
import math
import torch
from . import inits  # noqa: F401


def _aggregate(msg, index, num_nodes, aggr):
    out = msg.new_zeros((num_nodes, ) + tuple(msg.shape[1:]))
    out.index_add_(0, index, msg)
    if aggr == 'mean':
        cnt = torch.bincount(index, minlength=num_nodes).clamp_(min=1)
        out = out / cnt.view(-1, *([1] * (msg.dim() - 1))).to(out.dtype)
    return out


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr='add', flow='source_to_target'):
        super().__init__()
        self.aggr = aggr
        self.flow = flow

    def propagate(self, edge_index, x, **kwargs):
        i, j = (1, 0) if self.flow == 'source_to_target' else (0, 1)
        msg = self.message(x_j=x[edge_index[j]], **kwargs)
        return _aggregate(msg, edge_index[i], x.size(0), self.aggr)

    def message(self, x_j):
        return x_j


def spline_basis(pseudo, kernel_size, degree=1, is_open=True):
    """-> basis [E, 2^dim], weight_index [E, 2^dim]  (degree-1 B-spline)."""
    assert degree == 1
    E, dim = pseudo.shape
    v = pseudo * (kernel_size - degree * int(is_open))
    lo = v.floor()
    frac = v - lo
    lo = lo.long()
    S = 2**dim
    basis = pseudo.new_ones(E, S)
    widx = torch.zeros(E, S, dtype=torch.long, device=pseudo.device)
    for s in range(S):
        off = 1
        for d in range(dim):
            bit = (s >> d) & 1
            f = frac[:, d]
            basis[:, s] = basis[:, s] * (f if bit else (1 - f))
            widx[:, s] += ((lo[:, d] + bit) % kernel_size) * off
            off *= kernel_size
    return basis, widx


class SplineConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, dim, kernel_size, is_open_spline=True,
                 degree=1, aggr='mean', root_weight=True, bias=True):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.dim, self.kernel_size, self.degree, self.aggr = dim, kernel_size, degree, aggr
        self.is_open = is_open_spline
        self.K = kernel_size**dim
        self.weight = torch.nn.Parameter(torch.empty(self.K, in_channels, out_channels))
        self.root = torch.nn.Parameter(torch.empty(in_channels, out_channels)) \
            if root_weight else None
        self.bias = torch.nn.Parameter(torch.empty(out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        bound = 1.0 / math.sqrt(self.in_channels * self.K)
        torch.nn.init.uniform_(self.weight, -bound, bound)
        if self.root is not None:
            torch.nn.init.uniform_(self.root, -bound, bound)
        if self.bias is not None:
            torch.nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x, edge_index, pseudo):
        src, dst = edge_index[0], edge_index[1]
        basis, widx = spline_basis(pseudo, self.kernel_size, self.degree, self.is_open)
        x_j = x[src]
        msg = x.new_zeros(src.numel(), self.out_channels)
        for s in range(basis.size(1)):
            w = self.weight[widx[:, s]]  # [E, in, out]
            msg = msg + basis[:, s:s + 1] * torch.bmm(x_j.unsqueeze(1), w).squeeze(1)
        out = _aggregate(msg, dst, x.size(0), self.aggr)
        if self.root is not None:
            out = out + x @ self.root
        if self.bias is not None:
            out = out + self.bias
        return out

    def __repr__(self):
        return '{}({}, {}, dim={})'.format(self.__class__.__name__, self.in_channels,
                                           self.out_channels, self.dim)


class GINConv(torch.nn.Module):
    def __init__(self, nn, eps=0.0, train_eps=False):
        super().__init__()
        self.nn = nn
        self.initial_eps = eps
        if train_eps:
            self.eps = torch.nn.Parameter(torch.tensor([float(eps)]))
        else:
            self.register_buffer('eps', torch.tensor([float(eps)]))

    def reset_parameters(self):
        inits.reset(self.nn)
        self.eps.data.fill_(self.initial_eps)

    def forward(self, x, edge_index):
        agg = _aggregate(x[edge_index[0]], edge_index[1], x.size(0), 'add')
        return self.nn((1 + self.eps) * x + agg)

    def __repr__(self):
        return '{}(nn={})'.format(self.__class__.__name__, self.nn)
