"""GIN backbone (reference: dgmc/models/gin.py:9-59; PyG GINConv is third-party).  Used by the
reference's tests as psi_1/psi_2 (test/models/test_dgmc.py:9-10).  The neighbour sum runs on the
K5 segmented-gather kernel; the MLPs on the per-node GEMM (ops.node_linear)."""
import torch

from .mlp import MLP
from .. import ops


class GINConv(torch.nn.Module):
    """nn((1 + eps) * x_i + sum_{j -> i} x_j), eps trainable and initialised to 0."""
    def __init__(self, nn, train_eps=True):
        super().__init__()
        self.nn = nn
        self.eps = torch.nn.Parameter(torch.zeros(1), requires_grad=train_eps)

    def reset_parameters(self):
        self.nn.reset_parameters()
        self.eps.data.zero_()

    def forward(self, x, edge_index):
        graph = ops.GraphCSR.of(edge_index, x.size(0))
        return self.nn((1 + self.eps) * x + ops.neighbor_sum(x, graph))

    def __repr__(self):
        return '{}(nn={})'.format(type(self).__name__, self.nn)


class GIN(torch.nn.Module):
    def __init__(self, in_channels, out_channels, num_layers, batch_norm=False, cat=True, lin=True):
        super().__init__()
        self.in_channels, self.num_layers = in_channels, num_layers
        self.batch_norm, self.cat, self.lin = batch_norm, cat, lin
        self.convs = torch.nn.ModuleList()
        width = in_channels
        for _ in range(num_layers):
            self.convs.append(GINConv(MLP(width, out_channels, 2, batch_norm, dropout=0.0)))
            width = out_channels
        feat = in_channels + num_layers * out_channels if cat else out_channels
        if lin:
            self.out_channels = out_channels
            self.final = torch.nn.Linear(feat, out_channels)
        else:
            self.out_channels = feat
        self.reset_parameters()

    def reset_parameters(self):
        for conv in self.convs:
            conv.reset_parameters()
        if self.lin:
            self.final.reset_parameters()

    def forward(self, x, edge_index, *args):
        feats = [x]
        for conv in self.convs:
            feats.append(conv(feats[-1], edge_index))
        out = torch.cat(feats, dim=-1) if self.cat else feats[-1]
        return ops.node_linear(out, self.final.weight, self.final.bias) if self.lin else out

    def __repr__(self):
        return '{}({}, {}, num_layers={}, batch_norm={}, cat={}, lin={})'.format(
            type(self).__name__, self.in_channels, self.out_channels, self.num_layers,
            self.batch_norm, self.cat, self.lin)
