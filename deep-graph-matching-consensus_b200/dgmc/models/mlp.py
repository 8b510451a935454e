"""MLP backbone (reference: dgmc/models/mlp.py:7-44).  Only ever used inside GIN, which only the
reference's unit tests use (same constructor, state_dict names and repr).  Its Linear layers run on
the shared per-node GEMM (ops.node_linear: tensor-core kernel on CUDA)."""
import torch
import torch.nn.functional as F

from .. import ops


class MLP(torch.nn.Module):
    def __init__(self, in_channels, out_channels, num_layers, batch_norm=False, dropout=0.0):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.num_layers, self.batch_norm, self.dropout = num_layers, batch_norm, dropout
        widths = [in_channels] + [out_channels] * num_layers
        self.lins = torch.nn.ModuleList(
            torch.nn.Linear(a, b) for a, b in zip(widths[:-1], widths[1:]))
        self.batch_norms = torch.nn.ModuleList(
            torch.nn.BatchNorm1d(out_channels) for _ in range(num_layers))
        self.reset_parameters()

    def reset_parameters(self):
        for module in list(self.lins) + list(self.batch_norms):
            module.reset_parameters()

    def forward(self, x, *args):
        last = self.num_layers - 1
        for i in range(self.num_layers):
            if i == last:
                x = F.dropout(x, p=self.dropout, training=self.training)
            x = ops.node_linear(x, self.lins[i].weight, self.lins[i].bias, relu=(i != last))
            if i != last:
                if self.batch_norm:
                    x = self.batch_norms[i](x)
        return x

    def __repr__(self):
        return '{}({}, {}, num_layers={}, batch_norm={}, dropout={})'.format(
            type(self).__name__, self.in_channels, self.out_channels, self.num_layers,
            self.batch_norm, self.dropout)
