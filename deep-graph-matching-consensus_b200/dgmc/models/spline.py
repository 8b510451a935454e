"""SplineCNN backbone (reference: dgmc/models/spline.py:8-61; PyG SplineConv and
torch_spline_conv are third-party) on the K5 spline kernels.

SplineConv(in, out, dim, kernel_size=5): degree-1 open B-spline kernel, mean aggregation, root
weight and bias.  Parameters: `weight [K^dim, in, out]`, `root [in, out]`, `bias [out]`
(PyG changed these names between releases; copy weights explicitly when converting checkpoints).
The B200 path transforms first — Y = x @ W for all K^dim kernels in one GEMM (ops.node_linear) — and then
runs a deterministic CSR gather over the in-edges weighted by the cached B-spline basis."""
import math

import torch
import torch.nn.functional as F

from .. import ops


class SplineConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels, dim, kernel_size=5):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.dim, self.kernel_size = dim, kernel_size
        self.n_kernels = kernel_size**dim
        self.weight = torch.nn.Parameter(torch.empty(self.n_kernels, in_channels, out_channels))
        self.root = torch.nn.Parameter(torch.empty(in_channels, out_channels))
        self.bias = torch.nn.Parameter(torch.empty(out_channels))
        self.reset_parameters()

    def reset_parameters(self):
        bound = 1.0 / math.sqrt(self.in_channels * self.n_kernels)
        for p in (self.weight, self.root, self.bias):
            torch.nn.init.uniform_(p, -bound, bound)

    def forward(self, x, edge_index, edge_attr):
        graph = ops.GraphCSR.of(edge_index, x.size(0))
        sb = ops.SplineBasis.of(edge_attr, self.kernel_size)
        c = self.out_channels
        pad = (-c) % 4  # kernels move float4 channel vectors
        weight, root, bias = self.weight, self.root, self.bias
        if pad:
            weight, root, bias = F.pad(weight, (0, pad)), F.pad(root, (0, pad)), F.pad(bias, (0, pad))
        wide = weight.permute(1, 0, 2).reshape(self.in_channels, self.n_kernels * (c + pad))
        y = ops.node_linear(x, wide, kn=True)  # [N, K * C]
        base = ops.node_linear(x, root, bias, kn=True)
        out = ops.spline_gather(y, base, graph, sb, self.n_kernels)
        return out[:, :c] if pad else out

    def __repr__(self):
        return '{}({}, {}, dim={})'.format(type(self).__name__, self.in_channels, self.out_channels,
                                           self.dim)


class SplineCNN(torch.nn.Module):
    def __init__(self, in_channels, out_channels, dim, num_layers, cat=True, lin=True, dropout=0.0):
        super().__init__()
        self.in_channels, self.dim, self.num_layers = in_channels, dim, num_layers
        self.cat, self.lin, self.dropout = cat, lin, dropout
        self.convs = torch.nn.ModuleList()
        width = in_channels
        for _ in range(num_layers):
            self.convs.append(SplineConv(width, out_channels, dim, kernel_size=5))
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

    def forward(self, x, edge_index, edge_attr, *args):
        feats = [x]
        for conv in self.convs:
            feats.append(F.relu(conv(feats[-1], edge_index, edge_attr)))
        out = torch.cat(feats, dim=-1) if self.cat else feats[-1]
        out = F.dropout(out, p=self.dropout, training=self.training)
        return ops.node_linear(out, self.final.weight, self.final.bias) if self.lin else out

    def __repr__(self):
        return ('{}({}, {}, dim={}, num_layers={}, cat={}, lin={}, dropout={})').format(
            type(self).__name__, self.in_channels, self.out_channels, self.dim, self.num_layers,
            self.cat, self.lin, self.dropout)
