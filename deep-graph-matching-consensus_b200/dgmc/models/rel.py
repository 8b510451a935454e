"""RelCNN backbone (reference: dgmc/models/rel.py:7-99) on the K5 aggregation kernel.

RelConv computes root(x_i) + mean_{j->i} lin1(x_j) + mean_{i->j} lin2(x_j) (rel.py:25-31).  The
mean commutes with the linear maps, so the B200 path aggregates the raw features in both flow
directions ([x | mean_in x | mean_out x], one sparse [3N, N] product, deterministic, no atomics)
and applies ONE tensor-core GEMM [N, 3*C_in] x [3*C_in, C_out] with the three weight matrices side
by side (`ops.node_linear`: tcgen05 3xTF32, bias + ReLU in the epilogue).  Where the shape allows
(C_in a multiple of 32, C_out <= 64: psi_2) both happen in ONE kernel, `ops.relconv_fused`: the
aggregated operand is gathered straight into the GEMM's shared-memory stage.  Parameter names (lin1, lin2, root, batch_norms, final) match the reference's
state_dict so checkpoints interchange."""
import torch
import torch.nn.functional as F

from .. import ops


class RelConv(torch.nn.Module):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.lin1 = torch.nn.Linear(in_channels, out_channels, bias=False)
        self.lin2 = torch.nn.Linear(in_channels, out_channels, bias=False)
        self.root = torch.nn.Linear(in_channels, out_channels)
        self.reset_parameters()

    def reset_parameters(self):
        for lin in (self.lin1, self.lin2, self.root):
            lin.reset_parameters()

    def forward(self, x, edge_index, relu=False, out=None, passthrough=False, relu_in=None, relu_out=None):
        """relu=True additionally applies the ReLU that RelCNN puts after every conv (rel.py:84-85),
        fused into the GEMM epilogue.  out: write the result into this column block of a layer
        buffer; passthrough: also return x as handed through the aggregation (ops._RelAggregatePass).
        relu_in / relu_out: ops.ReluOwner of the layer that produced x / of this layer's output (RelCNN
        chains them so that a layer's ReLU backward runs inside the next layer's aggregation backward)."""
        graph = ops.GraphCSR.of(edge_index, x.size(0))
        if ops.relconv_fused_ok(x, graph.rel, self.out_channels):
            # aggregation and GEMM in one kernel (dgmc_relconv_tc)
            weight = ops.region_cached(('relconv', id(self)), lambda: torch.cat(
                [self.root.weight, self.lin1.weight, self.lin2.weight], dim=1))
            y, alias = ops.relconv_fused(x, weight, self.root.bias, graph.rel, relu=relu,
                                         key=('relconv', id(self)), out=out)
            return (y, alias) if passthrough else y
        if passthrough:
            stacked, alias = ops.rel_aggregate_pass(x, graph, relu_in)
        else:
            stacked = ops.rel_aggregate(x, graph)  # [N, 3 C_in] = [x | mean_in | mean_out]
        weight = ops.region_cached(('relconv', id(self)), lambda: torch.cat(
            [self.root.weight, self.lin1.weight, self.lin2.weight], dim=1))
        y = ops.node_linear(stacked, weight, self.root.bias, relu=relu, key=('relconv', id(self)), out=out,
                            relu_owner=relu_out if relu else None)
        return (y, alias) if passthrough else y

    def __repr__(self):
        return '{}({}, {})'.format(type(self).__name__, self.in_channels, self.out_channels)


class RelCNN(torch.nn.Module):
    def __init__(self, in_channels, out_channels, num_layers, batch_norm=False, cat=True, lin=True,
                 dropout=0.0):
        super().__init__()
        self.in_channels, self.num_layers = in_channels, num_layers
        self.batch_norm, self.cat, self.lin, self.dropout = batch_norm, cat, lin, dropout
        self.convs = torch.nn.ModuleList()
        self.batch_norms = torch.nn.ModuleList()
        width = in_channels
        for _ in range(num_layers):
            self.convs.append(RelConv(width, out_channels))
            self.batch_norms.append(torch.nn.BatchNorm1d(out_channels))
            width = out_channels
        feat = in_channels + num_layers * out_channels if cat else out_channels
        if lin:
            self.out_channels = out_channels
            self.final = torch.nn.Linear(feat, out_channels)
        else:
            self.out_channels = feat
        self.reset_parameters()

    def reset_parameters(self):
        for module in list(self.convs) + list(self.batch_norms):
            module.reset_parameters()
        if self.lin:
            self.final.reset_parameters()

    def _direct_cat(self, x):
        """The layers can write straight into the concatenated [N, in + L*out] buffer: no batch norm
        or dropout between a layer and the concatenation, float4-aligned column blocks."""
        return (self.cat and not self.batch_norm and not (self.training and self.dropout > 0)
                and x.is_cuda and x.dtype == torch.float32 and x.dim() == 2
                and self.in_channels % 4 == 0 and self.convs[0].out_channels % 4 == 0)

    def feature_buffer(self, n, device):
        """An empty [n, in + L*out] table for `forward(..., buffer=)`: a caller that PRODUCES the input
        (DGMC writes the indicators and their transport) can write it straight into the leading column
        block `buffer[:, :in_channels]` and save the copy.  None when the layers cannot write into one
        concatenated table (see _direct_cat)."""
        probe = torch.empty(0, self.in_channels, dtype=torch.float32, device=device)
        if not self._direct_cat(probe):
            return None
        return torch.empty(n, self.in_channels + self.num_layers * self.convs[0].out_channels,
                           dtype=torch.float32, device=device)

    def _forward_direct(self, x, edge_index, features_only=False, buffer=None):
        c_in, c_out = self.in_channels, self.convs[0].out_channels
        if buffer is not None and x.data_ptr() == buffer.data_ptr() and x.stride(0) == buffer.stride(0) \
                and x.size(0) == buffer.size(0) and buffer.size(1) == c_in + self.num_layers * c_out:
            buf, cur = buffer, x                      # x already IS the leading column block
        else:
            buf = torch.empty(x.size(0), c_in + self.num_layers * c_out, dtype=torch.float32, device=x.device)
            cur = ops.copy_into(x, buf, 0)
        parts, off = [], c_in
        owner = None                      # ReLU owner of `cur` (the layer input is not a ReLU output)
        fused = ops.relconv_fused_ok(cur, ops.GraphCSR.of(edge_index, cur.size(0)).rel, c_out)
        for conv in self.convs:
            # `cur` (a layer output) is consumed by the next layer's aggregation node only — through which
            # its share of the concatenation's gradient passes as well — so that node can apply the ReLU
            # backward of the layer that produced it
            nxt = None if fused else ops.ReluOwner()
            cur, alias = conv(cur, edge_index, relu=True, out=buf[:, off:off + c_out], passthrough=True,
                              relu_in=owner, relu_out=nxt)
            owner = nxt
            parts.append(alias)
            off += c_out
        parts.append(cur)
        # the gradient of `out` comes from this package's own GEMM backward (the final Lin, or the
        # caller's fused map when features_only) unless the features are the module's output
        out = ops.join_slices(buf, self.lin, *parts)
        if features_only or not self.lin:
            return out
        return ops.node_linear(out, self.final.weight, self.final.bias, key=('final', id(self)))

    def forward(self, x, edge_index, *args, features_only=False, buffer=None):
        """features_only: return the input of the final Lin (the concatenated layer outputs) instead
        of applying it — DGMC folds that Lin into the first consensus-MLP layer (two linear maps
        with nothing in between).  buffer: see feature_buffer()."""
        assert not features_only or self.lin
        if self._direct_cat(x):
            return self._forward_direct(x, edge_index, features_only, buffer)
        feats = [x]
        for conv, norm in zip(self.convs, self.batch_norms):
            h = conv(feats[-1], edge_index, relu=True)
            if self.batch_norm:
                h = norm(h)
            feats.append(F.dropout(h, p=self.dropout, training=self.training))
        out = torch.cat(feats, dim=-1) if self.cat else feats[-1]
        if features_only or not self.lin:
            return out
        return ops.node_linear(out, self.final.weight, self.final.bias, key=('final', id(self)))

    def __repr__(self):
        return ('{}({}, {}, num_layers={}, batch_norm={}, cat={}, lin={}, dropout={})').format(
            type(self).__name__, self.in_channels, self.out_channels, self.num_layers,
            self.batch_norm, self.cat, self.lin, self.dropout)
