"""DGMC — drop-in for `dgmc.models.DGMC` (reference: dgmc/models/dgmc.py:32-319) whose
correspondence-and-consensus path runs on hand-written sm_100a kernels (libdgmc_b200).

What stays in Python, on purpose: the module surface (constructor, mutable `num_steps` / `k` /
`detach` read at call time, parameter names `psi_1.* psi_2.* mlp.0.* mlp.2.*`), the random draws
(`torch.randint` for negatives, one `torch.randn((B, N_s, R_in))` per consensus step, in the
reference's order and shapes so a seed reproduces the reference's stream on the same device),
and the per-node library GEMMs.  Everything index-driven or row-wise is a kernel:

  dgmc.py:85-94    __top_k__            -> K1  ops.topk_*            (scores never hit HBM)
  dgmc.py:198-201  gather + mul + sum   -> K2  ops.gather_dot
  dgmc.py:202..225 softmax over k       -> K3  ops.softmax_k / ops.masked_softmax
  dgmc.py:209-212  scatter_add          -> K4  ops.transport         (by-target CSR gather)
  dgmc.py:219-223  gather, sub, mlp(D)  -> K6  ops.diff_mlp          (first Linear hoisted)
  dgmc.py:163-165  h_s h_t^T + softmax  ->     ops.dense_scores      (tcgen05, softmax in the epilogue)
  dgmc.py:170      S^T r_s (dense)      ->     ops.dense_transport   (small batched products)
  dgmc.py:178-179  dense D and mlp(D)   ->     ops.dense_consensus

There is no CPU path: CPU tensors raise `DgmcNativeError`.
"""
import torch
import torch.nn.functional as F
from torch.nn import Linear, ReLU, Sequential

from .. import ops

EPS = 1e-8


class BatchLayout(object):
    """Where the nodes of a batch of graphs sit in the padded [B, N_max] layout (PyG to_dense_batch as
    called at dgmc.py:154-155): `slot` [sum N] = flat padded position of every real node, `mask`
    [B, N_max].  Everything downstream un-pads / re-pads with `slot` (index_select / index_copy: no
    boolean indexing, hence no nonzero and no host synchronisation per consensus step).

    Only two numbers need the host: the number of graphs and the largest graph.  They are taken, in
    this order, from a hint attached to the batch vector by the collation — `BatchLayout.hint(batch,
    num_graphs, max_nodes)`, which dgmc.utils.collate_pairs does on the host, where the sizes are known
    before the data is copied to the device — from the cache (same batch tensor as last time), or from
    the device (two synchronisations, as in the reference)."""
    _cache = {}

    def __init__(self, batch, num_graphs, n_max):
        self.B, self.n_max = num_graphs, n_max
        ones = torch.ones_like(batch)
        sizes = torch.zeros(num_graphs, dtype=batch.dtype, device=batch.device).scatter_add_(0, batch, ones)
        first = torch.cumsum(sizes, 0) - sizes
        self.slot = batch * n_max + (torch.arange(batch.numel(), device=batch.device) - first[batch])
        mask = torch.zeros(num_graphs * n_max, dtype=torch.bool, device=batch.device)
        mask[self.slot] = True
        self.mask = mask.view(num_graphs, n_max)

    @staticmethod
    def hint(batch, num_graphs, max_nodes):
        batch.__dgmc_layout__ = (int(num_graphs), int(max_nodes))
        return batch

    @classmethod
    def of(cls, batch):
        key = (batch.data_ptr(), batch.numel(), batch._version, str(batch.device))
        hit = cls._cache.get(key)
        if hit is not None and hit[0] is batch:
            return hit[1]
        known = getattr(batch, '__dgmc_layout__', None)
        if known is None:
            num_graphs = int(batch[-1]) + 1                                   # host sync
            ones = torch.ones_like(batch)
            sizes = torch.zeros(num_graphs, dtype=batch.dtype, device=batch.device).scatter_add_(0, batch, ones)
            known = (num_graphs, int(sizes.max()))                            # host sync
        if len(cls._cache) > 64:
            cls._cache.clear()
        layout = cls(batch, *known)
        cls._cache[key] = (batch, layout)
        return layout


ops._EXTRA_CACHES.append(BatchLayout._cache)


def _pad_batch(x, batch):
    """[sum N, C] -> ([B, N_max, C] zero-padded, BatchLayout or None).  batch=None means one graph: a
    view, no copy."""
    if batch is None:
        return x.unsqueeze(0), None
    lay = BatchLayout.of(batch)
    dense = x.new_zeros(lay.B * lay.n_max, x.size(-1)).index_copy(0, lay.slot, x)
    return dense.view(lay.B, lay.n_max, -1), lay


def _rows(x, lay):
    """un-pad: [B, N, R] -> [sum N, R]  (to_sparse, dgmc.py:22-23)."""
    flat = x.reshape(-1, x.size(-1))
    return flat if lay is None else flat.index_select(0, lay.slot)


def _repad(x, lay):
    """[sum N, R] -> [B * N_max, R], zeros in the padding (to_dense, dgmc.py:26-29)."""
    if lay is None:
        return x
    return x.new_zeros(lay.B * lay.n_max, x.size(-1)).index_copy(0, lay.slot, x)


def _sparse_coo(coords, values, size):
    """Un-coalesced COO tensor as the reference builds it (dgmc.py:234-240), without torch's
    one-off 'invariant checks are implicitly disabled' warning."""
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore', UserWarning)
        return torch.sparse_coo_tensor(coords, values, size, requires_grad=values.requires_grad,
                                       check_invariants=False)


class DGMC(torch.nn.Module):
    r"""Deep Graph Matching Consensus: local matching with :obj:`psi_1`, then :obj:`num_steps`
    neighbourhood-consensus refinements with :obj:`psi_2`.  Same arguments, attributes, outputs
    and metrics as the reference class (dgmc.py:64-78); see the module docstring for what runs
    where.  :obj:`psi_2` must expose :obj:`in_channels` and :obj:`out_channels`."""

    TENSOR_CORE_MIN_PAIRS = 1 << 22  # N_s * N_t below this: the one-warp-per-row exact kernel

    def __init__(self, psi_1, psi_2, num_steps, k=-1, detach=False):
        super().__init__()
        self.psi_1, self.psi_2 = psi_1, psi_2
        self.num_steps, self.k, self.detach = num_steps, k, detach
        self.backend = 'auto'  # kept for API compatibility (KeOps backend switch, dgmc.py:72)
        width = psi_2.out_channels
        self.mlp = Sequential(Linear(width, width), ReLU(), Linear(width, 1))

    def reset_parameters(self):
        self.psi_1.reset_parameters()
        self.psi_2.reset_parameters()
        for layer in self.mlp:
            if hasattr(layer, 'reset_parameters'):
                layer.reset_parameters()

    # ---------------------------------------------------------------------------------------
    def __top_k__(self, x_s, x_t):
        """[B, N_s, C], [B, N_t, C] -> int64 [B, N_s, k]: per source row the columns of the k
        largest inner products, descending (dgmc.py:85-94).  The [N_s, N_t] score matrix is never
        written to memory."""
        k = self.k
        if k > x_t.size(1):
            raise RuntimeError('k=%d exceeds the number of target nodes %d' % (k, x_t.size(1)))
        x_s, x_t = x_s.detach(), x_t.detach()
        if ops.topk_tensor_core_ok(x_s, x_t, k) and \
                x_s.size(1) * x_t.size(1) >= self.TENSOR_CORE_MIN_PAIRS:
            return ops.topk_tensor_core(x_s, x_t, k).long()
        return ops.topk_simt(x_s.contiguous(), x_t.contiguous(), k).long()

    def __include_gt__(self, S_idx, s_mask, y):
        """Make sure y[1] is among the candidates of row y[0]: rows that miss it get their LAST
        slot overwritten (dgmc.py:96-112).  The reference does this with a masked_scatter that is
        only right when y[0] is ascending and duplicate-free; for such y
        this gives the same result, with static shapes (no nonzero / host sync, so the step can be
        captured in a CUDA graph).  For other y each listed row simply receives ITS ground truth."""
        B, n_s, k = S_idx.shape
        rows, cols = y[0], y[1]
        flat = S_idx.reshape(B * n_s, k)
        if isinstance(s_mask, BatchLayout):   # flat real-node index -> padded row, no synchronisation
            rows = s_mask.slot[rows]
        elif s_mask is not None:              # the reference's signature: a boolean mask [B, N_s]
            rows = torch.nonzero(s_mask.view(-1)).view(-1)[rows]
        missing = (flat[rows] != cols.view(-1, 1)).all(dim=-1)
        out = flat.clone()
        out[rows, k - 1] = torch.where(missing, cols, flat[rows, k - 1])
        return out.view(B, n_s, k)

    # ---------------------------------------------------------------------------------------
    def _hoisted_mlp(self):
        """First consensus-MLP layer as per-node maps, padded to a multiple of 4 channels."""
        w1, b1 = self.mlp[0].weight, self.mlp[0].bias
        w2, b2 = self.mlp[2].weight.view(-1), self.mlp[2].bias
        pad = (-w1.size(0)) % 4
        if pad:
            w1, b1, w2 = F.pad(w1, (0, 0, 0, pad)), F.pad(b1, (0, pad)), F.pad(w2, (0, pad))
        return w1, b1, w2, b2

    def _psi_2_then_lin1(self, x0, merged, w1, b1, n_source, buffer=None):
        """P = Lin1(psi_2(x0)) on the merged node table, with b1 on the source rows only.  When
        psi_2 ends in its own Lin (RelCNN(lin=True)), the two linear maps have nothing in between:
        P = feats (W1 W_f)^T + W1 b_f (+ b1), one GEMM instead of two (and one input-gradient and
        one weight-gradient GEMM less per consensus step).  W1 b_f is the same for source and target
        rows and cancels in Lin1(o_s - o_t) (dgmc.py:219-221) — which is also why the reference's
        gradient of psi_2.final.bias is exactly zero; it is kept in the graph with weight 0 so that
        the parameter still receives that zero gradient."""
        from .rel import RelCNN
        psi = self.psi_2
        if isinstance(psi, RelCNN) and psi.lin:
            feats = psi(x0, merged.edge_index, merged.edge_attr, features_only=True, buffer=buffer)
            w_eff, b_eff = ops.region_cached(('lin1_final', id(self)), lambda: (
                w1 @ psi.final.weight, b1 + 0.0 * (w1 @ psi.final.bias)))
            return ops.node_linear(feats, w_eff, b_eff, key=('lin1_final', id(self)), bias_rows=n_source)
        o = psi(x0, merged.edge_index, merged.edge_attr)
        return ops.node_linear(o, w1, b1, key=('mlp0', id(self)), bias_rows=n_source)

    def forward(self, x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t, edge_attr_t,
                batch_t, y=None):
        # psi_2 and the consensus MLP are applied num_steps times with the same weights: their
        # weight gradients are reduced once per backward pass (ops.WGradGroup)
        with ops.deferred_wgrad():
            return self._forward(x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t,
                                 edge_attr_t, batch_t, y)

    def _forward(self, x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t, edge_attr_t,
                 batch_t, y=None):
        """-> (S_0, S_L): dense [sum N_s, N_t] tensors for k < 1, else un-coalesced sparse COO
        tensors carrying `.__idx__` [sum N_s, k_eff] int64 and `.__val__` [sum N_s, k_eff]
        (gradients flow through `__val__`), exactly as dgmc.py:114-244."""
        h_s = self.psi_1(x_s, edge_index_s, edge_attr_s)
        h_t = self.psi_1(x_t, edge_index_t, edge_attr_t)
        if self.detach:
            h_s, h_t = h_s.detach(), h_t.detach()

        h_s, s_lay = _pad_batch(h_s, batch_s)
        h_t, t_lay = _pad_batch(h_t, batch_t)
        s_mask = None if s_lay is None else s_lay.mask
        t_mask = None if t_lay is None else t_lay.mask
        assert h_s.size(0) == h_t.size(0), 'Encountered unequal batch-sizes'
        (B, N_s, C), N_t = h_s.size(), h_t.size(1)
        R_in = self.psi_2.in_channels
        w1, b1, w2, b2 = self._hoisted_mlp()
        num_steps = self.num_steps or 0

        merged = ops.MergedPair.of(self.psi_2, self.training, edge_index_s, edge_index_t,
                                   edge_attr_s, edge_attr_t, x_s.size(0), x_t.size(0))

        def consensus_nodes(r_s, r_t):
            """psi_2 on both graphs, then the hoisted first MLP layer -> (P_s, P_t) flat."""
            r_s, r_t = _rows(r_s, s_lay), _rows(r_t, t_lay)
            if merged is not None:
                # psi_2 shares its weights between the two graphs: run it ONCE on their disjoint
                # union (same arithmetic per node, half the launches, twice the GEMM height)
                o = self.psi_2(torch.cat([r_s, r_t], dim=0), merged.edge_index, merged.edge_attr)
                p = ops.node_linear(o, w1, key=('mlp0', id(self)))
                # b1 after re-padding: the reference applies mlp[0] to padded source rows too (their
                # o_s is 0, so they see relu(b1 - W1 o_t), dgmc.py:217-223)
                return _repad(p[:r_s.size(0)], s_lay) + b1, _repad(p[r_s.size(0):], t_lay)
            o_s = self.psi_2(r_s, edge_index_s, edge_attr_s)
            o_t = self.psi_2(r_t, edge_index_t, edge_attr_t)
            o_s, o_t = _repad(o_s, s_lay), _repad(o_t, t_lay)
            return (ops.node_linear(o_s, w1, b1, key=('mlp0s', id(self))),
                    ops.node_linear(o_t, w1, key=('mlp0t', id(self))))

        if self.k < 1:
            # ------ dense variant (dgmc.py:161-183) ------
            sm = torch.ones(B, N_s, dtype=torch.uint8, device=h_s.device) if s_mask is None \
                else s_mask.to(torch.uint8)
            tm = torch.ones(B, N_t, dtype=torch.uint8, device=h_s.device) if t_mask is None \
                else t_mask.to(torch.uint8)
            small = ops.dense_scores_ok(h_s, h_t)
            if small:
                # graphs of up to 128 nodes (every reference configuration): scores on the tensor cores
                # with the masked softmax in the epilogue, transport as small exact-fp32 batched products
                S_hat, S = ops.dense_scores(h_s, h_t, sm, tm)
            else:
                S_hat = h_s @ h_t.transpose(-1, -2)
                S = ops.masked_softmax(S_hat, sm, tm)
            S_0 = _rows(S, s_lay)
            for step in range(num_steps):
                if step > 0:
                    S = ops.masked_softmax(S_hat, sm, tm)
                r_s = torch.randn((B, N_s, R_in), dtype=h_s.dtype, device=h_s.device)
                r_t = ops.dense_transport(S, r_s) if small else S.transpose(-1, -2) @ r_s
                p_s, p_t = consensus_nodes(r_s, r_t)
                S_hat = ops.dense_consensus(S_hat, p_s.view(B, N_s, -1), p_t.view(B, N_t, -1), w2,
                                            b2, sm, tm)
            S_L = _rows(ops.masked_softmax(S_hat, sm, tm), s_lay)
            return S_0, S_L

        # ------ sparse variant (dgmc.py:184-244) ------
        S_idx = self.__top_k__(h_s, h_t)
        if self.training and y is not None:
            rnd_size = (B, N_s, min(self.k, N_t - self.k))
            S_rnd_idx = torch.randint(N_t, rnd_size, dtype=torch.long, device=S_idx.device)
            S_idx = torch.cat([S_idx, S_rnd_idx], dim=-1)
            S_idx = self.__include_gt__(S_idx, s_mask if s_lay is None else s_lay, y)

        cand = ops.Candidates(S_idx, N_t)
        k = cand.k
        S_hat = ops.gather_dot(h_s.reshape(B * N_s, C), h_t.reshape(B * N_t, C), cand)
        S_0 = _rows(ops.softmax_k(S_hat), s_lay)
        one_pair = merged is not None and s_mask is None and t_mask is None
        from .rel import RelCNN
        direct_psi_2 = one_pair and isinstance(self.psi_2, RelCNN) and self.psi_2.lin
        for _ in range(num_steps):
            S = ops.softmax_k(S_hat)
            if one_pair:
                # one un-padded pair: psi_2 and the hoisted first MLP layer on the merged node table
                # [source rows ; target rows], which K6 reads (and whose gradient its backward writes) as
                # one tensor.  The indicators are drawn, and the transport is written, straight into
                # that table (no concatenation).  b1 belongs to the source rows: added by the GEMM
                # epilogue, its gradient summed by the weight-gradient kernel.
                # ... and that table is the leading column block of psi_2's concatenated feature
                # table where psi_2 can work in one (RelCNN(cat=True, lin=True) without batch norm /
                # dropout): psi_2 then does not copy its input either
                fb = self.psi_2.feature_buffer(B * N_s + cand.targets, h_s.device) \
                    if direct_psi_2 and h_s.dtype == torch.float32 else None
                x0 = fb[:, :R_in] if fb is not None else \
                    torch.empty(B * N_s + cand.targets, R_in, dtype=h_s.dtype, device=h_s.device)
                top = x0[:B * N_s]
                r_s = torch.randn((B, N_s, R_in), dtype=h_s.dtype, device=h_s.device, out=top.view(B, N_s, R_in))
                if r_s.data_ptr() != top.data_ptr():      # a replayed draw (tests): copy it in
                    top.copy_(r_s.reshape(B * N_s, R_in))
                r_t = ops.transport(S, top, cand, out=x0[B * N_s:])
                p = self._psi_2_then_lin1(ops.join_rows(x0, top, r_t), merged, w1, b1, B * N_s, buffer=fb)
                S_hat = ops.diff_mlp_merged(S_hat, p, B * N_s, None, w2, b2, cand)
                continue
            r_s = torch.randn((B, N_s, R_in), dtype=h_s.dtype, device=h_s.device)
            r_t = ops.transport(S, r_s.view(B * N_s, R_in), cand)
            p_s, p_t = consensus_nodes(r_s, r_t)
            S_hat = ops.diff_mlp(S_hat, p_s, p_t, w2, b2, cand)
        S_L = _rows(ops.softmax_k(S_hat), s_lay)
        S_idx = _rows(S_idx, s_lay)

        # wrap as the reference does (dgmc.py:228-244)
        n_rows = x_s.size(0)
        row = torch.arange(n_rows, device=S_idx.device).view(-1, 1).expand(-1, k)
        coords = torch.stack([row.reshape(-1), S_idx.reshape(-1)], dim=0)
        size = torch.Size([n_rows, N_t])
        outs = []
        for val in (S_0, S_L):
            sp = _sparse_coo(coords, val.reshape(-1), size)
            sp.__idx__, sp.__val__ = S_idx, val
            outs.append(sp)
        return outs[0], outs[1]

    # ---------------------------------------------------------------------------------------
    def loss(self, S, y, reduction='mean'):
        """Negative log-likelihood of the ground-truth entries (dgmc.py:246-267)."""
        assert reduction in ['none', 'mean', 'sum']
        if not S.is_sparse:
            val = S[y[0], y[1]]
        else:
            assert S.__idx__ is not None and S.__val__ is not None
            if reduction != 'none' and S.__val__.is_cuda and not (torch.is_grad_enabled() and
                                                                   S.__val__.requires_grad):
                # evaluation: one kernel over the ground-truth pairs, no gathers / masks
                nll, hits, _, _ = ops.sparse_metrics(S.__idx__, S.__val__, y, eps=EPS)
                return nll.sum() if reduction == 'sum' else nll.sum() / hits.sum()
            hit = S.__idx__[y[0]] == y[1].view(-1, 1)
            if reduction != 'none':
                # same value as indexing with the boolean mask (every hit entry counts once,
                # ground truths absent from the candidate list are dropped), but with static
                # shapes: no nonzero, no host sync
                nll = -torch.log(S.__val__[y[0]] + EPS)
                total = (nll * hit).sum()
                return total if reduction == 'sum' else total / hit.sum()
            val = S.__val__[y[0]][hit]
        nll = -torch.log(val + EPS)
        return nll if reduction == 'none' else getattr(torch, reduction)(nll)

    def acc(self, S, y, reduction='mean'):
        """Share (or count) of rows y[0] whose arg-max is y[1] (dgmc.py:269-288)."""
        assert reduction in ['mean', 'sum']
        if not S.is_sparse:
            pred = S[y[0]].argmax(dim=-1)
        else:
            assert S.__idx__ is not None and S.__val__ is not None
            if S.__val__.is_cuda:
                correct = int(ops.sparse_metrics(S.__idx__, S.__val__, y)[2].sum().item())
                return correct / y.size(1) if reduction == 'mean' else correct
            pred = S.__idx__[y[0], S.__val__[y[0]].argmax(dim=-1)]
        correct = (pred == y[1]).sum().item()
        return correct / y.size(1) if reduction == 'mean' else correct

    def hits_at_k(self, k, S, y, reduction='mean'):
        """Share (or count) of rows y[0] with y[1] among their k best (dgmc.py:290-311)."""
        assert reduction in ['mean', 'sum']
        if not S.is_sparse:
            pred = S[y[0]].argsort(dim=-1, descending=True)[:, :k]
        else:
            assert S.__idx__ is not None and S.__val__ is not None
            if S.__val__.is_cuda:
                correct = int(ops.sparse_metrics(S.__idx__, S.__val__, y, top_k=k)[3].sum().item())
                return correct / y.size(1) if reduction == 'mean' else correct
            order = S.__val__[y[0]].argsort(dim=-1, descending=True)[:, :k]
            pred = torch.gather(S.__idx__[y[0]], -1, order)
        correct = (pred == y[1].view(-1, 1)).sum().item()
        return correct / y.size(1) if reduction == 'mean' else correct

    def __repr__(self):
        return ('{}(\n'
                '    psi_1={},\n'
                '    psi_2={},\n'
                '    num_steps={}, k={}\n)').format(type(self).__name__, self.psi_1, self.psi_2,
                                                    self.num_steps, self.k)
