"""CPU ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A functional, dtype-generic (fp32 / fp64) restatement in plain PyTorch-on-CPU of the
reference's correspondence-and-consensus path, written so it can travel to the GPU box
(where /root/reference does not exist).  Every function cites the reference lines it
follows (paths relative to /root/reference/).

Who may import this file: `tests/`, `__graft_entry__.smoke()`, and `bench.py`'s
`cpu_baseline` / `--impl reference` legs.  The product package
(`deep-graph-matching-consensus_b200/`) never imports it and has no CPU fallback.

Parity status: PINNED.  `oracle/make_golden.py` imports the UNMODIFIED reference package
(through the third-party stand-ins in `oracle/shims/`) in the build container, records its
inputs, random draws, outputs, losses and gradients into `tests/golden/*.pt`, and
`tests/test_oracle_golden.py` checks this restatement against those vectors (and, when
/root/reference is present, against the live reference).  What stays unpinned by the
reference itself: KeOps argKmin (never run by its CI) and the numeric values
of RelConv / SplineConv / GINConv, whose arithmetic lives in third-party PyG code that is
not in the tree; those follow the published semantics restated in `oracle/shims/`.

Model description: a "psi spec" is a dict
    {'kind': 'rel'|'spline'|'gin', 'in': int, 'out': int, 'layers': int,
     'batch_norm': bool, 'cat': bool, 'lin': bool, 'dropout': float, 'dim': int}
and parameters come as a flat {name: tensor} dict using the reference's state_dict names
(`psi_1.convs.0.lin1.weight`, `mlp.0.weight`, ...).  Random draws are INJECTED (`draws`)
rather than generated, so that oracle, reference and CUDA path see identical indicator
functions and negatives whatever device or dtype they run in.  The functions are device-agnostic
plain torch: tests may also evaluate them in fp32 ON the GPU (ATen / cuBLAS arithmetic) to measure
the reference formulas' own fp32 noise on that hardware (parity mode (i)).
"""
import math

import torch
import torch.nn.functional as F

EPS = 1e-8  # dgmc.py:12


# --------------------------------------------------------------------------------------
# third-party semantics used by the path (restated; see oracle/shims for the stand-ins)
# --------------------------------------------------------------------------------------
def segment_mean(msg, index, num_nodes):
    """PyG MessagePassing(aggr='mean') reduction: sum / max(count, 1)  (rel.py:9, [3p])."""
    acc = msg.new_zeros((num_nodes, msg.size(1)))
    acc.index_add_(0, index, msg)
    deg = torch.bincount(index, minlength=num_nodes).clamp(min=1).to(msg.dtype)
    return acc / deg.unsqueeze(1)


def pad_batch(x, batch):
    """torch_geometric.utils.to_dense_batch as called at dgmc.py:154-155.

    -> (dense [B, N_max, C] zero-filled, mask [B, N_max] bool)."""
    if batch is None:
        return x.unsqueeze(0), torch.ones(1, x.size(0), dtype=torch.bool, device=x.device)
    B = int(batch.max()) + 1
    sizes = torch.bincount(batch, minlength=B)
    n_max = int(sizes.max())
    start = torch.cumsum(sizes, 0) - sizes
    local = torch.arange(batch.numel(), device=x.device) - start[batch]
    dense = x.new_zeros(B, n_max, x.size(1))
    mask = torch.zeros(B, n_max, dtype=torch.bool, device=x.device)
    dense[batch, local] = x
    mask[batch, local] = True
    return dense, mask


# --------------------------------------------------------------------------------------
# psi backbones
# --------------------------------------------------------------------------------------
def _linear(x, p, name, bias=True):
    w = p[name + '.weight']
    b = p.get(name + '.bias') if bias else None
    return F.linear(x, w, b)


def relconv(x, edge_index, p, pre):
    """RelConv.forward, rel.py:25-31: root(x) + mean_{j->i} lin1(x_j) + mean_{i->j} lin2(x_j)."""
    src, dst = edge_index[0], edge_index[1]
    n = x.size(0)
    a = F.linear(x, p[pre + '.lin1.weight'])
    b = F.linear(x, p[pre + '.lin2.weight'])
    incoming = segment_mean(a[src], dst, n)  # flow source_to_target (rel.py:27-28)
    outgoing = segment_mean(b[dst], src, n)  # flow target_to_source (rel.py:29-30)
    return _linear(x, p, pre + '.root') + incoming + outgoing


def _batch_norm(x, p, pre, training):
    return F.batch_norm(x, p.get(pre + '.running_mean'), p.get(pre + '.running_var'),
                        p[pre + '.weight'], p[pre + '.bias'], training=training,
                        momentum=0.0, eps=1e-5)


def _drop(x, rate, training, dropout_fn):
    if not training or rate == 0.0:
        return x
    if dropout_fn is None:
        raise ValueError('oracle: training-mode dropout needs an injected dropout_fn')
    return dropout_fn(x, rate)


def relcnn(x, edge_index, p, pre, spec, training=False, dropout_fn=None):
    """RelCNN.forward, rel.py:80-92."""
    feats = [x]
    for layer in range(spec['layers']):
        h = F.relu(relconv(feats[-1], edge_index, p, '%s.convs.%d' % (pre, layer)))
        if spec.get('batch_norm', False):
            h = _batch_norm(h, p, '%s.batch_norms.%d' % (pre, layer), training)
        feats.append(_drop(h, spec.get('dropout', 0.0), training, dropout_fn))
    out = torch.cat(feats, dim=-1) if spec.get('cat', True) else feats[-1]
    return _linear(out, p, pre + '.final') if spec.get('lin', True) else out


def spline_basis(pseudo, kernel_size):
    """Degree-1 open B-spline basis of torch_spline_conv, as SplineConv(kernel_size=5) uses it
    (spline.py:21, [3p]).  -> basis [E, 2^dim], weight index [E, 2^dim]."""
    E, dim = pseudo.shape
    pos = pseudo * (kernel_size - 1)
    cell = pos.floor()
    t = pos - cell
    cell = cell.long()
    n_corner = 1 << dim
    basis = pseudo.new_ones(E, n_corner)
    widx = torch.zeros(E, n_corner, dtype=torch.long)
    for corner in range(n_corner):
        stride = 1
        for d in range(dim):
            up = (corner >> d) & 1
            basis[:, corner] = basis[:, corner] * (t[:, d] if up else 1 - t[:, d])
            widx[:, corner] += ((cell[:, d] + up) % kernel_size) * stride
            stride *= kernel_size
    return basis, widx


def splineconv(x, edge_index, pseudo, weight, root, bias, kernel_size=5):
    """PyG SplineConv forward ([3p], constructed at spline.py:21): mean_j sum_s b_s x_j W[idx_s]
    + x root + bias.  Computed transform-first: Y = x W for all K^dim kernels, then gather."""
    src, dst = edge_index[0], edge_index[1]
    n, K = x.size(0), weight.size(0)
    basis, widx = spline_basis(pseudo, kernel_size)
    y = torch.einsum('ni,kio->nko', x, weight)  # [N, K, out]
    msg = x.new_zeros(src.numel(), weight.size(2))
    for corner in range(basis.size(1)):
        msg = msg + basis[:, corner:corner + 1] * y[src, widx[:, corner]]
    out = segment_mean(msg, dst, n)
    return out + x @ root + bias


def splinecnn(x, edge_index, edge_attr, p, pre, spec, training=False, dropout_fn=None):
    """SplineCNN.forward, spline.py:44-54.  Parameter names: convs.i.{weight,root,bias}."""
    feats = [x]
    for layer in range(spec['layers']):
        c = '%s.convs.%d' % (pre, layer)
        feats.append(F.relu(splineconv(feats[-1], edge_index, edge_attr, p[c + '.weight'],
                                       p[c + '.root'], p[c + '.bias'])))
    out = torch.cat(feats, dim=-1) if spec.get('cat', True) else feats[-1]
    out = _drop(out, spec.get('dropout', 0.0), training, dropout_fn)
    return _linear(out, p, pre + '.final') if spec.get('lin', True) else out


def mlp_stack(x, p, pre, num_layers, batch_norm, training):
    """MLP.forward, mlp.py:31-39 (dropout is 0 wherever the reference builds one, gin.py:22)."""
    for i in range(num_layers):
        x = _linear(x, p, '%s.lins.%d' % (pre, i))
        if i < num_layers - 1:
            x = F.relu(x)
            if batch_norm:
                x = _batch_norm(x, p, '%s.batch_norms.%d' % (pre, i), training)
    return x


def gin(x, edge_index, p, pre, spec, training=False):
    """GIN.forward, gin.py:44-53; GINConv ([3p]): nn((1+eps) x + sum_j x_j)."""
    src, dst = edge_index[0], edge_index[1]
    feats = [x]
    for layer in range(spec['layers']):
        c = '%s.convs.%d' % (pre, layer)
        h = feats[-1]
        agg = h.new_zeros(h.shape).index_add_(0, dst, h[src])
        h = (1 + p[c + '.eps']) * h + agg
        feats.append(mlp_stack(h, p, c + '.nn', 2, spec.get('batch_norm', False), training))
    out = torch.cat(feats, dim=-1) if spec.get('cat', True) else feats[-1]
    return _linear(out, p, pre + '.final') if spec.get('lin', True) else out


def psi(spec, x, edge_index, edge_attr, p, pre, training=False, dropout_fn=None):
    kind = spec['kind']
    if kind == 'rel':
        return relcnn(x, edge_index, p, pre, spec, training, dropout_fn)
    if kind == 'spline':
        return splinecnn(x, edge_index, edge_attr, p, pre, spec, training, dropout_fn)
    if kind == 'gin':
        return gin(x, edge_index, p, pre, spec, training)
    raise ValueError(kind)


def psi_out_channels(spec):
    """out_channels attribute logic shared by rel.py:65-70, spline.py:30-35, gin.py:32-37."""
    if spec.get('lin', True):
        return spec['out']
    return spec['in'] + spec['layers'] * spec['out'] if spec.get('cat', True) else spec['out']


# --------------------------------------------------------------------------------------
# correspondence + consensus
# --------------------------------------------------------------------------------------
def topk_rows(h_s, h_t, k, chunk=2048, dtype=None):
    """DGMC.__top_k__ (dgmc.py:85-94): indices of the k largest <h_s[b,i], h_t[b,j]> per row,
    descending.  The reference's non-KeOps route materialises [B,N_s,N_t,C]; this is the same
    arithmetic done per row chunk as a matmul (SURVEY.md §8c "at-scale caveat").  `dtype`
    lets tests ask for an fp64 index oracle."""
    hs = h_s if dtype is None else h_s.to(dtype)
    ht = h_t if dtype is None else h_t.to(dtype)
    out = []
    for lo in range(0, hs.size(1), chunk):
        scores = hs[:, lo:lo + chunk] @ ht.transpose(1, 2)
        out.append(scores.topk(k, dim=2)[1])
    return torch.cat(out, dim=1)


def include_gt(S_idx, s_mask, y):
    """DGMC.__include_gt__ (dgmc.py:96-112), including its ordering quirk: rows that miss their
    ground truth are visited in ascending flat order and consume `y[1][missing]` in y order."""
    B, N_s, k = S_idx.shape
    row, col = y[0], y[1]
    flat = S_idx[s_mask]  # [sum N_s, k]
    missing = (flat[row] != col.unsqueeze(1)).all(dim=1)
    needs = torch.zeros(flat.size(0), dtype=torch.bool)
    needs[row] = missing
    padded = torch.zeros(B, N_s, dtype=torch.bool)
    padded[s_mask] = needs
    targets = torch.nonzero(padded.view(-1)).view(-1)  # ascending padded-row order
    values = col[missing][:targets.numel()]
    out = S_idx.clone().view(B * N_s, k)
    out[targets, k - 1] = values
    return out.view(B, N_s, k)


def _masked_softmax(logits, mask):
    """masked_softmax, dgmc.py:15-19."""
    z = logits.masked_fill(~mask, float('-inf')).softmax(dim=-1)
    return z.masked_fill(~mask, 0)


def _consensus_mlp(d, p):
    """self.mlp = Lin(R,R) . ReLU . Lin(R,1), dgmc.py:74-78."""
    hidden = F.relu(F.linear(d, p['mlp.0.weight'], p['mlp.0.bias']))
    return F.linear(hidden, p['mlp.2.weight'], p['mlp.2.bias']).squeeze(-1)


def _unpad(x, mask):
    return x[mask]  # to_sparse, dgmc.py:22-23


def _repad(x, mask):
    out = x.new_zeros(tuple(mask.shape) + (x.size(-1), ))  # to_dense, dgmc.py:26-29
    out[mask] = x
    return out


def dgmc_forward(spec1, spec2, p, x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t,
                 edge_attr_t, batch_t, num_steps, k=-1, y=None, training=False, detach=False,
                 draws=None, dropout_fn=None, topk_dtype=None, S_idx=None):
    """DGMC.forward, dgmc.py:114-244.

    draws: {'r_s': [tensor [B,N_s,R_in]] * num_steps, 'neg': tensor [B,N_s,m] or None}
    Returns {'S_0','S_L'} as dense [sum N_s, N_t] tensors (k < 1) or, in the sparse variant,
    {'idx': [sum N_s, k_eff] int64, 'S_0': [sum N_s, k_eff], 'S_L': ..., 'N_t': int}
    (the `__idx__` / `__val__` side-cars of dgmc.py:236-242).  `S_idx` lets a test impose a
    candidate set (e.g. the one the CUDA top-k picked) instead of recomputing top-k.
    """
    draws = draws or {}
    h_s = psi(spec1, x_s, edge_index_s, edge_attr_s, p, 'psi_1', training, dropout_fn)
    h_t = psi(spec1, x_t, edge_index_t, edge_attr_t, p, 'psi_1', training, dropout_fn)
    if detach:
        h_s, h_t = h_s.detach(), h_t.detach()  # dgmc.py:152
    h_s, s_mask = pad_batch(h_s, batch_s)
    h_t, t_mask = pad_batch(h_t, batch_t)
    assert h_s.size(0) == h_t.size(0)
    B, N_s, _ = h_s.shape
    N_t = h_t.size(1)

    def run_psi2(r_s, r_t):
        o_s = psi(spec2, _unpad(r_s, s_mask), edge_index_s, edge_attr_s, p, 'psi_2', training)
        o_t = psi(spec2, _unpad(r_t, t_mask), edge_index_t, edge_attr_t, p, 'psi_2', training)
        return _repad(o_s, s_mask), _repad(o_t, t_mask)

    if k < 1:  # dense variant, dgmc.py:161-183
        pair_mask = s_mask.unsqueeze(2) & t_mask.unsqueeze(1)
        logits = h_s @ h_t.transpose(1, 2)
        S_0 = _masked_softmax(logits, pair_mask)[s_mask]
        for step in range(num_steps):
            S = _masked_softmax(logits, pair_mask)
            r_s = draws['r_s'][step].to(h_s.dtype)
            r_t = S.transpose(1, 2) @ r_s
            o_s, o_t = run_psi2(r_s, r_t)
            d = o_s.unsqueeze(2) - o_t.unsqueeze(1)
            logits = logits + _consensus_mlp(d, p).masked_fill(~pair_mask, 0)
        S_L = _masked_softmax(logits, pair_mask)[s_mask]
        return {'S_0': S_0, 'S_L': S_L}

    # sparse variant, dgmc.py:184-244
    if S_idx is None:
        S_idx = topk_rows(h_s, h_t, k, dtype=topk_dtype)
        if training and y is not None:  # dgmc.py:190-195
            neg = draws.get('neg')
            if neg is not None and neg.numel() > 0:
                S_idx = torch.cat([S_idx, neg], dim=-1)
            S_idx = include_gt(S_idx, s_mask, y)
    k_eff = S_idx.size(-1)
    rows_b = torch.arange(B, device=h_s.device).view(B, 1, 1)
    cand = h_t[rows_b, S_idx]  # [B, N_s, k_eff, C]   (gather, dgmc.py:198-200)
    logits = (h_s.unsqueeze(2) * cand).sum(dim=-1)  # dgmc.py:201
    S_0 = logits.softmax(dim=-1)[s_mask]
    for step in range(num_steps):
        S = logits.softmax(dim=-1)
        r_s = draws['r_s'][step].to(h_s.dtype)
        contrib = (r_s.unsqueeze(2) * S.unsqueeze(3)).reshape(B, N_s * k_eff, -1)
        r_t = r_s.new_zeros(B, N_t, r_s.size(-1))
        r_t.scatter_add_(1, S_idx.reshape(B, N_s * k_eff, 1).expand_as(contrib), contrib)
        o_s, o_t = run_psi2(r_s, r_t)
        d = o_s.unsqueeze(2) - o_t[rows_b, S_idx]  # dgmc.py:219-222
        logits = logits + _consensus_mlp(d, p)
    S_L = logits.softmax(dim=-1)[s_mask]
    return {'idx': S_idx[s_mask], 'S_0': S_0, 'S_L': S_L, 'N_t': N_t}


def sparse_to_dense(idx, val, n_cols):
    """What `.to_dense()` of the un-coalesced COO tensor of dgmc.py:229-240 yields
    (duplicate (row, col) entries add up)."""
    out = val.new_zeros(idx.size(0), n_cols)
    out.scatter_add_(1, idx, val)
    return out


# --------------------------------------------------------------------------------------
# loss / metrics  (dgmc.py:246-311)
# --------------------------------------------------------------------------------------
def loss(S, y, reduction='mean', idx=None):
    """DGMC.loss, dgmc.py:246-267.  Dense: S [rows, N_t].  Sparse: pass S=val and idx."""
    if idx is None:
        val = S[y[0], y[1]]
    else:
        hit = idx[y[0]] == y[1].unsqueeze(1)
        val = S[y[0]][hit]
    nll = -torch.log(val + EPS)
    return nll if reduction == 'none' else getattr(torch, reduction)(nll)


def acc(S, y, reduction='mean', idx=None):
    """DGMC.acc, dgmc.py:269-288."""
    if idx is None:
        pred = S[y[0]].argmax(dim=-1)
    else:
        pred = idx[y[0], S[y[0]].argmax(dim=-1)]
    correct = int((pred == y[1]).sum())
    return correct / y.size(1) if reduction == 'mean' else correct


def hits_at_k(k, S, y, reduction='mean', idx=None):
    """DGMC.hits_at_k, dgmc.py:290-311."""
    if idx is None:
        pred = S[y[0]].argsort(dim=-1, descending=True)[:, :k]
    else:
        order = S[y[0]].argsort(dim=-1, descending=True)[:, :k]
        pred = torch.gather(idx[y[0]], -1, order)
    correct = int((pred == y[1].unsqueeze(1)).sum())
    return correct / y.size(1) if reduction == 'mean' else correct


# --------------------------------------------------------------------------------------
# parameter initialisation helper (used by tests / bench to build random-init models
# without the reference): same distributions as torch.nn.Linear.reset_parameters.
# --------------------------------------------------------------------------------------
def init_linear(gen, out_f, in_f, bias=True, dtype=torch.float32):
    bound = 1.0 / math.sqrt(in_f)
    w = (torch.rand(out_f, in_f, generator=gen, dtype=dtype) * 2 - 1) * bound
    b = (torch.rand(out_f, generator=gen, dtype=dtype) * 2 - 1) * bound if bias else None
    return w, b


def init_psi(gen, spec, pre, dtype=torch.float32):
    p = {}
    c_in = spec['in']
    for layer in range(spec['layers']):
        c = '%s.convs.%d' % (pre, layer)
        if spec['kind'] == 'rel':
            p[c + '.lin1.weight'], _ = init_linear(gen, spec['out'], c_in, False, dtype)
            p[c + '.lin2.weight'], _ = init_linear(gen, spec['out'], c_in, False, dtype)
            p[c + '.root.weight'], p[c + '.root.bias'] = init_linear(gen, spec['out'], c_in,
                                                                     True, dtype)
        elif spec['kind'] == 'spline':
            K = 5**spec['dim']
            bound = 1.0 / math.sqrt(c_in * K)
            p[c + '.weight'] = (torch.rand(K, c_in, spec['out'], generator=gen, dtype=dtype) * 2
                                - 1) * bound
            p[c + '.root'] = (torch.rand(c_in, spec['out'], generator=gen, dtype=dtype) * 2
                              - 1) * bound
            p[c + '.bias'] = (torch.rand(spec['out'], generator=gen, dtype=dtype) * 2 - 1) * bound
        else:
            raise ValueError(spec['kind'])
        c_in = spec['out']
    if spec.get('lin', True):
        f_in = spec['in'] + spec['layers'] * spec['out'] if spec.get('cat', True) else spec['out']
        p[pre + '.final.weight'], p[pre + '.final.bias'] = init_linear(gen, spec['out'], f_in,
                                                                       True, dtype)
    return p


def init_dgmc(seed, spec1, spec2, dtype=torch.float32):
    gen = torch.Generator().manual_seed(seed)
    p = {}
    p.update(init_psi(gen, spec1, 'psi_1', dtype))
    p.update(init_psi(gen, spec2, 'psi_2', dtype))
    R = psi_out_channels(spec2)
    p['mlp.0.weight'], p['mlp.0.bias'] = init_linear(gen, R, R, True, dtype)
    p['mlp.2.weight'], p['mlp.2.bias'] = init_linear(gen, 1, R, True, dtype)
    return p
