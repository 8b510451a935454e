"""Per-kernel parity: each C-ABI entry point against a plain PyTorch restatement of the reference
op it replaces, on seeded inputs (fp32; tolerances written at each assert; index outputs
bit-exact).  Runs on the GPU box only."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = 'cuda'


def _gen(seed):
    return torch.Generator(device='cpu').manual_seed(seed)


def _randn(g, *shape):
    return torch.randn(*shape, generator=g).to(DEV)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('rows,k', [(1, 1), (7, 5), (1000, 20), (333, 50), (64, 130)])
def test_softmax_k(rows, k):
    from dgmc import ops
    g = _gen(rows + k)
    x = (_randn(g, rows, k) * 3).requires_grad_()
    dy = _randn(g, rows, k)
    y = ops.softmax_k(x)
    y.backward(dy)
    xr = x.detach().double().requires_grad_()
    yr = xr.softmax(-1)
    yr.backward(dy.double())
    assert torch.allclose(y.double(), yr, rtol=1e-5, atol=1e-7)
    assert torch.allclose(x.grad.double(), xr.grad, rtol=1e-4, atol=1e-7)


def test_masked_softmax_ragged():
    from dgmc import ops
    g = _gen(3)
    B, ns, nt = 5, 9, 37
    x = _randn(g, B, ns, nt).requires_grad_()
    sm = (torch.rand(B, ns, generator=g) > 0.3).to(DEV)
    tm = (torch.rand(B, nt, generator=g) > 0.3).to(DEV)
    tm[:, 0] = True
    y = ops.masked_softmax(x, sm.to(torch.uint8), tm.to(torch.uint8))
    dy = _randn(g, B, ns, nt)
    y.backward(dy)
    mask = sm.unsqueeze(2) & tm.unsqueeze(1)
    xr = x.detach().double().requires_grad_()
    yr = xr.masked_fill(~mask, float('-inf')).softmax(-1).masked_fill(~mask, 0)  # dgmc.py:15-19
    yr = torch.nan_to_num(yr, nan=0.0)
    keep = sm.unsqueeze(2).expand_as(yr)
    (yr * keep).backward(dy.double())
    assert torch.allclose(y.double(), yr * keep, rtol=1e-5, atol=1e-7)
    assert torch.allclose(x.grad.double(), torch.nan_to_num(xr.grad), rtol=1e-4, atol=1e-7)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('nnz,n_seg', [(0, 5), (1, 1), (1000, 7), (5000, 300), (70000, 19572),
                                       (200000, 1 << 17)])
def test_csr_build_is_a_stable_sort(nnz, n_seg):
    from dgmc import ops
    g = _gen(nnz + n_seg)
    keys = torch.randint(n_seg, (nnz, ), generator=g, dtype=torch.int32).to(DEV)
    if nnz > 100:  # a hub segment
        keys[::3] = n_seg // 2
    ptr, perm = ops.csr_build(keys, n_seg)
    order = torch.sort(keys.long(), stable=True)[1]
    assert torch.equal(perm.long(), order)
    counts = torch.bincount(keys.long(), minlength=n_seg)
    expect = torch.cat([torch.zeros(1, dtype=torch.long, device=DEV), counts.cumsum(0)])
    assert torch.equal(ptr.long(), expect)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('c', [8, 32, 64, 100, 128, 256, 300, 384, 512, 7])
def test_gather_dot_and_backward(c):
    from dgmc import ops
    g = _gen(c)
    m, nt, k = 200, 150, 12
    a = _randn(g, m, c).requires_grad_()
    b = _randn(g, nt, c).requires_grad_()
    idx = torch.randint(nt, (1, m, k), generator=g).to(DEV)
    cand = ops.Candidates(idx, nt)
    out = ops.gather_dot(a, b, cand)
    go = _randn(g, m, k)
    out.backward(go)
    ar, br = a.detach().double().requires_grad_(), b.detach().double().requires_grad_()
    ref = (ar.unsqueeze(1) * br[idx[0]]).sum(-1)  # dgmc.py:198-201
    ref.backward(go.double())
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=1e-5)
    assert torch.allclose(a.grad.double(), ar.grad, rtol=1e-5, atol=1e-5)
    assert torch.allclose(b.grad.double(), br.grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize('B,r', [(1, 32), (1, 64), (3, 8), (2, 128), (1, 6)])
def test_transport_matches_scatter_add(B, r):
    from dgmc import ops
    g = _gen(B * 100 + r)
    ns, nt, k = 90, 70, 9
    S = torch.rand(B * ns, k, generator=g).to(DEV).requires_grad_()
    r_s = _randn(g, B * ns, r).requires_grad_()
    idx = torch.randint(nt, (B, ns, k), generator=g).to(DEV)
    cand = ops.Candidates(idx, nt)
    r_t = ops.transport(S, r_s, cand)
    go = _randn(g, B * nt, r)
    r_t.backward(go)
    Sr, rr = S.detach().double().requires_grad_(), r_s.detach().double().requires_grad_()
    contrib = (rr.view(B, ns, 1, r) * Sr.view(B, ns, k, 1)).reshape(B, ns * k, r)
    ref = torch.zeros(B, nt, r, dtype=torch.double, device=DEV)
    ref = ref.scatter_add(1, idx.view(B, ns * k, 1).expand(-1, -1, r), contrib)  # dgmc.py:209-212
    ref.view(B * nt, r).backward(go.double())
    assert torch.allclose(r_t.double(), ref.view(B * nt, r), rtol=1e-5, atol=1e-6)
    assert torch.allclose(S.grad.double(), Sr.grad, rtol=1e-5, atol=1e-5)
    assert torch.allclose(r_s.grad.double(), rr.grad, rtol=1e-5, atol=1e-5)
    # determinism: by-target CSR order is fixed
    assert torch.equal(ops.transport(S, r_s, ops.Candidates(idx, nt)), r_t)


@pytest.mark.parametrize('B,r', [(1, 64), (1, 32), (2, 8), (1, 128), (1, 256)])
def test_diff_mlp_matches_reference_mlp(B, r):
    from dgmc import ops
    g = _gen(B + r)
    ns, nt, k = 80, 60, 11
    lin1 = torch.nn.Linear(r, r).to(DEV)
    lin2 = torch.nn.Linear(r, 1).to(DEV)
    o_s = _randn(g, B * ns, r).requires_grad_()
    o_t = _randn(g, B * nt, r).requires_grad_()
    s_hat = _randn(g, B * ns, k).requires_grad_()
    idx = torch.randint(nt, (B, ns, k), generator=g).to(DEV)
    cand = ops.Candidates(idx, nt)
    ps = torch.addmm(lin1.bias, o_s, lin1.weight.t())
    pt = o_t @ lin1.weight.t()
    out = ops.diff_mlp(s_hat, ps, pt, lin2.weight.view(-1), lin2.bias, cand)
    go = _randn(g, B * ns, k)
    out.backward(go)
    got = [t.grad.clone() for t in (s_hat, o_s, o_t, lin1.weight, lin1.bias, lin2.weight, lin2.bias)]
    # reference formulation, dgmc.py:219-223, in fp64
    d = lambda t: t.detach().double().requires_grad_()  # noqa: E731
    s2, os2, ot2 = d(s_hat), d(o_s), d(o_t)
    w1, b1, w2, b2 = d(lin1.weight), d(lin1.bias), d(lin2.weight), d(lin2.bias)
    D = os2.view(B, ns, 1, r) - ot2.view(B, nt, r)[torch.arange(B, device=DEV).view(B, 1, 1), idx]
    hid = torch.relu(D @ w1.t() + b1)
    ref = s2 + (hid @ w2.t() + b2).squeeze(-1).view(B * ns, k)
    ref.backward(go.double())
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=2e-5)
    for a, b_ in zip(got, (s2, os2, ot2, w1, b1, w2, b2)):
        scale = b_.grad.abs().max().item() + 1e-12
        assert (a.double() - b_.grad).abs().max().item() <= 2e-5 * scale + 1e-6


def test_dense_consensus_matches_reference_mlp():
    from dgmc import ops
    g = _gen(11)
    B, ns, nt, r = 4, 13, 17, 16
    lin1 = torch.nn.Linear(r, r).to(DEV)
    lin2 = torch.nn.Linear(r, 1).to(DEV)
    o_s = _randn(g, B, ns, r).requires_grad_()
    o_t = _randn(g, B, nt, r).requires_grad_()
    s_hat = _randn(g, B, ns, nt).requires_grad_()
    sm = (torch.rand(B, ns, generator=g) > 0.2).to(DEV)
    tm = (torch.rand(B, nt, generator=g) > 0.2).to(DEV)
    ps = o_s @ lin1.weight.t() + lin1.bias
    pt = o_t @ lin1.weight.t()
    out = ops.dense_consensus(s_hat, ps, pt, lin2.weight.view(-1), lin2.bias, sm.to(torch.uint8),
                              tm.to(torch.uint8))
    go = _randn(g, B, ns, nt)
    out.backward(go)
    got = [t.grad.clone() for t in (s_hat, o_s, o_t, lin1.weight, lin1.bias, lin2.weight, lin2.bias)]
    d = lambda t: t.detach().double().requires_grad_()  # noqa: E731
    s2, os2, ot2 = d(s_hat), d(o_s), d(o_t)
    w1, b1, w2, b2 = d(lin1.weight), d(lin1.bias), d(lin2.weight), d(lin2.bias)
    mask = sm.unsqueeze(2) & tm.unsqueeze(1)
    D = os2.unsqueeze(2) - os2.new_zeros(1) - ot2.unsqueeze(1)
    upd = (torch.relu(D @ w1.t() + b1) @ w2.t() + b2).squeeze(-1)
    ref = s2 + upd.masked_fill(~mask, 0)  # dgmc.py:178-179
    ref.backward(go.double())
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=2e-5)
    for a, b_ in zip(got, (s2, os2, ot2, w1, b1, w2, b2)):
        scale = b_.grad.abs().max().item() + 1e-12
        assert (a.double() - b_.grad).abs().max().item() <= 2e-5 * scale + 1e-6


# ---------------------------------------------------------------------------------------------
def _mean_agg(x, src, dst, n):
    out = torch.zeros(n, x.size(1), dtype=x.dtype, device=x.device).index_add_(0, dst, x[src])
    deg = torch.bincount(dst, minlength=n).clamp(min=1).to(x.dtype)
    return out / deg.unsqueeze(1)


@pytest.mark.parametrize('c', [4, 32, 64, 300, 10])
def test_rel_aggregate(c):
    from dgmc import ops
    g = _gen(c)
    n, e = 500, 2100
    x = _randn(g, n, c).requires_grad_()
    ei = torch.randint(n, (2, e), generator=g).to(DEV)
    ei[:, :50] = 3  # self-loops on a hub + isolated nodes elsewhere
    graph = ops.GraphCSR.of(ei, n)
    cat = ops.rel_aggregate(x, graph)
    go = _randn(g, n, 3 * c)
    cat.backward(go)
    xr = x.detach().double().requires_grad_()
    ref = torch.cat([xr, _mean_agg(xr, ei[0], ei[1], n), _mean_agg(xr, ei[1], ei[0], n)], 1)
    ref.backward(go.double())
    assert torch.allclose(cat.double(), ref, rtol=1e-5, atol=1e-6)
    assert torch.allclose(x.grad.double(), xr.grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize('c', [64, 256])
def test_hub_segments_reduce_identically_on_every_path(c):
    """The second pass of the balanced gathers hands a hub (a segment cut into more than 18 pieces) to
    its whole CTA, or — when the CTA's list is full — to the hub's own lane group, which adds in the
    same association: with the list switched off the result must be bit-identical, and right."""
    from dgmc import native, ops
    g = _gen(c)
    n, e = 3000, 60000
    dst = (torch.rand(e, generator=g).pow(4) * n).long().clamp_(max=n - 1)      # hubs of up to ~9,000 entries
    src = torch.randint(n, (e, ), generator=g)
    keys = dst.to(torch.int32).to(DEV)
    ptr, perm, seg_of = ops.csr_build(keys, n, True)
    assert int((ptr[1:] - ptr[:-1]).max()) > 2000
    x = _randn(g, n, c)
    w = torch.rand(e, generator=g).to(DEV)
    srcs = src.to(torch.int32).to(DEV)[perm.long()].contiguous()
    ws = w[perm.long()].contiguous()
    outs = []
    try:
        for cap in (64, 0, 1):
            native.call('dgmc_set_fixup_hub_list', cap)
            outs.append(ops.segment_gather(x, n, srcs, ptr=ptr, w=ws, seg_of=seg_of).clone())
    finally:
        native.call('dgmc_set_fixup_hub_list', 64)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    ref = torch.zeros(n, c, dtype=torch.float64, device=DEV).index_add_(
        0, dst.to(DEV), x.double()[src.to(DEV)] * w.double().unsqueeze(1))
    assert torch.allclose(outs[0].double(), ref, rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize('dim,cin,cout', [(2, 16, 32), (3, 16, 32), (1, 8, 8), (2, 12, 10)])
def test_splineconv_matches_oracle(dim, cin, cout):
    from dgmc.models.spline import SplineConv
    from oracle import dgmc_oracle as O
    g = _gen(dim * 100 + cin)
    n, e = 100, 400
    x = _randn(g, n, cin).requires_grad_()
    ei = torch.randint(n, (2, e), generator=g).to(DEV)
    pseudo = torch.rand(e, dim, generator=g).to(DEV)
    pseudo[0] = 1.0  # closed end of the open spline: cell index wraps (torch_spline_conv modulo)
    pseudo[1] = 0.0
    conv = SplineConv(cin, cout, dim).to(DEV)
    out = conv(x, ei, pseudo)
    go = _randn(g, n, cout)
    out.backward(go)
    xr = x.detach().double().cpu().requires_grad_()
    w, r, b = [p.detach().double().cpu().requires_grad_() for p in (conv.weight, conv.root, conv.bias)]
    ref = O.splineconv(xr, ei.cpu(), pseudo.double().cpu(), w, r, b)
    ref.backward(go.double().cpu())
    assert torch.allclose(out.double().cpu(), ref, rtol=1e-5, atol=1e-5)
    assert torch.allclose(x.grad.double().cpu(), xr.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(conv.weight.grad.double().cpu(), w.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(conv.root.grad.double().cpu(), r.grad, rtol=1e-4, atol=1e-5)
    assert torch.allclose(conv.bias.grad.double().cpu(), b.grad, rtol=1e-4, atol=1e-5)


def test_relcnn_matches_oracle_all_variants():
    from itertools import product
    from dgmc.models import RelCNN
    from oracle import dgmc_oracle as O
    g = _gen(5)
    x = _randn(g, 100, 16)
    ei = torch.randint(100, (2, 400), generator=g).to(DEV)
    for cat, lin in product([False, True], [False, True]):  # test/models/test_rel.py:15-20
        model = RelCNN(16, 32, 2, False, cat, lin, 0.0).to(DEV)
        out = model(x, ei)
        assert out.size() == (100, 16 + 2 * 32 if not lin and cat else 32)
        assert out.size() == (100, model.out_channels)
        spec = dict(kind='rel', out=32, layers=2, cat=cat, lin=lin, dropout=0.0)
        spec['in'] = 16
        p = {'psi.' + k: v.detach().double().cpu() for k, v in model.state_dict().items()}
        ref = O.relcnn(x.double().cpu(), ei.cpu(), p, 'psi', spec)
        assert torch.allclose(out.double().cpu(), ref, rtol=1e-5, atol=1e-5)
    bn = RelCNN(16, 32, num_layers=2, batch_norm=True, cat=True, lin=True, dropout=0.5).to(DEV)
    assert repr(bn) == 'RelCNN(16, 32, num_layers=2, batch_norm=True, cat=True, lin=True, dropout=0.5)'
    assert repr(bn.convs[0]) == 'RelConv(16, 32)'
    assert bn(x, ei).size() == (100, 32)


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('B,ns,nt,c,k', [(1, 4, 4, 16, 4), (2, 50, 70, 32, 10), (1, 300, 500, 256, 25),
                                         (1, 33, 40, 7, 40), (3, 20, 64, 64, 64)])
def test_topk_simt_exact(B, ns, nt, c, k):
    from dgmc import ops
    g = _gen(ns * nt + k)
    hs, ht = _randn(g, B, ns, c), _randn(g, B, nt, c)
    idx, score = ops.topk_simt(hs, ht, k, want_scores=True)
    full = hs.double() @ ht.double().transpose(1, 2)
    ref_v, ref_i = full.topk(k, dim=2)
    # index sets bit-exact wherever the k-th / (k+1)-th gap exceeds 1e-6 relative (north_star)
    picked = torch.gather(full, 2, idx.long())
    assert torch.allclose(picked, ref_v, rtol=1e-6, atol=1e-6)
    if k < nt:
        kth = full.topk(k + 1, dim=2)[0]
        safe = (kth[..., k - 1] - kth[..., k]) > 1e-6 * kth[..., k - 1].abs()
        same = (idx.long().sort(-1)[0] == ref_i.sort(-1)[0]).all(-1)
        assert bool(same[safe].all())
    assert torch.allclose(score.double(), picked, rtol=1e-5, atol=1e-5)
    assert bool((score[..., :-1] >= score[..., 1:]).all())  # descending


# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize('r', [8, 64, 128])
def test_balanced_kernels_with_hub_targets(r):
    """Half of all candidates point at one hub target, many targets are empty: the entry-parallel
    kernels must agree with the one-group-per-segment kernels and with scatter_add."""
    from dgmc import native, ops
    g = _gen(r)
    ns, nt, k = 700, 300, 13
    idx = torch.randint(nt // 2, (1, ns, k), generator=g)
    idx[0, :, ::2] = 7  # the hub
    idx = idx.to(DEV)
    cand = ops.Candidates(idx, nt)
    tptr, perm, seg_of = cand.csr
    S = torch.rand(ns, k, generator=g).to(DEV)
    r_s = _randn(g, ns, r)
    plain = torch.empty(nt, r, device=DEV)
    native.call('dgmc_segment_gather', tptr, 0, perm, k, perm, S, None, None, r_s, r, plain, r, nt, r, 0,
                native.stream_ptr())
    bal = ops.segment_gather(r_s, nt, perm, ptr=tptr, src_div=k, w_index=perm, w=S.view(-1), seg_of=seg_of)
    ref = torch.zeros(nt, r, dtype=torch.double, device=DEV).index_add_(
        0, idx.view(-1), (r_s.double().unsqueeze(1) * S.double().unsqueeze(2)).view(-1, r))
    assert torch.allclose(bal.double(), ref, rtol=1e-5, atol=1e-4)
    # the one-group-per-segment kernel adds the hub's ~4,500 terms into ONE fp32 accumulator, the
    # balanced one in blocks of 32: they differ by fp32 summation order only
    assert torch.allclose(bal, plain, rtol=1e-4, atol=2e-3)
    assert torch.equal(bal, ops.segment_gather(r_s, nt, perm, ptr=tptr, src_div=k, w_index=perm,
                                               w=S.view(-1), seg_of=seg_of))  # deterministic
    # accumulate into a strided column slice
    wide = torch.ones(nt, 3 * r, device=DEV)
    ops.segment_gather(r_s, nt, perm, ptr=tptr, src_div=k, w_index=perm, w=S.view(-1), seg_of=seg_of,
                       out=wide[:, r:2 * r], accumulate=True)
    assert torch.allclose(wide[:, r:2 * r].double(), ref + 1, rtol=1e-5, atol=1e-4)
    assert bool((wide[:, :r] == 1).all()) and bool((wide[:, 2 * r:] == 1).all())
    # K6 backward, target side: both entry points
    grad, ps, pt, w2 = _randn(g, ns, k), _randn(g, ns, r), _randn(g, nt, r), _randn(g, r)
    a, b = torch.empty(nt, r, device=DEV), torch.empty(nt, r, device=DEV)
    native.call('dgmc_diff_mlp_bwd_targets', grad, ps, pt, tptr, perm, w2, a, nt, k, r, native.stream_ptr())
    ws = torch.empty(native.query('dgmc_balanced_workspace', perm.numel(), r), dtype=torch.uint8, device=DEV)
    native.call('dgmc_diff_mlp_bwd_targets_balanced', grad, ps, pt, tptr, seg_of, perm, w2, b, nt,
                perm.numel(), k, r, ws, ws.numel(), native.stream_ptr())
    scale = a.abs().max().item()
    assert (a - b).abs().max().item() <= 2e-5 * scale + 1e-6


@pytest.mark.parametrize('k,m,n', [(39000, 64, 192), (5000, 64, 256), (3000, 8, 24), (2500, 33, 70),
                                   (100000, 64, 64)])
@pytest.mark.parametrize('tensor_core', [False, True])
def test_node_linear_wgrad_matches_autograd(k, m, n, tensor_core, monkeypatch):
    from dgmc import ops
    monkeypatch.setattr(ops, 'WGRAD_TENSOR_CORE', tensor_core)  # fp32 FFMA split-K / tcgen05 3xTF32
    g = _gen(k + m + n)
    x = _randn(g, k, n).requires_grad_()
    w = (_randn(g, m, n) * 0.1).requires_grad_()
    b = _randn(g, m).requires_grad_()
    go = _randn(g, k, m)
    out = ops.node_linear(x, w, b)
    out.backward(go)
    xr, wr, br = [t.detach().double().requires_grad_() for t in (x, w, b)]
    ref = torch.nn.functional.linear(xr, wr, br)
    ref.backward(go.double())
    assert torch.allclose(out.double(), ref, rtol=1e-5, atol=1e-5)
    for a, r_ in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        scale = r_.abs().max().item()
        assert (a.double() - r_).abs().max().item() <= 2e-5 * scale + 1e-6
    # strided input (a column slice) and no bias
    wide = _randn(g, k, n + 8)
    xs = wide[:, 4:4 + n].detach().requires_grad_()
    w2 = w.detach().clone().requires_grad_()
    ops.node_linear(xs, w2).backward(go)
    ref_w = go.double().t() @ wide[:, 4:4 + n].double()
    assert (w2.grad.double() - ref_w).abs().max().item() <= 2e-5 * ref_w.abs().max().item() + 1e-6


# ---------------------------------------------------------------------------------------------
def test_relu_bwd_strided():
    from dgmc import native
    g = _gen(21)
    wide_g, wide_y = _randn(g, 3000, 96), _randn(g, 3000, 160)
    gs, ys = wide_g[:, 32:80], wide_y[:, 64:112]           # column blocks: pitch != width
    out = torch.empty(3000, 48, device=DEV)
    native.call('dgmc_relu_bwd', native.RowStrided(gs), gs.stride(0), native.RowStrided(ys), ys.stride(0),
                out, out.stride(0), 3000, 48, native.stream_ptr())
    assert torch.equal(out, torch.where(ys > 0, gs, torch.zeros_like(gs)))


@pytest.mark.parametrize('lin', [True, False])
def test_relcnn_direct_cat_matches_layerwise(lin, monkeypatch):
    """RelCNN writing its layers straight into the concatenated buffer (and accumulating the two
    gradients of every layer input inside the gather kernel) against the plain layer-by-layer path:
    same kernels on the same values, so forward is bit-identical and gradients agree to rounding."""
    from dgmc.models import RelCNN
    g = _gen(31)
    n, c_in, c_out = 3000, 16, 32
    x0 = _randn(g, n, c_in)
    ei = torch.randint(n, (2, 4 * n), generator=g).to(DEV)
    go = _randn(g, n, c_out if lin else c_in + 3 * c_out)
    torch.manual_seed(3)
    model = RelCNN(c_in, c_out, 3, cat=True, lin=lin).to(DEV)

    def run():
        x = x0.clone().requires_grad_()
        model.zero_grad()
        y = model(x, ei)
        y.backward(go)
        return y.detach(), x.grad.clone(), [p.grad.clone() for p in model.parameters() if p.grad is not None]

    assert model._direct_cat(x0)
    y1, dx1, gp1 = run()
    monkeypatch.setattr(RelCNN, '_direct_cat', lambda self, x: False)
    y2, dx2, gp2 = run()
    assert torch.equal(y1, y2)
    assert torch.allclose(dx1, dx2, rtol=1e-5, atol=1e-7)
    assert len(gp1) == len(gp2)
    for a, b in zip(gp1, gp2):
        assert torch.allclose(a, b, rtol=1e-5, atol=1e-6 * b.abs().max().item())


def test_sparse_metrics_kernel_matches_the_reference_formulas():
    """loss / acc / hits_at_k on the side-cars in one kernel (dgmc_sparse_metrics) against the oracle's
    restatement of dgmc.py:246-311, with the cases the reference formulas treat specially: ground truths
    missing from the list (dropped from the mean), duplicated candidates (every copy counts)."""
    from dgmc import ops
    from dgmc.models import DGMC, GIN
    from oracle import dgmc_oracle as O
    g = torch.Generator().manual_seed(17)
    rows, k, n_t, pairs = 500, 20, 300, 260
    idx = torch.randint(n_t, (rows, k), generator=g)
    idx[:, 5] = idx[:, 2]                                    # a duplicated candidate in every row
    val = torch.rand(rows, k, generator=g).softmax(-1)       # distinct values: argsort order is unambiguous
    y_row = torch.sort(torch.randperm(rows, generator=g)[:pairs])[0]
    y_col = idx[y_row, torch.randint(k, (pairs, ), generator=g)]
    y_col[::7] = (idx[y_row[::7]].max(dim=1)[0] + 1) % n_t   # some ground truths are not listed at all
    y = torch.stack([y_row, y_col])
    model = DGMC(GIN(4, 4, 1), GIN(4, 4, 1), num_steps=0, k=k)
    S = torch.sparse_coo_tensor(torch.zeros(2, 0, dtype=torch.long), torch.zeros(0), (rows, n_t)).cuda()
    S.__idx__, S.__val__ = idx.cuda(), val.cuda()
    yc = y.cuda()
    with torch.no_grad():
        for red in ('mean', 'sum'):
            assert torch.allclose(model.loss(S, yc, red).cpu(), O.loss(val, y, red, idx=idx), rtol=1e-5)
    assert model.acc(S, yc) == pytest.approx(O.acc(val, y, idx=idx))
    assert model.acc(S, yc, 'sum') == O.acc(val, y, 'sum', idx=idx)
    for top in (1, 3, 10, 20):
        assert model.hits_at_k(top, S, yc) == pytest.approx(O.hits_at_k(top, val, y, idx=idx)), top
    nll, hits, correct, in_top = ops.sparse_metrics(S.__idx__, S.__val__, yc, top_k=k)
    assert torch.equal(hits.cpu(), (idx[y_row] == y_col.view(-1, 1)).sum(1).float())
    assert torch.equal(in_top, hits)                          # top_k = k: every listed copy is in
    # with gradients the loss goes through differentiable ops and must agree with the kernel's value
    S.__val__ = val.cuda().requires_grad_()
    loss = model.loss(S, yc)
    loss.backward()
    assert torch.allclose(loss.detach().cpu(), O.loss(val, y, idx=idx), rtol=1e-5)
    assert S.__val__.grad is not None


@pytest.mark.parametrize('B,n_s,n_t,c', [(512, 20, 20, 256), (1, 100, 100, 64), (3, 7, 5, 8), (5, 128, 128, 32),
                                          (70, 17, 23, 36), (2, 128, 3, 300)])
def test_dense_scores_tensor_core_matches_fp64(B, n_s, n_t, c):
    """S_hat = h_s h_t^T and S_0 = masked_softmax(S_hat) (dgmc.py:163-165) from the block-diagonal
    tcgen05 kernel against fp64, ragged masks, and the adjoints through the small batched products.
    Tolerance: fp32 level (3xTF32) — 2e-6 of the largest score, 1e-5 relative on S_0 (+1e-7)."""
    from dgmc import ops
    g = torch.Generator().manual_seed(B + n_s + c)
    sizes_s = torch.randint(max(1, n_s // 2), n_s + 1, (B, ), generator=g)
    sizes_t = torch.randint(max(1, n_t // 2), n_t + 1, (B, ), generator=g)
    sizes_s[0], sizes_t[0] = n_s, n_t
    sm = (torch.arange(n_s).view(1, -1) < sizes_s.view(-1, 1))
    tm = (torch.arange(n_t).view(1, -1) < sizes_t.view(-1, 1))
    h_s = torch.randn(B, n_s, c, generator=g) * sm.unsqueeze(-1)          # padded rows are zero (to_dense_batch)
    h_t = torch.randn(B, n_t, c, generator=g) * tm.unsqueeze(-1)
    g_hat, g_0 = torch.randn(B, n_s, n_t, generator=g), torch.randn(B, n_s, n_t, generator=g)

    def ref(dtype):
        a, b = h_s.to(dtype).requires_grad_(), h_t.to(dtype).requires_grad_()
        s_hat = a @ b.transpose(-1, -2)
        mask = sm.unsqueeze(2) & tm.unsqueeze(1)
        s_0 = s_hat.masked_fill(~mask, float('-inf')).softmax(-1).masked_fill(~mask, 0)
        s_0 = torch.nan_to_num(s_0, nan=0.0)
        ((s_hat * g_hat.to(dtype)).sum() + (s_0 * g_0.to(dtype)).sum()).backward()
        return s_hat.detach(), s_0.detach(), a.grad, b.grad

    r = ref(torch.float64)
    a, b = h_s.cuda().requires_grad_(), h_t.cuda().requires_grad_()
    assert ops.dense_scores_ok(a, b)
    s_hat, s_0 = ops.dense_scores(a, b, sm.cuda().to(torch.uint8), tm.cuda().to(torch.uint8))
    ((s_hat * g_hat.cuda()).sum() + (s_0 * g_0.cuda()).sum()).backward()
    scale = r[0].abs().max().item()
    assert (s_hat.detach().double().cpu() - r[0]).abs().max().item() <= 2e-6 * scale
    err0 = (s_0.detach().double().cpu() - r[1]).abs()
    assert bool((err0 <= 1e-5 * r[1].abs() + 4e-6 * scale * r[1].abs() + 1e-7).all())   # d softmax = S * d logit
    for got, want in ((a.grad, r[2]), (b.grad, r[3])):
        assert (got.double().cpu() - want).abs().max().item() <= 1e-5 * want.abs().max().item()


def test_small_batched_products_match_bmm():
    from dgmc import ops
    g = torch.Generator().manual_seed(4)
    for B, m, k, n in ((512, 20, 20, 128), (3, 100, 100, 32), (7, 5, 128, 6), (2, 128, 9, 301)):
        a = torch.randn(B, m, k, generator=g).cuda()
        at = torch.randn(B, k, m, generator=g).cuda()
        bm = torch.randn(B, k, n, generator=g).cuda()
        bt = torch.randn(B, n, k, generator=g).cuda()
        for got, want in ((ops._bmm_nn(a, bm), a.double() @ bm.double()),
                          (ops._bmm_nn(at, bm, trans_a=True), at.double().transpose(1, 2) @ bm.double()),
                          (ops._bmm_nt(a, bt), a.double() @ bt.double().transpose(1, 2))):
            assert (got.double() - want).abs().max().item() <= 2e-6 * want.abs().max().item() * max(1, k ** 0.5)
