"""K1 tensor-core top-k (dgmc_topk_tc) against an fp64 restatement of DGMC.__top_k__
(dgmc.py:85-94, via oracle.topk_rows).  Index sets must be bit-exact wherever the k-th/(k+1)-th
score gap exceeds 1e-6 relative (north_star); returned scores are exact fp32 (rtol 1e-5)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _planted(gen, n_s, n_t, c):
    hs = torch.randn(n_s, c, generator=gen)
    ht = torch.randn(n_t, c, generator=gen)
    m = min(n_s, n_t)
    perm = torch.randperm(n_t, generator=gen)[:m]
    ht[perm] = hs[:m] + 0.5 * torch.randn(m, c, generator=gen)
    return hs.cuda(), ht.cuda()


def _check(hs, ht, k, idx, score):
    from oracle import dgmc_oracle as O
    n_t = ht.size(0)
    full = hs.double() @ ht.double().t()
    ref_v, ref_i = full.topk(min(k + 1, n_t), dim=1)
    picked = torch.gather(full, 1, idx.long())
    assert int(idx.min()) >= 0 and int(idx.max()) < n_t
    assert torch.allclose(score.double(), picked, rtol=1e-5, atol=1e-4)
    assert bool((score[:, :-1] >= score[:, 1:]).all())
    if k < n_t:
        gap_ok = (ref_v[:, k - 1] - ref_v[:, k]) > 1e-6 * ref_v[:, k - 1].abs()
    else:
        gap_ok = torch.ones(hs.size(0), dtype=torch.bool, device=hs.device)
    same = (idx.long().sort(1)[0] == ref_i[:, :k].sort(1)[0]).all(1)
    bad = int((~same & gap_ok).sum())
    assert bad == 0, '%d of %d rows differ from the exact top-%d set' % (bad, hs.size(0), k)
    # oracle's own (fp64) index path agrees too
    oi = O.topk_rows(hs.cpu()[None], ht.cpu()[None], k, dtype=torch.float64)[0]
    same2 = (idx.cpu().long().sort(1)[0] == oi.sort(1)[0]).all(1)
    assert int((~same2 & gap_ok.cpu()).sum()) == 0


@pytest.mark.parametrize('n_s,n_t,c,k', [
    (300, 500, 256, 10),      # one CTA per tile, partial row block, partial column tile
    (128, 128, 64, 1),
    (1000, 1500, 64, 25),     # KP = 32
    (700, 900, 300, 10),      # C padded to 320 -> MH = 1
    (513, 2049, 512, 16),     # widest C
    (4000, 6000, 256, 10),    # stream-K runs cross row blocks (several partial lists per row)
    (19388, 19572, 256, 10),  # BASELINE.json configs[2] shape
])
def test_topk_tc_matches_exact(n_s, n_t, c, k):
    from dgmc import ops
    g = torch.Generator().manual_seed(n_s + n_t + c + k)
    hs, ht = _planted(g, n_s, n_t, c)
    assert ops.topk_tensor_core_ok(hs[None], ht[None], k)
    idx, score = ops.topk_tensor_core(hs[None], ht[None], k, want_scores=True)
    _check(hs, ht, k, idx[0], score[0])


def test_topk_tc_agrees_with_simt_kernel():
    from dgmc import ops
    g = torch.Generator().manual_seed(5)
    hs, ht = _planted(g, 640, 777, 128)
    a = ops.topk_tensor_core(hs[None], ht[None], 10)
    b = ops.topk_simt(hs[None], ht[None], 10)
    assert torch.equal(a.sort(-1)[0], b.sort(-1)[0])


def test_unsupported_shapes_are_reported():
    from dgmc import native, ops
    hs, ht = torch.randn(1, 200, 600).cuda(), torch.randn(1, 300, 600).cuda()
    assert not ops.topk_tensor_core_ok(hs, ht, 10)        # C > 512
    assert not ops.topk_tensor_core_ok(hs[..., :64], ht[..., :64], 40)   # k > 26
    with pytest.raises(native.DgmcNativeError, match='unsupported shape'):
        ops.topk_tensor_core(hs, ht, 10)
