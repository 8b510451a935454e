"""Edge cases of the kernels and of the module surface: empty edge lists, isolated nodes, k = 1 and
k = N_t, zero-row calls, candidate lists that all point at one target, non-multiple-of-4 widths."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = 'cuda'


def test_relcnn_on_graph_without_edges():
    from dgmc.models import RelCNN
    torch.manual_seed(0)
    model = RelCNN(12, 16, 2).to(DEV)
    x = torch.randn(50, 12, device=DEV, requires_grad=True)
    empty = torch.zeros(2, 0, dtype=torch.long, device=DEV)
    out = model(x, empty)
    out.sum().backward()
    # without edges RelConv reduces to its root map (rel.py:31 with empty aggregations)
    h = x.detach()
    feats = [h]
    for conv in model.convs:
        h = torch.relu(conv.root(feats[-1]))
        feats.append(h)
    ref = model.final(torch.cat(feats, -1))
    assert torch.allclose(out, ref, rtol=1e-5, atol=1e-6)
    assert torch.isfinite(x.grad).all()


def test_zero_row_calls_are_noops():
    from dgmc import ops
    empty = torch.empty(0, 7, device=DEV)
    assert ops.softmax_k(empty).shape == (0, 7)
    ptr, perm = ops.csr_build(torch.empty(0, dtype=torch.int32, device=DEV), 5)
    assert ptr.tolist() == [0] * 6 and perm.numel() == 0
    out = ops.segment_gather(torch.randn(4, 8, device=DEV), 5, perm, ptr=ptr,
                             seg_of=torch.empty(1, dtype=torch.int32, device=DEV))
    assert out.shape == (5, 8) and bool((out == 0).all())


@pytest.mark.parametrize('k', [1, 37])
def test_sparse_forward_k_extremes(k):
    from dgmc.models import DGMC, RelCNN
    torch.manual_seed(1)
    n_s, n_t = 29, 37
    model = DGMC(RelCNN(6, 10, 1), RelCNN(6, 6, 1), num_steps=2, k=k).to(DEV).eval()
    x_s, x_t = torch.randn(n_s, 6, device=DEV), torch.randn(n_t, 6, device=DEV)
    e_s = torch.randint(n_s, (2, 60), device=DEV)
    e_t = torch.randint(n_t, (2, 80), device=DEV)
    with torch.no_grad():
        S_0, S_L = model(x_s, e_s, None, None, x_t, e_t, None, None)
    assert S_L.__idx__.shape == (n_s, k)
    assert torch.allclose(S_L.__val__.sum(-1), torch.ones(n_s, device=DEV), atol=1e-5)
    if k == n_t:  # every column is a candidate: the sparse result is the dense one (test_dgmc.py:46-60)
        model.k = -1
        torch.manual_seed(5)
        with torch.no_grad():
            D_0, _ = model(x_s, e_s, None, None, x_t, e_t, None, None)
        assert torch.allclose(S_0.to_dense(), D_0, atol=1e-5)
    with pytest.raises(RuntimeError):
        model.k = n_t + 1
        model(x_s, e_s, None, None, x_t, e_t, None, None)


def test_all_candidates_point_at_one_target():
    from dgmc import ops
    g = torch.Generator().manual_seed(3)
    ns, nt, k, r = 500, 40, 6, 12      # r = 12: float4 rows, but not a power-of-two lane group
    idx = torch.full((1, ns, k), 17, dtype=torch.long, device=DEV)
    cand = ops.Candidates(idx, nt)
    S = torch.rand(ns, k, generator=g).to(DEV)
    r_s = torch.randn(ns, r, generator=g).to(DEV)
    r_t = ops.transport(S, r_s, cand)
    ref = (r_s.double() * S.double().sum(1, keepdim=True)).sum(0)
    assert torch.allclose(r_t[17].double(), ref, rtol=1e-5, atol=1e-4)
    mask = torch.ones(nt, dtype=torch.bool, device=DEV)
    mask[17] = False
    assert bool((r_t[mask] == 0).all())


def test_include_gt_rows_keep_their_own_ground_truth():
    from dgmc.models import DGMC, GIN
    model = DGMC(GIN(4, 4, 1), GIN(4, 4, 1), num_steps=0, k=2)
    S_idx = torch.tensor([[[0, 1], [2, 3], [4, 5], [6, 7]]], device=DEV)
    y = torch.tensor([[0, 2, 3], [1, 9, 6]], device=DEV)     # row 0 and row 3 already contain theirs
    out = model.__include_gt__(S_idx, None, y)
    assert out.tolist() == [[[0, 1], [2, 3], [4, 9], [6, 7]]]


def test_batched_forward_backward_makes_no_host_synchronisation():
    """with the sizes attached by the collation (dgmc.utils.collate_pairs) a batched
    call — dense or sparse, forward, loss and backward — never waits for the device: no boolean
    indexing, no .item(), no bincount (checked with torch's sync debug mode)."""
    from dgmc.models import DGMC, SplineCNN
    from dgmc.utils.data import PairData, collate_pairs
    g = torch.Generator().manual_seed(1)

    def graph(n):
        return dict(x=torch.randn(n, 16, generator=g), ei=torch.randint(n, (2, 4 * n), generator=g),
                    ea=torch.rand(4 * n, 2, generator=g))
    pairs = []
    for _ in range(12):
        n = int(torch.randint(5, 12, (1, ), generator=g))
        a, b = graph(n), graph(n + 1)
        pairs.append(PairData(x_s=a['x'], edge_index_s=a['ei'], edge_attr_s=a['ea'], x_t=b['x'],
                              edge_index_t=b['ei'], edge_attr_t=b['ea'], y=torch.randperm(n + 1, generator=g)[:n],
                              num_nodes=None))
    data = collate_pairs(pairs).to('cuda')
    y = torch.stack([torch.arange(data.y.numel(), device='cuda'), data.y])
    torch.manual_seed(0)
    for k in (-1, 4):
        model = DGMC(SplineCNN(16, 32, 2, 2, cat=False, dropout=0.5), SplineCNN(8, 8, 2, 2), num_steps=3, k=k).cuda().train()
        args = (data.x_s, data.edge_index_s, data.edge_attr_s, data.x_s_batch, data.x_t, data.edge_index_t,
                data.edge_attr_t, data.x_t_batch)
        model(*args, y)[1]           # warm-up: lazy one-time set-up (kernel attributes, cuBLAS handles)
        torch.cuda.synchronize()
        torch.cuda.set_sync_debug_mode('error')
        try:
            S_0, S_L = model(*args, y)
            loss = model.loss(S_0, y) + model.loss(S_L, y)
            loss.backward()
        finally:
            torch.cuda.set_sync_debug_mode('default')
        assert torch.isfinite(loss).item()
