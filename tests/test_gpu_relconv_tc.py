"""Fused RelConv kernel (dgmc_relconv_tc: aggregation gathered straight into the tensor-core GEMM's
operand stage) against an fp64 restatement of RelConv.forward (rel.py:25-31) written with plain
torch ops, forward and all gradients; and against the two-kernel route it replaces.

Tolerance: the kernel is 3xTF32 (fp32-level, like dgmc_linear_tc): max error <= 2e-6 of the largest
output entry against fp64, gradients <= 1e-5 of their largest entry."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fused_on():
    """The fused route is opt-in (it is slower than the two-kernel route, see ops.RELCONV_FUSED)."""
    from dgmc import ops
    old, ops.RELCONV_FUSED = ops.RELCONV_FUSED, True
    yield
    ops.RELCONV_FUSED = old


def _graph(n, e, g, hub=0):
    ei = torch.randint(n, (2, e), generator=g)
    if hub:   # one node with `hub` in-edges and `hub` out-edges: its tile overflows the staged entry lists
        extra = torch.randint(n, (hub, ), generator=g)
        ei = torch.cat([ei, torch.stack([extra, torch.full((hub, ), 5)]),
                        torch.stack([torch.full((hub, ), 5), extra])], dim=1)
    return ei


def _reference(x, ei, w_root, w1, w2, bias, relu):
    """rel.py:25-31 in the given dtype: root(x) + mean_{j->i} lin1(x_j) + mean_{i->j} lin2(x_j)."""
    n = x.size(0)
    src, dst = ei[0], ei[1]

    def mean(msg, index):
        acc = torch.zeros(n, msg.size(1), dtype=x.dtype, device=x.device).index_add_(0, index, msg)
        deg = torch.bincount(index, minlength=n).clamp(min=1).to(x.dtype)
        return acc / deg.unsqueeze(1)
    y = x @ w_root.t() + bias + mean((x @ w1.t())[src], dst) + mean((x @ w2.t())[dst], src)
    return torch.relu(y) if relu else y


@pytest.mark.parametrize('n,c,c_out,edges,hub,relu', [
    (100, 32, 16, 300, 0, True),          # less than one tile, padded output channels
    (128, 64, 64, 600, 0, False),
    (1000, 64, 64, 4000, 0, True),
    (1000, 32, 40, 4000, 0, True),        # one 32-column slice, output channels not a multiple of 32
    (4737, 64, 64, 19000, 0, False),      # 149 groups of 32 rows: one CTA gets a second, partial tile
    (3000, 64, 64, 9000, 2500, True),     # hub tile: operator entries read from global memory
    (38960, 64, 64, 165556, 0, True),     # psi_2 on the merged DBP15K pair: 305 tiles on 148 CTAs
    (40000, 32, 32, 160000, 0, True),     # reference default rnd_dim = 32 (examples/dbp15k.py:12)
])
def test_relconv_fused_matches_fp64_and_unfused(n, c, c_out, edges, hub, relu):
    from dgmc import ops
    g = torch.Generator().manual_seed(n + c)
    ei = _graph(n, edges, g, hub)
    x = torch.randn(n, c, generator=g)
    w_root, w1, w2 = (torch.randn(c_out, c, generator=g) / c ** 0.5 for _ in range(3))
    bias = torch.randn(c_out, generator=g)
    gy, gp = torch.randn(n, c_out, generator=g), torch.randn(n, c, generator=g)

    def run(fn, dtype, dev):
        xs = x.to(dev, dtype).requires_grad_()
        ws = [t.to(dev, dtype).requires_grad_() for t in (w_root, w1, w2)]
        b = bias.to(dev, dtype).requires_grad_()
        y, alias = fn(xs, ws, b)
        ((y * gy.to(dev, dtype)).sum() + (alias * gp.to(dev, dtype)).sum()).backward()
        return [y.detach()] + [t.grad for t in [xs] + ws + [b]]

    ref = run(lambda xs, ws, b: (_reference(xs, ei, ws[0], ws[1], ws[2], b, relu), xs), torch.float64, 'cpu')

    graph = ops.GraphCSR(ei.cuda(), n)
    rel = graph.rel   # build the operator (and what only the fused kernel reads) now: not part of the layer
    rel.t_ptr, rel.f_order, rel.t_order

    def fused(xs, ws, b):
        assert ops.relconv_fused_ok(xs, graph.rel, c_out)
        return ops.relconv_fused(xs, torch.cat(ws, dim=1), b, graph.rel, relu=relu)

    def unfused(xs, ws, b):
        stacked, alias = ops.rel_aggregate_pass(xs, graph)
        return ops.node_linear(stacked, torch.cat(ws, dim=1), b, relu=relu), alias

    launches = ops.native.lib.dgmc_launch_count()
    got = run(fused, torch.float32, 'cuda')
    # forward kernel, ReLU backward, backward kernel, weight-gradient partial + reduce, two weight splits
    assert ops.native.lib.dgmc_launch_count() - launches <= 8
    old = run(unfused, torch.float32, 'cuda')
    names = ['y', 'dx', 'dw_root', 'dw1', 'dw2', 'db']
    for name, a, o, r in zip(names, got, old, ref):
        scale = r.abs().max().item()
        err = (a.double().cpu() - r).abs().max().item()
        err_old = (o.double().cpu() - r).abs().max().item()
        tol = 2e-6 if name == 'y' else 1e-5
        # ReLU makes y's gradient discontinuous: an output within rounding of 0 may flip its mask
        assert err <= max(tol * scale, 2 * err_old), (name, err, err_old, scale)


def test_relconv_fused_into_column_block_and_eval_mode():
    """out= a column block of a wider buffer (RelCNN writes its layers straight into the concatenated
    feature buffer), no gradient requested -> no operand rows stored."""
    from dgmc import ops
    g = torch.Generator().manual_seed(3)
    n, c = 777, 64
    ei = _graph(n, 3000, g)
    x = torch.randn(n, c, generator=g).cuda()
    w = (torch.randn(c, 3 * c, generator=g) / c ** 0.5).cuda()
    b = torch.randn(c, generator=g).cuda()
    graph = ops.GraphCSR(ei.cuda(), n)
    buf = torch.full((n, 3 * c), 7.0, device='cuda')
    with torch.no_grad():
        y, _ = ops.relconv_fused(x, w, b, graph.rel, relu=True, out=buf[:, c:2 * c])
        ref = ops.node_linear(ops.rel_aggregate(x, graph), w, b, relu=True)
    assert y.data_ptr() == buf[:, c:2 * c].data_ptr()
    assert torch.allclose(buf[:, c:2 * c], ref, rtol=1e-5, atol=1e-5)
    assert bool((buf[:, :c] == 7.0).all()) and bool((buf[:, 2 * c:] == 7.0).all())   # neighbours untouched


def test_relcnn_uses_the_fused_kernel_and_matches_the_unfused_route():
    from dgmc import ops
    from dgmc.models import RelCNN
    g = torch.Generator().manual_seed(9)
    n = 5000
    ei = _graph(n, 20000, g).cuda()
    x = torch.randn(n, 64, generator=g).cuda()
    torch.manual_seed(0)
    psi = RelCNN(64, 64, 3).cuda()
    grads = {}
    for flag in (True, False):
        ops.RELCONV_FUSED = flag
        try:
            psi.zero_grad()
            xs = x.clone().requires_grad_()
            with ops.deferred_wgrad():
                out = psi(xs, ei) + psi(2 * xs, ei)       # two applications: grouped weight gradients
            out.square().sum().backward()
            grads[flag] = [out.detach(), xs.grad] + [p.grad.clone() for p in psi.parameters() if p.grad is not None]
        finally:
            ops.RELCONV_FUSED = True   # the autouse fixture restores the default afterwards
    for a, b in zip(grads[True], grads[False]):
        assert (a - b).abs().max().item() <= 2e-5 * b.abs().max().item() + 1e-7
