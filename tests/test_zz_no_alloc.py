"""dgmc_diff_mlp_bwd_rows_noalloc — the entry point for callers that cannot pass a workspace
(a serving process that owns the whole device through its own arena, or a loop where cudaMalloc
is banned because it serialises the device).  It must give what dgmc_diff_mlp_bwd_rows gives with
a caller-allocated workspace.  Called straight through the C ABI, twice in a row with different
inputs, so that whatever the first call left in the library's built-in scratch is there when the
second one runs.  Collected last on purpose (file name): a run that stops at the first failure
has already executed every other test."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(native, seed, m, n_t, k, r):
    g = torch.Generator().manual_seed(seed)
    dev = 'cuda'
    grad = torch.randn(m, k, generator=g).to(dev)
    ps = torch.randn(m, r, generator=g).to(dev)
    pt = torch.randn(n_t, r, generator=g).to(dev)
    idx = torch.randint(n_t, (m, k), generator=g, dtype=torch.int32).to(dev)
    w2 = torch.randn(r, generator=g).to(dev)
    stream = native.stream_ptr()

    ws = torch.empty(native.query('dgmc_diff_mlp_bwd_workspace', m, r), dtype=torch.uint8, device=dev)
    want = [torch.empty(m, r, device=dev), torch.empty(r, device=dev), torch.empty(1, device=dev)]
    native.call('dgmc_diff_mlp_bwd_rows', grad, ps, pt, idx, w2, *want, m, k, r, ws, ws.numel(), stream)

    got = [torch.full((m, r), float('nan'), device=dev), torch.full((r, ), float('nan'), device=dev),
           torch.full((1, ), float('nan'), device=dev)]
    native.call('dgmc_diff_mlp_bwd_rows_noalloc', grad, ps, pt, idx, w2, *got, m, k, r, stream)
    torch.cuda.synchronize()
    return want, got


def test_noalloc_matches_workspace_variant_on_two_calls_in_a_row():
    from dgmc import native
    # tolerance of tests/test_gpu_kernels.py::test_diff_mlp_matches_reference_mlp (2e-5 of the
    # largest entry); the two entry points run the same kernels, so they normally agree exactly
    for seed, (m, n_t, k, r) in enumerate([(3000, 500, 20, 64), (1777, 300, 7, 128)]):
        want, got = _both(native, seed, m, n_t, k, r)
        for a, b in zip(want, got):
            scale = a.abs().max().item() + 1e-12
            assert torch.isfinite(b).all()
            assert (a - b).abs().max().item() <= 2e-5 * scale + 1e-6


def test_noalloc_rejects_sizes_beyond_its_scratch():
    from dgmc import native
    dev = 'cuda'
    m, k, r = 8, 3, 132
    z = lambda *s: torch.zeros(*s, device=dev)  # noqa: E731
    idx = torch.zeros(m, k, dtype=torch.int32, device=dev)
    with pytest.raises(native.DgmcNativeError, match='exceeds the built-in scratch'):
        native.call('dgmc_diff_mlp_bwd_rows_noalloc', z(m, k), z(m, r), z(4, r), idx, z(r), z(m, r),
                    z(r), z(1), m, k, r, native.stream_ptr())
