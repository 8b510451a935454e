"""dgmc_linear_tc (tcgen05 kind::tf32 with hi/lo splitting) against fp64: the per-node GEMM behind
RelConv / SplineConv / Lin (rel.py:27-30, spline.py:40-47).  The bar is "fp32-level": the error
against the fp64 product must stay within 3x that of the library's IEEE fp32 SGEMM on the same
operands (plus a floor of 2e-7 of the largest output), at every shape — including long reductions,
where a single truncating accumulator would drift (see linear_tc.cu)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = 'cuda'


def _run(x, w, bias, relu, transposed):
    from dgmc.native import RowStrided, call, query, stream_ptr
    rows, k_in = x.shape
    n_out = w.shape[1] if transposed else w.shape[0]
    assert query('dgmc_linear_tc_supported', rows, n_out, k_in, x.stride(0), x.data_ptr())
    ws = torch.empty(query('dgmc_linear_tc_workspace', n_out, k_in), dtype=torch.uint8, device=DEV)
    call('dgmc_linear_tc_split_weight', w, w.stride(0), int(transposed), n_out, k_in, ws, ws.numel(),
         stream_ptr())
    y = torch.full((rows, n_out), float('nan'), device=DEV)
    call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, bias, int(relu), y, n_out, rows, n_out, k_in,
         stream_ptr())
    return y


@pytest.mark.parametrize('rows,n_out,k_in,relu,transposed,bias', [
    (1, 1, 4, False, False, True),          # smallest
    (257, 50, 72, True, False, True),       # ragged rows / columns, n_out not a multiple of 4
    (130, 130, 36, False, True, False),     # two column tiles, second nearly empty; k pad
    (35000, 64, 192, True, False, True),    # psi_2 RelConv at the DBP15K shape
    (35000, 64, 256, False, True, False),   # its input gradient
    (20000, 256, 900, True, False, True),   # psi_1 RelConv (300-d features, three aggregates)
    (20000, 256, 1068, False, False, True), # psi_1 final Lin
    (20000, 768, 256, False, True, False),  # psi_1 input gradient
    (4000, 6400, 256, False, True, False),  # SplineConv transform (25 kernels x 256)
    (3000, 96, 4096, False, False, True),   # long reduction
])
def test_linear_tc_fp32_level(rows, n_out, k_in, relu, transposed, bias):
    g = torch.Generator().manual_seed(rows + n_out + k_in)
    x = torch.randn(rows, k_in, generator=g).to(DEV)
    w = (torch.randn(k_in, n_out, generator=g) if transposed else torch.randn(n_out, k_in, generator=g)).to(DEV)
    b = torch.randn(n_out, generator=g).to(DEV) if bias else None
    y = _run(x, w, b, relu, transposed)
    w_kn = w if transposed else w.t()
    ref = x.double() @ w_kn.double()
    lib = x @ w_kn
    if bias:
        ref, lib = ref + b.double(), lib + b
    if relu:
        ref, lib = ref.relu(), lib.relu()
    assert torch.isfinite(y).all()
    scale = ref.abs().max().item()
    err = (y.double() - ref).abs().max().item()
    err_lib = (lib.double() - ref).abs().max().item()
    assert err <= 3 * err_lib + 2e-7 * scale, (err / scale, err_lib / scale)


def test_linear_tc_strided_input_and_zero_padding():
    """x is a column slice of a wider buffer (pitch != k_in); the columns next to it hold NaN, which
    must not leak in through the k padding (TMA zero-fills beyond k_in)."""
    g = torch.Generator().manual_seed(5)
    rows, n_out, k_in = 1000, 40, 44
    wide = torch.full((rows, 64), float('nan'), device=DEV)
    wide[:, 4:4 + k_in] = torch.randn(rows, k_in, generator=g).to(DEV)
    x = wide[:, 4:4 + k_in]
    w = torch.randn(n_out, k_in, generator=g).to(DEV)
    y = _run(x, w, None, False, False)
    ref = x.double() @ w.double().t()
    assert (y.double() - ref).abs().max().item() <= 2e-6 * ref.abs().max().item()


def test_linear_tc_rejects_misaligned():
    from dgmc.native import DgmcNativeError, RowStrided, call, query, stream_ptr
    x = torch.randn(100, 30, device=DEV)          # pitch of 30 floats: not 16-byte aligned
    assert not query('dgmc_linear_tc_supported', 100, 8, 30, x.stride(0), x.data_ptr())
    ws = torch.empty(query('dgmc_linear_tc_workspace', 8, 30), dtype=torch.uint8, device=DEV)
    y = torch.empty(100, 8, device=DEV)
    with pytest.raises(DgmcNativeError, match='unsupported'):
        call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, None, 0, y, 8, 100, 8, 30, stream_ptr())


@pytest.mark.parametrize('kn', [False, True])
@pytest.mark.parametrize('relu', [False, True])
@pytest.mark.parametrize('rows,n_in,n_out', [(5000, 96, 72),      # small weight: grouped split-K path
                                             (3000, 300, 260),    # > 64K weights: chunks of 128 channels
                                             (600, 40, 24)])      # few rows: library weight gradient
def test_node_linear_autograd(kn, relu, rows, n_in, n_out):
    """ops.node_linear (forward + input gradient on the tensor-core kernel, weight gradient on the
    tensor-core split-K kernel or the library) against fp64 autograd, both weight layouts."""
    from dgmc import ops
    g = torch.Generator().manual_seed(11)
    x = torch.randn(rows, n_in, generator=g).to(DEV).requires_grad_()
    w = (torch.randn(n_in, n_out, generator=g) * 0.1 if kn else torch.randn(n_out, n_in, generator=g) * 0.1).to(DEV).requires_grad_()
    b = torch.randn(n_out, generator=g).to(DEV).requires_grad_()
    go = torch.randn(rows, n_out, generator=g).to(DEV)
    y = ops.node_linear(x, w, b, relu=relu, kn=kn)
    y.backward(go)
    xr, wr, br = [t.detach().double().requires_grad_() for t in (x, w, b)]
    ref = xr @ (wr if kn else wr.t()) + br
    if relu:
        ref = ref.relu()
    ref.backward(go.double())
    assert (y.double() - ref).abs().max().item() <= 2e-6 * ref.abs().max().item()
    for a, r_ in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert (a.double() - r_).abs().max().item() <= 2e-5 * r_.abs().max().item() + 1e-6


@pytest.mark.parametrize('rows,n_in,n_out,tc_wgrad', [(5000, 64, 64, True),    # tensor-core forward + weight gradient
                                                      (5000, 64, 64, False),   # FFMA weight gradient
                                                      (700, 32, 16, True)])    # few rows: library weight gradient
def test_node_linear_bias_on_leading_rows(rows, n_in, n_out, tc_wgrad, monkeypatch):
    """bias_rows: the bias (first consensus-MLP layer on the merged node table) belongs to the
    source rows only — forward and bias gradient against fp64 autograd."""
    from dgmc import ops
    monkeypatch.setattr(ops, 'WGRAD_TENSOR_CORE', tc_wgrad)
    g = torch.Generator().manual_seed(17)
    n_b = rows // 3
    x = torch.randn(rows, n_in, generator=g).to(DEV).requires_grad_()
    w = (torch.randn(n_out, n_in, generator=g) * 0.1).to(DEV).requires_grad_()
    b = torch.randn(n_out, generator=g).to(DEV).requires_grad_()
    go = torch.randn(rows, n_out, generator=g).to(DEV)
    y = ops.node_linear(x, w, b, bias_rows=n_b)
    y.backward(go)
    xr, wr, br = [t.detach().double().requires_grad_() for t in (x, w, b)]
    ref = xr @ wr.t()
    ref = torch.cat([ref[:n_b] + br, ref[n_b:]], dim=0)
    ref.backward(go.double())
    assert (y.double() - ref).abs().max().item() <= 2e-6 * ref.abs().max().item()
    for a, r_ in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert (a.double() - r_).abs().max().item() <= 2e-5 * r_.abs().max().item() + 1e-6
