"""Launches the tensor-core GEMM kernels (dgmc_linear_tc, dgmc_wgrad_tc_partial) at the shapes of the
DBP15K training step, once each after a warm-up, so that
    ncu --set full --clock-control none --import-source on -k regex:'linear_tc_kernel|wgrad_tc_kernel' \
        -o gpurun_out/<name> python profiles/gemm_tc_shapes.py
captures one launch per shape.  Not part of the product or the tests."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), '..',
                                'deep-graph-matching-consensus_b200'))
from dgmc.native import RowStrided, call, query, stream_ptr  # noqa: E402

DEV = 'cuda'
torch.manual_seed(0)

# (rows, n_out, k_in, relu, transposed): psi_2 RelConv fwd, its input gradient, psi_1 RelConv fwd
for rows, n_out, k_in, relu, trans in [(38960, 64, 192, True, False), (38960, 192, 64, False, True),
                                       (19572, 256, 900, True, False)]:
    x = torch.randn(rows, k_in, device=DEV)
    w = torch.randn(k_in, n_out, device=DEV) * 0.05 if trans else torch.randn(n_out, k_in, device=DEV) * 0.05
    b = torch.randn(n_out, device=DEV)
    ws = torch.empty(query('dgmc_linear_tc_workspace', n_out, k_in), dtype=torch.uint8, device=DEV)
    call('dgmc_linear_tc_split_weight', w, w.stride(0), int(trans), n_out, k_in, ws, ws.numel(), stream_ptr())
    y = torch.empty(rows, n_out, device=DEV)
    for _ in range(2):
        call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, b, int(relu), y, n_out, rows, n_out, k_in,
             stream_ptr())
    torch.cuda.synchronize()

# (k nodes, m out, n in): psi_2 RelConv weight gradient
for k, m, n in [(38960, 64, 192)]:
    g = torch.randn(k, m, device=DEV)
    x = torch.randn(k, n, device=DEV)
    splits = query('dgmc_wgrad_tc_splits', k, m, n)
    partial = torch.empty(splits, m * n, device=DEV)
    pb = torch.empty(splits, m, device=DEV)
    for _ in range(2):
        call('dgmc_wgrad_tc_partial', g, g.stride(0), RowStrided(x), x.stride(0), k, m, n, partial, pb,
             stream_ptr())
    torch.cuda.synchronize()
