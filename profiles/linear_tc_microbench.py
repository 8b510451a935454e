"""Micro-benchmark: dgmc_linear_tc (tcgen05 3xTF32) against the library SGEMM at the GEMM shapes of the
DBP15K- and Pascal-shaped steps: error against fp64 and device time per launch inside a replayed CUDA
graph (host launch overhead excluded).  python profiles/linear_tc_microbench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc.native import call, query, stream_ptr, RowStrided
torch.manual_seed(0)
dev = 'cuda'
def run(x, w, b, relu, trans=False):
    rows, k_in = x.shape
    n_out = w.shape[1] if trans else w.shape[0]
    ws = torch.empty(query('dgmc_linear_tc_workspace', n_out, k_in), dtype=torch.uint8, device=dev)
    call('dgmc_linear_tc_split_weight', w, w.stride(0), int(trans), n_out, k_in, ws, ws.numel(), stream_ptr())
    y = torch.empty(rows, n_out, device=dev)
    call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, b, int(relu), y, n_out, rows, n_out, k_in, stream_ptr())
    return y, ws
for rows, n_out, k_in, relu, trans in [(257, 50, 72, True, False), (38960, 64, 192, True, False), (38960, 192, 64, False, True), (35000, 64, 192, True, False), (35000, 64, 256, False, True), (35000, 192, 64, False, True), (35000, 256, 64, False, True),
                                       (34500, 256, 900, True, False), (34500, 256, 1068, False, False), (34500, 768, 256, False, True),
                                       (10000, 6400, 256, False, False)]:
    x = torch.randn(rows, k_in, device=dev)
    w = torch.randn(k_in, n_out, device=dev) * 0.05 if trans else torch.randn(n_out, k_in, device=dev) * 0.05
    b = torch.randn(n_out, device=dev)
    y, ws = run(x, w, b, relu, trans)
    torch.cuda.synchronize()
    wm = w if trans else w.t()
    ref64 = x.double() @ wm.double() + b.double()
    lib = torch.addmm(b, x, wm)
    if relu: ref64, lib = ref64.relu(), lib.relu()
    scale = ref64.abs().max().item()
    e_tc = (y.double() - ref64).abs().max().item() / scale
    e_lib = (lib.double() - ref64).abs().max().item() / scale
    def t(fn, n=20):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(n): fn()
        g.replay(); torch.cuda.synchronize()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); c.record(); torch.cuda.synchronize()
        return a.elapsed_time(c) / n * 1e3
    y2 = torch.empty_like(y)
    t_tc = t(lambda: call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, b, int(relu), y2, n_out, rows, n_out, k_in, stream_ptr()))
    t_lib = t(lambda: torch._addmm_activation(b, x, wm, use_gelu=False) if relu else torch.addmm(b, x, wm))
    fl = 2.0 * rows * n_out * k_in
    print('rows %6d n_out %4d k_in %4d relu %d trans %d: err tc %.2e lib %.2e | tc %.1f us (%.0f TF/s eff) lib %.1f us (%.0f TF/s)' % (
        rows, n_out, k_in, relu, trans, e_tc, e_lib, t_tc, fl / t_tc / 1e6, t_lib, fl / t_lib / 1e6), flush=True)
