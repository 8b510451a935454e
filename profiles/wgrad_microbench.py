"""Micro-benchmark: fp32 FFMA split-K weight gradient (wgrad.cu) against the tcgen05 3xTF32 kernel
(wgrad_tc.cu) at the DBP15K psi_2 shape (K = 39,000 nodes, 64 x 192 output).  python profiles/wgrad_microbench.py"""
import sys, torch
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'deep-graph-matching-consensus_b200'))
from dgmc import ops, native
g = torch.randn(39000, 64, device='cuda'); x = torch.randn(39000, 192, device='cuda')
def run(tc):
    fs, fp = ('dgmc_wgrad_tc_splits', 'dgmc_wgrad_tc_partial') if tc else ('dgmc_wgrad_splits', 'dgmc_wgrad_partial')
    sp = native.query(fs, 39000, 64, 192)
    part = torch.empty(sp, 64 * 192, device='cuda'); pb = torch.empty(sp, 64, device='cuda')
    dw = torch.empty(64, 192, device='cuda'); db = torch.empty(64, device='cuda')
    def once():
        native.call(fp, g, 64, x, 192, 39000, 64, 192, part, pb, native.stream_ptr())
        native.call('dgmc_wgrad_reduce', part, pb, sp, 64, 192, dw, db, native.stream_ptr())
    for _ in range(3): once()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50): once()
    b.record(); torch.cuda.synchronize()
    ref = g.double().t() @ x.double()
    return a.elapsed_time(b) / 50 * 1e3, (dw.double() - ref).abs().max().item() / ref.abs().max().item(), sp
for tc in (False, True):
    print('tensor_core=%s: %.1f us per wgrad (partial+reduce), max rel err %.2e, splits %d' % ((tc,) + run(tc)))
