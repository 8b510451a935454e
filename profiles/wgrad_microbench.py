"""Micro-benchmark: tcgen05 3xTF32 weight gradient (wgrad_tc.cu) against the fp32 FFMA split-K kernel
(wgrad.cu): error against fp64 and device time per launch inside a replayed CUDA graph.
python profiles/wgrad_microbench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc.native import call, query, stream_ptr, RowStrided
dev = 'cuda'
torch.manual_seed(0)
def run(kind, g, x, want_b=True):
    k, m = g.shape; n = x.shape[1]
    splits = query('dgmc_wgrad_tc_splits' if kind == 'tc' else 'dgmc_wgrad_splits', k, m, n)
    partial = torch.empty(splits, m * n, device=dev); pb = torch.empty(splits, m, device=dev)
    dw = torch.empty(m, n, device=dev); db = torch.empty(m, device=dev)
    def fn():
        call('dgmc_wgrad_tc_partial' if kind == 'tc' else 'dgmc_wgrad_partial', g, g.stride(0), RowStrided(x), x.stride(0), k, m, n, partial, pb, stream_ptr())
    def red():
        call('dgmc_wgrad_reduce', partial, pb, splits, m, n, dw, db, stream_ptr())
    fn(); red(); torch.cuda.synchronize()
    def t(f, reps=20):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for _ in range(reps): f()
        gr.replay(); torch.cuda.synchronize()
        a, c = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); gr.replay(); c.record(); torch.cuda.synchronize()
        return a.elapsed_time(c) / reps * 1e3
    return dw, db, t(fn), t(red), splits
for k, m, n in [(3000, 8, 24), (38960, 64, 192), (38960, 64, 256), (38960, 64, 64), (100000, 128, 300), (5000, 36, 72)]:
    g = torch.randn(k, m, device=dev); x = torch.randn(k, n, device=dev)
    ref_w = g.double().t() @ x.double(); ref_b = g.double().sum(0)
    for kind in ('ffma', 'tc'):
        dw, db, tp, tr, splits = run(kind, g, x)
        ew = (dw.double() - ref_w).abs().max().item() / ref_w.abs().max().item()
        eb = (db.double() - ref_b).abs().max().item() / ref_b.abs().max().item()
        print('k %6d m %3d n %3d %-4s splits %4d: err dw %.2e db %.2e | partial %.1f us reduce %.1f us' % (k, m, n, kind, splits, ew, eb, tp, tr), flush=True)
