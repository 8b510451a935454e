"""K1 (dgmc_topk_tc: cast + tcgen05 GEMM/select + fp32 merge) at the DBP15K shape and at k = 25: device
time per call inside a replayed CUDA graph, and exactness against an fp64 top-k on sampled rows.
python profiles/topk_microbench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc import ops

torch.manual_seed(0)
n_s, n_t, c = 19388, 19572, 256
h_s, h_t = torch.randn(1, n_s, c, device='cuda'), torch.randn(1, n_t, c, device='cuda')


def t(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    g.replay()
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps * 1e3


for k in (10, 25):
    us = t(lambda: ops.topk_tensor_core(h_s, h_t, k))
    idx = ops.topk_tensor_core(h_s, h_t, k)[0]
    rows = torch.arange(0, n_s, 41, device='cuda')
    exact = h_s[0, rows].double() @ h_t[0].double().t()
    top_v, top_i = exact.topk(k + 1, dim=1)
    safe = (top_v[:, k - 1] - top_v[:, k]) > 1e-6 * top_v[:, k - 1].abs()
    same = (idx[rows].long().sort(-1)[0] == top_i[:, :k].sort(-1)[0]).all(-1)
    print('k=%d: %.1f us per call = %.0f TFLOP/s; exact on %d / %d sampled rows outside the tie band' % (
        k, us, 2.0 * n_s * n_t * c / us / 1e6, int(same[safe].sum()), int(safe.sum())), flush=True)
