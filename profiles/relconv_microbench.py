"""Micro-benchmark: the fused RelConv kernel (dgmc_relconv_tc) against the chain it replaces
(balanced aggregation + fix-up, then dgmc_linear_tc) on the merged DBP15K pair (38,960 nodes,
165,556 edges, 64 -> 64 channels), forward and backward operator; device time per launch inside a
replayed CUDA graph.   python profiles/relconv_microbench.py [nodes edges c]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc import ops

n, e, c = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (38960, 165556, 64)
torch.manual_seed(0)
ei = torch.randint(n, (2, e), device='cuda')
graph = ops.GraphCSR(ei, n)
rel = graph.rel
x = torch.randn(n, c, device='cuda')
w = torch.randn(c, 3 * c, device='cuda') * 0.05
b = torch.randn(c, device='cuda')
split = ops._split_weight(w, False)
y = torch.empty(n, c, device='cuda')
stacked = torch.empty(n, 3 * c, device='cuda')


def t(fn, reps=20):
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


with torch.no_grad():
    print('nodes %d edges %d c %d' % (n, e, c))
    print('fused forward (eval, no operand rows stored)  %.1f us' % t(
        lambda: ops._relconv_tc(rel.f_ptr, rel.f_src, rel.f_w, rel.f_order, x, split, b, True, None, None, y, c)))
    print('fused forward (training, operand rows stored)  %.1f us' % t(
        lambda: ops._relconv_tc(rel.f_ptr, rel.f_src, rel.f_w, rel.f_order, x, split, b, True, None, stacked, y, c)))
    print('fused backward operator (+ addend, in place)   %.1f us' % t(
        lambda: ops._relconv_tc(rel.t_ptr, rel.t_src, rel.t_w, rel.t_order, x, split, None, False, y, None, y, c)))
    t_agg = t(lambda: ops.rel_aggregate(x, graph))
    st = ops.rel_aggregate(x, graph)
    t_lin = t(lambda: ops.node_linear(st, w, b, relu=True))
    print('unfused: aggregation + fix-up %.1f us, linear_tc (incl. weight split) %.1f us' % (t_agg, t_lin))
