"""Micro-benchmark: the RelConv aggregation (K5, `[3N, N]` sparse product and its transpose) on the
entry-parallel balanced kernel (+ fix-up launch) against the one-lane-group-per-segment kernel, on the
merged DBP15K pair (38,960 nodes, 165,556 uniformly drawn edges) and on a hub-heavy edge list of the
same size (destination ~ Zipf); device time per launch inside a replayed CUDA graph.
    python profiles/k5_segment_vs_balanced.py [nodes edges c]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc import ops

n, e, c = (int(v) for v in sys.argv[1:4]) if len(sys.argv) >= 4 else (38960, 165556, 64)


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


def run(name, ei):
    rel = ops.GraphCSR(ei, n).rel
    x = torch.randn(n, c, device='cuda')
    g3 = torch.randn(3 * n, c, device='cuda')
    out_f = torch.empty(3 * n, c, device='cuda')
    out_b = torch.empty(n, c, device='cuda')
    longest_f = int((rel.f_ptr[1:] - rel.f_ptr[:-1]).max())
    longest_b = int((rel.b_ptr[1:] - rel.b_ptr[:-1]).max())
    print('%s: longest segment forward %d, transpose %d' % (name, longest_f, longest_b))
    with torch.no_grad():
        for label, seg_f, seg_b in (('balanced + fix-up', rel.f_seg, rel.b_seg), ('group per segment', None, None)):
            tf = t(lambda: ops.segment_gather(x, 3 * n, rel.f_src, ptr=rel.f_ptr, w=rel.f_w, out=out_f, seg_of=seg_f))
            tb = t(lambda: ops.segment_gather(g3, n, rel.b_src, ptr=rel.b_ptr, w=rel.b_w, out=out_b, seg_of=seg_b))
            print('  %-18s forward %.1f us   transpose %.1f us' % (label, tf, tb))


torch.manual_seed(0)
print('nodes %d edges %d c %d' % (n, e, c))
run('uniform', torch.randint(n, (2, e), device='cuda'))
zipf = (torch.rand(e, device='cuda').pow(4) * n).long().clamp_(max=n - 1)       # a few hundred-fold hubs
run('hub-heavy', torch.stack([torch.randint(n, (e, ), device='cuda'), zipf]))
