"""Launches the K5 RelConv aggregation (balanced_main_kernel + balanced_fixup_kernel) on the merged
DBP15K-shaped pair (38,960 nodes, 165,556 edges, 64 channels), forward and transposed, so that
    ncu --set full --clock-control none --import-source on -k regex:balanced_ -o gpurun_out/<name> \
        python profiles/k5_shape.py
captures them.  Not part of the product or the tests."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))
import bench  # noqa: E402
from dgmc import ops  # noqa: E402

w = bench.WORKLOAD
inputs = bench.make_inputs(w)
e_s, e_t = inputs['e_s'].cuda(), inputs['e_t'].cuda()
e_m = torch.cat([e_s, e_t + w['n_s']], dim=1)
n = w['n_s'] + w['n_t']
graph = ops.GraphCSR.of(e_m, n)
x = torch.randn(n, w['rnd_dim'], device='cuda').requires_grad_()
for _ in range(2):
    y = ops.rel_aggregate(x, graph)
    y.backward(torch.ones_like(y))
torch.cuda.synchronize()
