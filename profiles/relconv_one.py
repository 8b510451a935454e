"""One warm-up + a few launches of the fused RelConv kernel at the merged DBP15K shape (for ncu)."""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'deep-graph-matching-consensus_b200'))
from dgmc import ops
n, e, c = 38960, 165556, 64
torch.manual_seed(0)
ei = torch.randint(n, (2, e), device='cuda')
graph = ops.GraphCSR(ei, n); rel = graph.rel
x = torch.randn(n, c, device='cuda'); w = torch.randn(c, 3 * c, device='cuda') * 0.05; b = torch.randn(c, device='cuda')
split = ops._split_weight(w, False); y = torch.empty(n, c, device='cuda'); st = torch.empty(n, 3 * c, device='cuda')
for _ in range(4):
    ops._relconv_tc(rel.f_ptr, rel.f_src, rel.f_w, rel.f_order, x, split, b, True, None, st, y, c)
torch.cuda.synchronize()
print('ok')
