"""Peer-memory collectives (csrc/symm.cu) against NCCL: correctness and device time per call inside a
replayed CUDA graph.   torchrun --nproc-per-node N profiles/symm_microbench.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
from dgmc.symm import PeerCollectives

rows, cols = 19572, 64
pc = PeerCollectives.of(None, dev, max(world * rows * cols * 4, 78288 * cols * 4))
assert pc is not None, 'peer memory unavailable'
torch.manual_seed(rank)
x = torch.randn(rows, cols, device=dev)

# ---- correctness ----
pc.scratch_in(rows, cols).copy_(x)
got = pc.all_reduce(rows, cols).clone()
ref = x.clone(); dist.all_reduce(ref)
err_ar = (got - ref).abs().max().item()
blk = x[:4000].contiguous()
got = pc.all_gather(blk).clone()
ref = torch.empty(world * 4000, cols, device=dev); dist.all_gather_into_tensor(ref, blk)
err_ag = (got - ref).abs().max().item()
big = torch.randn(world * 4000, cols, device=dev)
pc.scratch_in(world * 4000, cols).copy_(big)
got = pc.reduce_scatter(4000, cols)
ref = torch.empty(4000, cols, device=dev); dist.reduce_scatter_tensor(ref, big)
err_rs = (got - ref).abs().max().item()
# every rank holds bit-identical all-reduce results
pc.scratch_in(rows, cols).copy_(x)
mine = pc.all_reduce(rows, cols).clone()
other = mine.clone(); dist.broadcast(other, 0)
same = bool(torch.equal(mine, other))
torch.cuda.synchronize()


def t(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode='thread_local'):
        for _ in range(reps):
            fn()
    g.replay(); torch.cuda.synchronize(); dist.barrier()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); z.record(); torch.cuda.synchronize()
    return a.elapsed_time(z) / reps * 1e3


y = x.clone()
res = {'all_reduce 5 MB: peer': t(lambda: pc.all_reduce(rows, cols)), 'nccl': t(lambda: dist.all_reduce(y)),
       'all_gather %d x 1.0 MB: peer' % world: t(lambda: pc.all_gather(blk)),
       'nccl ': t(lambda: dist.all_gather_into_tensor(ref_ag, blk)) if False else None}
ref_ag = torch.empty(world * 4000, cols, device=dev)
res['nccl '] = t(lambda: dist.all_gather_into_tensor(ref_ag, blk))
rs_out = torch.empty(4000, cols, device=dev)
res['reduce_scatter: peer'] = t(lambda: pc.reduce_scatter(4000, cols))
res['nccl  '] = t(lambda: dist.reduce_scatter_tensor(rs_out, big))
sweep = []
for r_ in (256, 2048, 19572, 78288):
    yy = torch.randn(r_, cols, device=dev)
    sweep.append((r_ * cols * 4 / 1e6, t(lambda: pc.all_reduce(r_, cols)), t(lambda: dist.all_reduce(yy)),
                  t(lambda: pc.all_gather(yy[:r_ // world * 1].contiguous())) if False else 0.0))
timeouts = pc.timeouts()
if rank == 0:
    print('multicast mapping:', pc.multicast is not None)
    print('world %d: max |err| all_reduce %.2g, all_gather %.2g, reduce_scatter %.2g; identical on all ranks: %s; '
          'timeouts %d' % (world, err_ar, err_ag, err_rs, same, timeouts))
    print('; '.join('%s %.1f us' % (k, v) for k, v in res.items()))
    print('all_reduce sweep (MB, peer us, nccl us):', ' | '.join('%.2f: %.1f / %.1f' % (a, b, c) for a, b, c, _ in sweep))
dist.barrier()
os._exit(0)
