"""Host-side logic of the row-sharded path on CPU with world_size = 2 (gloo): the row partition,
the all-reduce autograd function (forward sum, backward sum), gradient synchronisation and the
sharded loss / metrics arithmetic.  The kernels themselves are covered by the -m gpu tests."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from dgmc.models import GIN, RowShardedDGMC
        from dgmc.models.sharded import all_reduce_sum, row_block
        torch.manual_seed(0)
        # --- all-reduce with gradient ---------------------------------------------------------
        x = torch.full((3, 2), float(rank + 1), requires_grad=True)
        y = all_reduce_sum(x)
        assert torch.equal(y, torch.full((3, 2), 3.0))
        (y * (rank + 1)).sum().backward()           # d/dy = rank + 1 on each rank -> summed: 3
        assert torch.equal(x.grad, torch.full((3, 2), 3.0))
        # --- sharded metrics against the unsharded formulas ------------------------------------
        n_s, n_t, k = 11, 9, 4
        g = torch.Generator().manual_seed(1)
        idx = torch.stack([torch.randperm(n_t, generator=g)[:k] for _ in range(n_s)])
        val = torch.rand(n_s, k, generator=g).softmax(-1)
        yy = torch.stack([torch.arange(0, n_s, 2), idx[torch.arange(0, n_s, 2), 1]])
        yy[1, 0] = (idx[0].max() + 1) % n_t if (idx[0].max() + 1) % n_t not in idx[0] else yy[1, 0]
        model = RowShardedDGMC(GIN(4, 4, 1), GIN(4, 4, 1), num_steps=0, k=k)
        lo, hi = row_block(n_s, rank, world)
        S = torch.sparse_coo_tensor(torch.zeros(2, 0, dtype=torch.long), torch.zeros(0), (hi - lo, n_t))
        S.__idx__, S.__val__, S.__rows__ = idx[lo:hi], val[lo:hi].clone().requires_grad_(), (lo, hi)
        share = model.loss(S, yy)
        total = model.global_loss(share).item()
        hit = idx[yy[0]] == yy[1].view(-1, 1)
        ref = (-torch.log(val[yy[0]] + 1e-8))[hit].mean().item()      # dgmc.py:264-267
        assert abs(total - ref) < 1e-6, (total, ref)
        pred = idx[yy[0], val[yy[0]].argmax(-1)]
        assert model.acc(S, yy) == pytest.approx((pred == yy[1]).sum().item() / yy.size(1))
        order = val[yy[0]].argsort(-1, descending=True)[:, :2]
        ref_hits = (torch.gather(idx[yy[0]], -1, order) == yy[1].view(-1, 1)).sum().item() / yy.size(1)
        assert model.hits_at_k(2, S, yy) == pytest.approx(ref_hits)
        # --- gradient synchronisation -------------------------------------------------------------
        for p in model.parameters():
            p.grad = torch.full_like(p, float(rank + 1))
        model.sync_gradients()
        assert all(torch.equal(p.grad, torch.full_like(p, 3.0)) for p in model.parameters())
        out.put((rank, 'ok'))
    except Exception as e:  # pragma: no cover
        import traceback
        out.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_row_block_partitions_all_rows():
    sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))
    from dgmc.models.sharded import row_block
    for n, w in ((19388, 8), (10, 3), (2, 4), (0, 2)):
        blocks = [row_block(n, r, w) for r in range(w)]
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(blocks, blocks[1:]))
        assert max(h - l for l, h in blocks) - min(h - l for l, h in blocks) <= 1


def test_sharded_host_logic_world_size_2():
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == 'ok', 'rank %d: %s' % (rank, msg)
