"""Row-sharded DGMC on 2 GPUs (NCCL) against the single-GPU path on the same seed: candidate sets
bit-exact, S_L / loss within the e2e parity tolerance, summed gradients within the gradient
tolerance of tests/test_gpu_e2e.py.  Skipped on boxes with one GPU (run it with gpurun --gpus 2)."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _problem(n_s=900, n_t=1000, feat=32):
    g = torch.Generator().manual_seed(7)
    x_s = torch.randn(n_s, feat, generator=g)
    perm = torch.randperm(n_t, generator=g)[:n_s]
    x_t = torch.randn(n_t, feat, generator=g)
    x_t[perm] = x_s + 0.3 * torch.randn(n_s, feat, generator=g)
    e_s = torch.randint(n_s, (2, 4 * n_s), generator=g)
    e_t = torch.randint(n_t, (2, 4 * n_t), generator=g)
    rows = torch.arange(0, n_s, 4)
    return x_s, e_s, x_t, e_t, torch.stack([rows, perm[rows]])


def _worker(rank, world, port, out):
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=torch.device('cuda', rank))
    try:
        from dgmc.models import DGMC, RelCNN, RowShardedDGMC
        x_s, e_s, x_t, e_t, y = [t.cuda() for t in _problem()]

        def build(cls):
            torch.manual_seed(0)
            # psi_1 with dropout: the sharded path must reproduce the single-process mask stream
            m = cls(RelCNN(32, 48, 2, dropout=0.5), RelCNN(16, 16, 2), num_steps=3, k=6).cuda().train()
            return m

        single = build(DGMC)
        torch.manual_seed(123)
        _, S1 = single(x_s, e_s, None, None, x_t, e_t, None, None, y)
        l1 = single.loss(S1, y)
        l1.backward()
        floor = 1e-5 * max(p.grad.abs().max().item() for p in single.parameters() if p.grad is not None)
        # psi_2 replicated / node-sharded with all-gathers; psi_1 replicated / node-sharded
        for shard_psi, shard_psi_1 in ((False, False), (True, False), (False, True), (True, True)):
            sharded = build(RowShardedDGMC)
            sharded.shard_psi, sharded.shard_psi_1 = shard_psi, shard_psi_1
            torch.manual_seed(123)
            _, S2 = sharded(x_s, e_s, None, None, x_t, e_t, None, None, y)
            share = sharded.loss(S2, y)
            share.backward()
            sharded.sync_gradients()
            lo, hi = S2.__rows__
            assert torch.equal(S2.__idx__, S1.__idx__[lo:hi]), 'candidate sets differ'
            ref = S1.__val__[lo:hi]
            assert (S2.__val__ - ref).abs().max().item() <= 1e-5 * ref.abs().max().item() + 1e-7
            assert abs(sharded.global_loss(share).item() - l1.item()) <= 1e-5 * abs(l1.item()) + 1e-7
            assert sharded.acc(S2, y) == pytest.approx(single.acc(S1, y))
            assert sharded.hits_at_k(3, S2, y) == pytest.approx(single.hits_at_k(3, S1, y))
            # psi_2.final.bias cancels in o_s - o_t: its true gradient is 0 and what is computed is
            # rounding noise, hence the floor relative to the largest gradient of the whole model
            for (n, p), q in zip(single.named_parameters(), sharded.parameters()):
                if p.grad is None:
                    assert q.grad is None
                    continue
                scale = p.grad.abs().max().item()
                assert (p.grad - q.grad).abs().max().item() <= 5e-4 * scale + floor, (shard_psi, shard_psi_1, n)
        from dgmc.symm import PeerCollectives
        pc = PeerCollectives._by_group.get(id(None))
        assert pc, 'peer-memory collectives were not used'
        assert pc.timeouts() == 0
        out.put((rank, 'ok'))
    except Exception:  # pragma: no cover
        import traceback
        out.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_row_sharded_matches_single_gpu():
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = 29600 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == 'ok', 'rank %d: %s' % (rank, msg)


def _collectives_worker(rank, world, port, out):
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))
    os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        from dgmc.symm import PeerCollectives
        pc = PeerCollectives.of(None, dev, 1 << 22)
        assert pc is not None, 'peer memory unavailable on this box'
        g = torch.Generator(device='cpu').manual_seed(100 + rank)
        for rows, cols in ((1, 4), (37, 8), (5000, 64)):        # all-reduce: sum, identical on all ranks
            x = torch.randn(rows, cols, generator=g).to(dev)
            for _ in range(3):                                   # back to back: the flag words are reused
                pc.scratch_in(rows, cols).copy_(x)
                got = pc.all_reduce(rows, cols).clone()
                ref = x.clone()
                dist.all_reduce(ref)
                assert torch.allclose(got, ref, rtol=1e-6, atol=1e-6)
                same = got.clone()
                dist.broadcast(same, 0)
                assert torch.equal(got, same), 'ranks hold different sums'
        # all-gather / reduce-scatter with two runs per rank (source block + target block, natural order)
        n_s, n_t, c = 1001, 777, 16
        from dgmc.models.sharded import row_block
        (a, b), (t0, t1) = row_block(n_s, rank, world), row_block(n_t, rank, world)
        runs = ((a, b - a), (n_s + t0, t1 - t0))
        max_rows = max((row_block(n_s, q, world)[1] - row_block(n_s, q, world)[0]) +
                       (row_block(n_t, q, world)[1] - row_block(n_t, q, world)[0]) for q in range(world))
        full = torch.randn(n_s + n_t, c, generator=torch.Generator().manual_seed(5)).to(dev)   # same on all ranks
        mine = torch.cat([full[a:b], full[n_s + t0:n_s + t1]])
        for _ in range(3):
            got = pc.all_gather(mine, runs, n_s + n_t, max_rows).clone()
            assert torch.equal(got, full)
        contrib = torch.randn(n_s + n_t, c, generator=g).to(dev)
        total = contrib.clone()
        dist.all_reduce(total)
        for _ in range(3):
            pc.scratch_in(n_s + n_t, c).copy_(contrib)
            got = pc.reduce_scatter(mine.size(0), c, runs, max_rows)
            want = torch.cat([total[a:b], total[n_s + t0:n_s + t1]])
            assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
        # equal blocks in rank order (the default form)
        blk = torch.randn(300, 8, generator=g).to(dev)
        got = pc.all_gather(blk).clone()
        ref = torch.empty(world * 300, 8, device=dev)
        dist.all_gather_into_tensor(ref, blk)
        assert torch.equal(got, ref)
        assert pc.timeouts() == 0
        out.put((rank, 'ok'))
    except Exception:  # pragma: no cover
        import traceback
        out.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_peer_memory_collectives_match_nccl():
    """csrc/symm.cu through dgmc.symm against torch.distributed (NCCL): sums, identical results on all
    ranks, two-run all-gather / reduce-scatter, repeated calls (self-resetting handshakes), no timeouts."""
    import torch.multiprocessing as mp
    ctx = mp.get_context('spawn')
    out = ctx.Queue()
    port = 29700 + os.getpid() % 2000
    procs = [ctx.Process(target=_collectives_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    results = [out.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, msg in results:
        assert msg == 'ok', 'rank %d: %s' % (rank, msg)
