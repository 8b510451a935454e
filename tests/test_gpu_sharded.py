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
