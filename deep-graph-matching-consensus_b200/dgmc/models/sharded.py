"""Source-row sharding of the sparse DGMC path across the GPUs of one node (no counterpart in the
reference, which is single-process: SURVEY.md §2.3, §8e).

One process per GPU, torch.distributed for the plumbing.  Every rank holds the whole pair (features,
edge lists, weights) and the same RNG seed; the correspondence structures — candidate lists, S_hat,
S, and the K1/K2/K3/K4/K6 kernels that touch them — exist only for the rank's block of source rows:

  * top-k, gather-dot, softmax, diff-MLP are row-independent: no communication;
  * the indicator transport r_t = S^T r_s is a sum over source rows: each rank reduces its rows into
    a partial [N_t, R] and ONE all-reduce (NCCL over NVLink) per consensus step completes it; its
    backward is the matching all-reduce of d r_t;
  * psi_1 / psi_2 run replicated on identical inputs (r_s is drawn at full shape from the shared
    seed on every rank, as the reference's RNG contract requires), so their outputs agree bit for bit
    and need no exchange; weight gradients are partial (each rank back-propagates its own rows'
    loss) and are summed once per step by `sync_gradients()`.

Results equal the single-GPU path up to the fp32 summation order of the all-reduce.
"""
import torch
import torch.distributed as dist

from .. import ops
from .dgmc import DGMC, EPS, _rows


def row_block(n_rows, rank, world):
    """Contiguous, balanced block [lo, hi) of `n_rows` owned by `rank`."""
    base, extra = divmod(n_rows, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class _AllReduceSum(torch.autograd.Function):
    """y = sum over ranks of x; the adjoint of a sum whose result every rank consumes in a
    replicated computation is again the sum of the ranks' gradients."""
    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        y = x.contiguous().clone()
        dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
        return y

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous().clone()
        dist.all_reduce(g, op=dist.ReduceOp.SUM, group=ctx.group)
        return g, None


def all_reduce_sum(x, group=None):
    return _AllReduceSum.apply(x, group)


class RowShardedDGMC(DGMC):
    """DGMC whose sparse correspondence path is sharded by source rows over `group` (default: the
    world group).  forward() returns this rank's rows only: sparse [n_local, N_t] tensors with
    `.__idx__`, `.__val__` and `.__rows__ = (lo, hi)`; `loss` / `acc` / `hits_at_k` take the GLOBAL
    ground truth `y` and reduce over all ranks."""

    def __init__(self, psi_1, psi_2, num_steps, k, detach=False, group=None):
        super().__init__(psi_1, psi_2, num_steps, k=k, detach=detach)
        assert k >= 1, 'row sharding applies to the sparse variant (k >= 1)'
        self.group = group

    def _world(self):
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def forward(self, x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t, edge_attr_t,
                batch_t, y=None):
        assert batch_s is None and batch_t is None, 'row sharding handles one graph pair'
        rank, world = self._world()
        h_s = self.psi_1(x_s, edge_index_s, edge_attr_s)
        h_t = self.psi_1(x_t, edge_index_t, edge_attr_t)
        if self.detach:
            h_s, h_t = h_s.detach(), h_t.detach()
        N_s, N_t, C = h_s.size(0), h_t.size(0), h_s.size(1)
        lo, hi = row_block(N_s, rank, world)
        n_loc = hi - lo
        R_in = self.psi_2.in_channels
        w1, b1, w2, b2 = self._hoisted_mlp()
        num_steps = self.num_steps or 0
        merged = ops.MergedPair.of(self.psi_2, self.training, edge_index_s, edge_index_t,
                                   edge_attr_s, edge_attr_t, N_s, N_t)

        h_loc = h_s[lo:hi]
        S_idx = self.__top_k__(h_loc.unsqueeze(0), h_t.unsqueeze(0))  # [1, n_loc, k]
        if self.training and y is not None:
            # draw at the reference's full shape so every rank sees the same stream, keep own rows
            rnd = torch.randint(N_t, (1, N_s, min(self.k, N_t - self.k)), dtype=torch.long,
                                device=S_idx.device)
            S_idx = torch.cat([S_idx, rnd[:, lo:hi]], dim=-1)
            # __include_gt__ (dgmc.py:96-112) for the rows this rank owns; ground truths of other
            # ranks are routed to a scratch row so that shapes stay static (no nonzero, no sync)
            mine = (y[0] >= lo) & (y[0] < hi)
            rows = torch.where(mine, y[0] - lo, torch.full_like(y[0], n_loc))
            flat = torch.cat([S_idx.view(n_loc, -1), S_idx.new_zeros(1, S_idx.size(-1))], dim=0)
            last = flat.size(1) - 1
            missing = (flat[rows] != y[1].view(-1, 1)).all(dim=-1)
            flat[rows, last] = torch.where(missing, y[1], flat[rows, last])
            S_idx = flat[:n_loc].view(1, n_loc, -1)

        cand = ops.Candidates(S_idx, N_t)
        k = cand.k
        S_hat = ops.gather_dot(h_loc, h_t, cand)
        S_0 = ops.softmax_k(S_hat)
        for _ in range(num_steps):
            S = ops.softmax_k(S_hat)
            r_s = torch.randn((1, N_s, R_in), dtype=h_s.dtype, device=h_s.device).view(N_s, R_in)
            r_t = all_reduce_sum(ops.transport(S, r_s[lo:hi], cand), self.group)
            if merged is not None:
                o = self.psi_2(torch.cat([r_s, r_t], dim=0), merged.edge_index, merged.edge_attr)
                p = ops.node_linear(o, w1)
                p_s, p_t = p[:N_s] + b1, p[N_s:]
            else:
                o_s = self.psi_2(r_s, edge_index_s, edge_attr_s)
                o_t = self.psi_2(r_t, edge_index_t, edge_attr_t)
                p_s, p_t = ops.node_linear(o_s, w1, b1), ops.node_linear(o_t, w1)
            S_hat = ops.diff_mlp(S_hat, p_s[lo:hi], p_t, w2, b2, cand)
        S_L = ops.softmax_k(S_hat)
        S_idx = S_idx.view(n_loc, k)

        row = torch.arange(n_loc, device=S_idx.device).view(-1, 1).expand(-1, k)
        coords = torch.stack([row.reshape(-1), S_idx.reshape(-1)], dim=0)
        outs = []
        for val in (S_0, S_L):
            sp = torch.sparse_coo_tensor(coords, val.reshape(-1), torch.Size([n_loc, N_t]),
                                         requires_grad=val.requires_grad, check_invariants=False)
            sp.__idx__, sp.__val__, sp.__rows__ = S_idx, val, (lo, hi)
            outs.append(sp)
        return outs[0], outs[1]

    # ---- metrics over all ranks ------------------------------------------------------------------
    def _local(self, S, y):
        lo, hi = S.__rows__
        mine = (y[0] >= lo) & (y[0] < hi)
        rows = (y[0] - lo).clamp(0, max(hi - lo - 1, 0))
        return mine, rows

    def loss(self, S, y, reduction='mean'):
        """Global NLL (dgmc.py:246-267) from row-sharded S.  The value returned on each rank is that
        rank's share (its rows' sum divided by the global count for 'mean'), so that the sum of the
        ranks' gradients is the gradient of the global loss; `global_loss(share)` adds the shares."""
        assert reduction in ['mean', 'sum']
        mine, rows = self._local(S, y)
        hit = (S.__idx__[rows] == y[1].view(-1, 1)) & mine.view(-1, 1)
        nll = -torch.log(S.__val__[rows] + EPS)
        total = (nll * hit).sum()
        if reduction == 'sum':
            return total
        count = hit.sum().to(torch.float32)
        dist.all_reduce(count, op=dist.ReduceOp.SUM, group=self.group)
        return total / count

    def global_loss(self, share):
        value = share.detach().clone()
        dist.all_reduce(value, op=dist.ReduceOp.SUM, group=self.group)
        return value

    def _count(self, correct_local, y, reduction):
        c = correct_local.to(torch.float32)
        dist.all_reduce(c, op=dist.ReduceOp.SUM, group=self.group)
        c = c.item()
        return c / y.size(1) if reduction == 'mean' else int(c)

    def acc(self, S, y, reduction='mean'):
        mine, rows = self._local(S, y)
        pred = S.__idx__[rows, S.__val__[rows].argmax(dim=-1)]
        return self._count(((pred == y[1]) & mine).sum(), y, reduction)

    def hits_at_k(self, k, S, y, reduction='mean'):
        mine, rows = self._local(S, y)
        order = S.__val__[rows].argsort(dim=-1, descending=True)[:, :k]
        pred = torch.gather(S.__idx__[rows], -1, order)
        return self._count(((pred == y[1].view(-1, 1)) & mine.view(-1, 1)).sum(), y, reduction)

    def sync_gradients(self):
        """Sum the (partial) parameter gradients over the ranks: one flat all-reduce per step."""
        grads = [p.grad for p in self.parameters() if p.grad is not None]
        if not grads:
            return
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
