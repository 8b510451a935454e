"""Source-row sharding of the sparse DGMC path across the GPUs of one node (no counterpart in the
reference, which is single-process: SURVEY.md §2.3, §8e).

One process per GPU, torch.distributed for the plumbing.  Every rank holds the whole pair (features,
edge lists, weights) and the same RNG seed; the correspondence structures — candidate lists, S_hat,
S, and the K1/K2/K3/K4/K6 kernels that touch them — exist only for the rank's block of source rows:

  * top-k, gather-dot, softmax, diff-MLP are row-independent: no communication;
  * the indicator transport r_t = S^T r_s is a sum over source rows: each rank reduces its rows into
    a partial [N_t, R] and ONE all-reduce (NCCL over NVLink) per consensus step completes it; its
    backward is the matching all-reduce of d r_t;
  * psi_2 (when it is this package's RelCNN without batch-norm statistics or dropout) is sharded by
    NODES: each rank owns a block of source nodes (the same block as its rows of S) and a block of
    target nodes, computes every RelConv layer only for those, and all-gathers the layer output
    (its neighbours live on other ranks); the hoisted MLP projection of the target side, P_t, is
    all-gathered once per step because the difference kernel gathers arbitrary target rows — the
    "all-gather of o_t" of the north star.  Backward of an all-gather is a reduce-scatter.
    Any other psi_2, and psi_1, run replicated on identical inputs (r_s is drawn at full shape from
    the shared seed on every rank, as the reference's RNG contract requires);
  * weight gradients are partial on every rank and are summed once per step by `sync_gradients()`.

Results equal the single-GPU path up to the fp32 summation order of the all-reduce.
"""
import copy

import torch
import torch.distributed as dist

from .. import ops
from .dgmc import DGMC, EPS, _sparse_coo


def row_block(n_rows, rank, world):
    """Contiguous, balanced block [lo, hi) of `n_rows` owned by `rank`."""
    base, extra = divmod(n_rows, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


# messages up to this size go through the library's own peer-memory kernels (dgmc.symm / csrc/symm.cu;
# latency-optimised), larger ones through NCCL's bandwidth-optimised ring
PEER_MAX_BYTES = 2304 << 20


def _peer(group, device, nbytes, numel):
    """PeerCollectives of the group for a message of nbytes, or None (CPU / gloo, no peer access,
    message too large or not a multiple of 4 floats)."""
    if device.type != 'cuda' or nbytes > PEER_MAX_BYTES or numel % 4 != 0 or numel == 0:
        return None
    from ..symm import PeerCollectives
    return PeerCollectives.of(group, device, nbytes)


def _sum_over_ranks(x, group, keep=True):
    """-> the sum of x over the ranks (x itself may be the peer scratch); keep=False: the result may be
    a view of the peer scratch that the caller consumes before its next collective."""
    pc = _peer(group, x.device, x.numel() * 4, x.numel()) if x.dtype == torch.float32 else None
    if pc is None:
        y = x.contiguous().clone()
        dist.all_reduce(y, op=dist.ReduceOp.SUM, group=group)
        return y
    n = x.numel()
    src = pc.scratch_in(1, n)
    if x.data_ptr() != src.data_ptr():       # the producer may already have written into the scratch
        src.view(x.shape).copy_(x)           # (one copy also when x is a strided block of a wider table)
    y = pc.all_reduce(1, n)
    return (y.clone() if keep else y).view(x.shape)


class _AllReduceSum(torch.autograd.Function):
    """y = sum over ranks of x; the adjoint of a sum whose result every rank consumes in a
    replicated computation is again the sum of the ranks' gradients."""
    @staticmethod
    def forward(ctx, x, group, keep):
        ctx.group = group
        return _sum_over_ranks(x, group, keep)

    @staticmethod
    def backward(ctx, g):
        return _sum_over_ranks(g, ctx.group), None, None


def _scratch_in(group, device, rows, c):
    pc = _peer(group, device, rows * c * 4, 4 * c)
    return None if pc is None else pc.scratch_in(rows, c)


def all_reduce_sum(x, group=None, keep=True):
    return _AllReduceSum.apply(x, group, keep)


def transport_partial_buffer(n_t, r, device, group):
    """Where ops.transport should write this rank's partial r_t so that the all-reduce needs no copy:
    the peer scratch when peer-memory collectives are in use, else None (a fresh tensor)."""
    pc = _peer(group, device, n_t * r * 4, n_t * r)
    return None if pc is None else pc.scratch_in(1, n_t * r).view(n_t, r)


class ShardedPair(object):
    """Node partition of the merged (source + target) graph over the ranks, and this rank's slice
    of the RelConv aggregation operator.

    Rank q owns the source nodes row_block(N_s, q) — the same rows of S it owns — and the target nodes
    row_block(N_t, q).  Node features keep the NATURAL merged order [source nodes ; target nodes]
    everywhere: a rank's part of a gathered [N_s + N_t, C] tensor is two runs of rows (`runs`), which
    the peer-memory all-gather / reduce-scatter kernels address directly, so there is no permutation on
    either side.  `rel` is the [3 n_loc, N_s + N_t] operator (see ops._RelAggregate) restricted to this
    rank's nodes.  Built once per (edge_index_s, edge_index_t) with data-dependent shapes (boolean
    masks), hence cached and not capturable: see RowShardedDGMC._node_shard."""
    _cache = {}

    def __init__(self, e_s, e_t, n_s, n_t, rank, world):
        dev = e_s.device
        self.world, self.rank, self.n_s, self.n_t, self.n = world, rank, n_s, n_t, n_s + n_t
        self.blocks_s = [row_block(n_s, q, world) for q in range(world)]
        self.blocks_t = [row_block(n_t, q, world) for q in range(world)]
        self.ns_max = max(b[1] - b[0] for b in self.blocks_s)
        self.nt_max = max(c[1] - c[0] for c in self.blocks_t)
        (lo_s, hi_s), (lo_t, hi_t) = self.blocks_s[rank], self.blocks_t[rank]
        self.ns_loc, self.nt_loc = hi_s - lo_s, hi_t - lo_t
        self.n_loc = self.ns_loc + self.nt_loc
        self.runs = ((lo_s, self.ns_loc), (n_s + lo_t, self.nt_loc))      # in a [n_s + n_t, C] tensor
        self.runs_t = ((lo_t, self.nt_loc), (0, 0))                       # in a [n_t, C] tensor
        self.max_rows = max((b[1] - b[0]) + (c[1] - c[0]) for b, c in zip(self.blocks_s, self.blocks_t))
        n = self.n
        local = torch.cat([torch.arange(lo_s, hi_s, device=dev), n_s + torch.arange(lo_t, hi_t, device=dev)])
        self.local = local
        loc_of = torch.full((n, ), -1, dtype=torch.long, device=dev)
        loc_of[local] = torch.arange(self.n_loc, device=dev)
        src = torch.cat([e_s[0], e_t[0] + n_s])
        dst = torch.cat([e_s[1], e_t[1] + n_s])
        inv_in = torch.bincount(dst, minlength=n).clamp_(min=1).to(torch.float32).reciprocal_()
        inv_out = torch.bincount(src, minlength=n).clamp_(min=1).to(torch.float32).reciprocal_()
        m_in, m_out = loc_of[dst] >= 0, loc_of[src] >= 0
        rows = torch.cat([3 * torch.arange(self.n_loc, device=dev), 3 * loc_of[dst[m_in]] + 1,
                          3 * loc_of[src[m_out]] + 2]).to(torch.int32)
        cols = torch.cat([local, src[m_in], dst[m_out]]).to(torch.int32)
        w = torch.cat([torch.ones(self.n_loc, device=dev), inv_in[dst[m_in]], inv_out[src[m_out]]])
        rel = type('RelSlice', (), {})()
        rel.n_out, rel.n_in = self.n_loc, n
        rel.f_ptr, perm, rel.f_seg = ops.csr_build(rows, 3 * rel.n_out, True)
        rel.f_src, rel.f_w = cols[perm.long()].contiguous(), w[perm.long()].contiguous()
        rel.b_ptr, perm, rel.b_seg = ops.csr_build(cols, rel.n_in, True)
        rel.b_src, rel.b_w = rows[perm.long()].contiguous(), w[perm.long()].contiguous()
        self.rel = rel

    @classmethod
    def of(cls, e_s, e_t, n_s, n_t, rank, world):
        key = (e_s.data_ptr(), e_t.data_ptr(), e_s._version, e_t._version, n_s, n_t, rank, world)
        hit = cls._cache.get(key)
        if ops.DYNAMIC_STRUCTURE[0]:
            return None          # the step must follow its inputs' edge lists: no cached structures
        if hit is None or hit[0] is not e_s or hit[1] is not e_t:
            if torch.cuda.is_current_stream_capturing():
                return None      # data-dependent shapes: cannot be built inside a CUDA-graph capture
            if len(cls._cache) > 16:
                cls._cache.clear()
            hit = (e_s, e_t, cls(e_s, e_t, n_s, n_t, rank, world))
            cls._cache[key] = hit
        return hit[2]


ops._EXTRA_CACHES.append(ShardedPair._cache)


def _pad_rows(x, n):
    return x if x.size(0) == n else torch.cat([x, x.new_zeros(n - x.size(0), x.size(1))], dim=0)


class _AllGatherNodes(torch.autograd.Function):
    """This rank's rows of a node-feature tensor -> the whole tensor in natural node order on every
    rank; adjoint: every rank keeps the sum over ranks of its own rows of the gradient (reduce-scatter).
    part = 'all': x [n_loc, C] -> [N_s + N_t, C]; part = 't': x [nt_loc, C] -> [N_t, C].
    keep: the result outlives the next collective (it is copied out of the peer scratch)."""
    @staticmethod
    def forward(ctx, x, sp, group, part, keep):
        ctx.sp, ctx.group, ctx.part = sp, group, part
        x = x.contiguous()
        c = x.size(1)
        runs, total, max_rows = (sp.runs, sp.n, sp.max_rows) if part == 'all' else (sp.runs_t, sp.n_t, sp.nt_max)
        pc = _peer(group, x.device, total * c * 4, 4 * c) if x.dtype == torch.float32 else None
        if pc is not None:
            full = pc.all_gather(x, runs, total, max_rows)
            return full.clone() if keep else full
        # torch.distributed: equal padded blocks, then the natural order by static slices
        if part == 'all':
            blk = torch.cat([_pad_rows(x[:sp.ns_loc], sp.ns_max), _pad_rows(x[sp.ns_loc:], sp.nt_max)], dim=0)
        else:
            blk = _pad_rows(x, sp.nt_max)
        m = blk.size(0)
        g = x.new_empty(sp.world * m, c)
        dist.all_gather_into_tensor(g, blk, group=group)
        pieces = []
        if part == 'all':
            pieces += [g[q * m:q * m + (b[1] - b[0])] for q, b in enumerate(sp.blocks_s)]
            pieces += [g[q * m + sp.ns_max:q * m + sp.ns_max + (t[1] - t[0])] for q, t in enumerate(sp.blocks_t)]
        else:
            pieces += [g[q * m:q * m + (t[1] - t[0])] for q, t in enumerate(sp.blocks_t)]
        return torch.cat(pieces, dim=0)

    @staticmethod
    def backward(ctx, g):
        sp, group, part = ctx.sp, ctx.group, ctx.part
        c = g.size(1)
        runs, max_rows = (sp.runs, sp.max_rows) if part == 'all' else (sp.runs_t, sp.nt_max)
        rows = runs[0][1] + runs[1][1]
        pc = _peer(group, g.device, g.numel() * 4, 4 * c) if g.dtype == torch.float32 else None
        if pc is not None:
            src = pc.scratch_in(g.size(0), c)
            if g.data_ptr() != src.data_ptr():      # the aggregation's backward may have written it there
                src.copy_(g)
            return pc.reduce_scatter(rows, c, runs, max_rows), None, None, None, None
        zeros = g.new_zeros
        blocks = []
        for q in range(sp.world):
            (a, b), (t0, t1) = sp.blocks_s[q], sp.blocks_t[q]
            if part == 'all':
                blocks += [g[a:b], zeros(sp.ns_max - (b - a), c), g[sp.n_s + t0:sp.n_s + t1],
                           zeros(sp.nt_max - (t1 - t0), c)]
            else:
                blocks += [g[t0:t1], zeros(sp.nt_max - (t1 - t0), c)]
        padded = torch.cat(blocks, dim=0)
        m = padded.size(0) // sp.world
        out = g.new_empty(m, c)
        dist.reduce_scatter_tensor(out, padded, op=dist.ReduceOp.SUM, group=group)
        if part == 'all':
            out = torch.cat([out[:sp.ns_loc], out[sp.ns_max:sp.ns_max + sp.nt_loc]], dim=0)
        else:
            out = out[:sp.nt_loc]
        return out, None, None, None, None


def all_gather_nodes(x, sp, group=None, part='all', keep=True):
    return _AllGatherNodes.apply(x, sp, group, part, keep)


def sharded_relcnn(psi, x_full, sp, group, masks=None):
    """RelCNN.forward (rel.py:80-92) for this rank's nodes only.  x_full: the layer-0 features of ALL
    nodes, natural merged order [N_s + N_t, C]; masks: per layer the dropout factors of this rank's
    nodes (see reference_dropout_masks) or None; returns [n_loc, out_channels]."""
    feats = [torch.cat([x_full[sp.runs[0][0]:sp.runs[0][0] + sp.ns_loc],
                        x_full[sp.runs[1][0]:sp.runs[1][0] + sp.nt_loc]], dim=0)]
    full = x_full
    # from the second layer on the aggregation's input is an all-gather, whose backward is a
    # reduce-scatter of the aggregation's input gradient: that gradient is written straight into the
    # buffer the reduce-scatter reads
    rel_g = copy.copy(sp.rel)
    rel_g.grad_buffer = lambda rows, c, device: _scratch_in(group, device, rows, c)
    for i, conv in enumerate(psi.convs):
        stacked = ops.rel_aggregate_op(full, sp.rel if i == 0 else rel_g)    # [n_loc, 3 C]
        weight = ops.region_cached(('relconv', id(conv)), lambda c=conv: torch.cat(
            [c.root.weight, c.lin1.weight, c.lin2.weight], dim=1))
        h = ops.node_linear(stacked, weight, conv.root.bias, relu=True, key=('relconv', id(conv)))
        if psi.batch_norm:
            # only reached in eval mode (MergedPair.mergeable): the running-statistics affine map
            # is per node, so the node shard applies it without communication (rel.py:86-87)
            h = psi.batch_norms[i](h)
        if masks is not None:
            h = h * masks[i]
        feats.append(h)
        if i + 1 < len(psi.convs):
            # consumed by the next layer's aggregation right away (which keeps no reference to it)
            full = all_gather_nodes(h, sp, group, 'all', keep=False)
    out = torch.cat(feats, dim=-1) if psi.cat else feats[-1]
    return ops.node_linear(out, psi.final.weight, psi.final.bias, key=('final', id(psi))) if psi.lin else out


def reference_dropout_masks(psi, n_s, n_t, rows, device):
    """The dropout factors (0 or 1 / (1 - p)) the single-process forward would draw for psi on the two
    graphs, for the merged-pair rows `rows`, per layer.  The reference runs psi_1 on the source graph
    (one F.dropout per layer, rel.py:88) and then on the target graph (dgmc.py:149-150); F.dropout's
    mask depends only on the generator state and the tensor's shape, so drawing on tensors of ones in
    that order consumes the stream exactly as the reference does — every rank draws the full masks
    from the shared seed and keeps its rows."""
    import torch.nn.functional as F
    if not (psi.training and psi.dropout > 0):
        return None
    width = psi.convs[0].out_channels
    per_graph = [[F.dropout(torch.ones(n, width, device=device), p=psi.dropout, training=True)
                  for _ in psi.convs] for n in (n_s, n_t)]
    return [torch.cat([per_graph[0][i], per_graph[1][i]], dim=0)[rows] for i in range(len(psi.convs))]


class RowShardedDGMC(DGMC):
    """DGMC whose sparse correspondence path is sharded by source rows over `group` (default: the
    world group).  forward() returns this rank's rows only: sparse [n_local, N_t] tensors with
    `.__idx__`, `.__val__` and `.__rows__ = (lo, hi)`; `loss` / `acc` / `hits_at_k` take the GLOBAL
    ground truth `y` and reduce over all ranks."""

    # node-sharding psi_2 adds 2 (L + 1) collectives per consensus step; below this many nodes
    # (source + target) they cost more than the replicated psi_2 they save
    SHARD_PSI_MIN_NODES = 150000

    def __init__(self, psi_1, psi_2, num_steps, k, detach=False, group=None, shard_psi='auto',
                 shard_psi_1='auto'):
        super().__init__(psi_1, psi_2, num_steps, k=k, detach=detach)
        assert k >= 1, 'row sharding applies to the sparse variant (k >= 1)'
        self.group = group
        # node-shard psi_2 (RelCNN) instead of running it replicated: True / False / 'auto' (by size)
        self.shard_psi = shard_psi
        # node-shard psi_1 (RelCNN without batch-norm statistics): each rank embeds its own source rows
        # and a block of target nodes, layer outputs are all-gathered, h_t is all-gathered once
        self.shard_psi_1 = shard_psi_1

    def _world(self):
        return dist.get_rank(self.group), dist.get_world_size(self.group)

    def _forward(self, x_s, edge_index_s, edge_attr_s, batch_s, x_t, edge_index_t, edge_attr_t,
                 batch_t, y=None):
        assert batch_s is None and batch_t is None, 'row sharding handles one graph pair'
        rank, world = self._world()
        from .rel import RelCNN
        N_s, N_t = x_s.size(0), x_t.size(0)
        lo, hi = row_block(N_s, rank, world)
        n_loc = hi - lo
        psi_1 = self.psi_1
        shard_1 = world > 1 and isinstance(psi_1, RelCNN) and psi_1.lin and edge_attr_s is None and \
            not (psi_1.batch_norm and psi_1.training) and self.shard_psi_1 in (True, 'auto')
        # the node partition is built with data-dependent shapes: it is cached per edge list and cannot
        # be (re)built inside a CUDA-graph capture — a step captured with structure_is_static=False
        # therefore runs psi replicated (ShardedPair.of returns None there)
        sp1 = ShardedPair.of(edge_index_s, edge_index_t, N_s, N_t, rank, world) if shard_1 else None
        if sp1 is not None:
            # psi_1 on this rank's nodes only (its rows of S + a block of target nodes); the neighbours'
            # features of every layer come from an all-gather, h_t from one more
            masks = reference_dropout_masks(psi_1, N_s, N_t, sp1.local, x_s.device)
            h_own = sharded_relcnn(psi_1, torch.cat([x_s, x_t], dim=0), sp1, self.group, masks)
            if self.detach:
                h_own = h_own.detach()
            h_loc = h_own[:sp1.ns_loc]
            h_t = all_gather_nodes(h_own[sp1.ns_loc:], sp1, self.group, 't')
        else:
            h_s = psi_1(x_s, edge_index_s, edge_attr_s)
            h_t = psi_1(x_t, edge_index_t, edge_attr_t)
            if self.detach:
                h_s, h_t = h_s.detach(), h_t.detach()
            h_loc = h_s[lo:hi]
        C = h_t.size(1)
        R_in = self.psi_2.in_channels
        w1, b1, w2, b2 = self._hoisted_mlp()
        num_steps = self.num_steps or 0
        merged = ops.MergedPair.of(self.psi_2, self.training, edge_index_s, edge_index_t,
                                   edge_attr_s, edge_attr_t, N_s, N_t)
        sp = None
        shard_psi = (N_s + N_t >= self.SHARD_PSI_MIN_NODES) if self.shard_psi == 'auto' else self.shard_psi
        if shard_psi and world > 1 and merged is not None and isinstance(self.psi_2, RelCNN):
            sp = ShardedPair.of(edge_index_s, edge_index_t, N_s, N_t, rank, world)

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
        from .rel import RelCNN
        direct_psi_2 = sp is None and merged is not None and isinstance(self.psi_2, RelCNN) and self.psi_2.lin \
            and h_t.dtype == torch.float32
        for _ in range(num_steps):
            S = ops.softmax_k(S_hat)
            # replicated psi_2 = RelCNN: the merged node table [r_s ; r_t] is the leading column block of
            # psi_2's concatenated feature table — the draw goes straight into it (same stream as a
            # contiguous draw), the all-reduced transport is copied in once; no cat, no input copy
            fb = self.psi_2.feature_buffer(N_s + N_t, h_t.device) if direct_psi_2 else None
            if fb is not None:
                x0 = fb[:, :R_in]
                top = x0[:N_s]
                r_s = torch.randn((1, N_s, R_in), dtype=h_t.dtype, device=h_t.device, out=top.view(1, N_s, R_in))
                if r_s.data_ptr() != top.data_ptr():      # a replayed draw (tests): copy it in
                    top.copy_(r_s.reshape(N_s, R_in))
                r_s = top
            else:
                r_s = torch.randn((1, N_s, R_in), dtype=h_t.dtype, device=h_t.device).view(N_s, R_in)
            # the partial transport goes straight into the buffer the peers' all-reduce kernels read
            part = transport_partial_buffer(N_t, R_in, h_t.device, self.group)
            # (r_t is consumed by the concatenation below before the next collective: no copy out)
            r_t = all_reduce_sum(ops.transport(S, r_s[lo:hi], cand, out=part), self.group, keep=False)
            if sp is not None:
                # node-sharded psi_2: every rank holds the layer-0 features of all nodes (r_s from the
                # shared seed, the all-reduced r_t) and computes the layers for its own nodes
                o_loc = sharded_relcnn(self.psi_2, torch.cat([r_s, r_t], dim=0), sp, self.group)   # [n_loc, R_out]
                p_loc = ops.node_linear(o_loc, w1, key=('mlp0', id(self)))
                p_s_loc = p_loc[:sp.ns_loc] + b1
                p_t = all_gather_nodes(p_loc[sp.ns_loc:], sp, self.group, 't')
                S_hat = ops.diff_mlp(S_hat, p_s_loc, p_t, w2, b2, cand)
                continue
            # psi_2 replicated on the merged pair (this package's RelCNN / SplineCNN without batch-norm
            # statistics or dropout); K6 on this rank's source rows of the merged node table
            assert merged is not None, 'RowShardedDGMC needs a psi_2 that runs on the merged pair'
            if fb is not None:
                table = ops.join_rows(x0, top, ops.copy_to(r_t, x0[N_s:]))
            else:
                table = torch.cat([r_s, r_t], dim=0)
            p = self._psi_2_then_lin1(table, merged, w1, b1, N_s, buffer=fb)
            S_hat = ops.diff_mlp_merged(S_hat, p, N_s, None, w2, b2, cand, rows=(lo, hi))
        S_L = ops.softmax_k(S_hat)
        S_idx = S_idx.view(n_loc, k)

        row = torch.arange(n_loc, device=S_idx.device).view(-1, 1).expand(-1, k)
        coords = torch.stack([row.reshape(-1), S_idx.reshape(-1)], dim=0)
        outs = []
        for val in (S_0, S_L):
            sp = _sparse_coo(coords, val.reshape(-1), torch.Size([n_loc, N_t]))
            sp.__idx__, sp.__val__, sp.__rows__ = S_idx, val, (lo, hi)
            outs.append(sp)
        return outs[0], outs[1]

    # ---- metrics over all ranks ------------------------------------------------------------------
    def _local(self, S, y):
        lo, hi = S.__rows__
        mine = (y[0] >= lo) & (y[0] < hi)
        # ground truths of other ranks are masked out by `mine`; point them at DISTINCT local rows:
        # thousands of duplicates of one clamped row made the backward of the row gather (a sorted,
        # serialised index_put accumulate) take 0.45 ms
        spread = torch.arange(y.size(1), device=y.device) % max(hi - lo, 1)
        rows = torch.where(mine, y[0] - lo, spread)
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
        pad = (-flat.numel()) % 4
        if pad:
            flat = torch.cat([flat, flat.new_zeros(pad)])
        flat = _sum_over_ranks(flat, self.group)
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
