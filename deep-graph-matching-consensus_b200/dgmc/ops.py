"""Host-side wrappers (torch.autograd.Function) around the libdgmc_b200 kernels.

PyTorch provides device memory, the current stream and the autograd tape; every numeric step
below runs in a hand-written sm_100a kernel reached through the C ABI (`native.call`), including
the per-node GEMMs (`node_linear`: tcgen05 3xTF32); the library GEMM is only the route for
operands TMA cannot address (rows or bases not 16-byte aligned) and for tiny row counts.
"""
import torch

from . import native
from .native import RowStrided, call, query, stream_ptr


def _f32(t):
    if t.dtype != torch.float32:
        raise native.DgmcNativeError('dgmc (B200): kernels compute in fp32, got %s' % t.dtype)
    return t.contiguous()


def _workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)


# --------------------------------------------------------------------------------------------
# CSR helpers
# --------------------------------------------------------------------------------------------
def csr_build(keys, n_seg, want_sorted=False):
    """keys int32 [nnz] in [0, n_seg) -> (ptr int32 [n_seg + 1], perm int32 [nnz]) — a stable sort;
    with want_sorted also keys[perm] (the segment id of every CSR slot, used by the balanced kernels)."""
    keys = keys.contiguous()
    assert keys.dtype == torch.int32
    nnz = keys.numel()
    ptr = torch.empty(n_seg + 1, dtype=torch.int32, device=keys.device)
    perm = torch.empty(nnz, dtype=torch.int32, device=keys.device)
    seg_of = torch.empty(max(nnz, 1), dtype=torch.int32, device=keys.device) if want_sorted else None
    ws = _workspace(query('dgmc_csr_build_workspace', nnz, n_seg), keys.device)
    call('dgmc_csr_build', keys, nnz, n_seg, ptr, perm, seg_of, ws, ws.numel(), stream_ptr())
    return (ptr, perm, seg_of) if want_sorted else (ptr, perm)


def _vec_ok(t):
    return t.size(1) % 4 == 0 and t.stride(0) % 4 == 0 and t.data_ptr() % 16 == 0


def segment_gather(x, out_rows, src, ptr=None, fixed_len=0, src_div=1, w_index=None, w=None,
                   src_scale=None, seg_scale=None, out=None, accumulate=False, seg_of=None):
    """out[s] = seg_scale[s] * sum_{e in seg(s)} w_e * x[src[e] // src_div]  (see dgmc_b200.h).
    With `seg_of` (csr_build(..., want_sorted=True)) and float4-aligned rows the load-balanced kernel is used."""
    assert x.dim() == 2 and x.stride(1) == 1
    r = x.size(1)
    if out is None:
        out = torch.empty(out_rows, r, dtype=torch.float32, device=x.device)
        accumulate = False
    assert out.stride(1) == 1 and out.size(0) == out_rows
    if out_rows == 0:
        return out
    if ptr is not None and seg_of is not None and _vec_ok(x) and _vec_ok(out):
        nnz = src.numel()
        ws = _workspace(query('dgmc_balanced_workspace', nnz, r), x.device)
        call('dgmc_segment_gather_balanced', ptr, seg_of, src, int(src_div), w_index, w, src_scale,
             seg_scale, RowStrided(x), x.stride(0), RowStrided(out), out.stride(0), out_rows, nnz, r,
             int(bool(accumulate)), ws, ws.numel(), stream_ptr())
        return out
    call('dgmc_segment_gather', ptr, int(fixed_len), src, int(src_div), w_index, w, src_scale,
         seg_scale, RowStrided(x), x.stride(0), RowStrided(out), out.stride(0), out_rows, r,
         int(bool(accumulate)), stream_ptr())
    return out


class GraphCSR(object):
    """In-/out-adjacency of one (batched) graph in CSR form, built once per edge_index and cached.

    in_ptr/in_nbr: for node i, the sources j of edges j -> i (PyG flow 'source_to_target',
    rel.py:27-28); out_ptr/out_nbr: the destinations of edges i -> j ('target_to_source',
    rel.py:29-30).  inv_in/inv_out = 1 / max(degree, 1) (mean aggregation).  *_edge: the edge
    id behind each CSR slot (SplineConv needs per-edge data)."""
    _cache = {}

    def __init__(self, edge_index, num_nodes):
        # edge lists now; the sorted adjacency arrays only when something reads them (SplineConv, GIN:
        # RelConv works from `rel`, which sorts its own operator)
        self.num_nodes, self.num_edges = num_nodes, edge_index.size(1)
        self.src = edge_index[0].to(torch.int32).contiguous()
        self.dst = edge_index[1].to(torch.int32).contiguous()
        self._lazy = True

    @classmethod
    def union(cls, edge_index, g_a, g_b):
        """GraphCSR of the disjoint union (edge_index = [a's edges, b's edges shifted]) that takes its
        RelConv operator from the parts (no sort)."""
        self = cls(edge_index, g_a.num_nodes + g_b.num_nodes)
        self._rel = _RelOperator.concat(g_a.rel, g_b.rel)
        return self

    _ADJACENCY = ('in_ptr', 'in_edge', 'in_seg', 'out_ptr', 'out_edge', 'out_seg', 'in_nbr', 'out_nbr',
                  'inv_in', 'inv_out')

    def __getattr__(self, name):
        # only reached for attributes that are not set yet: build the adjacency arrays on first use
        if name in GraphCSR._ADJACENCY and self.__dict__.get('_lazy'):
            self.__dict__['_lazy'] = False
            self._build_adjacency()
            return self.__dict__[name]
        raise AttributeError(name)

    def _build_adjacency(self):
        src, dst, num_nodes = self.src, self.dst, self.num_nodes
        self.in_ptr, self.in_edge, self.in_seg = csr_build(dst, num_nodes, True)
        self.out_ptr, self.out_edge, self.out_seg = csr_build(src, num_nodes, True)
        self.in_nbr = src[self.in_edge.long()].contiguous()
        self.out_nbr = dst[self.out_edge.long()].contiguous()
        deg_in = (self.in_ptr[1:] - self.in_ptr[:-1]).clamp_(min=1).to(torch.float32)
        deg_out = (self.out_ptr[1:] - self.out_ptr[:-1]).clamp_(min=1).to(torch.float32)
        self.inv_in, self.inv_out = deg_in.reciprocal_(), deg_out.reciprocal_()

    @property
    def rel(self):
        """CSR of the [3N, N] RelConv aggregation matrix (see _RelAggregate) and of its transpose."""
        if getattr(self, '_rel', None) is None:
            self._rel = _RelOperator(self)
        return self._rel

    @classmethod
    def seed(cls, edge_index, num_nodes, graph):
        key = (edge_index.data_ptr(), tuple(edge_index.shape), edge_index._version, num_nodes,
               str(edge_index.device))
        cls._cache[key] = (edge_index, graph)

    @classmethod
    def of(cls, edge_index, num_nodes):
        key = (edge_index.data_ptr(), tuple(edge_index.shape), edge_index._version, num_nodes,
               str(edge_index.device))
        hit = cls._cache.get(key)
        if hit is None or hit[0] is not edge_index:
            if len(cls._cache) > 64:
                cls._cache.clear()
            hit = (edge_index, cls(edge_index, num_nodes))
            cls._cache[key] = hit
        return hit[1]


def _tile_order(ptr, n, tile=32):
    """For each group of `tile` consecutive nodes the group-local rows by descending neighbour-list
    length (segments 3i+1 and 3i+2 of the three-segments-per-node CSR `ptr`): the fused RelConv kernel
    gives rows that are adjacent in this order to the same warp.  int32 [ceil(n / tile) * tile]."""
    length = (ptr[3::3] - ptr[1::3])[:n]
    tiles = (n + tile - 1) // tile
    padded = torch.full((tiles * tile, ), -1, dtype=torch.int32, device=ptr.device)
    padded[:n] = length
    return padded.view(tiles, tile).argsort(dim=1, descending=True, stable=True).to(torch.int32).view(-1).contiguous()


class _RelOperator(object):
    # row orders for the fused RelConv kernel: computed on first use (the default route never asks)
    @property
    def f_order(self):
        if getattr(self, '_f_order', None) is None:
            self._f_order = _tile_order(self.f_ptr, self.n_out)
        return self._f_order

    @property
    def t_order(self):
        if getattr(self, '_t_order', None) is None:
            self._t_order = _tile_order(self.t_ptr, self.n_out)
        return self._t_order

    def _t(self):
        if getattr(self, '_t_built', None) is None:
            self._t_built = self._t_make()
        return self._t_built

    t_ptr = property(lambda self: self._t()[0])
    t_src = property(lambda self: self._t()[1])
    t_w = property(lambda self: self._t()[2])

    @classmethod
    def concat(cls, a, b):
        """Operator of the disjoint union of two graphs (b's nodes numbered after a's) from the two
        operators: block-diagonal, so the CSRs are concatenations with shifted indices — no sort."""
        self = cls.__new__(cls)
        self.n_out = self.n_in = a.n_out + b.n_out
        n_a = a.n_out

        def join(ptr_a, ptr_b, src_a, src_b, w_a, w_b, seg_a, seg_b, src_shift, seg_shift):
            nnz_a = src_a.numel()
            return (torch.cat([ptr_a[:-1], ptr_b + nnz_a]), torch.cat([src_a, src_b + src_shift]),
                    torch.cat([w_a, w_b]), torch.cat([seg_a[:nnz_a], seg_b + seg_shift]))
        # forward operator [3N, N]: segments = 3 rows per node, sources = node ids
        self.f_ptr, self.f_src, self.f_w, self.f_seg = join(a.f_ptr, b.f_ptr, a.f_src, b.f_src, a.f_w, b.f_w,
                                                            a.f_seg, b.f_seg, n_a, 3 * n_a)
        # its transpose [N, 3N]: segments = node ids, sources = rows of the forward operator
        self.b_ptr, self.b_src, self.b_w, self.b_seg = join(a.b_ptr, b.b_ptr, a.b_src, b.b_src, a.b_w, b.b_w,
                                                            a.b_seg, b.b_seg, 3 * n_a, n_a)
        # operator of the input gradient, again [3N, N] (see __init__): joined on first use
        self._t_make = lambda: join(a.t_ptr, b.t_ptr, a.t_src, b.t_src, a.t_w, b.t_w,
                                    a.t_ptr[:0], b.t_ptr[:0], n_a, 0)[:3]
        return self

    def __init__(self, graph):
        n, dev = graph.num_nodes, graph.src.device
        self.n_out = self.n_in = n
        src, dst = graph.src, graph.dst
        node = torch.arange(n, dtype=torch.int32, device=dev)
        rows = torch.cat([3 * node, 3 * dst + 1, 3 * src + 2])          # segment of A
        cols = torch.cat([node, src, dst])
        # 1 / max(degree, 1) from edge counts (integer histograms: exact, order-independent)
        # (scatter_add on integers rather than bincount, which sizes its result with a host sync)
        src_l, dst_l = src.long(), dst.long()
        one = torch.ones(src.numel(), dtype=torch.int32, device=dev)
        deg_in = torch.zeros(n, dtype=torch.int32, device=dev).scatter_add_(0, dst_l, one)
        deg_out = torch.zeros(n, dtype=torch.int32, device=dev).scatter_add_(0, src_l, one)
        inv_in = deg_in.clamp_(min=1).to(torch.float32).reciprocal_()
        inv_out = deg_out.clamp_(min=1).to(torch.float32).reciprocal_()
        w = torch.cat([torch.ones(n, device=dev), inv_in[dst_l], inv_out[src_l]])
        self.f_ptr, perm, self.f_seg = csr_build(rows, 3 * n, True)
        self.f_src, self.f_w = cols[perm.long()].contiguous(), w[perm.long()].contiguous()
        self.b_ptr, perm, self.b_seg = csr_build(cols, n, True)
        self.b_src, self.b_w = rows[perm.long()].contiguous(), w[perm.long()].contiguous()
        # The input gradient of the layer in the SAME form (three segments per node), for the fused
        # kernel: with z = [x | A_in x | A_out x] W^T,  dx = [g | P g | Q g] [W_root ; W_1 ; W_2]  where
        # (P g)[j] = sum_{j -> i} g[i] / deg_in(i)  and  (Q g)[j] = sum_{i -> j} g[i] / deg_out(i):
        # row 3j+1 lists the out-neighbours of j weighted by THEIR in-degree, row 3j+2 its in-neighbours
        # weighted by their out-degree.
        # (built on first use: only the fused RelConv kernel reads it)
        def make():
            t_rows = torch.cat([3 * node, 3 * src + 1, 3 * dst + 2])
            t_cols = torch.cat([node, dst, src])
            t_ptr, t_perm = csr_build(t_rows, 3 * n)
            return t_ptr, t_cols[t_perm.long()].contiguous(), w[t_perm.long()].contiguous()
        self._t_make = make


class MergedPair(object):
    """Disjoint union of the source and the target graph (target node ids shifted by N_s), so that
    a weight-sharing psi can process both in one pass.  Only used when that is arithmetically
    identical to two separate calls AND draws no random numbers: psi must be this package's
    RelCNN / SplineCNN without active batch-norm statistics or dropout."""
    _cache = {}

    def __init__(self, e_s, e_t, ea_s, ea_t, n_s, n_t=None):
        self.edge_index = torch.cat([e_s, e_t + n_s], dim=1)
        self.edge_attr = None if ea_s is None else torch.cat([ea_s, ea_t], dim=0)
        if n_t is not None:
            # the union's RelConv operator is the concatenation of the parts' (usually already built
            # for psi_1): seed the GraphCSR cache so that psi_2 does not sort the merged edge list
            g = GraphCSR.union(self.edge_index, GraphCSR.of(e_s, n_s), GraphCSR.of(e_t, n_t))
            GraphCSR.seed(self.edge_index, n_s + n_t, g)

    @staticmethod
    def mergeable(psi, training):
        from .models.rel import RelCNN
        from .models.spline import SplineCNN
        if not isinstance(psi, (RelCNN, SplineCNN)):
            return False
        if training and (getattr(psi, 'batch_norm', False) or psi.dropout > 0):
            return False
        return True

    @classmethod
    def of(cls, psi, training, e_s, e_t, ea_s, ea_t, n_s, n_t):
        if not cls.mergeable(psi, training) or (ea_s is None) != (ea_t is None):
            return None
        tensors = [t for t in (e_s, e_t, ea_s, ea_t) if t is not None]
        key = tuple((t.data_ptr(), tuple(t.shape), t._version) for t in tensors) + (n_s, n_t)
        hit = cls._cache.get(key)
        if hit is None or any(a is not b for a, b in zip(hit[0], tensors)):
            if len(cls._cache) > 64:
                cls._cache.clear()
            hit = (tensors, cls(e_s, e_t, ea_s, ea_t, n_s, n_t if ea_s is None else None))
            cls._cache[key] = hit
        return hit[1]


_EXTRA_CACHES = []   # structure caches of other modules (the node partition of the sharded path)

# Set while a step is warmed up / captured that must rebuild every graph structure from its inputs on
# each replay (GraphedTrainStep(structure_is_static=False)): structures whose construction has
# data-dependent shapes (the node partition of the sharded path) are then not used at all.
DYNAMIC_STRUCTURE = [False]


class dynamic_structure(object):
    def __enter__(self):
        self.prev, DYNAMIC_STRUCTURE[0] = DYNAMIC_STRUCTURE[0], True

    def __exit__(self, *exc):
        DYNAMIC_STRUCTURE[0] = self.prev


def clear_caches():
    """Forget cached graph structures (CSR, spline bases, merged pairs, node partitions)."""
    GraphCSR._cache.clear()
    SplineBasis._cache.clear()
    MergedPair._cache.clear()
    for cache in _EXTRA_CACHES:
        cache.clear()


class Candidates(object):
    """The sparse correspondence structure of one forward pass: S_idx (dgmc.py:186-195) as flat
    int32 target rows plus its by-target CSR (dgmc_csr_build), shared by K2/K4/K6 and their
    backward kernels.  idx_local: [B, N_s, k] int (graph-local columns)."""
    def __init__(self, idx_local, n_t):
        B, n_s, k = idx_local.shape
        self.B, self.n_s, self.n_t, self.k = B, n_s, n_t, k
        self.m, self.targets = B * n_s, B * n_t
        flat = idx_local.to(torch.int32)
        if B > 1:
            flat = flat + (torch.arange(B, device=flat.device, dtype=torch.int32) * n_t).view(B, 1, 1)
        self.idx = flat.reshape(self.m, k).contiguous()
        self._csr = None

    @property
    def csr(self):
        if self._csr is None:
            self._csr = csr_build(self.idx.view(-1), self.targets, True)
        return self._csr


# --------------------------------------------------------------------------------------------
# K3
# --------------------------------------------------------------------------------------------
class _SoftmaxK(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = _f32(x)
        y = torch.empty_like(x)
        call('dgmc_softmax_k_fwd', x, y, x.numel() // x.size(-1), x.size(-1), stream_ptr())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        y, = ctx.saved_tensors
        dy = _f32(dy)
        dx = torch.empty_like(y)
        call('dgmc_softmax_k_bwd', y, dy, dx, y.numel() // y.size(-1), y.size(-1), stream_ptr())
        return dx


def softmax_k(x):
    """Row softmax over the last dim (k candidates), dgmc.py:202,205,225."""
    return _SoftmaxK.apply(x)


class _MaskedSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, s_mask, t_mask):
        x = _f32(x)
        B, n_s, n_t = x.shape
        y = torch.empty_like(x)
        call('dgmc_masked_softmax_fwd', x, s_mask, t_mask, y, B, n_s, n_t, stream_ptr())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        y, = ctx.saved_tensors
        B, n_s, n_t = y.shape
        dx = torch.empty_like(y)
        call('dgmc_masked_softmax_bwd', y, _f32(dy), dx, B, n_s, n_t, stream_ptr())
        return dx, None, None


def masked_softmax(x, s_mask_u8, t_mask_u8):
    """masked_softmax of dgmc.py:15-19 with S_mask = s_mask (x) t_mask; masks are uint8."""
    return _MaskedSoftmax.apply(x, s_mask_u8, t_mask_u8)


# --------------------------------------------------------------------------------------------
# K2
# --------------------------------------------------------------------------------------------
def _gather_dot_raw(a, b, idx):
    """a, b: fp32 rows with unit column stride (any row pitch)."""
    m, k = idx.shape
    out = torch.empty(m, k, dtype=torch.float32, device=a.device)
    call('dgmc_gather_dot_fwd', RowStrided(a), a.stride(0), RowStrided(b), b.stride(0), idx, out, m, k,
         a.size(1), stream_ptr())
    return out


class _GatherDot(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, cand):
        a, b = _f32(a), _f32(b)
        ctx.cand = cand
        ctx.save_for_backward(a, b)
        return _gather_dot_raw(a, b, cand.idx)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        cand, g = ctx.cand, _f32(g)
        da = db = None
        if ctx.needs_input_grad[0]:  # da[m] = sum_j g[m,j] b[idx[m,j]]
            da = segment_gather(b, cand.m, cand.idx.view(-1), fixed_len=cand.k, w=g.view(-1))
        if ctx.needs_input_grad[1]:  # db[t] = sum_{e -> t} g[e] a[e // k]   (by-target CSR)
            tptr, perm, seg_of = cand.csr
            db = segment_gather(a, cand.targets, perm, ptr=tptr, src_div=cand.k, w_index=perm,
                                w=g.view(-1), seg_of=seg_of)
        return da, db, None


def gather_dot(a, b, cand):
    """S_hat[m, j] = <a[m], b[idx[m, j]]>  (dgmc.py:198-201).  a [M, C], b [B*N_t, C]."""
    return _GatherDot.apply(a, b, cand)


# --------------------------------------------------------------------------------------------
# K4
# --------------------------------------------------------------------------------------------
class _Transport(torch.autograd.Function):
    @staticmethod
    def forward(ctx, S, r_s, cand, out):
        S, r_s = _f32(S), _rows_f32(r_s)       # r_s may be a row block of a wider table (row pitch > R)
        tptr, perm, seg_of = cand.csr
        ctx.cand = cand
        ctx.save_for_backward(S, r_s)
        return segment_gather(r_s, cand.targets, perm, ptr=tptr, src_div=cand.k, w_index=perm,
                              w=S.view(-1), seg_of=seg_of, out=out)

    @staticmethod
    def backward(ctx, g):
        S, r_s = ctx.saved_tensors
        # g may be a row block of a wider gradient buffer (psi_2's concatenated features): both kernels
        # below take a row pitch, so no contiguous copy
        cand, g = ctx.cand, _rows_f32(g)
        dS = dr = None
        if ctx.needs_input_grad[0]:  # dS[m, j] = <r_s[m], g[idx[m, j]]>
            dS = _gather_dot_raw(r_s, g, cand.idx)
        if ctx.needs_input_grad[1]:  # dr_s[m] = sum_j S[m, j] g[idx[m, j]]
            dr = segment_gather(g, cand.m, cand.idx.view(-1), fixed_len=cand.k, w=S.view(-1))
        return dS, dr, None, None


def transport(S, r_s, cand, out=None):
    """r_t[t] = sum_{(m,j): idx[m,j] = t} S[m,j] r_s[m]  (scatter_add at dgmc.py:209-212),
    computed as a deterministic by-target segmented gather.  -> [B*N_t, R].  `out`: write the result
    there (the row-sharded path hands in the peer-memory buffer its all-reduce reads)."""
    return _Transport.apply(S, r_s, cand, out)


# --------------------------------------------------------------------------------------------
# K6
# --------------------------------------------------------------------------------------------
class _DiffMLP(torch.autograd.Function):
    @staticmethod
    def forward(ctx, s_hat, ps, pt, w2, b2, cand):
        s_hat, ps, pt, w2, b2 = _f32(s_hat), _f32(ps), _f32(pt), _f32(w2), _f32(b2)
        out = torch.empty_like(s_hat)
        call('dgmc_diff_mlp_fwd', s_hat, ps, pt, cand.idx, w2, b2, out, cand.m, cand.k, ps.size(1),
             stream_ptr())
        ctx.cand = cand
        ctx.save_for_backward(ps, pt, w2)
        return out

    @staticmethod
    def backward(ctx, g):
        ps, pt, w2 = ctx.saved_tensors
        cand, g = ctx.cand, _f32(g)
        r = ps.size(1)
        d_ps = torch.empty_like(ps)
        d_w2 = torch.empty_like(w2)
        d_b2 = torch.empty(1, dtype=torch.float32, device=g.device)
        ws = _workspace(query('dgmc_diff_mlp_bwd_workspace', cand.m, r), g.device)
        call('dgmc_diff_mlp_bwd_rows', g, ps, pt, cand.idx, w2, d_ps, d_w2, d_b2, cand.m, cand.k, r,
             ws, ws.numel(), stream_ptr())
        tptr, perm, seg_of = cand.csr
        d_pt = torch.empty_like(pt)
        ws2 = _workspace(query('dgmc_balanced_workspace', perm.numel(), r), g.device)
        call('dgmc_diff_mlp_bwd_targets_balanced', g, ps, pt, tptr, seg_of, perm, w2, d_pt,
             cand.targets, perm.numel(), cand.k, r, ws2, ws2.numel(), stream_ptr())
        return g, d_ps, d_pt, d_w2, d_b2, None


class _DiffMLPMerged(torch.autograd.Function):
    """_DiffMLP on ONE node table p = [source rows ; target rows] (psi_2 ran on the merged pair) with
    the first-layer bias b1 added to the source rows here (b1 = None: the GEMM already did).  The
    backward writes d_ps and d_pt into the row blocks of a single d_p, so autograd does not have to
    zero-pad and add two slice gradients (two fills, two copies and an add of [N_s + N_t, R] per
    consensus step).  rows = (lo, hi): the source rows this call owns (row sharding); the other
    source rows get a zero gradient."""
    @staticmethod
    def forward(ctx, s_hat, p, b1, w2, b2, cand, n_s, lo, hi):
        s_hat, p, w2, b2 = _f32(s_hat), _f32(p), _f32(w2), _f32(b2)
        ps = p[lo:hi] + b1 if b1 is not None else p[lo:hi]
        pt = p[n_s:]
        out = torch.empty_like(s_hat)
        call('dgmc_diff_mlp_fwd', s_hat, ps, pt, cand.idx, w2, b2, out, cand.m, cand.k, ps.size(1),
             stream_ptr())
        ctx.cand, ctx.rows = cand, (n_s, lo, hi)
        ctx.save_for_backward(ps, pt, w2)
        return out

    @staticmethod
    def backward(ctx, g):
        ps, pt, w2 = ctx.saved_tensors
        cand, (n_s, lo, hi), g = ctx.cand, ctx.rows, _f32(g)
        r = ps.size(1)
        d_p = torch.empty(n_s + pt.size(0), r, dtype=torch.float32, device=g.device)
        if lo > 0:
            d_p[:lo].zero_()
        if hi < n_s:
            d_p[hi:n_s].zero_()
        d_ps, d_pt = d_p[lo:hi], d_p[n_s:]
        d_w2 = torch.empty_like(w2)
        d_b2 = torch.empty(1, dtype=torch.float32, device=g.device)
        ws = _workspace(query('dgmc_diff_mlp_bwd_workspace', cand.m, r), g.device)
        call('dgmc_diff_mlp_bwd_rows', g, ps, pt, cand.idx, w2, d_ps, d_w2, d_b2, cand.m, cand.k, r,
             ws, ws.numel(), stream_ptr())
        tptr, perm, seg_of = cand.csr
        ws2 = _workspace(query('dgmc_balanced_workspace', perm.numel(), r), g.device)
        call('dgmc_diff_mlp_bwd_targets_balanced', g, ps, pt, tptr, seg_of, perm, w2, d_pt,
             cand.targets, perm.numel(), cand.k, r, ws2, ws2.numel(), stream_ptr())
        d_b1 = d_ps.sum(0) if ctx.needs_input_grad[2] else None   # b1 is None: no gradient asked
        return g, d_p, d_b1, d_w2, d_b2, None, None, None, None


def diff_mlp_merged(s_hat, p, n_s, b1, w2, b2, cand, rows=None):
    """diff_mlp(s_hat, p[lo:hi] + b1, p[n_s:], w2, b2, cand) for a merged node table p [n_s + n_t, R4];
    rows = (lo, hi) defaults to all n_s source rows."""
    lo, hi = (0, n_s) if rows is None else rows
    assert cand.m == hi - lo and p.size(0) == n_s + cand.targets
    return _DiffMLPMerged.apply(s_hat, p, b1, w2, b2, cand, n_s, lo, hi)


def diff_mlp(s_hat, ps, pt, w2, b2, cand):
    """S_hat + Lin2(ReLU(Lin1(o_s - o_t[S_idx]))) with Lin1 hoisted to the nodes (dgmc.py:219-223).
    ps [M, R4], pt [B*N_t, R4] (R4 = R padded to a multiple of 4), w2 [R4], b2 [1]."""
    return _DiffMLP.apply(s_hat, ps, pt, w2, b2, cand)


# --------------------------------------------------------------------------------------------
# per-node linear map with a split-K weight gradient
# --------------------------------------------------------------------------------------------
WGRAD_TENSOR_CORE = True


class WGradGroup(object):
    """All applications of ONE weight inside one forward pass (the consensus loop applies psi_2 and
    the MLP num_steps times).  Their backward passes each write partial tiles into a shared buffer;
    only the LAST one to run reduces them and hands autograd the summed gradient — the others
    return None (= no contribution), so every parameter receives one gradient per step instead
    of num_steps, and one reduction kernel instead of num_steps."""
    def __init__(self):
        self.calls = 0          # forward applications
        self.done = 0           # backward passes seen
        self.partial = self.partial_b = None
        self.splits = 0
        self.tc = None          # kernel family (tensor core / FFMA) that sized `partial`
        self.written = None     # per application: partial tiles written
        self.split_fwd = self.split_bwd = None   # hi / lo copies of the weight for dgmc_linear_tc


_DEFER = [None]  # {key: WGradGroup} while a deferred-wgrad region is active


class deferred_wgrad(object):
    """Context manager around a region that applies the same Linear maps repeatedly and whose
    every application takes part in the backward pass."""
    def __enter__(self):
        self.prev, _DEFER[0] = _DEFER[0], {}
        return self

    def __exit__(self, *exc):
        _DEFER[0] = self.prev


def region_cached(key, make):
    """Value shared by all uses inside the current deferred_wgrad() region (e.g. the concatenated
    RelConv weight, identical for every consensus step); computed on the spot outside a region."""
    region = _DEFER[0]
    if region is None or not torch.is_grad_enabled():
        return make()
    slot = ('cache', ) + tuple(key)
    if slot not in region:
        region[slot] = make()
    return region[slot]


LINEAR_TENSOR_CORE = True   # forward / input-gradient GEMMs on dgmc_linear_tc (tcgen05, 3xTF32)


def _linear_tc_ok(x, n_out, k_in):
    return LINEAR_TENSOR_CORE and x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 \
        and x.stride(1) == 1 and x.size(0) >= 1 \
        and bool(query('dgmc_linear_tc_supported', x.size(0), n_out, k_in, x.stride(0), x.data_ptr()))


def _split_weight(weight, transposed, holder=None, slot=None):
    """hi / lo split (K-major, padded) of a weight for dgmc_linear_tc; `transposed`: weight is
    [k_in, n_out].  Cached on `holder` (a WGradGroup: one forward pass of one weight)."""
    if holder is not None and getattr(holder, slot) is not None:
        return getattr(holder, slot)
    weight = weight.contiguous()
    k_in, n_out = (weight.size(0), weight.size(1)) if transposed else (weight.size(1), weight.size(0))
    ws = torch.empty(query('dgmc_linear_tc_workspace', n_out, k_in), dtype=torch.uint8,
                     device=weight.device)
    call('dgmc_linear_tc_split_weight', weight, weight.stride(0), int(transposed), n_out, k_in, ws,
         ws.numel(), stream_ptr())
    if holder is not None:
        setattr(holder, slot, ws)
    return ws


def _linear_tc(x, split, bias, relu, n_out, k_in, out=None, bias_rows=None):
    """`out`: optional [rows, n_out] row-strided destination (e.g. a column block of a wider buffer);
    `bias_rows`: add the bias to the first bias_rows rows only."""
    y = torch.empty(x.size(0), n_out, dtype=torch.float32, device=x.device) if out is None else out
    if bias_rows is None or bias is None:
        call('dgmc_linear_tc', RowStrided(x), x.stride(0), split, bias, int(relu), RowStrided(y), y.stride(0),
             x.size(0), n_out, k_in, stream_ptr())
    else:
        call('dgmc_linear_tc_rows', RowStrided(x), x.stride(0), split, bias, int(bias_rows), int(relu),
             RowStrided(y), y.stride(0), x.size(0), n_out, k_in, stream_ptr())
    return y


def _wgrad_tc_chunked(a, b):
    """dw[i, j] = sum_k a[k, i] b[k, j] on dgmc_wgrad_tc_partial, 128 columns of `a` per call
    (the kernel's limit on the operand that becomes the UMMA N dimension).  None if TMA cannot read
    the operands (alignment)."""
    k, m, n = a.size(0), a.size(1), b.size(1)
    if a.stride(1) != 1 or b.stride(1) != 1:
        return None
    if not query('dgmc_wgrad_tc_supported', min(m, 128), n, a.stride(0), b.stride(0), a.data_ptr(),
                 b.data_ptr()):
        return None
    dw = torch.empty(m, n, dtype=torch.float32, device=a.device)
    splits = query('dgmc_wgrad_tc_splits', k, min(m, 128), n)
    partial = torch.empty(splits, min(m, 128) * n, dtype=torch.float32, device=a.device)
    for c0 in range(0, m, 128):
        mc = min(128, m - c0)
        call('dgmc_wgrad_tc_partial', RowStrided(a[:, c0:c0 + mc]), a.stride(0), RowStrided(b), b.stride(0),
             k, mc, n, partial, None, stream_ptr())
        call('dgmc_wgrad_reduce', partial, None, splits, mc, n, dw[c0:c0 + mc], None, stream_ptr())
    return dw


class ReluOwner(object):
    """Who applies the backward of a layer's ReLU.  node_linear(relu=True, relu_owner=o) leaves it to the
    consumer of its output when that consumer announces, during the forward pass, that it will zero the
    gradient where the output is <= 0 (o.masked = True: the transposed aggregation of the next RelConv
    does it inside its store) — one elementwise kernel less per layer and consensus step."""
    def __init__(self):
        self.masked = False


class _NodeLinear(torch.autograd.Function):
    """kn = False: weight [out, in] (torch Linear); kn = True: weight [in, out] (SplineConv)."""
    @staticmethod
    def forward(ctx, x, weight, bias, relu, group, kn, out, bias_rows, relu_owner=None):
        ctx.relu_owner = relu_owner
        k_in, n_out = (weight.size(0), weight.size(1)) if kn else (weight.size(1), weight.size(0))
        if _linear_tc_ok(x, n_out, k_in):
            y = _linear_tc(x, _split_weight(weight, kn, group, 'split_fwd'), bias, relu, n_out, k_in, out,
                           bias_rows)
        else:
            w_kn = weight if kn else weight.t()
            if bias_rows is not None and bias is not None:
                y = x @ w_kn
                y[:bias_rows] += bias
                if relu:
                    y = torch.relu_(y)
            elif relu:  # bias + ReLU in the library GEMM's epilogue
                y = torch._addmm_activation(bias, x, w_kn, use_gelu=False)
            else:
                y = torch.addmm(bias, x, w_kn) if bias is not None else x @ w_kn
            if out is not None:
                out.copy_(y)
                y = out
        if relu:
            ctx.save_for_backward(x, weight, y)
        else:
            ctx.save_for_backward(x, weight)
        ctx.has_bias, ctx.relu, ctx.group, ctx.kn = bias is not None, relu, group, kn
        ctx.bias_rows = bias_rows if bias is not None else None
        if group is not None:
            ctx.slot = group.calls
            group.calls += 1
        return y

    @staticmethod
    def backward(ctx, g):
        if ctx.relu:
            x, weight, y = ctx.saved_tensors
            if ctx.relu_owner is not None and ctx.relu_owner.masked:
                pass                                   # the consumer of y has already zeroed g where y <= 0
            elif g.dim() == 2 and g.stride(1) == 1 and y.stride(1) == 1 and _vec_ok(g) and _vec_ok(y) \
                    and g.dtype == torch.float32:
                gm = torch.empty(g.size(0), g.size(1), dtype=torch.float32, device=g.device)
                call('dgmc_relu_bwd', RowStrided(g), g.stride(0), RowStrided(y), y.stride(0), gm, gm.stride(0),
                     g.size(0), g.size(1), stream_ptr())   # g * (y > 0), row-strided operands
                g = gm
            else:
                g = torch.ops.aten.threshold_backward(g, y, 0)
        else:
            x, weight = ctx.saved_tensors
        # (a gradient that arrives as a column block of a wider table stays where it is when the kernels
        # below can read it with its row pitch)
        g = _rows_f32(g) if (g.dim() == 2 and _vec_ok(g)) else _f32(g)
        kn = ctx.kn
        k_in, n_out = (weight.size(0), weight.size(1)) if kn else (weight.size(1), weight.size(0))
        k = x.size(0)
        dx = None
        if ctx.needs_input_grad[0]:
            if g.stride(1) != 1:
                g = g.contiguous()
            if _linear_tc_ok(g, k_in, n_out):  # dx = g W: the same kernel on the transposed weight
                dx = _linear_tc(g, _split_weight(weight, not kn, ctx.group, 'split_bwd'), None, False,
                                k_in, n_out)
            else:
                dx = g @ (weight.t() if kn else weight)
        dw, db = _weight_grad(g, x, weight, kn, ctx.has_bias, ctx.bias_rows, ctx.group,
                              getattr(ctx, 'slot', 0), ctx.needs_input_grad[1],
                              ctx.has_bias and ctx.needs_input_grad[2])
        return dx, dw, db, None, None, None, None, None, None


def _weight_grad(g, x, weight, kn, has_bias, bias_rows, grp, slot, want_w, want_b):
    """dW (and db) of y = x W^T (+ b) from the output gradient g [nodes, out] and the GEMM operand
    x [nodes, in]; with a WGradGroup the partial tiles of all applications of the weight are reduced
    once, by the last backward pass to run (the others return None)."""
    dw = db = None
    if not (want_w or want_b):
        return None, None
    k_in, n_out = (weight.size(0), weight.size(1)) if kn else (weight.size(1), weight.size(0))
    k = x.size(0)
    if kn or weight.numel() > 64 * 1024 or k < 2048:
        # not a "tiny output, long reduction" shape: no grouping across consensus steps; the
        # tensor-core split-K kernel in chunks of 128 output channels when TMA can read the
        # operands, else an ordinary library GEMM
        if want_w:
            a, b = (x, g) if kn else (g, x)          # dw[i, j] = sum_k a[k, i] b[k, j]
            dw = _wgrad_tc_chunked(a, b) if (WGRAD_TENSOR_CORE and k >= 2048) else None
            if dw is None:
                dw = a.t() @ b
        if want_b:
            db = g.sum(0) if bias_rows is None else g[:bias_rows].sum(0)
        return dw, db
    m, n = n_out, k_in
    # tcgen05 3xTF32 kernel on TMA-loaded MN-major tiles when the operands meet TMA's alignment
    # rules (WGRAD_TENSOR_CORE), else the fp32 FFMA split-K kernel
    tc = WGRAD_TENSOR_CORE and bool(query('dgmc_wgrad_tc_supported', m, n, g.stride(0), x.stride(0),
                                          g.data_ptr(), x.data_ptr()))
    rows_bias = bias_rows is not None      # bias gradient over the leading rows only
    calls, slot = (1, 0) if grp is None else (grp.calls, slot)
    holder = grp if grp is not None else WGradGroup()
    if holder.partial is None:
        holder.tc = tc
        holder.splits = query('dgmc_wgrad_tc_splits' if tc else 'dgmc_wgrad_splits', k, m, n)
        holder.partial = torch.empty(calls * holder.splits, m * n, dtype=torch.float32,
                                     device=g.device)
        holder.partial_b = torch.empty(calls * holder.splits, m, dtype=torch.float32,
                                       device=g.device) if has_bias else None
        holder.written = [False] * calls
    elif holder.tc != tc:
        # every application of a weight must write the same number of partial tiles: an operand
        # that the tensor-core kernel cannot read (alignment) is copied instead of switching kernel
        if holder.tc:
            g, x = g.contiguous().clone(), x.contiguous().clone()
        tc = holder.tc
    fn_partial = 'dgmc_wgrad_tc_partial' if tc else 'dgmc_wgrad_partial'
    off = slot * holder.splits
    pb = None if holder.partial_b is None else holder.partial_b[off:off + holder.splits]
    if rows_bias and tc:
        call('dgmc_wgrad_tc_partial_rows', RowStrided(g), g.stride(0), RowStrided(x), x.stride(0), k, m, n,
             int(bias_rows), holder.partial[off:off + holder.splits], pb, stream_ptr())
    else:
        call(fn_partial, RowStrided(g), g.stride(0), RowStrided(x), x.stride(0), k, m, n,
             holder.partial[off:off + holder.splits], pb, stream_ptr())
        if rows_bias:   # the FFMA kernel sums all rows: overwrite its bias partials
            pb.zero_()
            pb[0] = g[:bias_rows].sum(0)
    holder.written[slot] = True
    holder.done += 1
    if holder.done == calls:  # last backward of this weight: reduce everything once
        assert all(holder.written), 'WGradGroup: an application of the weight ran its backward twice'
        dw = torch.empty_like(weight)
        db = torch.empty(m, dtype=torch.float32, device=g.device) if has_bias else None
        call('dgmc_wgrad_reduce', holder.partial, holder.partial_b, calls * holder.splits, m, n,
             dw, db, stream_ptr())
        holder.partial = holder.partial_b = None
        holder.split_fwd = holder.split_bwd = None
        holder.done = 0  # a second backward over a retained graph starts afresh
    return dw, db


def node_linear(x, weight, bias=None, relu=False, key=None, kn=False, out=None, bias_rows=None,
                relu_owner=None):
    """relu?(x [nodes, in] @ W + bias) with W = weight[out, in]^T (kn=False) or weight[in, out]
    (kn=True).  Forward and input gradient run on the tensor-core kernel dgmc_linear_tc (3xTF32,
    bias and ReLU in its epilogue) whenever x meets TMA's alignment rules, else on the library GEMM;
    the weight / bias gradient (a reduction over all nodes with a tiny output) runs on the split-K
    kernel dgmc_wgrad when the output is small, else stays with the library.  `out`: write the
    result into this row-strided [nodes, out] view (a column block of a layer buffer) instead of a
    fresh tensor.  `bias_rows`: the bias belongs to the first bias_rows nodes only."""
    if x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 \
            and weight.dtype == torch.float32 and (bias is not None or not relu):
        group = None
        if key is not None and _DEFER[0] is not None and torch.is_grad_enabled():
            group = _DEFER[0].setdefault((key, x.size(0), tuple(weight.shape)), WGradGroup())
        return _NodeLinear.apply(x, weight, bias, relu, group, kn, out, bias_rows, relu_owner)
    assert out is None and bias_rows is None, 'node_linear(out=..., bias_rows=...) are CUDA fp32 features'
    y = torch.addmm(bias, x, weight) if (kn and bias is not None) else (
        x @ weight if kn else torch.nn.functional.linear(x, weight, bias))
    return torch.relu(y) if relu else y


# --------------------------------------------------------------------------------------------
# K5: RelConv mean aggregation in both flow directions
# --------------------------------------------------------------------------------------------
class _RelAggregate(torch.autograd.Function):
    """[x | mean_in x | mean_out x] as ONE sparse product A x with A of shape [3N, N]: row 3i is
    e_i, row 3i+1 the in-neighbour mean of node i, row 3i+2 its out-neighbour mean — the row-major
    [3N, C] result IS the [N, 3C] operand of the layer GEMM.  Backward is A^T g with g viewed as
    [3N, C].  Both directions therefore cost one balanced segmented gather (+ its fix-up)."""
    @staticmethod
    def forward(ctx, x, rel):
        x = _f32(x)
        c = x.size(1)
        cat = torch.empty(rel.n_out, 3 * c, dtype=torch.float32, device=x.device)
        segment_gather(x, 3 * rel.n_out, rel.f_src, ptr=rel.f_ptr, w=rel.f_w,
                       out=cat.view(3 * rel.n_out, c), seg_of=rel.f_seg)
        ctx.rel = rel
        return cat

    @staticmethod
    def backward(ctx, g):
        rel = ctx.rel
        g = _f32(g)
        c = g.size(1) // 3
        # the node-sharded path hands in the buffer its reduce-scatter reads (no copy in between)
        buf = getattr(rel, 'grad_buffer', None)
        out = buf(rel.n_in, c, g.device) if buf is not None else None
        dx = segment_gather(g.view(3 * rel.n_out, c), rel.n_in, rel.b_src, ptr=rel.b_ptr, w=rel.b_w,
                            seg_of=rel.b_seg, out=out)
        return dx, None


def _rows_f32(t):
    """fp32 2-D tensor with unit column stride (row pitch may exceed the width)."""
    if t.dtype != torch.float32:
        raise native.DgmcNativeError('dgmc (B200): kernels compute in fp32, got %s' % t.dtype)
    return t if (t.dim() == 2 and t.stride(1) == 1) else t.contiguous()


class _RelAggregatePass(torch.autograd.Function):
    """_RelAggregate that also hands its input through as a second output.  A layer input that
    feeds both the aggregation and the final concatenation then has ONE consumer in the autograd
    graph, and the backward adds the two gradients inside the gather kernel (accumulate mode, in
    place in the concatenation's gradient block) instead of in a separate add kernel."""
    @staticmethod
    def forward(ctx, x, rel, relu_owner=None):
        x = _rows_f32(x)
        c = x.size(1)
        cat = torch.empty(rel.n_out, 3 * c, dtype=torch.float32, device=x.device)
        segment_gather(x, 3 * rel.n_out, rel.f_src, ptr=rel.f_ptr, w=rel.f_w,
                       out=cat.view(3 * rel.n_out, c), seg_of=rel.f_seg)
        ctx.rel = rel
        # x is the ReLU output of the previous layer and this node is its only consumer: its backward
        # zeroes the gradient where x <= 0 (see ReluOwner)
        ctx.relu_x = None
        if relu_owner is not None and ctx.needs_input_grad[0]:
            relu_owner.masked = True
            ctx.relu_x = x
        return cat, x.view_as(x)

    @staticmethod
    def backward(ctx, g, g_pass):
        if not ctx.needs_input_grad[0]:
            return None, None, None
        rel = ctx.rel
        g = _f32(g)
        c = g.size(1) // 3
        g_pass = _rows_f32(g_pass)
        if not _vec_ok(g_pass):
            g_pass = g_pass.contiguous()
        y = ctx.relu_x
        g3 = g.view(3 * rel.n_out, c)
        if y is not None and c <= 128 and y.stride(0) == g_pass.stride(0) and _vec_ok(y) \
                and rel.n_in * g_pass.stride(0) < 2 ** 31 and rel.b_src.numel() > 0:
            # one kernel: g_pass = (g_pass + A^T g) * [y > 0]
            ws = _workspace(query('dgmc_balanced_workspace', rel.b_src.numel(), c), g.device)
            call('dgmc_segment_gather_balanced_relu', rel.b_ptr, rel.b_seg, rel.b_src, rel.b_w, g3, c,
                 RowStrided(g_pass), g_pass.stride(0), RowStrided(y), rel.n_in, rel.b_src.numel(), c, ws,
                 ws.numel(), stream_ptr())
            return g_pass, None, None
        dx = segment_gather(g3, rel.n_in, rel.b_src, ptr=rel.b_ptr, w=rel.b_w,
                            seg_of=rel.b_seg, out=g_pass, accumulate=True)
        if y is not None:                      # the promise of the forward pass, unfused
            dx = torch.ops.aten.threshold_backward(dx, y, 0)
        return dx, None, None


def rel_aggregate_pass(x, graph, relu_owner=None):
    """-> (rel_aggregate(x, graph), x); see _RelAggregatePass.  relu_owner: x is the output of a
    node_linear(relu=True, relu_owner=relu_owner) that nothing else consumes."""
    return _RelAggregatePass.apply(x, graph.rel, relu_owner)


# RelConv layers on the fused gather + tensor-core GEMM kernel (dgmc_relconv_tc).  OFF by default: it is
# parity-tested but measured SLOWER than the two-kernel route at every shape tried (merged DBP15K pair,
# 64 -> 64 channels: 45 us against 16 + 17 us, profiles/README.md "Fused RelConv"): one CTA per SM leaves
# 16 gather warps where the stand-alone aggregation kernel has 40, and the gather is latency- and
# issue-bound.
RELCONV_FUSED = False


def relconv_fused_ok(x, rel, n_out):
    """Whether dgmc_relconv_tc can run this layer: this package's full-graph operator (node i's own
    entry is column i), C_in = 32 or 64, at most 64 output channels, TMA-addressable rows."""
    return (RELCONV_FUSED and x.is_cuda and x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
            and getattr(rel, 't_ptr', None) is not None and rel.n_in == rel.n_out == x.size(0)
            and bool(query('dgmc_relconv_tc_supported', x.size(0), x.size(1), n_out, x.stride(0), x.data_ptr())))


def _relconv_tc(ptr, src, w, order, x, split, bias, relu, addend, stacked, y, n_out):
    call('dgmc_relconv_tc', ptr, src, w, order, RowStrided(x), x.stride(0), split, bias, int(relu),
         None if addend is None else RowStrided(addend), 0 if addend is None else addend.stride(0),
         None if stacked is None else RowStrided(stacked), 0 if stacked is None else stacked.stride(0),
         RowStrided(y), y.stride(0), x.size(0), x.size(1), n_out, stream_ptr())
    return y


class _RelConvFused(torch.autograd.Function):
    """One RelConv layer, relu?(root(x) + mean_in lin1(x) + mean_out lin2(x)) (rel.py:25-31), as ONE
    kernel: the aggregated operand [x | mean_in x | mean_out x] is gathered straight into the GEMM's
    shared-memory stage (dgmc_relconv_tc).  weight = [root | lin1 | lin2] side by side, [C_out, 3 C_in].
    Second output: x handed through (see _RelAggregatePass) so that a layer input that also feeds the
    concatenation has one consumer; its gradient is added in the epilogue of the backward kernel.

    Backward: the same kernel on the transposed operator and the block-transposed weight,
    dx = [g | P g | Q g] [W_root ; W_1 ; W_2] (ops._RelOperator), the weight gradient from the operand
    rows the forward kernel stored (`stacked`)."""
    @staticmethod
    def forward(ctx, x, weight, bias, rel, relu, group, out):
        c_out, c_in = weight.size(0), x.size(1)
        n = x.size(0)
        split = _split_weight(weight, False, group, 'split_fwd')
        want_w = ctx.needs_input_grad[1] or ctx.needs_input_grad[2]
        stacked = torch.empty(n, 3 * c_in, dtype=torch.float32, device=x.device) if want_w else None
        y = torch.empty(n, c_out, dtype=torch.float32, device=x.device) if out is None else out
        _relconv_tc(rel.f_ptr, rel.f_src, rel.f_w, rel.f_order, x, split, bias, relu, None, stacked, y, c_out)
        ctx.rel, ctx.relu, ctx.group, ctx.has_bias = rel, relu, group, bias is not None
        ctx.save_for_backward(weight, stacked, y if relu else None)
        if group is not None:
            ctx.slot = group.calls
            group.calls += 1
        ctx.set_materialize_grads(False)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, g, g_pass):
        weight, stacked, y = ctx.saved_tensors
        rel = ctx.rel
        c_out, c_in = weight.size(0), weight.size(1) // 3
        if g is None:   # only the handed-through input was used
            return g_pass, None, None, None, None, None, None
        g = _rows_f32(g)
        if ctx.relu:
            gm = torch.empty(g.size(0), g.size(1), dtype=torch.float32, device=g.device)
            if _vec_ok(g) and _vec_ok(y):
                call('dgmc_relu_bwd', RowStrided(g), g.stride(0), RowStrided(y), y.stride(0), gm, gm.stride(0),
                     g.size(0), g.size(1), stream_ptr())
                g = gm
            else:
                g = torch.ops.aten.threshold_backward(g.contiguous(), y.contiguous(), 0)
        dx = None
        if ctx.needs_input_grad[0]:
            if g_pass is not None:
                g_pass = _rows_f32(g_pass)
                if not _vec_ok(g_pass):
                    g_pass = g_pass.contiguous()
            if bool(query('dgmc_relconv_tc_supported', g.size(0), c_out, c_in, g.stride(0), g.data_ptr())):
                # W' [C_in, 3 C_out]: block p of W' is block p of the weight, transposed
                def make():
                    w_t = weight.view(c_out, 3, c_in).permute(2, 1, 0).reshape(c_in, 3 * c_out).contiguous()
                    return _split_weight(w_t, False)
                grp = ctx.group
                if grp is not None:
                    if grp.split_bwd is None:
                        grp.split_bwd = make()
                    split = grp.split_bwd
                else:
                    split = make()
                dx = g_pass if g_pass is not None else torch.empty(g.size(0), c_in, dtype=torch.float32,
                                                                   device=g.device)
                _relconv_tc(rel.t_ptr, rel.t_src, rel.t_w, rel.t_order, g, split, None, False, g_pass, None, dx, c_in)
            else:
                d_st = g @ weight                                            # [N, 3 C_in]
                dx = segment_gather(d_st.view(3 * rel.n_out, c_in), rel.n_in, rel.b_src, ptr=rel.b_ptr,
                                    w=rel.b_w, seg_of=rel.b_seg, out=g_pass, accumulate=g_pass is not None)
        elif g_pass is not None:
            dx = g_pass
        dw, db = _weight_grad(g, stacked, weight, False, ctx.has_bias, None, ctx.group,
                              getattr(ctx, 'slot', 0), ctx.needs_input_grad[1],
                              ctx.has_bias and ctx.needs_input_grad[2]) if stacked is not None else (None, None)
        return dx, dw, db, None, None, None, None


def relconv_fused(x, weight, bias, rel, relu=False, key=None, out=None):
    """-> (layer output, x handed through); see _RelConvFused."""
    group = None
    if key is not None and _DEFER[0] is not None and torch.is_grad_enabled():
        group = _DEFER[0].setdefault((key, x.size(0), tuple(weight.shape)), WGradGroup())
    return _RelConvFused.apply(x, weight, bias, rel, relu, group, out)


class _CopyInto(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, buf, offset):
        view = buf[:, offset:offset + x.size(1)]
        view.copy_(x)
        return view

    @staticmethod
    def backward(ctx, g):
        return g, None, None


class _CopyTo(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, dst):
        # through .data: the table's version counter must not move — other row blocks of the same table
        # (disjoint from dst) may already be saved for a backward pass
        dst.data.copy_(x)
        return dst.view_as(dst)

    @staticmethod
    def backward(ctx, g):
        return g, None


def copy_to(x, dst):
    """dst[...] = x for a pre-allocated block `dst` of a larger table; returns dst (gradient: identity)."""
    return _CopyTo.apply(x, dst)


def copy_into(x, buf, offset):
    """buf[:, offset : offset + x.size(1)] = x; returns that column block (gradient: identity)."""
    return _CopyInto.apply(x, buf, offset)


class _JoinSlices(torch.autograd.Function):
    @staticmethod
    def forward(ctx, buf, own_grad, *parts):
        off = 0
        for part in parts:  # the parts must already BE the column blocks of buf, in order
            assert part.data_ptr() == buf.data_ptr() + 4 * off and part.stride(0) == buf.stride(0)
            off += part.size(1)
        assert off == buf.size(1)
        ctx.widths, ctx.own_grad = [part.size(1) for part in parts], own_grad
        return buf.view_as(buf)

    @staticmethod
    def backward(ctx, g):
        if not ctx.own_grad or g.stride(1) != 1:
            g = g.clone(memory_format=torch.contiguous_format)  # consumers accumulate into their block in place
        outs, off = [], 0
        for width in ctx.widths:
            outs.append(g[:, off:off + width])
            off += width
        return (None, None) + tuple(outs)


def join_slices(buf, own_grad, *parts):
    """torch.cat(parts, dim=-1) when every part was written straight into its column block of `buf`
    (node_linear(out=...), copy_into): no copy in the forward pass, column views of the gradient in
    the backward pass.  own_grad: the gradient of the result is produced by this package's own
    GEMM backward and may be modified in place by the consumers of the parts; otherwise it is cloned
    first."""
    return _JoinSlices.apply(buf, own_grad, *parts)


class _JoinRows(torch.autograd.Function):
    @staticmethod
    def forward(ctx, buf, top, bottom):
        n = top.size(0)
        assert top.data_ptr() == buf.data_ptr() and bottom.data_ptr() == buf[n:].data_ptr() \
            and n + bottom.size(0) == buf.size(0)
        ctx.n = n
        return buf.view_as(buf)

    @staticmethod
    def backward(ctx, g):
        return None, g[:ctx.n], g[ctx.n:]


def join_rows(buf, top, bottom):
    """torch.cat([top, bottom], dim=0) when both were written straight into `buf` (top = buf[:n],
    bottom = buf[n:]): no copy in the forward pass, row views of the gradient in the backward pass."""
    return _JoinRows.apply(buf, top, bottom)


def rel_aggregate(x, graph):
    """-> [N, 3C] = [x | mean_{j->i} x_j | mean_{i->j} x_j]  (the two propagate calls of
    rel.py:27-30, taken BEFORE the linear maps: mean(lin(x_j)) == lin(mean(x_j)))."""
    return _RelAggregate.apply(x, graph.rel)


def rel_aggregate_op(x, rel):
    """Same with an explicit aggregation operator `rel` ([3 n_out, n_in] sparse matrix: f_* its CSR,
    b_* the CSR of its transpose) — used by the node-sharded backbone, whose operator maps the
    all-gathered features to this rank's nodes."""
    return _RelAggregate.apply(x, rel)


class _NeighborSum(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, graph):
        ctx.graph = graph
        x = _f32(x)
        return segment_gather(x, x.size(0), graph.in_nbr, ptr=graph.in_ptr, seg_of=graph.in_seg)

    @staticmethod
    def backward(ctx, g):
        graph = ctx.graph
        g = _f32(g)
        return segment_gather(g, g.size(0), graph.out_nbr, ptr=graph.out_ptr,
                              seg_of=graph.out_seg), None


def neighbor_sum(x, graph):
    """sum_{j -> i} x_j (GINConv's aggregation, gin.py:23 [3p])."""
    return _NeighborSum.apply(x, graph)


# --------------------------------------------------------------------------------------------
# K5: SplineConv
# --------------------------------------------------------------------------------------------
class SplineBasis(object):
    """(basis, widx) of one edge_attr tensor, computed once and cached (edge attributes are static
    across the consensus steps and, in the reference scripts, across epochs)."""
    _cache = {}

    def __init__(self, pseudo, kernel_size):
        pseudo = _f32(pseudo)
        if pseudo.dim() == 1:
            pseudo = pseudo.view(-1, 1)
        E, dim = pseudo.shape
        self.n_corners = 1 << dim
        self.basis = torch.empty(E, self.n_corners, dtype=torch.float32, device=pseudo.device)
        self.widx = torch.empty(E, self.n_corners, dtype=torch.int32, device=pseudo.device)
        call('dgmc_spline_basis', pseudo, E, dim, kernel_size, self.basis, self.widx, stream_ptr())

    @classmethod
    def of(cls, pseudo, kernel_size):
        key = (pseudo.data_ptr(), tuple(pseudo.shape), pseudo._version, kernel_size,
               str(pseudo.device))
        hit = cls._cache.get(key)
        if hit is None or hit[0] is not pseudo:
            if len(cls._cache) > 64:
                cls._cache.clear()
            hit = (pseudo, cls(pseudo, kernel_size))
            cls._cache[key] = hit
        return hit[1]


class _SplineGather(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y, base, graph, sb, n_kernels):
        """y [N, K*C] -> base + mean-aggregated spline messages [N, C]."""
        y = _f32(y)
        n, c = base.shape
        out = _f32(base).clone()
        call('dgmc_spline_gather_fwd', y, graph.in_ptr, graph.in_nbr, graph.in_edge, sb.basis,
             sb.widx, sb.n_corners, graph.inv_in, out, out.stride(0), n, n_kernels, c, 1,
             stream_ptr())
        ctx.graph, ctx.sb, ctx.n_kernels = graph, sb, n_kernels
        return out

    @staticmethod
    def backward(ctx, g):
        graph, sb, K = ctx.graph, ctx.sb, ctx.n_kernels
        g = _f32(g)
        n, c = g.shape
        dy = torch.empty(n, K * c, dtype=torch.float32, device=g.device)
        call('dgmc_spline_gather_bwd', g, g.stride(0), graph.out_ptr, graph.out_nbr, graph.out_edge,
             sb.basis, sb.widx, sb.n_corners, graph.inv_in, dy, n, K, c, stream_ptr())
        return dy, g, None, None, None


def spline_gather(y, base, graph, sb, n_kernels):
    """base[i] + (1/deg_i) sum_{e -> i} sum_s basis[e,s] * y[src_e, widx[e,s]]  (PyG SplineConv)."""
    return _SplineGather.apply(y, base, graph, sb, n_kernels)


# --------------------------------------------------------------------------------------------
# loss / acc / hits_at_k on the sparse side-cars
# --------------------------------------------------------------------------------------------
def sparse_metrics(idx, val, y, top_k=1, eps=1e-8):
    """One kernel over the ground-truth pairs y [2, G] of a sparse correspondence (idx int64 / val fp32
    [rows, k]) -> device tensors (nll [G], hits [G], correct [G], in_top [G]); see dgmc_sparse_metrics."""
    idx, val = idx.contiguous(), _f32(val.detach())
    y_row, y_col = y[0].contiguous(), y[1].contiguous()
    assert idx.dtype == torch.int64 and y_row.dtype == torch.int64 and idx.shape == val.shape
    pairs = y_row.numel()
    out = torch.zeros(4, max(pairs, 1), dtype=torch.float32, device=val.device)
    call('dgmc_sparse_metrics', idx, val, y_row, y_col, pairs, idx.size(1), int(top_k), float(eps),
         out[0], out[1], out[2], out[3], stream_ptr())
    return out[0, :pairs], out[1, :pairs], out[2, :pairs], out[3, :pairs]


# --------------------------------------------------------------------------------------------
# dense branch
# --------------------------------------------------------------------------------------------
def _bmm_nn(a, bm, trans_a=False):
    """out[b] = op(a[b]) @ bm[b] for small per-graph matrices (dgmc_small_bmm_nn), exact fp32."""
    a, bm = _f32(a), _f32(bm)
    batch = a.size(0)
    m, k = (a.size(2), a.size(1)) if trans_a else (a.size(1), a.size(2))
    n = bm.size(2)
    assert bm.size(1) == k and bm.size(0) == batch
    out = torch.empty(batch, m, n, dtype=torch.float32, device=a.device)
    call('dgmc_small_bmm_nn', a, int(trans_a), bm, out, batch, m, k, n, stream_ptr())
    return out


def _bmm_nt(a, bm):
    """out[b] = a[b] @ bm[b]^T  (a [B, m, k], bm [B, n, k])."""
    a, bm = _f32(a), _f32(bm)
    batch, m, k = a.shape
    n = bm.size(1)
    out = torch.empty(batch, m, n, dtype=torch.float32, device=a.device)
    call('dgmc_small_bmm_nt', a, bm, out, batch, m, n, k, stream_ptr())
    return out


def dense_small_ok(n_s, n_t):
    return n_s <= 128 and n_t <= 128


class _DenseTransport(torch.autograd.Function):
    """r_t = S^T r_s per graph (dgmc.py:170): S [B, N_s, N_t], r_s [B, N_s, R] -> [B, N_t, R]."""
    @staticmethod
    def forward(ctx, S, r_s):
        ctx.save_for_backward(S, r_s)
        return _bmm_nn(S, r_s, trans_a=True)

    @staticmethod
    def backward(ctx, g):
        S, r_s = ctx.saved_tensors
        dS = _bmm_nt(r_s, g) if ctx.needs_input_grad[0] else None        # dS[i, j] = <r_s[i], g[j]>
        dr = _bmm_nn(S, g) if ctx.needs_input_grad[1] else None           # dr_s = S g
        return dS, dr


def dense_transport(S, r_s):
    return _DenseTransport.apply(S, r_s)


class _DenseScores(torch.autograd.Function):
    """(S_hat, S_0) = (h_s h_t^T, masked_softmax(S_hat)) per graph (dgmc.py:163-165) on the tensor cores
    (dgmc_dense_scores_tc); adjoints through the small batched products."""
    @staticmethod
    def forward(ctx, h_s, h_t, sm, tm):
        h_s, h_t = _f32(h_s), _f32(h_t)
        B, n_s, c = h_s.shape
        n_t = h_t.size(1)
        s_hat = torch.empty(B, n_s, n_t, dtype=torch.float32, device=h_s.device)
        s_0 = torch.empty_like(s_hat)
        ws = _workspace(query('dgmc_dense_scores_tc_workspace', B, n_t, c) + 128, h_s.device)
        ws = ws[(-ws.data_ptr()) % 128:]
        call('dgmc_dense_scores_tc', h_s, h_t, sm, tm, s_hat, s_0, B, n_s, n_t, c, ws, ws.numel(), stream_ptr())
        ctx.save_for_backward(h_s, h_t, s_0)
        return s_hat, s_0

    @staticmethod
    def backward(ctx, g_hat, g_0):
        h_s, h_t, s_0 = ctx.saved_tensors
        B, n_s, n_t = s_0.shape
        g = None
        if g_0 is not None:   # through the softmax: dx = y (dy - <y, dy>), zero outside the mask
            g = torch.empty_like(s_0)
            call('dgmc_masked_softmax_bwd', s_0, _f32(g_0), g, B, n_s, n_t, stream_ptr())
        if g_hat is not None:
            g = _f32(g_hat) if g is None else g + g_hat
        if g is None:
            return None, None, None, None
        d_hs = _bmm_nn(g, h_t) if ctx.needs_input_grad[0] else None                    # dS h_t
        d_ht = _bmm_nn(g, h_s, trans_a=True) if ctx.needs_input_grad[1] else None      # dS^T h_s
        return d_hs, d_ht, None, None


def dense_scores_ok(h_s, h_t):
    B, n_s, c = h_s.shape
    return h_s.is_cuda and h_s.dtype == torch.float32 and dense_small_ok(n_s, h_t.size(1)) and c % 4 == 0


def dense_scores(h_s, h_t, sm, tm):
    """-> (S_hat, S_0) of the dense variant; h_s [B, N_s, C], h_t [B, N_t, C] zero-padded batches, masks uint8."""
    return _DenseScores.apply(h_s.contiguous(), h_t.contiguous(), sm, tm)


class _DenseConsensus(torch.autograd.Function):
    @staticmethod
    def forward(ctx, s_hat, ps, pt, w2, b2, sm, tm):
        s_hat, ps, pt, w2, b2 = _f32(s_hat), _f32(ps), _f32(pt), _f32(w2), _f32(b2)
        B, n_s, n_t = s_hat.shape
        out = torch.empty_like(s_hat)
        call('dgmc_dense_consensus_fwd', s_hat, ps, pt, w2, b2, sm, tm, out, B, n_s, n_t,
             ps.size(-1), stream_ptr())
        ctx.save_for_backward(ps, pt, w2, sm, tm)
        return out

    @staticmethod
    def backward(ctx, g):
        ps, pt, w2, sm, tm = ctx.saved_tensors
        g = _f32(g)
        B, n_s, n_t = g.shape
        r = ps.size(-1)
        d_ps, d_pt, d_w2 = torch.empty_like(ps), torch.empty_like(pt), torch.empty_like(w2)
        d_b2 = torch.empty(1, dtype=torch.float32, device=g.device)
        ws = _workspace(query('dgmc_dense_consensus_bwd_workspace', r), g.device)
        call('dgmc_dense_consensus_bwd', g, ps, pt, w2, sm, tm, d_ps, d_pt, d_w2, d_b2, B, n_s, n_t,
             r, ws, ws.numel(), stream_ptr())
        return g, d_ps, d_pt, d_w2, d_b2, None, None


def dense_consensus(s_hat, ps, pt, w2, b2, sm, tm):
    """S_hat + mlp(o_s[:, :, None] - o_t[:, None]).masked_fill(~S_mask, 0)  (dgmc.py:178-179)
    with the first Linear hoisted: ps [B, N_s, R4], pt [B, N_t, R4]."""
    return _DenseConsensus.apply(s_hat, ps, pt, w2, b2, sm, tm)


# --------------------------------------------------------------------------------------------
# K1
# --------------------------------------------------------------------------------------------
def topk_tensor_core_ok(h_s, h_t, k):
    """Whether the tcgen05 streaming top-k kernel supports this shape (dgmc_b200.h)."""
    _, n_s, c = h_s.shape
    n_t = h_t.size(1)
    return n_s >= 128 and n_t >= 128 and bool(query('dgmc_topk_tc_supported', n_s, n_t, c, k))


def topk_tensor_core(h_s, h_t, k, want_scores=False):
    """h_s [B, N_s, C], h_t [B, N_t, C] -> int32 [B, N_s, k] via the tcgen05/TMA kernel (one call
    per graph pair; sparse matching is single-pair in every reference script)."""
    h_s, h_t = _f32(h_s), _f32(h_t)
    B, n_s, c = h_s.shape
    n_t = h_t.size(1)
    idx = torch.empty(B, n_s, k, dtype=torch.int32, device=h_s.device)
    score = torch.empty(B, n_s, k, dtype=torch.float32, device=h_s.device) if want_scores else None
    ws = _workspace(query('dgmc_topk_tc_workspace', n_s, n_t, c, k) + 256, h_s.device)
    off = (-ws.data_ptr()) % 256
    ws = ws[off:]
    for b in range(B):
        call('dgmc_topk_tc', h_s[b], c, h_t[b], c, n_s, n_t, c, k, idx[b],
             None if score is None else score[b], ws, ws.numel(), stream_ptr())
    return (idx, score) if want_scores else idx


def topk_simt(h_s, h_t, k, want_scores=False):
    """Exact fp32 top-k for small pairs: h_s [B, N_s, C], h_t [B, N_t, C] -> int32 [B, N_s, k]."""
    h_s, h_t = _f32(h_s), _f32(h_t)
    B, n_s, c = h_s.shape
    n_t = h_t.size(1)
    idx = torch.empty(B, n_s, k, dtype=torch.int32, device=h_s.device)
    score = torch.empty(B, n_s, k, dtype=torch.float32, device=h_s.device) if want_scores else None
    call('dgmc_topk_simt', h_s, h_t, B, n_s, n_t, c, k, idx, score, stream_ptr())
    return (idx, score) if want_scores else idx
