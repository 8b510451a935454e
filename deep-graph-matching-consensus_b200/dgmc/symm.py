"""Collectives of the row-sharded path over NVLink peer memory (no counterpart in the reference).

`PeerCollectives` owns one symmetric allocation per process group (torch.distributed._symmetric_memory
does the allocation and the exchange of the peer mappings — plumbing) and runs the library's own
kernels on it (csrc/symm.cu: dgmc_sym_all_reduce / all_gather / reduce_scatter): every rank reads and
writes its peers' buffers directly (2 GPUs: 22 us per 5 MB all-reduce where NCCL takes 43 us, 14-17 us per
all-gather / reduce-scatter of 1 MB blocks against 19-21 us), and every
rank receives bit-identical sums.  Layout of the allocation: [flag words | input scratch | output
scratch]; results are copied out (or consumed) before the next call reuses the scratch — every call opens
and closes with a handshake, which is what makes that safe across ranks (also across replays of a
captured graph).

Used when the group's ranks are CUDA devices of one node with peer access; otherwise (CPU / gloo tests,
no NVLink) `PeerCollectives.of()` returns None and the callers use torch.distributed collectives.
"""
import ctypes
import os

import torch
import torch.distributed as dist

from . import native
from .native import call, query, stream_ptr

ENABLED = os.environ.get('DGMC_PEER_COLLECTIVES', '1') != '0'


class PeerCollectives(object):
    _by_group = {}

    def __init__(self, group, device, scratch_bytes):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.pad_bytes = (int(query('dgmc_sym_pad_bytes')) + 255) // 256 * 256
        self.scratch_bytes = (int(scratch_bytes) + 255) // 256 * 256
        # [flag words | input scratch | output scratch]
        total = self.pad_bytes + 2 * self.scratch_bytes
        self.buf = symm_mem.empty(total, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group=dist.group.WORLD if group is None else group)
        assert self.handle.world_size == self.world and self.handle.rank == self.rank
        self.bases = (ctypes.c_void_p * self.world)(*[int(p) for p in self.handle.buffer_ptrs])
        self.in_off, self.out_off = self.pad_bytes, self.pad_bytes + self.scratch_bytes
        # NVLink-SHARP multicast mapping of the allocation (0 when the fabric has none).  Opt-in: on two
        # GPUs the switch reduction measured SLOWER than peer loads (26.8 vs 22.5 us per 5 MB all-reduce)
        mc = int(getattr(self.handle, 'multicast_ptr', 0) or 0)
        self.multicast = ctypes.c_void_p(mc) if (mc and os.environ.get('DGMC_PEER_MULTIMEM', '0') == '1') else None
        torch.cuda.synchronize(device)
        dist.barrier(group)     # every rank's flag words are zero before anyone signals

    # ---- scratch views (this rank's copy) ------------------------------------------------------
    def _view(self, off, rows, cols):
        n = rows * cols * 4
        assert n <= self.scratch_bytes
        return self.buf[off:off + n].view(torch.float32).view(rows, cols)

    def scratch_in(self, rows, cols):
        """Where the NEXT all_reduce / reduce_scatter call reads this rank's contribution."""
        return self._view(self.in_off, rows, cols)

    def scratch_out(self, rows, cols):
        return self._view(self.out_off, rows, cols)

    # ---- collectives -------------------------------------------------------------------------------
    def all_reduce(self, rows, cols):
        """Sum over ranks of what every rank has written into scratch_in(rows, cols) -> scratch_out view
        (valid until the next collective of this object)."""
        n = rows * cols
        assert n % 4 == 0
        call('dgmc_sym_all_reduce', self.bases, self.multicast, self.rank, self.world, self.in_off, self.out_off, 0, n,
             stream_ptr())
        return self.scratch_out(rows, cols)

    def all_gather(self, x, runs=None, total_rows=None, max_rows=None):
        """x [n, c] (any memory) -> scratch_out view of the gathered tensor.  Default: equal blocks in rank
        order, [world * n, c].  runs = ((row0, n0), (row1, n1)): this rank's rows are rows row0.. and
        row1.. of a [total_rows, c] tensor (x holds them in that order); max_rows: the largest n0 + n1
        over the ranks."""
        x = x.contiguous()
        c = x.size(1)
        assert x.dtype == torch.float32 and c % 4 == 0
        if runs is None:
            runs, total_rows, max_rows = ((self.rank * x.size(0), x.size(0)), (0, 0)), self.world * x.size(0), x.size(0)
        (r0, n0), (r1, n1) = runs
        assert n0 + n1 == x.size(0) and total_rows * c * 4 <= self.scratch_bytes
        # both handshakes, always: the C ABI lets a caller that alternates between two output regions skip
        # the opening one, but a replayed CUDA graph repeats the region choice made at capture time, so
        # alternation cannot be guaranteed across replays (an odd number of calls per graph breaks it)
        call('dgmc_sym_all_gather', self.bases, self.rank, self.world, x, self.out_off, 0, r0 * c, n0 * c,
             r1 * c, n1 * c, max_rows * c, 1, stream_ptr())
        return self.scratch_out(total_rows, c)

    def reduce_scatter(self, rows, cols, runs=None, max_rows=None):
        """Every rank has written the whole tensor into scratch_in -> this rank's summed part, a fresh
        tensor [rows, cols] (default: block `rank` of `world` equal blocks; else its `runs`)."""
        assert cols % 4 == 0
        if runs is None:
            runs, max_rows = ((self.rank * rows, rows), (0, 0)), rows
        (r0, n0), (r1, n1) = runs
        assert n0 + n1 == rows
        out = torch.empty(rows, cols, dtype=torch.float32, device=self.buf.device)
        call('dgmc_sym_reduce_scatter', self.bases, self.rank, self.world, self.in_off, out, 0, r0 * cols,
             n0 * cols, r1 * cols, n1 * cols, max_rows * cols, 1, stream_ptr())     # both handshakes, see all_gather
        return out

    def timeouts(self):
        return int(native.lib.dgmc_sym_timeouts())

    # ---- one object per process group, grown (collectively) when a larger message shows up --------
    @classmethod
    def of(cls, group, device, need_bytes):
        """The group's PeerCollectives with at least need_bytes of scratch per direction, or None when
        peer memory is not available.  Every rank calls this with the same need_bytes (the message
        shapes of the sharded path are the same on all ranks), so growing is itself collective."""
        if not ENABLED or not dist.is_initialized() or device.type != 'cuda' or \
                dist.get_backend(group) != 'nccl' or dist.get_world_size(group) > 8:
            return None
        key = id(group)
        cur = cls._by_group.get(key)
        if cur is False:
            return None
        if cur is None or cur.scratch_bytes < need_bytes:
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError('dgmc: the peer-memory scratch must be sized before CUDA-graph capture '
                                   '(run one eager step first)')
            try:
                size = max(int(need_bytes), 1 << 24)
                cur = cls(group, device, size if cur is None else max(size, 2 * cur.scratch_bytes))
            except Exception as exc:   # no peer access / symmetric memory unsupported on this box
                if os.environ.get('DGMC_PEER_COLLECTIVES') == 'require':
                    raise
                import warnings
                warnings.warn('dgmc: peer-memory collectives unavailable (%s); using torch.distributed' % exc)
                cur = False
            cls._by_group[key] = cur
        return cur or None
