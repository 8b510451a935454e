#!/usr/bin/env python
"""bench.py — DGMC forward+backward steps/s at the DBP15K shape (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One step = what examples/dbp15k.py:37-47 does in its consensus phase (:68-69): zero_grad,
forward(train mode, detach=True, k=10 (+10 negatives, ground truth included), num_steps=10),
loss(S_L, train_y), backward, Adam step.  Workload = BASELINE.json configs[2]: a synthetic
DBP15K zh_en-shaped pair (19,388 x 19,572 nodes, 300-d features, RelCNN 300->256 L3 / RelCNN
64->64 L3).  Prints ONE JSON line (contract in the task statement).  `value` is measured with
inputs resident in HBM; `e2e` through the public API with pinned HOST inputs copied every step
and the loss read back.  `--impl reference` times the CPU oracle port of the reference
(the unmodified reference package from oracle/_ref — copied there, git-ignored, by oracle/make_ref.py
in the build container — else the oracle port oracle/dgmc_oracle.py).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, 'deep-graph-matching-consensus_b200')
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

# stdout carries exactly ONE JSON line: native libraries (NCCL prints its version banner there)
# write to file descriptor 1 directly, so point fd 1 at stderr and keep the real stdout aside.
_REAL_STDOUT = os.fdopen(os.dup(1), 'w')
os.dup2(2, 1)

import torch  # noqa: E402


def emit(obj):
    _REAL_STDOUT.write(json.dumps(obj) + '\n')
    _REAL_STDOUT.flush()

WORKLOAD = dict(name='DBP15K zh_en-shaped synthetic pair', n_s=19388, n_t=19572, feat=300,
                e_s=70414, e_t=95142, dim=256, rnd_dim=64, layers=3, k=10, num_steps=10,
                train_pairs=4500)
METRIC = 'DGMC fwd+bwd steps/s at DBP15K shape'


def make_inputs(w, seed=0):
    """Seeded synthetic pair with planted structure: the first n_s target nodes are a permuted,
    noisy copy of the source nodes, so top-k score gaps look like an alignment task rather than
    pure-noise ties."""
    g = torch.Generator().manual_seed(seed)
    x_s = torch.randn(w['n_s'], w['feat'], generator=g)
    perm = torch.randperm(w['n_t'], generator=g)[:w['n_s']]
    x_t = torch.randn(w['n_t'], w['feat'], generator=g)
    x_t[perm] = x_s + 0.5 * torch.randn(w['n_s'], w['feat'], generator=g)
    e_s = torch.randint(w['n_s'], (2, w['e_s']), generator=g)
    e_t = torch.randint(w['n_t'], (2, w['e_t']), generator=g)
    rows = torch.sort(torch.randperm(w['n_s'], generator=g)[:w['train_pairs']])[0]
    y = torch.stack([rows, perm[rows]])
    return dict(x_s=x_s, e_s=e_s, x_t=x_t, e_t=e_t, y=y)


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    FIELDS = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
              'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
              'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.FIELDS,
                 '--format=csv,noheader,nounits', '-lms', '100'], stdout=subprocess.PIPE,
                stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(',')])

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        sm.sort()
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': max(mx), 'reasons': sorted(reasons),
                'samples': len(sm)}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the CPU oracle port of the reference
# ------------------------------------------------------------------------------------------------
def oracle_specs(w):
    s1 = dict(kind='rel', out=w['dim'], layers=w['layers'], batch_norm=False, cat=True, lin=True,
              dropout=0.5)
    s1['in'] = w['feat']
    s2 = dict(kind='rel', out=w['rnd_dim'], layers=w['layers'], batch_norm=False, cat=True, lin=True,
              dropout=0.0)
    s2['in'] = w['rnd_dim']
    return s1, s2


REF_DIR = os.path.join(ROOT, 'oracle', '_ref')


def _reference_step_fn(w, inputs):
    """One training step of the UNMODIFIED reference package (oracle/_ref/dgmc, copied from
    /root/reference by oracle/make_ref.py), imported in a way that cannot collide with the product
    package of the same name: examples/dbp15k.py:37-47 — zero_grad, forward(train, y), loss(S_L),
    backward, Adam.  torch_geometric / torch_scatter come from the stand-ins in oracle/shims/ (not
    installable offline).  One deviation, which SURVEY.md §8(c) prescribes: DGMC.__top_k__'s non-KeOps
    route materialises [N_s, N_t, C] = 388.6 GB at this shape, so it is replaced by the same arithmetic
    done per row chunk (matmul + topk)."""
    import importlib
    saved_path, saved_mods = list(sys.path), {k: v for k, v in sys.modules.items() if k == 'dgmc' or k.startswith('dgmc.')}
    for k in saved_mods:
        del sys.modules[k]
    sys.path[:0] = [REF_DIR, os.path.join(ROOT, 'oracle', 'shims')]
    try:
        ref = importlib.import_module('dgmc')
        assert os.path.realpath(ref.__file__).startswith(os.path.realpath(REF_DIR)), ref.__file__
        from dgmc.models import DGMC, RelCNN
    finally:
        sys.path[:] = saved_path

    def chunked_top_k(self, x_s, x_t):
        out = []
        for lo in range(0, x_s.size(1), 2048):
            out.append((x_s[:, lo:lo + 2048] @ x_t.transpose(-1, -2)).topk(self.k, dim=2)[1])
        return torch.cat(out, dim=1)
    DGMC.__top_k__ = chunked_top_k
    torch.manual_seed(0)
    psi_1 = RelCNN(w['feat'], w['dim'], w['layers'], batch_norm=False, cat=True, lin=True, dropout=0.5)
    psi_2 = RelCNN(w['rnd_dim'], w['rnd_dim'], w['layers'], batch_norm=False, cat=True, lin=True, dropout=0.0)
    model = DGMC(psi_1, psi_2, num_steps=w['num_steps'], k=w['k'], detach=True)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    import warnings

    def one_step():
        opt.zero_grad()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            _, S_L = model(inputs['x_s'], inputs['e_s'], None, None, inputs['x_t'], inputs['e_t'], None, None,
                           inputs['y'])
            loss = model.loss(S_L, inputs['y'])
        loss.backward()
        opt.step()
        return loss.item()
    # leave the reference's modules out of sys.modules again: the GPU arm imports the product's `dgmc`
    for k in [k for k in sys.modules if k == 'dgmc' or k.startswith('dgmc.')]:
        del sys.modules[k]
    sys.modules.update(saved_mods)
    return one_step


def _port_step_fn(w, inputs):
    """The oracle port (oracle/dgmc_oracle.py) of the same step — used when oracle/_ref is absent."""
    import torch.nn.functional as F
    from oracle import dgmc_oracle as O
    s1, s2 = oracle_specs(w)
    params = {k: v.requires_grad_(True) for k, v in O.init_dgmc(0, s1, s2).items()}
    opt = torch.optim.Adam(list(params.values()), lr=1e-3)
    gen = torch.Generator().manual_seed(1)

    def one_step():
        opt.zero_grad()
        neg = torch.randint(w['n_t'], (1, w['n_s'], min(w['k'], w['n_t'] - w['k'])), generator=gen)
        draws = {'neg': neg, 'r_s': [torch.randn(1, w['n_s'], w['rnd_dim'], generator=gen)
                                     for _ in range(w['num_steps'])]}
        out = O.dgmc_forward(s1, s2, params, inputs['x_s'], inputs['e_s'], None, None,
                             inputs['x_t'], inputs['e_t'], None, None, w['num_steps'], w['k'],
                             y=inputs['y'], training=True, detach=True, draws=draws,
                             dropout_fn=lambda x, p: F.dropout(x, p, True))
        loss = O.loss(out['S_L'], inputs['y'], idx=out['idx'])
        loss.backward()
        opt.step()
        return loss.item()
    return one_step


def cpu_reference_steps(w, inputs, steps, warmup, budget_s):
    """Times the reference's CPU implementation of the step on all host threads: the unmodified
    reference package from oracle/_ref when it is there (kind 'reference'), else the oracle port (kind
    'port').  Returns (steps_per_s, steps_timed, cores, seconds_per_step, warmups, kind, what)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    if os.path.isdir(os.path.join(REF_DIR, 'dgmc')):
        one_step, kind = _reference_step_fn(w, inputs), 'reference'
        what = ('unmodified reference package (oracle/_ref/dgmc, sha256 %s) through the torch_geometric / '
                'torch_scatter stand-ins of oracle/shims, __top_k__ row-chunked (matmul + topk)' %
                open(os.path.join(REF_DIR, 'DIGEST')).read().strip()[:12])
    else:
        one_step, kind = _port_step_fn(w, inputs), 'port'
        what = 'oracle port of the reference (oracle/dgmc_oracle.py: torch CPU ops, row-chunked matmul + topk)'

    t0 = time.perf_counter()
    one_step()  # warm-up (also sizes the sample)
    first = time.perf_counter() - t0
    extra = max(0, min(warmup - 1, int(budget_s * 0.2 / max(first, 1e-3))))
    for _ in range(extra):
        one_step()
    n = max(1, min(steps, int(budget_s * 0.8 / max(first, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        one_step()
    dt = (time.perf_counter() - t0) / n
    return 1.0 / dt, n, cores, dt, 1 + extra, kind, what


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def build_model(w, device, sharded=False):
    from dgmc.models import DGMC, RelCNN, RowShardedDGMC
    torch.manual_seed(0)
    psi_1 = RelCNN(w['feat'], w['dim'], w['layers'], batch_norm=False, cat=True, lin=True, dropout=0.5)
    psi_2 = RelCNN(w['rnd_dim'], w['rnd_dim'], w['layers'], batch_norm=False, cat=True, lin=True,
                   dropout=0.0)
    cls = RowShardedDGMC if sharded else DGMC
    model = cls(psi_1, psi_2, num_steps=w['num_steps'], k=w['k'], detach=True).to(device)
    if sharded and os.environ.get('DGMC_SHARD_PSI'):      # study switch: 0 / 1 instead of 'auto'
        model.shard_psi = bool(int(os.environ['DGMC_SHARD_PSI']))
    model.train()
    return model


def train_step(model, opt, d):
    opt.zero_grad(set_to_none=True)
    _, S_L = model(d['x_s'], d['e_s'], None, None, d['x_t'], d['e_t'], None, None, d['y'])
    loss = model.loss(S_L, d['y'])
    loss.backward()
    if hasattr(model, 'sync_gradients'):
        model.sync_gradients()
    opt.step()
    return loss


def timed(fn, iters, flush=None):
    """Average device time (ms) of fn over `iters` launches, CUDA events on the current stream;
    `flush` (a >L2 buffer) is rewritten before every launch so each one starts with a cold L2."""
    total = 0.0
    fn()  # untimed: lazy one-time set-up (cached CSRs, kernel attributes) must not be measured
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        total += a.elapsed_time(b)
    return total / iters


def timed_graph(fn, iters, flush):
    """Like timed(), for calls whose host side (tensor-map encoding, autograd) takes longer than the
    kernels: [flush, fn] and [flush] are captured as CUDA graphs and the difference of their replay
    times is the device time of fn with a cold L2."""
    def capture(body):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            body()
        return g

    def replay_ms(g):
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(iters):
            g.replay()
        b.record()
        b.synchronize()
        return a.elapsed_time(b) / iters
    fn()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        g1 = capture(lambda: (flush.zero_(), fn()))
        g0 = capture(lambda: flush.zero_())
        t1, t0 = replay_ms(g1), replay_ms(g0)
    torch.cuda.current_stream().wait_stream(side)
    return max(t1 - t0, 1e-4)


# ------------------------------------------------------------------------------------------------
# roofline: the hot kernels timed alone, their DRAM traffic read from the committed ncu captures
# ------------------------------------------------------------------------------------------------
def load_ncu_traffic(tag):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch, parsed at run time from the newest
    `profiles/*ncu_full_hot_<tag>.csv` (`ncu --set full` raw-page export of `python bench.py
    --hot-kernels`, which launches every hot kernel once).  -> ({kernel function name: [bytes per
    launch, in launch order]}, file name) or ({}, None) when no capture is committed."""
    import csv
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*ncu_full_hot_%s.csv' % tag)))
    if not files:
        return {}, None
    rows = list(csv.reader(open(files[-1])))
    head, units = rows[0], rows[1]
    col = {name: head.index(name) for name in ('Kernel Name', 'dram__bytes_read.sum', 'dram__bytes_write.sum')}
    scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    out = {}
    for r in rows[2:]:
        if len(r) <= max(col.values()):
            continue
        name = re.sub(r'^void\s+', '', r[col['Kernel Name']])
        name = re.split(r'[<(]', name)[0].split('::')[-1]
        total = sum(float(r[col[c]]) * scale[units[col[c]]] for c in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
        out.setdefault(name, []).append(total)
    return out, os.path.basename(files[-1])


def hot_kernel_cases(w, model, e_t, e_s=None, with_k1=True):
    """The hot kernels of one training step on step-shaped data.  Each case: what to launch, how many
    times a step launches it, its algorithmic work per launch (`bytes_min` — every
    operand read once, every result written once — when the gathered table fits in L2, else the bytes
    the gathers really request), and which kernel functions one launch consists of."""
    from dgmc import ops
    dev = e_t.device
    n_s, n_t, C, R, k = w['n_s'], w['n_t'], w['dim'], w['rnd_dim'], w['k']
    n_g = n_s                         # source nodes of the graph (psi_2 runs on all of them on every rank)
    n_s = w.get('cand_rows', n_s)     # rows of S this rank owns (row sharding): K1-K4 and K6 work on these
    k_eff = k + min(k, n_t - k)
    L, T = w['layers'], w['num_steps']
    l2 = torch.cuda.get_device_properties(dev).L2_cache_size
    g = torch.Generator(device='cpu').manual_seed(3)

    def rnd(*shape):
        return torch.randn(*shape, generator=g, dtype=torch.float32).to(dev)

    h_s, h_t = rnd(1, n_s, C), rnd(1, n_t, C)
    idx = torch.randint(n_t, (1, n_s, k_eff), generator=g).to(dev)
    cand = ops.Candidates(idx, n_t)
    cand.csr
    nnz = n_s * k_eff
    S = torch.rand(n_s, k_eff, generator=g).to(dev)
    r_s, p_s, p_t = rnd(n_s, R), rnd(n_s, R), rnd(n_t, R)
    w2, b2 = rnd(R), torch.zeros(1, device=dev)
    # K5 as the step runs it: psi_2 aggregates on the MERGED pair (source + target graph), once per
    # layer in the forward and once (the transposed operator, same kernel) in the backward pass
    if e_s is not None:
        e_m = torch.cat([e_s, e_t + n_g], dim=1)
        n_m, k5_name, k5_calls = n_g + n_t, 'merged pair', 2 * L * T
    else:
        e_m, n_m, k5_name, k5_calls = e_t, n_t, 'target graph', L * T
    graph_m = ops.GraphCSR.of(e_m, n_m)
    graph_m.rel
    x_m = rnd(n_m, R)
    E = e_m.size(1)
    cat, wl, bl = rnd(n_m, 3 * R), rnd(R, 3 * R) * 0.05, rnd(R)
    gy = rnd(n_m, R)

    def pick(table_bytes, b_min, b_gathered):
        return (b_gathered, 'gathered') if table_bytes > l2 else (b_min, 'min')

    cases = []

    def add(key, name, fn, calls, work, bound, basis, kernels, graphed=False, survey_logical=None, subtract=None):
        cases.append(dict(key=key, name=name, fn=fn, calls=calls, work=work, bound=bound, basis=basis,
                          kernels=kernels, graphed=graphed, survey_logical=survey_logical, subtract=subtract))

    if with_k1:
        add('K1', 'K1 top-k (cast + tcgen05 GEMM/select + fp32 merge)', lambda: model.__top_k__(h_s, h_t), 1,
            2.0 * n_s * n_t * C, 'tensor', '2*N_s*N_t*C flop', ['cast_bf16_kernel', 'topk_tc_kernel', 'topk_merge_kernel'])
    wk, b = pick(n_t * 4 * C, (n_s + n_t) * 4 * C + 8 * nnz, 4 * C * n_s * (1 + k_eff) + 8 * nnz)
    add('K2', 'K2 gather_dot', lambda: ops.gather_dot(h_s[0], h_t[0], cand), 1, wk, 'hbm', b, ['gather_dot_kernel'],
        survey_logical=4 * C * n_s * (1 + k_eff) + 8 * nnz)
    add('K3', 'K3 softmax_k fwd', lambda: ops.softmax_k(S), T + 2, 8 * nnz, 'hbm', 'min', ['softmax_k_fwd_kernel'])
    # K4 is a by-target GATHER: per entry it reads (perm, seg, S) and one r_s row, and writes each r_t row
    # once.  SURVEY.md §8(d)'s `logical` formula models a scatter with read-modify-write atomics
    # (4R N_s (1 + 2 k_eff)), i.e. twice the row traffic this kernel has; it is reported beside the rate.
    wk, b = pick(n_s * 4 * R, 8 * nnz + 4 * R * (n_s + n_t), 12 * nnz + 4 * R * nnz + 4 * R * n_t)
    add('K4', 'K4 transport (balanced segmented gather by target)', lambda: ops.transport(S, r_s, cand), T, wk,
        'hbm', b, ['balanced_lean_kernel', 'balanced_fixup_kernel', 'balanced_main_kernel'],
        survey_logical=8 * nnz + 4 * R * n_s * (1 + 2 * k_eff))
    wk, b = pick(n_m * 4 * R, 8 * n_m * R + 8 * E + 8 * n_m, 12 * n_m * R + (n_m + 2 * E) * (4 * R + 12))
    add('K5', 'K5 rel_aggregate ([3N,N] sparse product, %s)' % k5_name, lambda: ops.rel_aggregate(x_m, graph_m),
        k5_calls, wk, 'hbm', b, ['balanced_lean_kernel', 'balanced_fixup_kernel', 'balanced_main_kernel'],
        survey_logical=8 * n_m * R + 2 * E * (4 * R + 4))
    wk, b = pick(n_t * 4 * R, 4 * R * (n_s + n_t) + 12 * nnz, 4 * R * n_s * (1 + k_eff) + 12 * nnz)
    add('K6', 'K6 diff_mlp fwd', lambda: ops.diff_mlp(S, p_s, p_t, w2, b2, cand), T, wk, 'hbm', b,
        ['diff_mlp_fwd_kernel'], survey_logical=4 * R * n_s * (1 + k_eff) + 12 * nnz)
    # psi_2 RelConv GEMM on the merged pair: [N_s + N_t, 3R] -> R, bias + ReLU (tcgen05 3xTF32).
    # Bound by HBM, not by the tensor pipe: 3 x 2*N*3R*R flop is 2.9 GFLOP (2 us at peak).
    add('G1', 'G1 linear_tc (psi_2 RelConv GEMM, %s)' % k5_name, lambda: ops.node_linear(cat, wl, bl, relu=True),
        (L + 1) * T, 4 * n_m * (3 * R + R), 'hbm', 'min', ['linear_tc_kernel'], graphed=True)
    # weight gradient dW [R, 3R] = G^T X over N nodes: split-K tcgen05 kernel + reduction of the partial
    # tiles, called as the backward pass calls them (no autograd thread in between)
    add('G2', 'G2 wgrad_tc + reduce (psi_2 RelConv weight gradient)', lambda: ops._wgrad_tc_chunked(gy, cat),
        (L + 1) * T, 4 * n_m * (3 * R + R), 'hbm', 'min', ['wgrad_tc_kernel', 'wgrad_reduce_kernel'], graphed=True)
    return cases


def kernel_rooflines(w, model, e_t, peaks, step_ms, iters=5, e_s=None, traffic_tag=None, with_k1=True):
    """Times each hot kernel alone (cold L2: a 256 MB buffer is rewritten before every launch)."""
    dev = e_t.device
    cases = hot_kernel_cases(w, model, e_t, e_s, with_k1)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    hbm, tf = peaks['hbm_gbs'], peaks['bf16_tflops']
    traffic, traffic_file = load_ncu_traffic(traffic_tag) if traffic_tag else ({}, None)
    cursor = {}   # kernel name -> launches of the capture already attributed (the capture is in case order)
    rows = []
    for c in cases:
        with torch.no_grad():
            ms = timed_graph(c['fn'], iters, flush) if c['graphed'] else timed(c['fn'], 1 if c['key'] == 'K1' and iters < 3 else iters, flush)
            if c['subtract'] is not None:
                ms = max(ms - timed_graph(c['subtract'], iters, flush), 1e-4)
        t_bytes, found = 0.0, False
        for kn in c['kernels']:
            if kn in traffic:
                i = cursor.get(kn, 0)
                n_launch = 2 if kn == 'cast_bf16_kernel' else 1
                vals = traffic[kn][i:i + n_launch]
                if vals:
                    t_bytes += sum(vals)
                    found = True
                    cursor[kn] = i + n_launch
        achieved = c['work'] / (ms * 1e-3) / (1e12 if c['bound'] == 'tensor' else 1e9)
        peak = tf if c['bound'] == 'tensor' else hbm
        rows.append({'kernel': c['name'], 'bound': c['bound'], 'achieved': round(achieved, 2), 'peak': peak,
                     'unit': 'TFLOP/s' if c['bound'] == 'tensor' else 'GB/s', 'frac': round(achieved / peak, 4),
                     'traffic': t_bytes if found else None, 'traffic_source': traffic_file if found else None,
                     'work_basis': c['basis'], 'work_per_launch': c['work'],
                     'survey_logical_gbs': round(c['survey_logical'] / (ms * 1e-3) / 1e9, 1) if c['survey_logical'] else None,
                     'ms_per_launch': round(ms, 4), 'launches_per_step': c['calls'],
                     'share_of_step': round(ms * c['calls'] / step_ms, 4) if step_ms else None})
    return rows


def run_hot_kernels(w, dev, tag, with_k1):
    """`--hot-kernels`: every hot kernel once (after one warm-up launch each) inside the NVTX range
    "hot", for `ncu --set full --nvtx --nvtx-include "hot/"`; the raw-page CSV export of that capture,
    committed as profiles/<round>_ncu_full_hot_<tag>.csv, is what `roofline*.traffic` is read from."""
    from dgmc.models import DGMC, RelCNN
    g = torch.Generator().manual_seed(11)
    e_s = torch.randint(w['n_s'], (2, w['e_s']), generator=g).to(dev)
    e_t = torch.randint(w['n_t'], (2, w['e_t']), generator=g).to(dev)
    model = DGMC(RelCNN(8, 8, 1), RelCNN(w['rnd_dim'], w['rnd_dim'], 1), num_steps=w['num_steps'], k=w['k']).to(dev)
    cases = hot_kernel_cases(w, model, e_t, e_s, with_k1)
    with torch.no_grad():
        for c in cases:
            c['fn']()
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_push('hot')
    with torch.no_grad():
        for c in cases:
            c['fn']()
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_pop()
    emit({'hot_kernels': [c['key'] for c in cases], 'tag': tag})


def pascal_step_bench(dev, steps, warmup):
    """`--pascal`: BASELINE.json configs[1] — PascalVOC-keypoint-shaped synthetic batch (512 graph
    pairs of 10-20 nodes, 1024-d features, 2-d pseudo-coordinates, SplineCNN psi_1 1024->256 L2 /
    psi_2 128->128 L2, dense S, 10 consensus steps, examples/pascal.py:46-76).  A parity-test
    configuration, timed here for the record (not the headline metric)."""
    from dgmc.models import DGMC, SplineCNN
    g = torch.Generator().manual_seed(2)
    B = 512

    def batch(sizes):
        xs, es, ats, bs, off = [], [], [], [], 0
        for i, n in enumerate(sizes):
            xs.append(torch.randn(n, 1024, generator=g))
            es.append(torch.randint(n, (2, 5 * n), generator=g) + off)
            ats.append(torch.rand(5 * n, 2, generator=g))
            bs.append(torch.full((n, ), i, dtype=torch.long))
            off += n
        return [t.to(dev) for t in (torch.cat(xs), torch.cat(es, 1), torch.cat(ats), torch.cat(bs))]

    from dgmc.models.dgmc import BatchLayout
    from dgmc.utils import GraphedTrainStep
    sizes = torch.randint(10, 21, (B, ), generator=g).tolist()
    x_s, e_s, a_s, b_s = batch(sizes)
    x_t, e_t, a_t, b_t = batch(sizes)
    # what dgmc.utils.collate_pairs attaches on the host: (number of graphs, largest graph)
    BatchLayout.hint(b_s, B, max(sizes))
    BatchLayout.hint(b_t, B, max(sizes))
    y = torch.stack([torch.arange(x_s.size(0)), torch.cat([torch.randperm(n, generator=g) for n in sizes])]).to(dev)
    torch.manual_seed(0)
    model = DGMC(SplineCNN(1024, 256, 2, 2, cat=False, dropout=0.5), SplineCNN(128, 128, 2, 2, cat=True),
                 num_steps=10).to(dev).train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)

    def step():
        opt.zero_grad(set_to_none=True)
        S_0, S_L = model(x_s, e_s, a_s, b_s, x_t, e_t, a_t, b_t)
        loss = model.loss(S_0, y) + model.loss(S_L, y)
        loss.backward()
        opt.step()
        return loss

    def timed_steps(fn):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            out = fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / steps, float(out)   # (a float: keeps no autograd graph alive)

    ms, loss = timed_steps(step)
    if os.environ.get('DGMC_BENCH_KINETO'):
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(3):
                step()
            torch.cuda.synchronize()
        rows = sorted(((e.key, e.count, e.device_time_total) for e in prof.key_averages()
                       if e.device_time_total > 0), key=lambda r: -r[2])
        print('kernel time per step: %.3f ms' % (sum(r[2] for r in rows) / 3e3), file=sys.stderr)
        for name, cnt, t in rows[:25]:
            print('%8.1f us/step %5d x  %s' % (t / 3, cnt // 3, name[:110]), file=sys.stderr)
    # the same step captured once and replayed (a fixed padded layout: batches are bucketed by shape)
    graphed = GraphedTrainStep(model, opt, x_s, e_s, a_s, b_s, x_t, e_t, a_t, b_t, y,
                               loss_fn=lambda S_0, S_L, yy: model.loss(S_0, yy) + model.loss(S_L, yy),
                               structure_is_static=True)
    ms_graph, loss_g = timed_steps(lambda: graphed(x_s, e_s, a_s, b_s, x_t, e_t, a_t, b_t, y))
    result = ({'pascal_shaped': {'graph_pairs': B, 'nodes': int(x_s.size(0)), 'edges': int(e_s.size(1)),
                            'ms_per_step': ms_graph, 'steps_per_s': 1e3 / ms_graph, 'pairs_per_s': B * 1e3 / ms_graph,
                            'eager_ms_per_step': ms, 'eager_pairs_per_s': B * 1e3 / ms,
                            'loss': loss_g, 'eager_loss': loss,
                            'execution': 'whole step replayed as one CUDA graph at a fixed padded layout (batches '
                                         'bucketed by shape); eager_*: the same step issued from Python, which the '
                                         'reference scripts do with a new batch shape every step'}})
    del graphed, model, opt
    sys.modules['dgmc.ops'].clear_caches()
    torch.cuda.empty_cache()
    return result['pascal_shaped']


def scale_workload(n):
    """BASELINE.json configs[4] shape: an N x N-node pair, 256-d embeddings, R = 64, k = 10 (k_eff = 20),
    5 out-edges per node — where the gathered tables no longer fit in L2 and HBM is the roofline."""
    return dict(n_s=n, n_t=n, dim=256, rnd_dim=64, k=10, layers=3, num_steps=10, e_s=5 * n, e_t=5 * n)


def scale_rooflines(n, dev, iters=3):
    from dgmc.models import DGMC, RelCNN
    w = scale_workload(n)
    g = torch.Generator().manual_seed(11)
    e_s = torch.randint(n, (2, 5 * n), generator=g).to(dev)
    e_t = torch.randint(n, (2, 5 * n), generator=g).to(dev)
    model = DGMC(RelCNN(8, 8, 1), RelCNN(64, 64, 1), num_steps=10, k=10).to(dev)
    tag = '1m' if n == 1000000 else None
    return kernel_rooflines(w, model, e_t, load_peaks(), None, iters=iters, e_s=e_s, traffic_tag=tag)


def scale_kernels(n, dev):
    """`--scale-kernels N`: the hot kernels alone on an N x N-node synthetic pair."""
    emit({'scale_kernels': {'n_s': n, 'n_t': n, 'C': 256, 'R': 64, 'k_eff': 20, 'edges': 5 * n},
          'roofline_all': scale_rooflines(n, dev)})


def sharded_parity(model, w, d, dev, rank, world):
    """Multi-GPU parity inside the run the driver sees: one un-graphed training-mode forward of the
    row-sharded model on all ranks and the same forward of the single-GPU model (same weights, same
    seed -> same dropout masks, negatives and indicators) on rank 0; the ranks' rows of S_L are gathered
    and compared.  Candidate lists must be identical; values agree up to the summation order of the
    r_t all-reduce (tolerance 1e-4 relative at this shape, see tests/test_gpu_bench_configs.py)."""
    import torch.distributed as dist
    from dgmc.models import DGMC
    from dgmc.models.sharded import row_block
    seed = 4321
    with torch.no_grad():
        torch.manual_seed(seed)
        _, S_sh = model(d['x_s'], d['e_s'], None, None, d['x_t'], d['e_t'], None, None, d['y'])
        share = model.loss(S_sh, d['y'])
        loss_sh = model.global_loss(share).item()
    n_s, k_eff = w['n_s'], S_sh.__idx__.size(1)
    rows_max = max(row_block(n_s, q, world)[1] - row_block(n_s, q, world)[0] for q in range(world))
    pad_v = torch.zeros(rows_max, k_eff, device=dev)
    pad_i = torch.full((rows_max, k_eff), -1, dtype=torch.long, device=dev)
    pad_v[:S_sh.__val__.size(0)] = S_sh.__val__
    pad_i[:S_sh.__idx__.size(0)] = S_sh.__idx__
    all_v = [torch.empty_like(pad_v) for _ in range(world)]
    all_i = [torch.empty_like(pad_i) for _ in range(world)]
    dist.all_gather(all_v, pad_v)
    dist.all_gather(all_i, pad_i)
    result = None
    if rank == 0:
        val = torch.cat([all_v[q][:row_block(n_s, q, world)[1] - row_block(n_s, q, world)[0]] for q in range(world)])
        idx = torch.cat([all_i[q][:row_block(n_s, q, world)[1] - row_block(n_s, q, world)[0]] for q in range(world)])
        single = DGMC(model.psi_1, model.psi_2, num_steps=model.num_steps, k=model.k, detach=model.detach)
        single.mlp = model.mlp
        single.train(model.training)
        with torch.no_grad():
            torch.manual_seed(seed)
            _, S_1 = single(d['x_s'], d['e_s'], None, None, d['x_t'], d['e_t'], None, None, d['y'])
            loss_1 = single.loss(S_1, d['y']).item()
        ref = S_1.__val__
        # Candidate lists: identical except where two scores tie within fp32 rounding — the sharded psi_1
        # sums its neighbourhoods in a different block order than the single-GPU kernel, so embeddings
        # differ in the last bit and a row whose k-th / (k+1)-th scores are that close may pick the other
        # one (BASELINE.json: index sets bit-exact where the gap exceeds 1e-6 relative).  Values are
        # compared on the rows whose lists agree.
        same = (idx == S_1.__idx__).all(dim=1)
        differing = int((~same).sum())
        rel_all = ((val - ref).abs() / ref.abs().clamp_min(1e-6))[same]
        rel = rel_all.max().item() if rel_all.numel() else 0.0
        q999 = torch.quantile(rel_all.flatten()[:8000000].float(), 0.999).item() if rel_all.numel() else 0.0
        # every differing row is CHECKED to be such a tie: the columns that only one of the two lists holds
        # must score within 2e-5 of each other (relative to the row's largest score), with the scores
        # recomputed here from the single-GPU embeddings (same seed -> same dropout masks)
        ties_verified, worst_gap = True, 0.0
        if differing:
            bad = (~same).nonzero().view(-1)[:256]
            with torch.no_grad():
                torch.manual_seed(seed)
                h_s = single.psi_1(d['x_s'], d['e_s'], None)
                h_t = single.psi_1(d['x_t'], d['e_t'], None)
                sc = h_s[bad] @ h_t.t()
            # (a row differs when a near-tie swaps two neighbours of its descending list, or swaps the
            # k-th best for the (k+1)-th: either way the two columns at a differing POSITION must tie)
            scale = sc.abs().max(dim=1, keepdim=True).values
            a, b = S_1.__idx__[bad], idx[bad]
            gaps = (sc.gather(1, a) - sc.gather(1, b)).abs() / scale
            worst_gap = gaps[a != b].max().item()
            ties_verified = worst_gap <= 2e-5
        allowed = max(8, int(2e-4 * n_s))
        result = {'checked': True, 'idx_equal': differing == 0, 'rows_with_other_candidates': differing,
                  'rows_allowed_in_tie_band': allowed, 'ties_verified': ties_verified,
                  'worst_tie_gap_rel': worst_gap, 'max_rel_err': rel, 'q999_rel_err': q999,
                  'loss_rel_err': abs(loss_sh - loss_1) / max(abs(loss_1), 1e-12), 'tolerance': 1e-4,
                  'what': 'S_L rows of all ranks (training-mode forward, shared seed) against the '
                          'single-GPU model on rank 0'}
        result['ok'] = differing <= allowed and ties_verified and q999 <= 1e-4 and \
            result['loss_rel_err'] <= 1e-4 and (differing > 0 or rel <= 1e-4)
    dist.barrier()
    return result


def scale_study(n, dev, rank, world, steps=3, warmup=2):
    """BASELINE.json configs[4] in every bench line: one whole training step (forward + loss + backward +
    Adam, issued eagerly: at this size the kernels are milliseconds long and a captured graph of the step
    would not fit beside anything else) on an n x n-node pair — the same model, 5 edges per node, n / 4
    training pairs — on the N GPUs of this run, max over ranks; at N > 1 with the parity check of the
    headline workload.  The driver's scaling table divides these numbers across its N = 1, 2, 4, 8 runs."""
    import torch.distributed as dist
    w = dict(WORKLOAD, name='synthetic %d x %d-node pair' % (n, n), n_s=n, n_t=n, e_s=5 * n, e_t=5 * n,
             train_pairs=n // 4)
    host = make_inputs(w, seed=5)
    d = {k: v.to(dev) for k, v in host.items()}
    del host
    model = build_model(w, dev, sharded=world > 1)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)
    torch.manual_seed(99)
    for _ in range(warmup):
        train_step(model, opt, d)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        loss = train_step(model, opt, d)
    b.record()
    torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / steps], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    out = {'workload': '%d x %d nodes, 300-d, RelCNN(300,256,L3)/RelCNN(64,64,L3), sparse k=10 (+10), 10 consensus '
                       'steps, %d train pairs, 5 edges per node; eager step' % (n, n, n // 4),
           'n_gpus': world, 'ms_per_step': ms.item(), 'steps_per_s': 1e3 / ms.item(), 'steps': steps,
           'warmup': warmup, 'peak_memory_gb': round(torch.cuda.max_memory_allocated(dev) / 2 ** 30, 1)}
    if world > 1:
        out['parity'] = sharded_parity(model, w, d, dev, rank, world)
    del model, opt, d, loss
    sys.modules['dgmc.ops'].clear_caches()
    torch.cuda.empty_cache()
    return out


def stage(msg):
    if os.environ.get('DGMC_BENCH_VERBOSE'):
        print('[bench rank %s] %s' % (os.environ.get('RANK', '0'), msg), file=sys.stderr, flush=True)


def load_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        p = json.load(open(path))
        return {'hbm_gbs': p['hbm_gbs'], 'bf16_tflops': p['bf16_tflops'],
                'bf16_tflops_sustained': p.get('bf16_tflops_sustained'), 'source': 'measured'}
    return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0,
            'source': 'fallback'}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--cpu-budget', type=float, default=None, help='seconds of CPU baseline work')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--k', type=int, default=0, help='override the sparsity k (BASELINE.json configs[3] uses 25)')
    ap.add_argument('--nodes', type=int, default=0, metavar='N',
                    help='scaling study: replace the DBP15K shape by an N x N-node pair (5 edges per '
                         'node, N/4 train pairs); not the headline configuration')
    ap.add_argument('--pascal', action='store_true', help='time the PascalVOC-shaped dense config and exit')
    ap.add_argument('--scale-kernels', type=int, default=0, metavar='N',
                    help='time the hot kernels alone on an N x N-node pair and exit')
    ap.add_argument('--hot-kernels', action='store_true',
                    help='launch every hot kernel once inside the NVTX range "hot" and exit (for ncu --set full '
                         '--nvtx --nvtx-include "hot/"); with --scale-kernels N on the N x N pair')
    ap.add_argument('--no-pascal', action='store_true', help='skip the PascalVOC-shaped block of the N=1 line')
    ap.add_argument('--no-scale-study', action='store_true', help='skip the large-pair step of the bench line')
    ap.add_argument('--scale-study-nodes', type=int, default=1000000, metavar='N')
    ap.add_argument('--no-roofline-1m', action='store_true', help='skip the 1M x 1M kernel sweep of the N=1 line')
    ap.add_argument('--profile-eager', action='store_true',
                    help='warm up, run --steps eager steps inside the NVTX range "timed_step" and exit '
                         '(for ncu --nvtx --nvtx-include "timed_step/")')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'b200' else args.warmup

    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))
    w = dict(WORKLOAD)
    if args.k:
        w['k'] = args.k
    if args.nodes:
        w.update(name='synthetic %d x %d-node pair (scaling study)' % (args.nodes, args.nodes),
                 n_s=args.nodes, n_t=args.nodes, e_s=5 * args.nodes, e_t=5 * args.nodes,
                 train_pairs=args.nodes // 4)
    config = {'workload': '%s: %d x %d nodes, %d-d, RelCNN(%d,%d,L%d)/RelCNN(%d,%d,L%d), sparse k=%d '
              '(+%d negatives), %d consensus steps, detach=True, %d train pairs; step = fwd + '
              'loss + bwd + Adam' % (w['name'], w['n_s'], w['n_t'], w['feat'], w['feat'], w['dim'],
                                     w['layers'], w['rnd_dim'], w['rnd_dim'], w['layers'], w['k'],
                                     w['k'], w['num_steps'], w['train_pairs']),
              'baseline_config': 'BASELINE.json configs[2]',
              # (the same dictionary in both arms; how each arm executes the workload is in `execution`)
              'l2': 'working set per step (inputs + saved activations, ~1 GB) exceeds the 126 MB L2; '
                    'per-kernel roofline timings flush L2 with a 256 MB write'}

    if args.impl == 'reference':
        if rank != 0:
            return 0
        inputs = make_inputs(w)
        budget = args.cpu_budget or 150.0
        sps, n, cores, dt, wu, kind, what = cpu_reference_steps(w, inputs, args.steps, args.warmup, budget)
        sample = '%d full steps of the same workload: %s; %d threads' % (n, what, cores)
        emit({
            'impl': 'reference', 'metric': METRIC, 'value': sps, 'unit': 'steps/s', 'n_gpus': 0,
            'steps': n, 'warmup': wu, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
            'config': config,
            'execution': {'parallelism': '%d host threads' % cores, 'step': what},
            'cpu_baseline': {'value': sps, 'unit': 'steps/s', 'cores': cores, 'kind': kind,
                             'sample': sample},
            'e2e': {'value': sps, 'unit': 'steps/s', 'h2d_bytes_per_step': 0,
                    'd2h_bytes_per_step': 0}})
        return 0

    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if args.hot_kernels:
        if args.scale_kernels:
            run_hot_kernels(scale_workload(args.scale_kernels), dev, '%d' % args.scale_kernels, with_k1=False)
        else:
            run_hot_kernels(w, dev, 'dbp15k', with_k1=True)
        return 0
    if args.scale_kernels:
        scale_kernels(args.scale_kernels, dev)
        return 0
    if args.pascal:
        emit({'pascal_shaped': pascal_step_bench(dev, args.steps, args.warmup)})
        return 0
    if world > 1:
        os.environ.setdefault('TORCH_NCCL_ASYNC_ERROR_HANDLING', '0')  # collectives get graph-captured
        dist.init_process_group('nccl', device_id=dev)
    from dgmc import native

    host = make_inputs(w)
    pinned = {k: v.pin_memory() for k, v in host.items()}
    d = {k: v.to(dev) for k, v in host.items()}
    model = build_model(w, dev, sharded=world > 1)
    torch.manual_seed(1234)  # every rank draws the same indicator / negative / dropout streams
    opt = torch.optim.Adam([p for p in model.parameters()], lr=1e-3, capturable=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if args.profile_eager:
        # backward on the calling thread: NVTX ranges (which ncu filters on) are per thread
        torch.autograd.set_multithreading_enabled(False)
        for _ in range(args.warmup):
            train_step(model, opt, d)
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_push('timed_step')
        for _ in range(args.steps):
            train_step(model, opt, d)
        torch.cuda.synchronize()
        torch.cuda.nvtx.range_pop()
        emit({'profile_eager_steps': args.steps})
        return 0

    parity = sharded_parity(model, w, d, dev, rank, world) if world > 1 else None
    torch.manual_seed(1234)

    from dgmc.utils import GraphedTrainStep
    args_of = lambda t: (t['x_s'], t['e_s'], None, None, t['x_t'], t['e_t'], None, None, t['y'])  # noqa: E731

    stage('eager warm-up')
    # ---- eager step (plain Python loop over the public API), for reference ------------------
    for _ in range(args.warmup):
        train_step(model, opt, d)
    barrier()
    launches0 = native.lib.dgmc_launch_count()
    n_eager = max(2, min(5, args.steps))
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(n_eager):
        train_step(model, opt, d)
    end.record()
    barrier()
    eager_ms = start.elapsed_time(end) / n_eager
    launches = (native.lib.dgmc_launch_count() - launches0) // n_eager

    # ---- resident inputs: the whole step captured in one CUDA graph -------------------------
    stage('capturing graph (static structure)')
    graphed = GraphedTrainStep(model, opt, *args_of(d), structure_is_static=True)
    stage('graph captured')
    for _ in range(args.warmup):
        graphed(*args_of(d))
    barrier()
    with ClockSampler(local) as clocks:
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        for _ in range(args.steps):
            graphed(*args_of(d))
        end.record()
        barrier()
    ms = torch.tensor([start.elapsed_time(end)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = ms.item() / args.steps
    value = 1e3 / ms_per_step  # one job: with N > 1 the source rows are sharded over the ranks

    if os.environ.get('DGMC_BENCH_KINETO'):
        # per-kernel device time of graph replays (warm caches, real overlap), via CUPTI
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            for _ in range(5):
                graphed(*args_of(d))
            torch.cuda.synchronize()
        rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0]
        rows.sort(key=lambda r: -r[2])
        tot = sum(r[2] for r in rows)
        print('kernel time per replay: %.3f ms over %d kernel names' % (tot / 5e3, len(rows)), file=sys.stderr)
        for name, cnt, t in rows[:40]:
            print('%8.1f us/replay %5d x  %s' % (t / 5, cnt // 5, name[:110]), file=sys.stderr)

    # ---- end-to-end: pinned host inputs copied every step (graph structure rebuilt inside the
    #      captured step), loss read back --------------------------------------------------------
    h2d = sum(v.numel() * v.element_size() for v in pinned.values())
    stage('capturing e2e graph')
    # (scaling studies on large pairs reuse the captured step: a second graph of a 1M x 1M step does not
    # fit beside the first; their edge lists are then fixed, features and ground truth copied per step)
    e2e_step = graphed if args.nodes else GraphedTrainStep(model, opt, *args_of(d), structure_is_static=False)
    stage('e2e graph captured')
    # Input pipelining: the host-to-device copy of step i+1 is issued (side stream, pinned memory)
    # right after step i is launched and before its loss is read, so it overlaps step i's kernels;
    # every step still copies its 49.5 MB from the host and reads its loss back.
    # Loss read-back is pipelined the same way: the 4-byte result of step i is copied to pinned host
    # memory behind the step (async), and the host reads it after step i+1 has been queued, so the
    # GPU never waits for the host between steps.  Every step's loss is read inside the timed region.
    host_loss = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    copied = [torch.cuda.Event(), torch.cuda.Event()]

    def run_e2e(n_steps):
        value = None
        for i in range(n_steps):
            loss_t = e2e_step(*args_of(pinned))
            host_loss[i % 2].copy_(loss_t.detach().view(1), non_blocking=True)
            copied[i % 2].record()
            e2e_step.prefetch(*args_of(pinned))
            if i > 0:
                copied[(i - 1) % 2].synchronize()
                value = host_loss[(i - 1) % 2].item()
        copied[(n_steps - 1) % 2].synchronize()
        return host_loss[(n_steps - 1) % 2].item() if n_steps else value

    run_e2e(3)
    barrier()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    loss_value = run_e2e(args.steps)
    end.record()
    barrier()
    ms2 = torch.tensor([start.elapsed_time(end)], device=dev)
    if world > 1:
        dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_value = 1e3 / (ms2.item() / args.steps)

    stage('timing done')
    study = None
    if not args.nodes and not args.k and not args.no_scale_study:
        del graphed, e2e_step
        sys.modules['dgmc.ops'].clear_caches()
        torch.cuda.empty_cache()
        study = scale_study(args.scale_study_nodes, dev, rank, world)
        stage('scale study done')
    if world > 1:  # all collective work is over: leave the group before rank 0's local epilogue
        dist.barrier()
        torch.cuda.synchronize()
        stage('last barrier passed')
        # Tearing the NCCL communicator down while captured graphs still reference it can block, and
        # a rank that vanishes early makes the others abort: the helper ranks wait (on the TCP store,
        # not on NCCL) until rank 0 has printed, then everybody leaves without a teardown.
        store = dist.distributed_c10d._get_default_store()
        if rank != 0:
            import datetime
            try:
                store.wait(['dgmc_bench_done'], datetime.timedelta(seconds=900))
            except Exception:
                pass
            sys.stderr.flush()
            os._exit(0)
    out = None
    if rank == 0:
        peaks = load_peaks()
        w_rl = dict(w)
        if world > 1:   # rank 0's share of the rows of S; psi_2 (K5, G1, G2) is replicated on the whole pair
            from dgmc.models.sharded import row_block
            lo, hi = row_block(w['n_s'], 0, world)
            w_rl['cand_rows'] = hi - lo
        if args.nodes and args.nodes > 200000:
            rl, roofline = [], None     # scaling study: whole-step numbers only (see --scale-kernels for the kernels)
        else:
            rl = kernel_rooflines(w_rl, model, d['e_t'], peaks, ms_per_step, e_s=d['e_s'],
                                  traffic_tag='dbp15k' if (not args.nodes and world == 1 and not args.k) else None)
            top = max(rl, key=lambda r: r['share_of_step'])
            roofline = {k_: top[k_] for k_ in ('bound', 'achieved', 'peak', 'unit', 'frac', 'traffic', 'traffic_source')}
            roofline.update(kernel=top['kernel'], peak_source=peaks['source'] + ' (burst: kernel timed alone)')
        out = {
            'metric': METRIC, 'value': value, 'unit': 'steps/s', 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': ms_per_step, 'higher_is_better': True,
            # one job = one graph pair whatever N is: its source rows are divided over the GPUs
            'scaling': 'strong', 'vs_baseline': None, 'dtype': 'f32',
            'data': 'synthetic',
            'config': config,
            'execution': dict(parallelism=('source rows of S sharded over %d GPUs; psi_1 node-sharded; psi_2 %s; '
                                                'collectives: the library\'s peer-memory kernels over NVLink '
                                                '(r_t all-reduce per consensus step, layer all-gathers)' %
                                                (world, 'node-sharded' if (w['n_s'] + w['n_t'] >= 150000 or
                                                                          os.environ.get('DGMC_SHARD_PSI') == '1')
                                                 else 'replicated')) if world > 1 else 'single GPU',
                              step='whole step (zero_grad+fwd+loss+bwd+Adam) replayed as one CUDA '
                                   'graph (dgmc.utils.GraphedTrainStep); eager_ms_per_step is the '
                                   'same step issued from Python'),
            'clocks': clocks.summary(),
            'e2e': {'value': e2e_value, 'unit': 'steps/s', 'h2d_bytes_per_step': h2d,
                    'd2h_bytes_per_step': 4},
            'gpu_launches': int(launches), 'final_loss': loss_value,
            'eager_ms_per_step': eager_ms,
            'roofline': roofline, 'roofline_all': rl,
        }
        if parity is not None:
            out['parity'] = parity
        if study is not None:
            out['scale_study'] = study
        if world == 1 and not args.nodes and not args.k and not args.no_pascal:
            # BASELINE.json configs[1] (dense variant, SplineCNN, 512 pairs of 10-20-node graphs): for the record
            out['pascal_shaped'] = pascal_step_bench(dev, 10, 3)
        if world == 1 and not args.nodes and not args.k and not args.no_roofline_1m:
            # BASELINE.json configs[4] shape, kernels alone: where HBM, not latency, is the roofline
            del model
            ops_mod = sys.modules['dgmc.ops']
            ops_mod.clear_caches()
            torch.cuda.empty_cache()
            out['roofline_1m'] = {'shape': '1,000,000 x 1,000,000 nodes, C=256, R=64, k_eff=20, 5 M edges per graph; '
                                           'kernels timed alone, cold L2',
                                  'kernels': scale_rooflines(1000000, dev, iters=2)}
        if world == 1 and not args.no_cpu_baseline and not args.nodes and not args.k:
            sps, n, cores, dt, _, kind, what = cpu_reference_steps(w, host, 2, 1, args.cpu_budget or 30.0)
            out['cpu_baseline'] = {
                'value': sps, 'unit': 'steps/s', 'cores': cores, 'kind': kind,
                'sample': '%d full step(s) of the same workload after 1 warm-up: %s; %.1f s/step' % (n, what, dt)}
        emit(out)
    if world > 1 and rank == 0 and study is not None and not study['parity']['ok']:
        parity = study['parity']
    if world > 1 and rank == 0 and parity is not None and not parity['ok']:
        print('bench: multi-GPU parity check FAILED: %s' % parity, file=sys.stderr, flush=True)
        store.set('dgmc_bench_done', '1')
        time.sleep(1.0)
        os._exit(3)
    if world > 1:
        store.set('dgmc_bench_done', '1')
        sys.stderr.flush()
        time.sleep(1.0)
        os._exit(0)
    return 0


if __name__ == '__main__':
    sys.exit(main())
