"""Golden-vector generator — TEST INFRASTRUCTURE, run in the BUILD CONTAINER only.

    python oracle/make_golden.py            # rewrites tests/golden/*.pt

Imports the UNMODIFIED reference package from /root/reference (a Python reference cannot
travel to the GPU box) through the third-party stand-ins in oracle/shims/, runs
`dgmc.models.DGMC` (dgmc.py:114-311) on small seeded inputs in fp32 and in fp64, and stores
inputs, parameters (reference state_dict names), the random draws the reference made
(`torch.randn` at dgmc.py:169/206, `torch.randint` at dgmc.py:192), outputs, losses and
parameter gradients.  tests/test_oracle_golden.py pins oracle/dgmc_oracle.py on these files;
the `-m gpu` parity tests check the CUDA path against them as well.
"""
import contextlib
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get('DGMC_REFERENCE', '/root/reference')
sys.path.insert(0, os.path.join(HERE, 'shims'))
sys.path.insert(0, REF)

import dgmc.models as ref_models  # noqa: E402  (the reference, unmodified)

OUT = os.path.join(ROOT, 'tests', 'golden')


class DrawTape(object):
    """Records (first run) or replays (later runs, any dtype) torch.randn / torch.randint."""
    def __init__(self):
        self.randn, self.randint, self.mode = [], [], 'record'
        self._i = self._j = 0

    @contextlib.contextmanager
    def active(self):
        real_randn, real_randint = torch.randn, torch.randint
        self._i = self._j = 0

        def randn(size, *a, dtype=None, device=None, **kw):
            if self.mode == 'record':
                t = real_randn(size, dtype=torch.float32)
                self.randn.append(t)
            else:
                t = self.randn[self._i]
                self._i += 1
            return t.to(dtype or torch.float32)

        def randint(high, size, *a, dtype=None, device=None, **kw):
            if self.mode == 'record':
                t = real_randint(high, size, dtype=torch.long)
                self.randint.append(t)
            else:
                t = self.randint[self._j]
                self._j += 1
            return t

        torch.randn, torch.randint = randn, randint
        try:
            yield self
        finally:
            torch.randn, torch.randint = real_randn, real_randint


def build_psi(spec):
    if spec['kind'] == 'rel':
        return ref_models.RelCNN(spec['in'], spec['out'], spec['layers'],
                                 batch_norm=spec.get('batch_norm', False), cat=spec.get('cat', True),
                                 lin=spec.get('lin', True), dropout=spec.get('dropout', 0.0))
    if spec['kind'] == 'spline':
        return ref_models.SplineCNN(spec['in'], spec['out'], spec['dim'], spec['layers'],
                                    cat=spec.get('cat', True), lin=spec.get('lin', True),
                                    dropout=spec.get('dropout', 0.0))
    if spec['kind'] == 'gin':
        return ref_models.GIN(spec['in'], spec['out'], spec['layers'],
                              batch_norm=spec.get('batch_norm', False), cat=spec.get('cat', True),
                              lin=spec.get('lin', True))
    raise ValueError(spec['kind'])


def run_reference(model, inp, y, tape, dtype, training, loss_on):
    model = model.to(dtype)
    model.train(training)
    model.zero_grad()

    def cast(t):
        return t.to(dtype) if (t is not None and t.is_floating_point()) else t

    args = [cast(inp[n]) for n in ('x_s', 'edge_index_s', 'edge_attr_s', 'batch_s', 'x_t',
                                   'edge_index_t', 'edge_attr_t', 'batch_t')]
    with tape.active():
        if y is not None and training and model.k >= 1:
            S_0, S_L = model(*args, y)
        else:
            S_0, S_L = model(*args)
    out = {}
    if S_0.is_sparse:
        out['idx'] = S_0.__idx__.clone()
        out['S_0'], out['S_L'] = S_0.__val__.detach().clone(), S_L.__val__.detach().clone()
        if S_0.size(0) * S_0.size(1) <= 4096:  # keep fixtures small
            out['S_0_dense'] = S_0.to_dense().detach().clone()
            out['S_L_dense'] = S_L.to_dense().detach().clone()
    else:
        out['S_0'], out['S_L'] = S_0.detach().clone(), S_L.detach().clone()
    if y is not None:
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            l0, lL = model.loss(S_0, y), model.loss(S_L, y)
            out['loss_0'], out['loss_L'] = l0.detach().clone(), lL.detach().clone()
            out['acc_L'] = torch.tensor(float(model.acc(S_L, y)))
            out['hits3_L'] = torch.tensor(float(model.hits_at_k(3, S_L, y)))
            if training:
                total = {'0': l0, 'L': lL, '0+L': l0 + lL}[loss_on]
                total.backward()
                out['grads'] = {n: q.grad.detach().clone() for n, q in model.named_parameters()
                                if q.grad is not None}
    return out


def make_case(name, spec1, spec2, inp, y, num_steps, k, training, detach=False, loss_on='L',
              seed=0):
    torch.manual_seed(seed)
    model = ref_models.DGMC(build_psi(spec1), build_psi(spec2), num_steps=num_steps, k=k,
                            detach=detach)
    params = {n: t.detach().clone() for n, t in model.state_dict().items()}
    tape = DrawTape()
    out32 = run_reference(model, inp, y, tape, torch.float32, training, loss_on)
    tape.mode = 'replay'
    model.load_state_dict(params)
    out64 = run_reference(model, inp, y, tape, torch.float64, training, loss_on)
    case = {
        'name': name, 'spec1': spec1, 'spec2': spec2, 'params': params, 'inputs': inp, 'y': y,
        'num_steps': num_steps, 'k': k, 'training': training, 'detach': detach,
        'loss_on': loss_on,
        'draws': {'r_s': tape.randn, 'neg': tape.randint[0] if tape.randint else None},
        'out32': out32, 'out64': out64,
    }
    os.makedirs(OUT, exist_ok=True)
    torch.save(case, os.path.join(OUT, name + '.pt'))
    size = os.path.getsize(os.path.join(OUT, name + '.pt'))
    print('%-28s %7.1f KB  S_L32-vs-64 max|d| = %.3g' %
          (name, size / 1024, (out32['S_L'].double() - out64['S_L']).abs().max()))


def rand_graph(gen, n, e):
    return torch.randint(n, (2, e), generator=gen)


def batched(gen, sizes, deg, feat, pseudo_dim=None):
    xs, eis, bs, attrs, off = [], [], [], [], 0
    for g, n in enumerate(sizes):
        xs.append(torch.randn(n, feat, generator=gen))
        ei = torch.randint(n, (2, n * deg), generator=gen)
        eis.append(ei + off)
        bs.append(torch.full((n, ), g, dtype=torch.long))
        if pseudo_dim:
            attrs.append(torch.rand(n * deg, pseudo_dim, generator=gen))
        off += n
    attr = torch.cat(attrs) if pseudo_dim else None
    return torch.cat(xs), torch.cat(eis, dim=1), attr, torch.cat(bs)


def main():
    gen = torch.Generator().manual_seed(20261106)

    # -- BASELINE.json configs[0]: dense, N=100 random graph pair with permuted target -----
    n = 100
    x = torch.randn(n, 32, generator=gen)
    ei = rand_graph(gen, n, 400)
    perm = torch.randperm(n, generator=gen)  # target node perm[i] <- source node i
    x_t = torch.empty_like(x)
    x_t[perm] = x
    inp = dict(x_s=x, edge_index_s=ei, edge_attr_s=None, batch_s=None, x_t=x_t,
               edge_index_t=perm[ei], edge_attr_t=None, batch_t=None)
    y = torch.stack([torch.arange(n), perm])
    rel1 = dict(kind='rel', **{'in': 32}, out=64, layers=3, batch_norm=False, cat=True, lin=True,
                dropout=0.0)
    rel2 = dict(kind='rel', **{'in': 32}, out=32, layers=3, batch_norm=False, cat=True, lin=True,
                dropout=0.0)
    make_case('cfg1_dense_rel_n100', rel1, rel2, inp, y, 10, -1, True, loss_on='0+L', seed=1)

    # -- ragged dense batch with SplineCNN (PascalVOC-like, tiny) ---------------------------
    xs, eis, eas, bs = batched(gen, [5, 7, 6], 3, 8, 2)
    xt, eit, eat, bt = batched(gen, [6, 7, 5], 3, 8, 2)
    inp = dict(x_s=xs, edge_index_s=eis, edge_attr_s=eas, batch_s=bs, x_t=xt, edge_index_t=eit,
               edge_attr_t=eat, batch_t=bt)
    y = torch.stack([torch.arange(18), torch.tensor([0, 1, 2, 3, 4, 6, 5, 4, 3, 2, 1, 0,
                                                     0, 1, 2, 3, 4, 0])])
    sp1 = dict(kind='spline', **{'in': 8}, out=16, dim=2, layers=2, cat=False, lin=True,
               dropout=0.0)
    sp2 = dict(kind='spline', **{'in': 8}, out=8, dim=2, layers=2, cat=True, lin=True, dropout=0.0)
    make_case('ragged_dense_spline', sp1, sp2, inp, y, 3, -1, True, loss_on='0+L', seed=2)

    # -- DBP15K-lite sparse pair, training with negatives + ground truth --------------------
    ns, nt = 300, 320
    x = torch.randn(ns, 24, generator=gen)
    x_t = torch.randn(nt, 24, generator=gen)
    match = torch.randperm(nt, generator=gen)[:ns]
    x_t[match] = x + 0.3 * torch.randn(ns, 24, generator=gen)
    inp = dict(x_s=x, edge_index_s=rand_graph(gen, ns, 1100), edge_attr_s=None, batch_s=None,
               x_t=x_t, edge_index_t=rand_graph(gen, nt, 1400), edge_attr_t=None, batch_t=None)
    rows = torch.sort(torch.randperm(ns, generator=gen)[:60])[0]
    y = torch.stack([rows, match[rows]])
    r1 = dict(kind='rel', **{'in': 24}, out=32, layers=2, batch_norm=False, cat=True, lin=True,
              dropout=0.0)
    r2 = dict(kind='rel', **{'in': 8}, out=8, layers=2, batch_norm=False, cat=True, lin=True,
              dropout=0.0)
    make_case('sparse_rel_train', r1, r2, inp, y, 3, 5, True, loss_on='L', seed=3)
    make_case('sparse_rel_train_detach', r1, r2, inp, y, 3, 5, True, detach=True, loss_on='L',
              seed=3)
    make_case('sparse_rel_eval', r1, r2, inp, y, 3, 5, False, seed=3)
    make_case('sparse_rel_steps0', r1, r2, inp, y, 0, 5, True, loss_on='L', seed=3)

    # -- the reference's own test fixture (test_dgmc.py:5-10,67-84): GIN, 2 x 4-node path ----
    x4 = torch.randn(4, 32, generator=gen)
    e4 = torch.tensor([[0, 1, 1, 2, 2, 3], [1, 0, 2, 1, 3, 2]])
    xb = torch.cat([x4, x4])
    eb = torch.cat([e4, e4 + 4], dim=1)
    bb = torch.tensor([0, 0, 0, 0, 1, 1, 1, 1])
    inp = dict(x_s=xb, edge_index_s=eb, edge_attr_s=None, batch_s=bb, x_t=xb, edge_index_t=eb,
               edge_attr_t=None, batch_t=bb)
    y = torch.stack([torch.arange(8), torch.arange(8) % 4])
    g1 = dict(kind='gin', **{'in': 32}, out=16, layers=2, batch_norm=False, cat=True, lin=True)
    g2 = dict(kind='gin', **{'in': 8}, out=8, layers=2, batch_norm=False, cat=True, lin=True)
    make_case('gin_batched_dense', g1, g2, inp, y, 1, -1, True, loss_on='0', seed=12345)
    make_case('gin_batched_sparse_kN', g1, g2, inp, y, 1, 4, True, loss_on='0', seed=12345)

    # -- ragged SPARSE batch with a RelCNN psi_2 and two consensus steps: the reference leaves padded
    #    source rows in play (they hold uniform S and scatter into real targets)
    #    and applies the consensus MLP to them, so the second step sees what they contributed.
    #    k = N_t(max) selects every column, which makes the candidate SETS independent of tie order.
    xs, eis, _, bs = batched(gen, [5, 7, 6], 3, 12)
    xt, eit, _, bt = batched(gen, [6, 7, 5], 3, 12)
    inp = dict(x_s=xs, edge_index_s=eis, edge_attr_s=None, batch_s=bs, x_t=xt, edge_index_t=eit,
               edge_attr_t=None, batch_t=bt)
    y = torch.stack([torch.arange(18), torch.tensor([0, 1, 2, 3, 4, 6, 5, 4, 3, 2, 1, 0,
                                                     0, 1, 2, 3, 4, 0])])
    q1 = dict(kind='rel', **{'in': 12}, out=16, layers=2, batch_norm=False, cat=True, lin=True,
              dropout=0.0)
    q2 = dict(kind='rel', **{'in': 8}, out=8, layers=2, batch_norm=False, cat=True, lin=True,
              dropout=0.0)
    make_case('ragged_sparse_rel_kN', q1, q2, inp, y, 2, 7, True, loss_on='L', seed=4)


if __name__ == '__main__':
    main()
