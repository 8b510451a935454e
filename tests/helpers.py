"""Shared helpers for the parity tests (test infrastructure)."""
import glob
import os

import torch

from oracle import dgmc_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def golden_names():
    return sorted(os.path.splitext(os.path.basename(f))[0]
                  for f in glob.glob(os.path.join(GOLDEN, '*.pt')))


def load_case(name):
    return torch.load(os.path.join(GOLDEN, name + '.pt'), weights_only=False)


def cast_tree(obj, dtype):
    if torch.is_tensor(obj):
        return obj.to(dtype) if obj.is_floating_point() else obj
    if isinstance(obj, dict):
        return {k: cast_tree(v, dtype) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [cast_tree(v, dtype) for v in obj]
    return obj


def run_oracle(case, dtype, with_grad=True, S_idx=None, topk_dtype=None):
    """Oracle forward (+ loss, + backward) on a golden case in `dtype`."""
    p = {k: v.to(dtype).clone().requires_grad_(with_grad and case['training'])
         if v.is_floating_point() else v for k, v in case['params'].items()}
    inp = cast_tree(case['inputs'], dtype)
    y = case['y']
    use_y = y if (case['training'] and case['k'] >= 1) else None
    out = O.dgmc_forward(case['spec1'], case['spec2'], p, inp['x_s'], inp['edge_index_s'],
                         inp['edge_attr_s'], inp['batch_s'], inp['x_t'], inp['edge_index_t'],
                         inp['edge_attr_t'], inp['batch_t'], case['num_steps'], case['k'], y=use_y,
                         training=case['training'], detach=case['detach'], draws=case['draws'],
                         S_idx=S_idx, topk_dtype=topk_dtype)
    idx = out.get('idx')
    res = dict(out)
    if y is not None:
        res['loss_0'] = O.loss(out['S_0'], y, idx=idx)
        res['loss_L'] = O.loss(out['S_L'], y, idx=idx)
        res['acc_L'] = O.acc(out['S_L'], y, idx=idx)
        res['hits3_L'] = O.hits_at_k(3, out['S_L'], y, idx=idx)
        if with_grad and case['training']:
            total = {'0': res['loss_0'], 'L': res['loss_L'],
                     '0+L': res['loss_0'] + res['loss_L']}[case['loss_on']]
            total.backward()
            res['grads'] = {k: v.grad for k, v in p.items()
                            if torch.is_tensor(v) and v.is_floating_point() and v.grad is not None}
    return res


def parity_bound(ref64, ref32, rel=1e-5, floor=1e-7):
    """|new - oracle64| <= max(rel*|oracle64|, 2*|oracle32 - oracle64|) + floor."""
    ref64 = ref64.double()
    return torch.maximum(rel * ref64.abs(), 2 * (ref32.double() - ref64).abs()) + floor


def assert_parity(new, ref64, ref32, rel=1e-5, floor=1e-7, what=''):
    err = (new.detach().double().cpu() - ref64.double()).abs()
    bound = parity_bound(ref64, ref32, rel, floor)
    worst = (err - bound).max().item()
    assert worst <= 0, '%s: exceeds parity bound by %.3g (max err %.3g)' % (what, worst,
                                                                        err.max().item())
