"""Pins oracle/dgmc_oracle.py on the golden vectors recorded from the UNMODIFIED reference
(oracle/make_golden.py).  CPU only."""
import os

import pytest
import torch

from helpers import golden_names, load_case, run_oracle
from oracle import dgmc_oracle as O

NAMES = golden_names()


def _close(a, b, rtol, atol, what):
    a, b = a.detach().double(), b.detach().double()
    assert a.shape == b.shape, what
    err = (a - b).abs().max().item() if a.numel() else 0.0
    assert torch.allclose(a, b, rtol=rtol, atol=atol), '%s: max|d| = %.3g' % (what, err)


def test_golden_files_present():
    assert len(NAMES) >= 8


@pytest.mark.parametrize('name', NAMES)
def test_oracle_fp64_matches_reference_fp64(name):
    case = load_case(name)
    ref = case['out64']
    got = run_oracle(case, torch.float64)
    if 'idx' in ref:
        assert torch.equal(got['idx'], ref['idx']), 'candidate indices differ'
    _close(got['S_0'], ref['S_0'], 1e-9, 1e-12, 'S_0')
    _close(got['S_L'], ref['S_L'], 1e-9, 1e-12, 'S_L')
    if 'loss_L' in ref:
        _close(got['loss_0'], ref['loss_0'], 1e-10, 1e-12, 'loss_0')
        _close(got['loss_L'], ref['loss_L'], 1e-10, 1e-12, 'loss_L')
        assert got['acc_L'] == pytest.approx(ref['acc_L'].item())
        assert got['hits3_L'] == pytest.approx(ref['hits3_L'].item())
    if 'grads' in ref:
        assert set(got['grads']) == set(ref['grads'])
        for k, g in ref['grads'].items():
            _close(got['grads'][k], g, 1e-7, 1e-11, 'grad ' + k)
    if 'S_L_dense' in ref:
        _close(O.sparse_to_dense(got['idx'], got['S_L'], got['N_t']), ref['S_L_dense'], 1e-9,
               1e-12, 'to_dense')


@pytest.mark.parametrize('name', NAMES)
def test_oracle_fp32_matches_reference_fp32(name):
    case = load_case(name)
    ref = case['out32']
    got = run_oracle(case, torch.float32)
    if 'idx' in ref:
        assert torch.equal(got['idx'], ref['idx'])
    _close(got['S_0'], ref['S_0'], 2e-5, 1e-7, 'S_0')
    _close(got['S_L'], ref['S_L'], 2e-5, 1e-7, 'S_L')
    if 'loss_L' in ref:
        _close(got['loss_L'], ref['loss_L'], 1e-5, 1e-6, 'loss_L')


def test_include_gt_known_answer():
    # the reference's one known-answer vector, test/models/test_dgmc.py:87-95
    S_idx = torch.tensor([[[0, 1], [1, 2]], [[1, 2], [0, 1]]])
    s_mask = torch.tensor([[True, False], [True, True]])
    y = torch.tensor([[0, 1], [0, 0]])
    assert O.include_gt(S_idx, s_mask, y).tolist() == [[[0, 1], [1, 2]], [[1, 0], [0, 1]]]


def test_dense_equals_sparse_k_equals_n():
    # the invariant the reference's tests rest on (test_dgmc.py:59-64,83-84)
    dense, sparse = load_case('gin_batched_dense'), load_case('gin_batched_sparse_kN')
    sparse['params'] = dense['params']
    sparse['draws'] = {'r_s': dense['draws']['r_s'], 'neg': None}
    a = run_oracle(dense, torch.float64, with_grad=False)
    b = run_oracle(sparse, torch.float64, with_grad=False)
    assert torch.allclose(a['S_0'], O.sparse_to_dense(b['idx'], b['S_0'], b['N_t']))
    assert torch.allclose(a['S_L'], O.sparse_to_dense(b['idx'], b['S_L'], b['N_t']))
    assert torch.allclose(a['loss_0'], b['loss_0'])


def test_steps0_SL_equals_S0():
    case = load_case('sparse_rel_steps0')
    got = run_oracle(case, torch.float32, with_grad=False)
    assert torch.equal(got['S_0'], got['S_L'])


@pytest.mark.skipif(not os.path.isdir('/root/reference/dgmc'),
                    reason='live reference only exists in the build container')
def test_live_reference_regenerates_same_golden(tmp_path):
    """Re-run the generator against the live reference and compare with the committed files."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = ("import sys; sys.argv=['x']; import runpy, torch, os;"
              "ns = runpy.run_path(os.path.join(%r, 'oracle', 'make_golden.py'));"
              "ns['make_case'].__globals__['OUT'] = %r; ns['main']()") % (root, str(tmp_path))
    subprocess.run([sys.executable, '-W', 'ignore', '-c', script], check=True, cwd=root,
                   capture_output=True)
    for name in NAMES:
        new = torch.load(os.path.join(str(tmp_path), name + '.pt'), weights_only=False)
        old = load_case(name)
        assert torch.allclose(new['out64']['S_L'], old['out64']['S_L'], rtol=1e-12, atol=0)
