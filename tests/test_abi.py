"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads without a GPU and
exports every symbol include/dgmc_b200.h declares; the host package mirrors the reference's
public surface; CPU tensors are refused (no fallback).  No compute calls here."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def native():
    import __graft_entry__ as entry
    entry.build()
    from dgmc import native
    return native


def test_library_exports_every_declared_symbol(native):
    header = open(os.path.join(ROOT, 'include', 'dgmc_b200.h')).read()
    declared = set(re.findall(r'\b(dgmc_[a-z0-9_]+)\s*\(', re.sub(r'/\*.*?\*/', ' ', header, flags=re.S)))
    assert len(declared) >= 15
    lib = ctypes.CDLL(native.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), 'missing export: ' + name
    assert declared == set(native.PROTOTYPES)
    assert lib.dgmc_abi_version() == 1


def test_no_torch_types_in_signatures(native):
    header = open(os.path.join(ROOT, 'include', 'dgmc_b200.h')).read()
    assert 'torch' not in header.lower().replace('pytorch', '').replace('torch_', '')
    assert 'extern "C"' in header


def test_reference_public_surface():
    import dgmc
    from dgmc.models import DGMC, GIN, MLP, RelCNN, SplineCNN
    from dgmc.utils import PairDataset, ValidPairDataset  # noqa: F401
    assert dgmc.__version__ == '1.0.0' and dgmc.DGMC is DGMC
    psi_1, psi_2 = GIN(32, 16, num_layers=2), GIN(8, 8, num_layers=2)
    model = DGMC(psi_1, psi_2, num_steps=1)
    assert repr(model) == (  # test/models/test_dgmc.py:17-25
        'DGMC(\n'
        '    psi_1=GIN(32, 16, num_layers=2, batch_norm=False, cat=True, lin=True),\n'
        '    psi_2=GIN(8, 8, num_layers=2, batch_norm=False, cat=True, lin=True),\n'
        '    num_steps=1, k=-1\n)')
    model.reset_parameters()
    assert [n for n, _ in model.mlp.named_parameters()] == ['0.weight', '0.bias', '2.weight', '2.bias']
    assert repr(MLP(16, 32, num_layers=2, batch_norm=True, dropout=0.5)) == \
        'MLP(16, 32, num_layers=2, batch_norm=True, dropout=0.5)'
    assert repr(SplineCNN(16, 32, dim=3, num_layers=2, cat=True, lin=True, dropout=0.5)) == \
        'SplineCNN(16, 32, dim=3, num_layers=2, cat=True, lin=True, dropout=0.5)'
    rel = RelCNN(16, 32, num_layers=2, batch_norm=True, cat=True, lin=True, dropout=0.5)
    assert sorted(rel.state_dict())[:4] == ['batch_norms.0.bias', 'batch_norms.0.num_batches_tracked',
                                            'batch_norms.0.running_mean', 'batch_norms.0.running_var']
    assert {'convs.0.lin1.weight', 'convs.0.lin2.weight', 'convs.0.root.weight', 'convs.0.root.bias',
            'final.weight', 'final.bias'} <= set(rel.state_dict())
    for cat in (False, True):
        for lin in (False, True):
            assert RelCNN(16, 32, 2, True, cat, lin, 0.5).out_channels == \
                (16 + 2 * 32 if not lin and cat else 32)
    assert DGMC(psi_1, psi_2, num_steps=None, k=10).num_steps is None  # examples/dbp15k.py:33


def test_state_dict_names_match_golden_reference_names():
    from helpers import load_case, build_product
    for name in ('sparse_rel_train', 'ragged_dense_spline', 'gin_batched_dense'):
        case = load_case(name)
        model = build_product(case, 'cpu')
        mine = {k for k in model.state_dict() if 'num_batches_tracked' not in k}
        theirs = {k for k in case['params'] if 'num_batches_tracked' not in k}
        assert mine == theirs, (mine ^ theirs)


def test_cpu_tensors_are_refused(native):
    from dgmc import ops
    with pytest.raises(native.DgmcNativeError):
        ops.softmax_k(torch.randn(4, 4))
    from dgmc.models import DGMC, RelCNN
    model = DGMC(RelCNN(8, 8, 1), RelCNN(4, 4, 1), num_steps=1, k=2)
    x, e = torch.randn(5, 8), torch.tensor([[0, 1, 2], [1, 2, 3]])
    with pytest.raises(native.DgmcNativeError):
        model(x, e, None, None, x, e, None, None)


def test_pair_datasets():
    # test/utils/test_data.py:6-74, with a minimal Data stand-in
    from dgmc.utils import PairDataset, ValidPairDataset

    class Data(object):
        def __init__(self, **kw):
            self.edge_attr = None
            self.__dict__.update(kw)

        def __repr__(self):
            return 'Data()'

    x = torch.randn(10, 16)
    ei = torch.randint(10, (2, 30))
    data = Data(x=x, edge_index=ei)
    ds = PairDataset([data, data], [data, data], sample=True)
    assert len(ds) == 2 and len(ds[0]) == 4
    assert repr(ds) == 'PairDataset([Data(), Data()], [Data(), Data()], sample=True)'
    ds = PairDataset([data, data], [data, data], sample=False)
    assert len(ds) == 4
    pair = ds[0]
    assert torch.equal(pair.x_s, x) and pair.edge_index_t.tolist() == ei.tolist()
    data = Data(x=x, edge_index=ei, y=torch.randperm(10))
    for sample, n in ((True, 2), (False, 4)):
        ds = ValidPairDataset([data, data], [data, data], sample=sample)
        assert len(ds) == n
        pair = ds[0]
        assert len(pair) == 5 and pair.y.tolist() == list(range(10))
    from dgmc.utils.data import collate_pairs
    batch = collate_pairs([ds[0], ds[1]])
    assert batch.x_s.shape == (20, 16) and batch.x_s_batch.tolist() == [0] * 10 + [1] * 10
    assert batch.edge_index_s[:, 30:].min() >= 10 and batch.y.tolist() == list(range(10)) * 2
