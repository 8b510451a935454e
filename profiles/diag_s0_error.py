"""Where the S_0 error at the full DBP15K shape comes from: psi_1 (tensor-core GEMMs) or the
candidate gather-dot.  python profiles/diag_s0_error.py  (GPU box)."""
import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'deep-graph-matching-consensus_b200')]
from oracle import dgmc_oracle as O
from dgmc import ops
from dgmc.models import RelCNN

g = torch.Generator().manual_seed(0)
n_s, n_t, feat, dim = 19388, 19572, 300, 256
x = torch.randn(n_s, feat, generator=g)
e = torch.randint(n_s, (2, 70414), generator=g)
spec = dict(kind='rel', out=dim, layers=3, batch_norm=False, cat=True, lin=True, dropout=0.0); spec['in'] = feat
p = O.init_psi(torch.Generator().manual_seed(1), spec, 'psi_1')
psi = RelCNN(feat, dim, 3)
psi.load_state_dict({k[len('psi_1.'):]: v for k, v in p.items()}, strict=False)
psi = psi.cuda().eval()
with torch.no_grad():
    h = psi(x.cuda(), e.cuda())
    h64 = O.relcnn(x.double(), e, {k: v.double() for k, v in p.items()}, 'psi_1', spec)
    h32 = O.relcnn(x, e, p, 'psi_1', spec)
    for name, flag in (('library GEMM', False), ):
        ops.LINEAR_TENSOR_CORE = flag
        h_lib = psi(x.cuda(), e.cuda())
        ops.LINEAR_TENSOR_CORE = True
    sc = h64.abs().max().item()
    print('psi_1 out: max|h| %.3g' % sc)
    print('  tensor-core path  max err / max|h| = %.3g' % ((h.double().cpu() - h64).abs().max().item() / sc))
    print('  library GEMM path max err / max|h| = %.3g' % ((h_lib.double().cpu() - h64).abs().max().item() / sc))
    print('  CPU fp32 oracle   max err / max|h| = %.3g' % ((h32.double() - h64).abs().max().item() / sc))
    # gather-dot on identical fp32 inputs
    idx = torch.randint(n_s, (1, n_s, 20), generator=g)
    cand = ops.Candidates(idx.cuda(), n_s)
    s = ops.gather_dot(h, h, cand)
    ref = (h.double().unsqueeze(1) * h.double()[idx[0].cuda()]).sum(-1)
    cpu32 = (h.cpu().unsqueeze(1) * h.cpu()[idx[0]]).sum(-1)
    print('gather_dot: max|logit| %.3g, kernel max abs err %.3g, torch CPU fp32 sum max abs err %.3g' % (
        ref.abs().max().item(), (s.double() - ref).abs().max().item(), (cpu32.double() - ref.cpu()).abs().max().item()))
