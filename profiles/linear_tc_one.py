"""One launch of dgmc_linear_tc per GEMM shape of the DBP15K step (for `ncu --set full -k
regex:linear_tc_kernel`): psi_2 forward (38,960 x 192 -> 64), its input gradient (64 -> 192), psi_1's
first layer (34,500 x 900 -> 256: the range-draining variant) and psi_1's input gradient (256 -> 768).
    python profiles/linear_tc_one.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                'deep-graph-matching-consensus_b200'))
from dgmc.native import call, query, stream_ptr, RowStrided

torch.manual_seed(0)
dev = 'cuda'
for rows, n_out, k_in, relu, trans in [(38960, 64, 192, True, False), (38960, 192, 64, False, True),
                                       (34500, 256, 900, True, False), (34500, 768, 256, False, True)]:
    x = torch.randn(rows, k_in, device=dev)
    w = torch.randn(k_in, n_out, device=dev) * 0.05 if trans else torch.randn(n_out, k_in, device=dev) * 0.05
    b = torch.randn(n_out, device=dev)
    ws = torch.empty(query('dgmc_linear_tc_workspace', n_out, k_in), dtype=torch.uint8, device=dev)
    call('dgmc_linear_tc_split_weight', w, w.stride(0), int(trans), n_out, k_in, ws, ws.numel(), stream_ptr())
    y = torch.empty(rows, n_out, device=dev)
    for _ in range(2):      # the second launch of each shape is the warm one
        call('dgmc_linear_tc', RowStrided(x), x.stride(0), ws, b, int(relu), y, n_out, rows, n_out, k_in, stream_ptr())
    torch.cuda.synchronize()
    print('rows %d n_out %d k_in %d done' % (rows, n_out, k_in))
