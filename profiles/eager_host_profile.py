"""Host-side profile (cProfile) of the eager DBP15K-shaped training step: where the Python / ctypes time
goes when the drop-in API is called without graph capture.  python profiles/eager_host_profile.py"""
import cProfile
import os
import pstats
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, 'deep-graph-matching-consensus_b200')]
sys.argv = ['bench.py']
import bench  # noqa: E402

w = dict(bench.WORKLOAD)
dev = torch.device('cuda', 0)
d = {k: v.to(dev) for k, v in bench.make_inputs(w).items()}
model = bench.build_model(w, dev)
opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)
for _ in range(5):
    bench.train_step(model, opt, d)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for _ in range(10):
    bench.train_step(model, opt, d)
host_issue = (time.perf_counter() - t0) / 10
torch.cuda.synchronize()
total = (time.perf_counter() - t0) / 10
print('host issue time %.2f ms/step, wall %.2f ms/step' % (host_issue * 1e3, total * 1e3), file=sys.stderr)
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    bench.train_step(model, opt, d)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr, stream=sys.stderr)
st.sort_stats('tottime').print_stats(28)
