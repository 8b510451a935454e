#!/usr/bin/env python
"""The reference's DBP15K training script (examples/dbp15k.py:1-76) on a synthetic DBP15K-shaped
pair, running on the B200 drop-in.  Same model, same two-phase schedule: first `--epochs1` epochs
with num_steps = 0 (psi_1 + top-k + S_0 only), then `--epochs2` epochs of neighbourhood consensus
with psi_1 detached.  There is no network in this environment, so the PyG `DBP15K` dataset is
replaced by a seeded generator: the target graph is a permuted copy of the source graph with noisy
features, 30 % of the correspondences are training pairs, the rest is the test set.

    python examples/dbp15k_synthetic.py [--nodes 19388] [--graphed]
"""
import argparse
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))

from dgmc.models import DGMC, RelCNN  # noqa: E402
from dgmc.utils import GraphedTrainStep  # noqa: E402

parser = argparse.ArgumentParser()
parser.add_argument('--nodes', type=int, default=19388)
parser.add_argument('--dim', type=int, default=256)
parser.add_argument('--rnd_dim', type=int, default=32)
parser.add_argument('--num_layers', type=int, default=3)
parser.add_argument('--num_steps', type=int, default=10)
parser.add_argument('--k', type=int, default=10)
parser.add_argument('--epochs1', type=int, default=100)
parser.add_argument('--epochs2', type=int, default=100)
parser.add_argument("--graphed", action="store_true", help="capture each phase in one CUDA graph")
parser.add_argument("--noise", type=float, default=4.0, help="feature noise of the target copies")
args = parser.parse_args()

device = 'cuda'
gen = torch.Generator().manual_seed(0)
n_s, n_t = args.nodes, args.nodes + 184
x1 = torch.randn(n_s, 300, generator=gen)
e1 = torch.randint(n_s, (2, int(3.6 * n_s)), generator=gen)
perm = torch.randperm(n_t, generator=gen)[:n_s]
x2 = torch.randn(n_t, 300, generator=gen)
x2[perm] = x1 + args.noise * torch.randn(n_s, 300, generator=gen)   # noisy copies of the source features
keep = torch.rand(e1.size(1), generator=gen) < 0.8                  # 80 % of the relations survive
e2 = torch.cat([perm[e1[:, keep]], torch.randint(n_t, (2, int(1.3 * n_s)), generator=gen)], dim=1)
rows = torch.randperm(n_s, generator=gen)
n_train = int(0.3 * n_s)
train_rows, test_rows = torch.sort(rows[:n_train])[0], torch.sort(rows[n_train:])[0]
train_y = torch.stack([train_rows, perm[train_rows]]).to(device)
test_y = torch.stack([test_rows, perm[test_rows]]).to(device)
x1, e1, x2, e2 = x1.to(device), e1.to(device), x2.to(device), e2.to(device)

psi_1 = RelCNN(x1.size(-1), args.dim, args.num_layers, batch_norm=False, cat=True, lin=True, dropout=0.5)
psi_2 = RelCNN(args.rnd_dim, args.rnd_dim, args.num_layers, batch_norm=False, cat=True, lin=True,
               dropout=0.0)
model = DGMC(psi_1, psi_2, num_steps=None, k=args.k).to(device)
optimizer = torch.optim.Adam(model.parameters(), lr=0.001, capturable=args.graphed)
inputs = (x1, e1, None, None, x2, e2, None, None, train_y)


def train():
    model.train()
    optimizer.zero_grad()
    _, S_L = model(*inputs)
    loss = model.loss(S_L, train_y)
    loss.backward()
    optimizer.step()
    return loss


@torch.no_grad()
def test():
    model.eval()
    _, S_L = model(x1, e1, None, None, x2, e2, None, None)
    return model.acc(S_L, test_y), model.hits_at_k(10, S_L, test_y)


print('Optimize initial feature matching...')
model.num_steps = 0
step = None
t0 = time.time()
for epoch in range(1, args.epochs1 + args.epochs2 + 1):
    if epoch == args.epochs1 + 1:
        print('Refine correspondence matrix...')
        model.num_steps = args.num_steps
        model.detach = True
        step = None
    if args.graphed:
        if step is None:
            model.train()
            step = GraphedTrainStep(model, optimizer, *inputs, structure_is_static=True)
        loss = step(*inputs)
    else:
        loss = train()
    if epoch % 10 == 0 or epoch > args.epochs1:
        if epoch % 10 == 0:
            hits1, hits10 = test()
            model.train()
            print(f'{epoch:03d}: Loss: {loss.item():.4f}, Hits@1: {hits1:.4f}, Hits@10: {hits10:.4f} '
                  f'[{time.time() - t0:.1f}s]')
