#!/usr/bin/env python
"""The reference's PascalVOC keypoint script (examples/pascal.py:1-109; WILLOW's examples/willow.py has
the same loop) on synthetic keypoint graphs, running on the B200 drop-in, THROUGH the pair-data
pipeline of dgmc.utils: ValidPairDataset(sample=True) -> collate_pairs (the `follow_batch=['x_s',
'x_t']` collation of pascal.py:42-43) -> DGMC(SplineCNN, SplineCNN, num_steps) dense matching ->
loss(S_0) + loss(S_L) -> Adam; evaluation = acc over sampled pairs (pascal.py:80-101).

There is no network here, so PascalVOCKeypoints + Delaunay + Cartesian are replaced by a seeded
generator with the same structure: per category a set of keypoint classes with a class "appearance"
vector; an instance shows a random subset (>= 3) of them at jittered 2-d positions, node feature =
appearance + noise (1024-d like the VGG16 features), edges = k nearest keypoints in both directions,
edge attribute = relative position mapped to [0, 1]^2 (T.Cartesian).

    python examples/pascal_synthetic.py [--epochs 5] [--batch_size 512]
"""
import argparse
import os
import random
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'deep-graph-matching-consensus_b200'))

from dgmc.models import DGMC, SplineCNN  # noqa: E402
from dgmc.utils import ValidPairDataset  # noqa: E402
from dgmc.utils.data import PairData, collate_pairs  # noqa: E402

parser = argparse.ArgumentParser()
parser.add_argument('--dim', type=int, default=256)
parser.add_argument('--rnd_dim', type=int, default=128)
parser.add_argument('--num_layers', type=int, default=2)
parser.add_argument('--num_steps', type=int, default=10)
parser.add_argument('--lr', type=float, default=0.001)
parser.add_argument('--batch_size', type=int, default=512)
parser.add_argument('--epochs', type=int, default=5)
parser.add_argument('--categories', type=int, default=4)
parser.add_argument('--graphs', type=int, default=512, help='instances per category and split')
parser.add_argument('--features', type=int, default=1024)
parser.add_argument('--noise', type=float, default=1.0)
args = parser.parse_args()

random.seed(0)
gen = torch.Generator().manual_seed(0)


class Graph(object):
    """What the reference's datasets yield per instance: x, edge_index, edge_attr, y (keypoint class)."""
    def __init__(self, x, edge_index, edge_attr, y):
        self.x, self.edge_index, self.edge_attr, self.y = x, edge_index, edge_attr, y


def make_category(n_classes, n_graphs):
    appearance = torch.randn(n_classes, args.features, generator=gen)
    layout = torch.rand(n_classes, 2, generator=gen)
    graphs = []
    for _ in range(n_graphs):
        n = int(torch.randint(max(3, n_classes // 2), n_classes + 1, (1, ), generator=gen))
        y = torch.randperm(n_classes, generator=gen)[:n]
        pos = layout[y] + 0.05 * torch.randn(n, 2, generator=gen)
        x = appearance[y] + args.noise * torch.randn(n, args.features, generator=gen)
        d = (pos.unsqueeze(1) - pos.unsqueeze(0)).norm(dim=-1)
        d.fill_diagonal_(float('inf'))
        nbr = d.topk(min(4, n - 1), largest=False)[1]                 # Delaunay stand-in: 4 nearest
        src = torch.arange(n).view(-1, 1).expand_as(nbr).reshape(-1)
        edge_index = torch.cat([torch.stack([src, nbr.reshape(-1)]), torch.stack([nbr.reshape(-1), src])], dim=1)
        rel = pos[edge_index[1]] - pos[edge_index[0]]
        edge_attr = (rel / (2 * rel.abs().max().clamp_min(1e-6)) + 0.5).clamp(0, 1)   # T.Cartesian
        graphs.append(Graph(x, edge_index, edge_attr, y))
    return graphs


train_sets, test_sets = [], []
for _ in range(args.categories):
    n_classes = int(torch.randint(10, 21, (1, ), generator=gen))
    graphs = make_category(n_classes, 2 * args.graphs)
    train_sets.append(ValidPairDataset(graphs[:args.graphs], graphs[:args.graphs], sample=True))
    test_sets.append(ValidPairDataset(graphs[args.graphs:], graphs[args.graphs:], sample=True))
train_dataset = torch.utils.data.ConcatDataset(train_sets)


def loader(dataset, batch_size, shuffle):
    order = list(range(len(dataset)))
    if shuffle:
        random.shuffle(order)
    for lo in range(0, len(order), batch_size):
        yield collate_pairs([dataset[i] for i in order[lo:lo + batch_size]])


device = 'cuda'
psi_1 = SplineCNN(args.features, args.dim, 2, args.num_layers, cat=False, dropout=0.5)
psi_2 = SplineCNN(args.rnd_dim, args.rnd_dim, 2, args.num_layers, cat=True, dropout=0.0)
model = DGMC(psi_1, psi_2, num_steps=args.num_steps).to(device)
optimizer = torch.optim.Adam(model.parameters(), lr=args.lr)


def generate_y(y_col):
    return torch.stack([torch.arange(y_col.size(0), device=device), y_col], dim=0)


def forward(data):
    return model(data.x_s, data.edge_index_s, data.edge_attr_s, data.x_s_batch, data.x_t,
                 data.edge_index_t, data.edge_attr_t, data.x_t_batch)


def train():
    model.train()
    total, pairs = 0.0, 0
    for data in loader(train_dataset, args.batch_size, True):
        optimizer.zero_grad()
        n_pairs = int(data.x_s_batch[-1]) + 1
        data = data.to(device)
        S_0, S_L = forward(data)
        y = generate_y(data.y)
        loss = model.loss(S_0, y)
        loss = model.loss(S_L, y) + loss if model.num_steps > 0 else loss
        loss.backward()
        optimizer.step()
        total += loss.item() * n_pairs
        pairs += n_pairs
    return total / pairs, pairs


@torch.no_grad()
def test(dataset):
    model.eval()
    correct = examples = 0
    for data in loader(dataset, args.batch_size, False):
        data = data.to(device)
        _, S_L = forward(data)
        y = generate_y(data.y)
        correct += model.acc(S_L, y, reduction='sum')
        examples += y.size(1)
    return correct / examples


print(model)
for epoch in range(1, args.epochs + 1):
    torch.cuda.synchronize()
    t0 = time.time()
    loss, pairs = train()
    torch.cuda.synchronize()
    dt = time.time() - t0
    accs = [100 * test(ds) for ds in test_sets]
    print('Epoch %02d  loss %.4f  %6.0f pairs/s  acc per category %s  mean %.1f' %
          (epoch, loss, pairs / dt, ' '.join('%.1f' % a for a in accs), sum(accs) / len(accs)), flush=True)
