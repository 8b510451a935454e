"""dgmc.utils.GraphedTrainStep: the captured step must consume fresh inputs on every replay, keep
training (loss falls on a planted alignment task) and agree with the eager step in expectation."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(seed, n_s=600, n_t=640, feat=24):
    g = torch.Generator().manual_seed(seed)
    x_s = torch.randn(n_s, feat, generator=g)
    perm = torch.randperm(n_t, generator=g)[:n_s]
    x_t = torch.randn(n_t, feat, generator=g)
    x_t[perm] = x_s + 0.1 * torch.randn(n_s, feat, generator=g)
    e_s = torch.randint(n_s, (2, 4 * n_s), generator=g)
    e_t = torch.randint(n_t, (2, 4 * n_t), generator=g)
    rows = torch.arange(0, n_s, 3)
    y = torch.stack([rows, perm[rows]])
    return [x_s, e_s, None, None, x_t, e_t, None, None, y]


@pytest.mark.parametrize('static_structure', [True, False])
def test_graphed_step_trains_and_reads_new_inputs(static_structure):
    from dgmc.models import DGMC, RelCNN
    from dgmc.utils import GraphedTrainStep
    torch.manual_seed(0)
    model = DGMC(RelCNN(24, 32, 2, dropout=0.0), RelCNN(8, 8, 2), num_steps=3, k=5).cuda().train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-2, capturable=True)
    host = [None if t is None else t.pin_memory() for t in _problem(1)]
    dev = [None if t is None else t.cuda() for t in host]
    step = GraphedTrainStep(model, opt, *dev, structure_is_static=static_structure)
    before = [p.detach().clone() for p in model.parameters()]
    losses = [step(*host).item() for _ in range(30)]
    assert all(torch.isfinite(torch.tensor(losses)))
    assert sum(losses[-5:]) < sum(losses[:5])          # it trains
    assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))
    if not static_structure:
        # a different pair of the same shape must give a different loss: inputs (features, edges,
        # ground truth) are really re-read and the graph structure rebuilt inside the replay
        other = [None if t is None else t.pin_memory() for t in _problem(2)]
        a = step(*host).item()
        b = step(*other).item()
        assert abs(a - b) > 1e-3
        # and the eager path on the same model/inputs gives a loss of the same size
        model.eval()
        with torch.no_grad():
            _, S_L = model(*[None if t is None else t.cuda() for t in other][:-1])
            eager = model.loss(S_L, other[-1].cuda()).item()
        assert abs(eager - b) < 0.5 * max(abs(b), 0.1) + 0.5


def test_prefetch_pipelines_inputs_without_changing_results():
    from dgmc.models import DGMC, RelCNN
    from dgmc.utils import GraphedTrainStep
    torch.manual_seed(0)
    model = DGMC(RelCNN(24, 32, 2, dropout=0.0), RelCNN(8, 8, 2), num_steps=2, k=5).cuda().eval()
    # eval-mode model + zero learning rate: the step's loss depends on the inputs only (up to the
    # random indicators), so feeding pair A / pair B alternately must give alternating losses
    opt = torch.optim.Adam(model.parameters(), lr=0.0, capturable=True)
    a = [None if t is None else t.pin_memory() for t in _problem(1)]
    b = [None if t is None else t.pin_memory() for t in _problem(2)]
    step = GraphedTrainStep(model, opt, *[None if t is None else t.cuda() for t in a],
                            structure_is_static=False)
    plain = [step(*(a if i % 2 == 0 else b)).item() for i in range(6)]
    piped = []
    loss = step(*a)
    for i in range(6):
        nxt = b if i % 2 == 0 else a
        step.prefetch(*nxt)
        piped.append(loss.item())
        loss = step(*nxt)
    mean = lambda v: sum(v) / len(v)  # noqa: E731
    assert abs(mean(plain[0::2]) - mean(piped[0::2])) < 0.05 * abs(mean(plain[0::2])) + 1e-3
    assert abs(mean(plain[1::2]) - mean(piped[1::2])) < 0.05 * abs(mean(plain[1::2])) + 1e-3
    assert abs(mean(plain[0::2]) - mean(plain[1::2])) > 1e-3   # the two pairs really differ
