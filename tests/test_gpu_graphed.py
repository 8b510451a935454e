"""dgmc.utils.GraphedTrainStep: the captured step must consume fresh inputs on every replay, keep
training (loss falls on a planted alignment task) and reproduce the eager step: same seed, same
weights -> same loss and same updated parameters (the replayed graph is what bench.py times)."""
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


@pytest.mark.parametrize('static_structure', [True, False])
def test_graph_replay_equals_eager_step(static_structure):
    """One eager training step and one graph replay from the same weights, optimizer state and
    Philox seed/offset: the loss and every updated parameter must agree to 1e-6 relative (they are
    the same kernels in the same order; torch's generator hands a captured graph the seed and offset
    current at replay time, so both runs draw the same dropout masks, negatives and indicators)."""
    import copy
    from dgmc.models import DGMC, RelCNN
    from dgmc.utils import GraphedTrainStep
    torch.manual_seed(0)
    model = DGMC(RelCNN(24, 32, 2, dropout=0.5), RelCNN(8, 8, 2), num_steps=3, k=5, detach=False).cuda().train()
    start = copy.deepcopy(model.state_dict())
    dev = [None if t is None else t.cuda() for t in _problem(3)]

    # eager, on a copy of the model (autograd ties a parameter's gradient accumulation to the stream
    # of its first backward pass; the capture below runs on a side stream)
    twin = copy.deepcopy(model)
    opt = torch.optim.Adam(twin.parameters(), lr=1e-2, capturable=True)
    torch.cuda.manual_seed(77)
    opt.zero_grad(set_to_none=True)
    _, S_L = twin(*dev)
    loss_eager = twin.loss(S_L, dev[-1])
    loss_eager.backward()
    opt.step()
    eager = {n: p.detach().clone() for n, p in twin.named_parameters()}
    loss_eager = loss_eager.item()
    del twin, opt, S_L

    # graphed: capture (its warm-up steps train), then put weights and optimizer state back
    opt = torch.optim.Adam(model.parameters(), lr=1e-2, capturable=True)
    step = GraphedTrainStep(model, opt, *dev, structure_is_static=static_structure)
    with torch.no_grad():
        for n, p in model.named_parameters():
            p.copy_(start[n])
        for state in opt.state.values():
            for v in state.values():
                if torch.is_tensor(v):
                    v.zero_()
    torch.cuda.manual_seed(77)
    loss_graph = step(*dev).item()
    assert abs(loss_graph - loss_eager) <= 1e-6 * abs(loss_eager), (loss_graph, loss_eager)
    for n, p in model.named_parameters():
        ref = eager[n]
        err = (p.detach() - ref).abs().max().item()
        assert err <= 1e-6 * ref.abs().max().item(), (n, err)
