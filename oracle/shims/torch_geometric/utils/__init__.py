"""Stand-in for torch_geometric.utils.to_dense_batch (reference call site dgmc.py:154-155)."""
import torch


def to_dense_batch(x, batch=None, fill_value=0, max_num_nodes=None):
    # batch=None: a single graph -> [1, N, C] and an all-true mask.
    if batch is None:
        mask = torch.ones(1, x.size(0), dtype=torch.bool, device=x.device)
        return x.unsqueeze(0), mask
    num_graphs = int(batch[-1]) + 1 if batch.numel() > 0 else 1
    counts = torch.bincount(batch, minlength=num_graphs)
    first = torch.cumsum(counts, 0) - counts
    n_max = int(counts.max()) if max_num_nodes is None else max_num_nodes
    slot = torch.arange(batch.numel(), device=x.device) - first[batch] + batch * n_max
    out = x.new_full((num_graphs * n_max, ) + tuple(x.shape[1:]), fill_value)
    out[slot] = x
    mask = torch.zeros(num_graphs * n_max, dtype=torch.bool, device=x.device)
    mask[slot] = True
    return out.view(num_graphs, n_max, *x.shape[1:]), mask.view(num_graphs, n_max)
