"""TEST INFRASTRUCTURE ONLY — stand-in for the third-party `torch_scatter` package.

The reference imports `scatter_add` at /root/reference/dgmc/models/dgmc.py:3 and calls
it at dgmc.py:212 as `scatter_add(src [B,M,R], index [B,M,1], dim=1, dim_size=N_t)`.
torch_scatter is not installed (and not installable offline), so the golden-vector
generator (`oracle/make_golden.py`) puts this directory on `sys.path` to import the
UNMODIFIED reference.  Semantics restated from the library's documentation: `index` is
broadcast against `src`, and `out[..., index[i], ...] += src[..., i, ...]` along `dim`.
Nothing in the product package may import this module.
"""
import torch


def scatter_add(src, index, dim=-1, out=None, dim_size=None):
    dim = dim if dim >= 0 else src.dim() + dim
    if index.dim() == 1 and src.dim() > 1:
        shape = [1] * src.dim()
        shape[dim] = -1
        index = index.view(shape)
    index = index.expand_as(src)
    if out is None:
        size = list(src.size())
        size[dim] = int(index.max()) + 1 if dim_size is None else dim_size
        out = src.new_zeros(size)
    return out.scatter_add_(dim, index, src)


def scatter_mean(src, index, dim=-1, out=None, dim_size=None):
    total = scatter_add(src, index, dim, out, dim_size)
    ones = torch.ones_like(src)
    count = scatter_add(ones, index, dim, None, total.size(dim)).clamp_(min=1)
    return total / count
