"""Stand-in for torch_scatter (semantics per SURVEY.md App. B): out length = dim_size or
index.max()+1; mean = sum / max(count, 1); max over an empty group = 0; scatter_max also
returns the arg (lowest source position among ties).  Test infrastructure only."""
import torch


def _size(index, dim_size):
    if dim_size is not None:
        return int(dim_size)
    return int(index.max()) + 1 if index.numel() else 0


def _bcast(index, src, dim):
    if dim < 0:
        dim += src.dim()
    if index.dim() == 1 and src.dim() > 1:
        shape = [1] * src.dim()
        shape[dim] = -1
        index = index.view(shape).expand_as(src)
    return index, dim


def scatter_add(src, index, dim=-1, out=None, dim_size=None):
    index, dim = _bcast(index, src, dim)
    shape = list(src.shape)
    shape[dim] = _size(index, dim_size)
    res = src.new_zeros(shape)
    return res.scatter_add(dim, index, src)


scatter_sum = scatter_add


def scatter_mean(src, index, dim=-1, out=None, dim_size=None):
    total = scatter_add(src, index, dim, dim_size=dim_size)
    ones = torch.ones_like(src)
    count = scatter_add(ones, index, dim, dim_size=dim_size).clamp(min=1)
    return total / count


def _scatter_ext(src, index, dim, dim_size, mode):
    index, dim = _bcast(index, src, dim)
    shape = list(src.shape)
    shape[dim] = _size(index, dim_size)
    res = src.new_zeros(shape)
    res = res.scatter_reduce(dim, index, src, reduce=mode, include_self=False)
    # arg: lowest source position that attains the extreme value
    pos = torch.arange(src.shape[dim], device=src.device)
    vshape = [1] * src.dim()
    vshape[dim] = -1
    pos = pos.view(vshape).expand_as(src)
    hit = res.gather(dim, index) == src
    big = src.shape[dim]
    cand = torch.where(hit, pos, torch.full_like(pos, big))
    arg = torch.full(shape, big, dtype=torch.long, device=src.device)
    arg = arg.scatter_reduce(dim, index, cand, reduce='amin', include_self=True)
    return res, arg


def scatter_max(src, index, dim=-1, out=None, dim_size=None):
    return _scatter_ext(src, index, dim, dim_size, 'amax')


def scatter_min(src, index, dim=-1, out=None, dim_size=None):
    return _scatter_ext(src, index, dim, dim_size, 'amin')


def scatter(src, index, dim=-1, out=None, dim_size=None, reduce='sum'):
    if reduce in ('sum', 'add'):
        return scatter_add(src, index, dim, dim_size=dim_size)
    if reduce == 'mean':
        return scatter_mean(src, index, dim, dim_size=dim_size)
    if reduce == 'max':
        return scatter_max(src, index, dim, dim_size=dim_size)[0]
    if reduce == 'min':
        return scatter_min(src, index, dim, dim_size=dim_size)[0]
    raise ValueError(reduce)
