"""Stand-in for `torch_scatter` 2.1.2 (absent in this image, environment.yaml:43 of the reference) in plain torch ops.

Used by BOTH tracker arms of the harness (this repository's kernels and REF-CUDA) so that the update operator's SoftAgg
(dpvo/blocks.py:41-43) and dpvo/ba.py's scatter_sum cost the same on either side; the fused replacement of SoftAgg is
wild-video-3d-reconstruction_b200/blocks.py and is measured separately.  Semantics follow the published torch_scatter API:
index is broadcast along `dim`, out-of-place, dim_size defaults to index.max() + 1."""
import torch


def _expand(index, src, dim):
    if dim < 0:
        dim += src.dim()
    if index.dim() == 1:
        shape = [1] * src.dim()
        shape[dim] = -1
        index = index.view(shape)
    elif index.dim() != src.dim():
        for _ in range(src.dim() - index.dim()):
            index = index.unsqueeze(-1)
    return index.expand_as(src), dim


def scatter_sum(src, index, dim=-1, out=None, dim_size=None):
    idx, dim = _expand(index, src, dim)
    if out is None:
        shape = list(src.shape)
        shape[dim] = int(dim_size) if dim_size is not None else (int(index.max()) + 1 if index.numel() else 0)
        out = torch.zeros(shape, dtype=src.dtype, device=src.device)
    return out.scatter_add_(dim, idx, src)


scatter_add = scatter_sum


def scatter_mean(src, index, dim=-1, out=None, dim_size=None):
    s = scatter_sum(src, index, dim, out, dim_size)
    cnt = scatter_sum(torch.ones_like(src), index, dim, None, s.shape[dim]).clamp_(min=1)
    return s / cnt


def scatter_max(src, index, dim=-1, out=None, dim_size=None):
    idx, dim = _expand(index, src, dim)
    shape = list(src.shape)
    shape[dim] = int(dim_size) if dim_size is not None else (int(index.max()) + 1 if index.numel() else 0)
    res = torch.full(shape, float("-inf") if src.is_floating_point() else torch.iinfo(src.dtype).min, dtype=src.dtype, device=src.device)
    res.scatter_reduce_(dim, idx, src, reduce="amax", include_self=True)
    # argmax: first position along dim whose value equals its group's maximum
    pos = torch.arange(src.shape[dim], device=src.device).view([-1 if d == dim else 1 for d in range(src.dim())]).expand_as(src)
    hit = torch.where(src == res.gather(dim, idx), pos, torch.full_like(pos, src.shape[dim]))
    arg = torch.full(shape, src.shape[dim], dtype=torch.long, device=src.device).scatter_reduce_(dim, idx, hit, reduce="amin", include_self=True)
    return res, arg


def scatter_softmax(src, index, dim=-1, eps=1e-12):
    idx, dim = _expand(index, src, dim)
    n = int(index.max()) + 1 if index.numel() else 0
    shape = list(src.shape)
    shape[dim] = n
    mx = torch.full(shape, float("-inf"), dtype=src.dtype, device=src.device).scatter_reduce_(dim, idx, src, reduce="amax", include_self=True)
    e = (src - mx.gather(dim, idx)).exp_()
    den = torch.zeros(shape, dtype=src.dtype, device=src.device).scatter_add_(dim, idx, e)
    return e / den.gather(dim, idx)
