"""Mirror of the reference's dpvo/blocks.py::SoftAgg (:31-48): same constructor, parameters (f, g, h -> a reference checkpoint's
`update.agg_kk.*` / `update.agg_ij.*` entries load unchanged) and forward(x, ix) semantics, with torch.unique + the two
torch_scatter calls (:41-43) replaced by this repository's `cuda_agg` module: a sort-free group plan and ONE kernel that does
scatter_softmax and scatter_sum together (csrc/agg.cu).  The Linear layers stay torch (cuBLAS)."""
import torch
import torch.nn as nn

from . import native_agg


def group_plan(ix, key_min=0, key_max=-1, g_capacity=None):
    """(jx, plan): jx == torch.unique(ix, return_inverse=True)[1]; plan feeds scatter_softmax_sum.  key_min / key_max (inclusive)
    bound the keys when the caller knows them (saves a reduction + host read); keys outside raise IndexError.
    g_capacity: a bound on the number of distinct keys; with it (and key bounds) nothing here waits for the device -- y gets
    g_capacity rows, the ones beyond the groups present are zero and never gathered by jx."""
    nosync = g_capacity is not None and key_max >= key_min
    jx, order, offsets, ng, G = native_agg().group_plan(ix, key_min, key_max, not nosync)
    return jx, (order, offsets, ng, int(g_capacity) if nosync else G)


def scatter_softmax_sum(fx, gx, plan):
    """scatter_sum(fx * scatter_softmax(gx, jx, dim=1), jx, dim=1) -> [B, G, D] (dpvo/blocks.py:42-43)"""
    order, offsets, ng, G = plan
    return native_agg().soft_agg(fx, gx, order, offsets, ng, G)


def _soft_agg_autograd(fx, gx, jx, G):
    """training path (gradients needed): the same arithmetic composed from differentiable torch ops on the device"""
    idx = jx.view(1, -1, 1).expand_as(gx)
    mx = torch.full((gx.shape[0], G, gx.shape[2]), float("-inf"), device=gx.device, dtype=gx.dtype).scatter_reduce(
        1, idx, gx.detach(), "amax", include_self=True)
    w = (gx - mx.gather(1, idx)).exp()
    den = torch.zeros_like(mx).scatter_add(1, idx, w)
    return torch.zeros_like(mx).scatter_add(1, idx, fx * (w / den.gather(1, idx)))


class SoftAgg(nn.Module):
    def __init__(self, dim=512, expand=True):
        super(SoftAgg, self).__init__()
        self.dim = dim
        self.expand = expand
        self.f = nn.Linear(self.dim, self.dim)
        self.g = nn.Linear(self.dim, self.dim)
        self.h = nn.Linear(self.dim, self.dim)

    def forward(self, x, ix, key_min=0, key_max=-1, g_capacity=None):
        # the optional bounds are hints (see group_plan); forward(x, ix) is the reference's call
        jx, plan = group_plan(ix, key_min, key_max, g_capacity if self.expand else None)
        fx, gx = self.f(x), self.g(x)
        if torch.is_grad_enabled() and (fx.requires_grad or gx.requires_grad):
            y = _soft_agg_autograd(fx, gx, jx, plan[3])
        else:
            y = scatter_softmax_sum(fx, gx, plan)
        if self.expand:
            return self.h(y)[:, jx]
        return self.h(y)
