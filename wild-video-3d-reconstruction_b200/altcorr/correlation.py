"""Mirror of the reference's dpvo/altcorr/correlation.py (same names, arguments and autograd behaviour) over
this repository's `cuda_corr` build.  The reference file itself also works unchanged against that module."""
import torch

from .. import native_modules

cuda_corr, _ = native_modules()


class CorrLayer(torch.autograd.Function):
    """corr(fmap1, fmap2, coords, ii, jj, radius, dropout): dpvo/altcorr/correlation.py:4-30"""

    @staticmethod
    def forward(ctx, fmap1, fmap2, coords, ii, jj, radius, dropout):
        ctx.save_for_backward(fmap1, fmap2, coords, ii, jj)
        ctx.radius, ctx.dropout = radius, dropout
        return cuda_corr.forward(fmap1, fmap2, coords, ii, jj, radius)[0]

    @staticmethod
    def backward(ctx, grad):
        fmap1, fmap2, coords, ii, jj = ctx.saved_tensors
        if ctx.dropout < 1:  # edge dropout of the reference's training path (:19-25)
            keep = torch.rand(len(ii), device=grad.device) < ctx.dropout
            coords, grad, ii, jj = coords[:, keep], grad[:, keep], ii[keep], jj[keep]
        g1, g2 = cuda_corr.backward(fmap1, fmap2, coords, ii, jj, grad, ctx.radius)
        return g1, g2, None, None, None, None, None


class PatchLayer(torch.autograd.Function):
    """integer-tap (2r+2)^2 crop around each point: dpvo/altcorr/correlation.py:33-49"""

    @staticmethod
    def forward(ctx, net, coords, radius):
        ctx.radius = radius
        ctx.save_for_backward(net, coords)
        return cuda_corr.patchify_forward(net, coords, radius)[0]

    @staticmethod
    def backward(ctx, grad):
        net, coords = ctx.saved_tensors
        return cuda_corr.patchify_backward(net, coords, grad, ctx.radius)[0], None, None


def patchify(net, coords, radius, mode='bilinear'):
    """extract (2r+1)^2 patches, bilinearly blended from the (2r+2)^2 crop (dpvo/altcorr/correlation.py:51-69)"""
    crop = PatchLayer.apply(net, coords, radius)
    if mode != 'bilinear':
        return crop
    frac = coords - coords.floor()
    dx = frac[..., 0][:, :, None, None, None]
    dy = frac[..., 1][:, :, None, None, None]
    d = 2 * radius + 1
    top = (1 - dx) * crop[..., :d, :d] + dx * crop[..., :d, 1:]
    bot = (1 - dx) * crop[..., 1:, :d] + dx * crop[..., 1:, 1:]
    return (1 - dy) * top + dy * bot


def corr(fmap1, fmap2, coords, ii, jj, radius=1, dropout=1):
    return CorrLayer.apply(fmap1, fmap2, coords, ii, jj, radius, dropout)


def corr_stacked(fmap1, pyramid, coords, ii, jj, radius=3, divs=(1, 4)):
    """DPVO.corr (dpvo/dpvo.py:325-333) as one call:

        torch.stack([corr(fmap1, pyramid[0], coords / divs[0], ii, jj, radius),
                     corr(fmap1, pyramid[1], coords / divs[1], ii, jj, radius)], -1).view(B, E, -1)

    The tracker's configuration (half maps, C = 128, 3x3 patches, radius 3, no gradient) runs both levels of every edge in one
    kernel that writes the stacked layout directly; anything else is composed from `corr` exactly as the reference does."""
    f0, f1l = pyramid
    B, E = coords.shape[0], coords.shape[1]
    fused = (fmap1.dtype == torch.float16 and f0.dtype == torch.float16 and f1l.dtype == torch.float16 and radius == 3
             and fmap1.shape[2] == 128 and tuple(fmap1.shape[3:]) == (3, 3) and f0.shape[1] == f1l.shape[1]
             and all(d > 0 and (d & (d - 1)) == 0 for d in (int(divs[0]), int(divs[1]))) and int(divs[0]) == divs[0] and int(divs[1]) == divs[1]
             and not (torch.is_grad_enabled() and any(t.requires_grad for t in (fmap1, f0, f1l))))
    if fused:
        return cuda_corr.forward_stacked(fmap1, f0, f1l, coords, ii, jj, radius, float(divs[0]), float(divs[1]))[0]
    c0 = corr(fmap1, f0, coords / divs[0], ii, jj, radius)
    c1 = corr(fmap1, f1l, coords / divs[1], ii, jj, radius)
    return torch.stack([c0, c1], -1).view(B, E, -1)
