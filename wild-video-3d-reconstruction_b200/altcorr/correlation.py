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
