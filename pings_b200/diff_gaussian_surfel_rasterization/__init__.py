"""Drop-in for the reference package `diff_gaussian_surfel_rasterization`
(submodules/diff-gaussian-surfel-rasterization-w-pose/diff_gaussian_surfel_rasterization/__init__.py):
same names, argument order, return tuples and error behaviour; the work is done by the
sm_100a kernels behind the C-ABI library (no CPU fallback).
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from .. import _lib as _L
from .. import _raster_c as _C


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta,
                        rho, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, theta, rho, raster_settings)


# Extension (no reference counterpart): GaussianRasterizer.param_grads_out, a flat fp32 buffer of 14*P elements that
# the backward writes the Gaussian-parameter gradients into ([means3D | colors | opacity | scales | rotations], the
# layout of pings_b200.parallel.FlatGradAllReduce), so a data-parallel caller all-reduces it without a gather copy.
_PENDING: dict = {}


class _RasterizeGaussians(torch.autograd.Function):
    """reference: __init__.py:48-178."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, theta, rho,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.prcppoint, rs.patch_bbox, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, rs.config)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                res = _C.rasterize_gaussians_surfel(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            res = _C.rasterize_gaussians_surfel(*args)
        num_rendered, color, normal, depth, opac, radii, contributions, geomBuffer, binningBuffer, imgBuffer = res
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.param_grads_out = _PENDING.pop("grads_out", None)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii, contributions)
        return color, normal, depth, opac, radii, contributions

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_normal, grad_out_depth, grad_out_opac, _, grad_out_contribution):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer,
         imgBuffer) = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.prcppoint, rs.patch_bbox, rs.tanfovx, rs.tanfovy,
                grad_out_color, grad_out_normal, grad_out_depth, grad_out_opac, sh, rs.sh_degree, rs.campos,
                geomBuffer, ctx.num_rendered, binningBuffer, imgBuffer, rs.debug, rs.config)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                res = _C.rasterize_gaussians_surfel_backward(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            res = _C.rasterize_gaussians_surfel_backward(*args, param_grads_out=ctx.param_grads_out)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations, grad_tau) = res
        # the kernel already reduced dL_dtau over surfels ([1,6]); same slicing as the reference
        # the kernel already reduced the pose gradient over the Gaussians to [1,6] (the reference sums [P,6] here)
        grad_tau = grad_tau.view(6) if grad_tau.numel() == 6 else torch.sum(grad_tau.view(-1, 6), dim=0)
        grad_theta = grad_tau[3:].view(1, -1)
        grad_rho = grad_tau[:3].view(1, -1)
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, grad_theta, grad_rho, None)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    patch_bbox: torch.Tensor
    prcppoint: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool
    config: torch.Tensor


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings
        self.param_grads_out = None   # optional flat gradient buffer (see _PENDING above)

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix, variant=_L.PG_VARIANT_SURFEL)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, theta=None, rho=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (
                (scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        theta = empty if theta is None else theta
        rho = empty if rho is None else rho
        if self.param_grads_out is not None:
            _PENDING["grads_out"] = self.param_grads_out
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations,
                                   cov3D_precomp, theta, rho, rs)
