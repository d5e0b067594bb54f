"""A recording stand-in for the rasterizer packages, used to pin `render()`'s post-processing numerically
(tests/golden/make_render_golden.py runs the REFERENCE render() over it; tests/test_renderer_gpu.py runs the
mirror over the same stand-in).  forward() returns fixed seeded images as autograd leaves, so the loss gradient
w.r.t. the rasterizer outputs (i.e. the backward of the post-processing) is observable."""
from __future__ import annotations

from typing import NamedTuple

import torch


class Settings3(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class SettingsS(NamedTuple):
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


def make_fake(surfel: bool, visible: torch.Tensor, seed: int, log: dict):
    """-> a GaussianRasterizer class.  `visible` is what markVisible returns; `log` receives the settings, the
    inputs and the output leaves of the last forward."""

    class FakeRasterizer(torch.nn.Module):
        def __init__(self, raster_settings):
            super().__init__()
            self.raster_settings = raster_settings
            log["settings"] = raster_settings

        def markVisible(self, positions):
            return visible.to(positions.device)

        def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                    cov3D_precomp=None, theta=None, rho=None):
            s = self.raster_settings
            H, W, P = s.image_height, s.image_width, means3D.shape[0]
            dev = means3D.device
            g = torch.Generator().manual_seed(seed)
            leaf = lambda t: t.to(dev).requires_grad_(True)
            color = leaf(torch.rand(3, H, W, generator=g))
            alpha = torch.rand(1, H, W, generator=g)
            alpha[:, : H // 4] *= 5e-4          # a band below min_alpha, so the visibility mask matters
            alpha = leaf(alpha)
            depth = leaf(torch.rand(1, H, W, generator=g) * 20.0 + 1.0)
            radii = (torch.arange(P) % 3).to(torch.int32).to(dev)
            log.update(inputs=dict(means3D=means3D, means2D=means2D, opacities=opacities, colors=colors_precomp,
                                   scales=scales, rotations=rotations, theta=theta, rho=rho))
            if surfel:
                normal = leaf(torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0))
                contrib = torch.rand(P, generator=g).to(dev)
                log["leaves"] = dict(color=color, normal=normal, depth=depth, alpha=alpha)
                return color * 1.0, normal * 1.0, depth * 1.0, alpha * 1.0, radii, contrib
            log["leaves"] = dict(color=color, depth=depth, alpha=alpha)
            # non-leaf outputs: the reference edits the 3DGS depth in place (GR:430-437)
            return color * 1.0, radii, depth * 1.0, alpha * 1.0, torch.zeros(P, dtype=torch.int32, device=dev)

    return FakeRasterizer


def render_loss(res: dict, seed: int) -> torch.Tensor:
    """A fixed scalar of every differentiable image in the result dict."""
    g = torch.Generator().manual_seed(seed + 17)
    loss = 0.0
    for k in ("render", "surf_depth", "rend_alpha", "rend_normal", "surf_normal"):
        v = res.get(k)
        if v is None:
            continue
        w = torch.randn(v.shape, generator=g).to(v.device)
        loss = loss + (v * w).sum()
    return loss
