"""Fused post-raster loss assembly: the photometric L1 and the masked depth L1
of the reference's GS mapping loss (utils/mapper.py:1239,1252-1267, `l1_loss` = mean |a - b|) as one
CUDA pass with its backward seeds, instead of ~15 eager full-image kernels."""
from __future__ import annotations

import torch

from . import _lib as L


class _PhotoDepthLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rgb, gt_rgb, depth, gt_depth, alpha, lambda_depth, depth_min, depth_max, min_alpha, inverse_depth):
        if not rgb.is_cuda:
            raise RuntimeError("pings_b200.losses: expected CUDA tensors (there is no CPU fallback)")
        lib = L.lib()
        dev = rgb.device
        H, W = rgb.shape[-2], rgb.shape[-1]
        f = lambda t: None if t is None else (t if (t.dtype is torch.float32 and t.is_contiguous()) else t.detach().to(torch.float32).contiguous())
        rgb_c, gt_c, d_c, gd_c, a_c = f(rgb), f(gt_rgb), f(depth), f(gt_depth), f(alpha)
        if d_c is None or gd_c is None:
            d_c = gd_c = a_c = None
        seed_rgb = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        seed_d = torch.empty(H, W, dtype=torch.float32, device=dev) if d_c is not None else None
        sums = torch.empty(3, dtype=torch.float64, device=dev)
        loss = torch.empty(3, dtype=torch.float32, device=dev)
        with L.device_guard(dev):
            L.check(lib.pg_photo_depth_loss_forward(H, W, rgb_c.data_ptr(), gt_c.data_ptr(), L.ptr(d_c), L.ptr(gd_c),
                                                    L.ptr(a_c), float(lambda_depth), float(depth_min), float(depth_max),
                                                    float(min_alpha), int(bool(inverse_depth)), seed_rgb.data_ptr(),
                                                    L.ptr(seed_d), sums.data_ptr(), loss.data_ptr(),
                                                    L.raw_stream(dev)), "photo/depth loss")
        ctx.save_for_backward(seed_rgb, seed_d, sums)
        ctx.meta = (H, W, float(lambda_depth), rgb.shape, None if depth is None else depth.shape)
        l_total, l_rgb, l_depth = loss.unbind(0)
        ctx.mark_non_differentiable(l_rgb, l_depth)
        return l_total, l_rgb, l_depth

    @staticmethod
    def backward(ctx, g_total, g_rgb, g_depth):
        seed_rgb, seed_d, sums = ctx.saved_tensors
        H, W, lam, rgb_shape, depth_shape = ctx.meta
        lib = L.lib()
        dev = seed_rgb.device
        up = g_total.detach().to(torch.float32).reshape(1).contiguous()
        d_rgb = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        want_d = seed_d is not None and ctx.needs_input_grad[2]
        d_depth = torch.empty(H, W, dtype=torch.float32, device=dev) if want_d else None
        with L.device_guard(dev):
            L.check(lib.pg_photo_depth_loss_backward(H, W, lam, up.data_ptr(), sums.data_ptr(), seed_rgb.data_ptr(),
                                                     L.ptr(seed_d), d_rgb.data_ptr(), L.ptr(d_depth),
                                                     L.raw_stream(dev)), "photo/depth loss bwd")
        return (d_rgb.reshape(rgb_shape), None, None if d_depth is None else d_depth.reshape(depth_shape), None, None,
                None, None, None, None, None)


def photo_depth_loss(rendered_rgb, gt_rgb, rendered_depth=None, gt_depth=None, rendered_alpha=None, lambda_depth=0.0,
                     depth_min=0.0, depth_max=float("inf"), min_alpha=0.0, inverse_depth=False):
    """Returns (loss, rgb_l1, depth_l1): loss = l1(rendered_rgb, gt_rgb) + lambda_depth * l1 over the valid depth
    pixels (utils/mapper.py:1252-1267; the alpha mask uses the detached alpha, as the reference does)."""
    return _PhotoDepthLoss.apply(rendered_rgb, gt_rgb, rendered_depth, gt_depth, rendered_alpha, lambda_depth, depth_min,
                                 depth_max, min_alpha, inverse_depth)
