"""Post-raster image ops of render() as CUDA kernels: depth2normal
(gaussian_splatting/utils/point_utils.py:83-148, ~25 eager full-image kernels in the reference) and the affine
exposure correction (GR:449-458), each one kernel forward and one backward behind the C-ABI
(pings_b200/csrc/pg_imgops.cuh)."""
from __future__ import annotations

import torch

from . import _lib as L


def _stream(dev) -> int:
    return L.raw_stream(dev)


class _Depth2Normal(torch.autograd.Function):
    @staticmethod
    def forward(ctx, depth, mask, alpha, cx, cy, fx, fy):
        if not depth.is_cuda:
            raise RuntimeError("pings_b200.imgops: expected CUDA tensors (there is no CPU fallback)")
        lib = L.lib()
        dev = depth.device
        H, W = depth.shape[-2], depth.shape[-1]
        d = depth.detach().to(torch.float32).contiguous()
        m = mask.detach().contiguous().view(torch.uint8)
        a = None if alpha is None else alpha.detach().to(torch.float32).contiguous()
        out = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        with L.device_guard(dev):
            L.check(lib.pg_depth2normal_forward(H, W, cx, cy, fx, fy, d.data_ptr(), m.data_ptr(), L.ptr(a), out.data_ptr(),
                                                _stream(dev)), "depth2normal")
        ctx.save_for_backward(d, m, a if a is not None else d.new_empty(0))
        ctx.k = (H, W, cx, cy, fx, fy, depth.shape, alpha is not None)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        lib = L.lib()
        d, m, a = ctx.saved_tensors
        H, W, cx, cy, fx, fy, shape, has_alpha = ctx.k
        g = g.to(torch.float32).contiguous()
        dd = torch.empty(H, W, dtype=torch.float32, device=d.device)
        with L.device_guard(d.device):
            L.check(lib.pg_depth2normal_backward(H, W, cx, cy, fx, fy, d.data_ptr(), m.data_ptr(),
                                                 a.data_ptr() if has_alpha else None, g.data_ptr(), dd.data_ptr(),
                                                 _stream(d.device)), "depth2normal backward")
        return dd.reshape(shape), None, None, None, None, None, None


def depth2normal(depth, mask, camera, img_scale: int = 1, alpha=None):
    """point_utils.py:83-148: normal map in the camera frame [3,H,W] from depth [1,H,W] and the visibility mask
    [1,H,W]; with `alpha` the result is multiplied by it (GR:381-382 multiplies by the detached rendered alpha)."""
    cx = float(camera.prcppoint[0]) * camera.image_width / img_scale
    cy = float(camera.prcppoint[1]) * camera.image_height / img_scale
    return _Depth2Normal.apply(depth, mask, alpha, cx, cy, float(camera.fx) / img_scale, float(camera.fy) / img_scale)


class _ExposureAffine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, mat, offset):
        lib = L.lib()
        dev = image.device
        H, W = image.shape[-2], image.shape[-1]
        img = image.detach().to(torch.float32).contiguous()
        m = mat.detach().to(torch.float32).contiguous()
        o = offset.detach().to(torch.float32).reshape(3).contiguous()
        out = torch.empty(3, H, W, dtype=torch.float32, device=dev)
        with L.device_guard(dev):
            L.check(lib.pg_exposure_affine_forward(H, W, img.data_ptr(), m.data_ptr(), o.data_ptr(), out.data_ptr(),
                                                   _stream(dev)), "exposure")
        ctx.save_for_backward(img, m)
        ctx.k = (H, W, offset.shape)
        return out

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        lib = L.lib()
        img, m = ctx.saved_tensors
        H, W, oshape = ctx.k
        g = g.to(torch.float32).contiguous()
        d_img = torch.empty_like(img) if ctx.needs_input_grad[0] else None
        d12 = torch.empty(12, dtype=torch.float32, device=img.device)
        with L.device_guard(img.device):
            L.check(lib.pg_exposure_affine_backward(H, W, img.data_ptr(), m.data_ptr(), g.data_ptr(), L.ptr(d_img),
                                                    d12.data_ptr(), _stream(img.device)), "exposure backward")
        return d_img, d12[:9].view(3, 3), d12[9:].reshape(oshape)


def exposure_affine(image, mat, offset):
    """GR:449-458: image [3,H,W] -> mat @ rgb + offset per pixel (exposure_mat [3,3], exposure_offset [1,3])."""
    return _ExposureAffine.apply(image, mat, offset)
