"""Drop-in for the reference package `fused_ssim` (submodules/fused-ssim/fused_ssim/__init__.py:8-42):
`fused_ssim(img1, img2, padding="same", train=True)` -> scalar, gradient w.r.t. img1 only.
The kernels are the sm_100a ones behind the C-ABI (pg_ssim_forward / pg_ssim_backward)."""
import ctypes as C  # noqa: F401

import torch

from .. import _lib as L

allowed_padding = ["same", "valid"]


def _stream(dev):
    return L.raw_stream(dev)


def fusedssim(C1, C2, img1, img2, train):
    """FS/ssim.cu:434-477: -> (ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12); the partials are empty when !train."""
    if not img1.is_cuda:
        raise RuntimeError("pings_b200.fused_ssim: expected CUDA tensors (there is no CPU fallback)")
    img1 = img1.detach().to(torch.float32).contiguous()
    img2 = img2.detach().to(torch.float32).contiguous()
    B, CH, H, W = img1.shape
    ssim_map = torch.empty_like(img1)          # the kernel writes every element (the reference zero-fills first)
    if train:
        p = [torch.empty_like(img1) for _ in range(3)]
    else:
        p = [torch.empty(0, device=img1.device) for _ in range(3)]
    with L.device_guard(img1.device):
        L.check(L.lib().pg_ssim_forward(B, CH, H, W, C1, C2, img1.data_ptr(), img2.data_ptr(), ssim_map.data_ptr(),
                                        L.ptr(p[0]), L.ptr(p[1]), L.ptr(p[2]), _stream(img1.device)), "ssim forward")
    return ssim_map, p[0], p[1], p[2]


def fusedssim_backward(C1, C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12):
    """FS/ssim.cu:482-516."""
    img1 = img1.detach().to(torch.float32).contiguous()
    img2 = img2.detach().to(torch.float32).contiguous()
    dL_dmap = dL_dmap.to(torch.float32).contiguous()
    B, CH, H, W = img1.shape
    out = torch.empty_like(img1)
    with L.device_guard(img1.device):
        L.check(L.lib().pg_ssim_backward(B, CH, H, W, C1, C2, img1.data_ptr(), img2.data_ptr(), dL_dmap.data_ptr(), 0.0,
                                         dm_dmu1.data_ptr(), dm_dsigma1_sq.data_ptr(), dm_dsigma12.data_ptr(),
                                         out.data_ptr(), _stream(img1.device)), "ssim backward")
    return out


class FusedSSIMMap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, C1, C2, img1, img2, padding="same", train=True):
        ssim_map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = fusedssim(C1, C2, img1, img2, train)
        if padding == "valid":
            ssim_map = ssim_map[:, :, 5:-5, 5:-5]
        ctx.save_for_backward(img1.detach(), img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        ctx.C1, ctx.C2, ctx.padding = C1, C2, padding
        return ssim_map

    @staticmethod
    def backward(ctx, opt_grad):
        img1, img2, dm_dmu1, dm_dsigma1_sq, dm_dsigma12 = ctx.saved_tensors
        dL_dmap = opt_grad
        if ctx.padding == "valid":
            dL_dmap = torch.zeros_like(img1)
            dL_dmap[:, :, 5:-5, 5:-5] = opt_grad
        grad = fusedssim_backward(ctx.C1, ctx.C2, img1, img2, dL_dmap, dm_dmu1, dm_dsigma1_sq, dm_dsigma12)
        return None, None, grad, None, None, None


class _FusedSSIMMean(torch.autograd.Function):
    """fused_ssim(..., padding="same") in one piece: the `.mean()` is reduced inside the forward kernel (the map is
    never written) and the backward takes the upstream scalar on the device - same value and gradient as
    FusedSSIMMap(...).mean(), 8 B/element less traffic and no expand/contiguous of the mean's gradient."""

    @staticmethod
    def forward(ctx, C1, C2, img1, img2, train):
        if not img1.is_cuda:
            raise RuntimeError("pings_b200.fused_ssim: expected CUDA tensors (there is no CPU fallback)")
        a = img1.detach().to(torch.float32).contiguous()
        b = img2.detach().to(torch.float32).contiguous()
        B, CH, H, W = a.shape
        need = bool(train) and ctx.needs_input_grad[2]
        p = [torch.empty_like(a) for _ in range(3)] if need else [None] * 3
        scratch = torch.empty(1, dtype=torch.float64, device=a.device)
        mean = torch.empty((), dtype=torch.float32, device=a.device)
        with L.device_guard(a.device):
            L.check(L.lib().pg_ssim_mean_forward(B, CH, H, W, C1, C2, a.data_ptr(), b.data_ptr(), L.ptr(p[0]), L.ptr(p[1]),
                                                 L.ptr(p[2]), scratch.data_ptr(), mean.data_ptr(), _stream(a.device)),
                    "ssim mean forward")
        if need:
            ctx.save_for_backward(a, b, *p)
        return mean

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, g):
        a, b, p0, p1, p2 = ctx.saved_tensors
        B, CH, H, W = a.shape
        up = g.detach().to(torch.float32).reshape(1).contiguous()
        out = torch.empty_like(a)
        with L.device_guard(a.device):
            L.check(L.lib().pg_ssim_mean_backward(B, CH, H, W, a.data_ptr(), b.data_ptr(), up.data_ptr(), p0.data_ptr(),
                                                  p1.data_ptr(), p2.data_ptr(), out.data_ptr(), _stream(a.device)),
                    "ssim mean backward")
        return None, None, out, None, None


def fused_ssim(img1, img2, padding="same", train=True):
    C1 = 0.01 ** 2
    C2 = 0.03 ** 2
    assert padding in allowed_padding
    img1 = img1.contiguous()
    if padding == "same" and img1.numel() > 0 and img1.dim() == 4:
        return _FusedSSIMMean.apply(C1, C2, img1, img2, train)
    map = FusedSSIMMap.apply(C1, C2, img1, img2, padding, train)
    return map.mean()
