"""TEST INFRASTRUCTURE — CPU restatement of SSIM with an 11x11 Gaussian window (sigma 1.5), zero
"same" padding, C1 = 0.01^2, C2 = 0.03^2: the formulation the reference's own test pins fused-ssim
to (submodules/fused-ssim/tests/test.py:14-54) and that lives in-tree at
gaussian_splatting/utils/loss_utils.py:182-219.  Separable fp64-capable torch ops; gradient by
autograd.  Pinned against tests/golden/ssim_small.npz (made by running the reference's
loss_utils.ssim, tests/golden/make_ssim_golden.py)."""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def gauss11(dtype=torch.float32) -> torch.Tensor:
    g = torch.tensor([math.exp(-((x - 5) ** 2) / (2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float64)
    return (g / g.sum()).to(dtype)


def _blur(x: torch.Tensor, g: torch.Tensor) -> torch.Tensor:
    C = x.shape[1]
    kx = g.view(1, 1, 1, 11).expand(C, 1, 1, 11)
    ky = g.view(1, 1, 11, 1).expand(C, 1, 11, 1)
    return F.conv2d(F.conv2d(x, kx, padding=(0, 5), groups=C), ky, padding=(5, 0), groups=C)


def ssim_map(img1: torch.Tensor, img2: torch.Tensor) -> torch.Tensor:
    g = gauss11(img1.dtype)
    mu1, mu2 = _blur(img1, g), _blur(img2, g)
    s1 = _blur(img1 * img1, g) - mu1 * mu1
    s2 = _blur(img2 * img2, g) - mu2 * mu2
    s12 = _blur(img1 * img2, g) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    return ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))


def ssim(img1, img2, padding="same"):
    m = ssim_map(img1, img2)
    if padding == "valid":
        m = m[:, :, 5:-5, 5:-5]
    return m.mean()


def ssim_and_grad(img1, img2, padding="same"):
    x = img1.detach().clone().requires_grad_(True)
    v = ssim(x, img2, padding)
    v.backward()
    return v.detach(), x.grad
