"""Generates tests/golden/ssim_small.npz by running the reference's own torch SSIM
(/root/reference/gaussian_splatting/utils/loss_utils.py:189-219) on CPU. Build container only."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_python as rp  # noqa: E402

rp._stub()
sys.path.insert(0, rp.REF)
from gaussian_splatting.utils.loss_utils import ssim as ref_ssim  # noqa: E402

g = torch.Generator().manual_seed(0)
out = {}
for name, shape in [("a", (2, 3, 40, 52)), ("b", (1, 1, 11, 70)), ("c", (1, 3, 37, 33))]:
    img1 = torch.rand(shape, generator=g, dtype=torch.float64).requires_grad_(True)
    img2 = (img1.detach() + 0.2 * torch.randn(shape, generator=g, dtype=torch.float64)).clamp(0, 1)
    v = ref_ssim(img1, img2)
    v.backward()
    out.update({f"{name}.img1": img1.detach().float().numpy(), f"{name}.img2": img2.float().numpy(),
                f"{name}.value": v.detach().numpy(), f"{name}.grad": img1.grad.numpy()})
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "ssim_small.npz"), **out)
print({k: v.shape for k, v in out.items()})
