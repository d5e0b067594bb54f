"""CPU: the SSIM oracle (oracle/ssim_cpu.py) against the golden vectors made with the reference's own
torch SSIM (fp64 values computed from fp32-rounded images)."""
import os

import numpy as np
import pytest
import torch

from oracle import ssim_cpu

Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ssim_small.npz"))


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_ssim_oracle_matches_reference(name):
    img1 = torch.from_numpy(Z[f"{name}.img1"]).double()
    img2 = torch.from_numpy(Z[f"{name}.img2"]).double()
    v, g = ssim_cpu.ssim_and_grad(img1, img2)
    # the golden was evaluated on the un-rounded fp64 images; fp32 rounding of the inputs moves it by ~1e-7
    assert abs(float(v) - float(Z[f"{name}.value"])) < 1e-6
    np.testing.assert_allclose(g.numpy(), Z[f"{name}.grad"], rtol=1e-3, atol=1e-7)
