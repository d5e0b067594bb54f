"""GPU parity of fused SSIM (through the C-ABI) against (a) the reference's own CUDA extension compiled
for sm_100a (oracle/_ref/fused_ssim_cuda_ref.so), (b) the fp64 CPU oracle (conv2d formulation the
reference's test pins to, FS/tests/test.py:82-91), (c) the golden vectors.  Bar: torch.isclose defaults
(rtol 1e-5, atol 1e-8) on the scalar like the reference's test; maps/gradients rtol 1e-4 atol 1e-6
(the reference is itself a --use_fast_math build)."""
import os

import numpy as np
import pytest
import torch

from oracle import ssim_cpu

pytestmark = pytest.mark.gpu
Z = np.load(os.path.join(os.path.dirname(__file__), "golden", "ssim_small.npz"))


def _pair(shape, seed):
    g = torch.Generator().manual_seed(seed)
    a = torch.rand(shape, generator=g)
    b = (a + 0.2 * torch.randn(shape, generator=g)).clamp(0, 1)
    return a.cuda(), b.cuda()


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_golden(name):
    from pings_b200.fused_ssim import fused_ssim
    img1 = torch.from_numpy(Z[f"{name}.img1"]).cuda().requires_grad_(True)
    img2 = torch.from_numpy(Z[f"{name}.img2"]).cuda()
    v = fused_ssim(img1, img2)
    v.backward()
    assert abs(float(v) - float(Z[f"{name}.value"])) < 2e-6
    np.testing.assert_allclose(img1.grad.cpu().numpy(), Z[f"{name}.grad"], rtol=2e-3, atol=1e-7)


@pytest.mark.parametrize("shape,padding", [((1, 3, 370, 1226), "same"), ((2, 5, 135, 241), "same"),
                                           ((1, 3, 129, 200), "valid"), ((1, 1, 7, 9), "same")])
def test_vs_reference_cuda_and_cpu_oracle(shape, padding):
    from oracle.ref_api import ref_ssim_module
    from pings_b200.fused_ssim import fused_ssim, fusedssim, fusedssim_backward
    img1, img2 = _pair(shape, 3)
    x = img1.clone().requires_grad_(True)
    v = fused_ssim(x, img2, padding=padding)
    v.backward()
    vo, go = ssim_cpu.ssim_and_grad(img1.cpu().double(), img2.cpu().double(), padding)
    assert torch.isclose(v.cpu().double(), vo, rtol=1e-5, atol=1e-8)
    scale = go.abs().max()
    assert ((x.grad.cpu().double() - go).abs().max() / scale) < 1e-4
    ref = ref_ssim_module()
    if ref is None:
        pytest.fail("oracle/_ref/fused_ssim_cuda_ref.so missing")
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m, a, b, c = fusedssim(C1, C2, img1, img2, True)
    mr, ar, br, cr = ref.fusedssim(C1, C2, img1, img2, True)
    torch.testing.assert_close(m, mr, rtol=1e-4, atol=1e-6)
    for t, tr in ((a, ar), (b, br), (c, cr)):
        assert ((t - tr).abs().max() / tr.abs().max()) < 1e-4
    dmap = torch.randn(shape, generator=torch.Generator().manual_seed(5)).cuda()
    g = fusedssim_backward(C1, C2, img1, img2, dmap, a, b, c)
    gr = ref.fusedssim_backward(C1, C2, img1, img2, dmap, ar, br, cr)
    assert ((g - gr).abs().max() / gr.abs().max()) < 1e-4
    # inference mode returns empty partials like the reference (FS/ssim.cu:458-460)
    m2, a2, _, _ = fusedssim(C1, C2, img1, img2, False)
    assert a2.numel() == 0 and torch.equal(m2, m)


def test_properties_full_size():
    """1080p: ssim(x, x) == 1 everywhere; symmetric in its arguments; gradient of ssim(x,x) w.r.t. x ~ 0."""
    from pings_b200.fused_ssim import fused_ssim, fusedssim
    img1, img2 = _pair((1, 3, 1080, 1920), 9)
    m, *_ = fusedssim(1e-4, 9e-4, img1, img1, False)
    assert float((m - 1).abs().max()) < 1e-5
    assert torch.isclose(fused_ssim(img1, img2, train=False), fused_ssim(img2, img1, train=False), rtol=1e-6, atol=1e-8)
    # non-contiguous rows slice as in utils/mapper.py:1227-1243
    a = img1[0][:, 0:-1, :]
    b = img2[0][:, 0:-1, :]
    v = fused_ssim(a[None], b[None])
    vo = ssim_cpu.ssim(a[None].cpu().double(), b[None].cpu().double())
    assert torch.isclose(v.cpu().double(), vo, rtol=1e-5, atol=1e-8)
