"""Fused photometric + depth loss (pings_b200.losses) against the reference's eager formulation
(utils/mapper.py:1239,1252-1267 with utils/loss-style l1 = mean |a - b|).  fp32 sums in a different
order: values within 1e-5 relative, gradients within 1e-6 absolute."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(rgb, gt, depth, gt_depth, alpha, lam, dmin, dmax, amin, inverse):
    loss = (rgb - gt).abs().mean()
    l_d = torch.zeros((), device=rgb.device)
    if depth is not None:
        valid = (gt_depth > dmin) & (gt_depth < dmax)
        if alpha is not None:
            valid = valid & (alpha.detach() > amin)
        g, r = gt_depth[valid], depth[valid]
        if g.numel():
            l_d = (1.0 / g - 1.0 / r).abs().mean() if inverse else (g - r).abs().mean()
        loss = loss + lam * l_d
    return loss, l_d


@pytest.mark.parametrize("inverse", [False, True])
@pytest.mark.parametrize("with_alpha", [False, True])
def test_photo_depth_loss_matches_eager(inverse, with_alpha):
    from pings_b200.losses import photo_depth_loss
    H, W = 370, 1226
    g = torch.Generator(device="cuda").manual_seed(3)
    rgb = torch.rand(3, H, W, device="cuda", generator=g, requires_grad=True)
    gt = torch.rand(3, H, W, device="cuda", generator=g)
    gt[:, :5] = rgb.detach()[:, :5]                      # exact zeros: sign(0) = 0
    depth = (torch.rand(1, H, W, device="cuda", generator=g) * 60 + 0.5).requires_grad_(True)
    gt_depth = torch.rand(1, H, W, device="cuda", generator=g) * 90
    alpha = torch.rand(1, H, W, device="cuda", generator=g) if with_alpha else None
    args = (0.3, 1.0, 70.0, 0.4, inverse)
    loss, l_rgb, l_d = photo_depth_loss(rgb, gt, depth, gt_depth, alpha, *args)
    (loss * 1.7).backward()
    rgb2, depth2 = rgb.detach().clone().requires_grad_(True), depth.detach().clone().requires_grad_(True)
    ref, ref_d = _ref(rgb2, gt, depth2, gt_depth, alpha, *args)
    (ref * 1.7).backward()
    assert abs(float(loss) - float(ref)) <= 1e-5 * abs(float(ref))
    assert abs(float(l_d) - float(ref_d)) <= 1e-5 * abs(float(ref_d)) + 1e-9
    assert float((rgb.grad - rgb2.grad).abs().max()) <= 1e-9
    assert float((depth.grad - depth2.grad).abs().max()) <= 1e-6 * float(depth2.grad.abs().max()) + 1e-12


def test_rgb_only_and_empty_mask():
    from pings_b200.losses import photo_depth_loss
    rgb = torch.rand(3, 33, 47, device="cuda", requires_grad=True)
    gt = torch.rand(3, 33, 47, device="cuda")
    loss, l_rgb, l_d = photo_depth_loss(rgb, gt)
    assert abs(float(loss) - float((rgb - gt).abs().mean())) < 1e-6 and float(l_d) == 0.0
    depth = torch.rand(1, 33, 47, device="cuda", requires_grad=True)
    loss, _, l_d = photo_depth_loss(rgb, gt, depth, torch.zeros(1, 33, 47, device="cuda"), None, 0.5, 1.0, 2.0)
    loss.backward()
    assert float(l_d) == 0.0 and float(depth.grad.abs().max()) == 0.0
