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


# ---- depth2normal / exposure kernels (pings_b200/imgops.py) ------------------------------------------------------
def _d2n_restatement(depth, mask, prcp, iw, ih, fx, fy, img_scale):
    """TEST ORACLE: eager restatement of gaussian_splatting/utils/point_utils.py:83-148 (the numeric pin against the
    reference itself is tests/golden/render_post.npz; this one adds autograd for the backward check)."""
    camD = depth.permute([1, 2, 0])
    m = mask.permute([1, 2, 0])
    H, W, _ = camD.shape
    h, w = torch.meshgrid(torch.arange(H, device=depth.device, dtype=torch.float32),
                          torch.arange(W, device=depth.device, dtype=torch.float32), indexing="ij")
    p = torch.stack([w - prcp[0] * iw / img_scale, h - prcp[1] * ih / img_scale], dim=-1) * camD
    p = p / torch.tensor([fx / img_scale, fy / img_scale], device=depth.device)
    cam = torch.cat([p, camD], -1)
    pp = torch.nn.functional.pad(cam[None], [0, 0, 1, 1, 1, 1], mode="replicate")
    mm = torch.nn.functional.pad(m[None].float(), [0, 0, 1, 1, 1, 1], mode="replicate")
    pc = pp[:, 1:-1, 1:-1] * mm[:, 1:-1, 1:-1]
    pu = (pp[:, :-2, 1:-1] - pc) * mm[:, :-2, 1:-1]
    pl = (pp[:, 1:-1, :-2] - pc) * mm[:, 1:-1, :-2]
    pb = (pp[:, 2:, 1:-1] - pc) * mm[:, 2:, 1:-1]
    pr = (pp[:, 1:-1, 2:] - pc) * mm[:, 1:-1, 2:]
    cr = torch.linalg.cross
    n = (cr(pu, pl) + cr(pr, pu) + cr(pb, pr) + cr(pl, pb))[0]
    n = torch.nn.functional.normalize(n, dim=-1)
    return (n * mm[0, 1:-1, 1:-1]).permute([2, 0, 1])


@pytest.mark.parametrize("H,W,scale", [(64, 96, 1), (37, 53, 2)])
def test_depth2normal_and_exposure_kernels(H, W, scale):
    from types import SimpleNamespace
    from pings_b200.imgops import depth2normal, exposure_affine
    g = torch.Generator().manual_seed(5)
    depth = (torch.rand(1, H, W, generator=g) * 20 + 1).cuda().requires_grad_(True)
    alpha = torch.rand(1, H, W, generator=g)
    alpha[:, : H // 4] *= 5e-4
    alpha = alpha.cuda()
    mask = alpha > 1e-3
    cam = SimpleNamespace(prcppoint=torch.tensor([0.48, 0.53]), image_width=W * scale, image_height=H * scale, fx=80.0, fy=75.0)
    out = depth2normal(depth, mask, cam, img_scale=scale, alpha=alpha)
    d2 = depth.detach().clone().requires_grad_(True)
    ref = _d2n_restatement(d2, mask, [0.48, 0.53], W * scale, H * scale, 80.0, 75.0, scale) * alpha
    bad = (out - ref).abs().amax(0) > 1e-4
    assert int(bad.sum()) == 0, (int(bad.sum()), torch.nonzero(bad)[:8].tolist(), out[:, bad][:, :4].tolist(), ref[:, bad][:, :4].tolist())
    torch.testing.assert_close(out, ref, rtol=1e-4, atol=5e-5)   # components near zero come from cancelling cross products
    w = torch.randn(3, H, W, generator=g).cuda()
    (out * w).sum().backward()
    (ref * w).sum().backward()
    # Pixels whose un-normalised normal is EXACTLY zero (a replicated border difference + a masked neighbour) are
    # excluded: F.normalize's eps turns the cotangent into g / 1e-12 there, the eager graph then adds +X and -X
    # (X ~ 1e12) into that pixel's depth gradient and whatever else was accumulated is absorbed - its value (0.0 on
    # CPU) is an artefact.  Our kernel drops the analytically cancelling pair and keeps the neighbours' contributions.
    with torch.no_grad():
        r0 = _d2n_restatement(d2.detach(), mask, [0.48, 0.53], W * scale, H * scale, 80.0, 75.0, scale)
        degenerate = (r0.abs().amax(0) == 0) & mask[0]
    keep = ~degenerate[None]
    assert int(degenerate.sum()) <= 4
    gmax = float(d2.grad[keep].abs().max())
    torch.testing.assert_close(depth.grad[keep], d2.grad[keep], rtol=1e-3, atol=1e-4 * gmax)
    assert bool(torch.isfinite(depth.grad).all()) and float(depth.grad[~keep].abs().max() if int(degenerate.sum()) else 0.0) <= gmax
    # affine exposure
    img = torch.rand(3, H, W, generator=g).cuda().requires_grad_(True)
    M = (torch.eye(3) + 0.1 * torch.randn(3, 3, generator=g)).cuda().requires_grad_(True)
    off = (0.05 * torch.randn(1, 3, generator=g)).cuda().requires_grad_(True)
    o1 = exposure_affine(img, M, off)
    img2, M2, off2 = (t.detach().clone().requires_grad_(True) for t in (img, M, off))
    o2 = (img2.permute(1, 2, 0).reshape(-1, 3) @ M2.T + off2).view(H, W, 3).permute(2, 0, 1)
    torch.testing.assert_close(o1, o2, rtol=1e-5, atol=1e-6)
    (o1 * w).sum().backward()
    (o2 * w).sum().backward()
    torch.testing.assert_close(img.grad, img2.grad, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(M.grad, M2.grad, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(off.grad, off2.grad, rtol=1e-4, atol=1e-3)
