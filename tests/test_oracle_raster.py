"""CPU checks of the plain-C raster restatement (oracle/cpu_raster.c): structural properties of the tile
lists and a finite-difference check of its analytic gradients.  (Its agreement with the reference CUDA
build is checked on the GPU box: tests/test_raster3_gpu.py::test_cpu_oracle_vs_reference_and_ours.)"""
import torch

from oracle.cpu_raster import CpuRaster3
from pings_b200 import synth


def _scene(P, W, H, seed, **kw):
    cam = synth.kitti_camera(seed, W, H)
    sc = synth.gaussians_in_view(cam, P, seed, **kw)
    view, full, proj, center = cam.matrices()
    return cam, sc, (view, full, proj)


def test_tile_lists_are_a_stable_depth_sorted_partition():
    P, W, H = 3000, 200, 120
    cam, sc, mats = _scene(P, W, H, 2)
    cr = CpuRaster3()
    r = cr.forward(sc, mats, W, H, cam.tanfovx, cam.tanfovy, torch.zeros(3))
    R, pl, rg = r["num_rendered"], r["point_list"].long(), r["ranges"].long()
    assert R == int(pl.numel()) and R > P
    depth = (torch.cat([sc["means3D"], torch.ones(P, 1)], 1) @ mats[0])[:, 2]
    covered = 0
    for t in range(rg.shape[0]):
        a, b = int(rg[t, 0]), int(rg[t, 1])
        if a == b:
            continue
        covered += b - a
        ids = pl[a:b]
        d = depth[ids]
        assert bool((d[1:] >= d[:-1] - 1e-4).all())          # front to back
        assert int(ids.unique().numel()) == b - a             # a Gaussian appears once per tile
        assert bool((r["radii"][ids] > 0).all())
    assert covered == R                                        # ranges partition the list
    # opacity + transmittance bookkeeping
    assert float(r["opacity"].min()) >= 0.0 and float(r["opacity"].max()) <= 1.0
    assert int(r["n_contrib"].max()) <= int((rg[:, 1] - rg[:, 0]).max())


def test_analytic_gradients_match_finite_differences():
    W, H, P = 96, 64, 150
    cam = synth.kitti_camera(4, W, H)
    sc = synth.gaussians_in_view(cam, P, 4, scale_lo=0.4, scale_hi=1.2)
    sc["opacities"] = sc["opacities"] * 0.15 + 0.02   # smooth scene: no saturation, few 1/255 cut-offs
    g = torch.Generator().manual_seed(0)
    wc, wd = torch.randn(3, H, W, generator=g), torch.randn(1, H, W, generator=g) * 0.05
    bg = torch.zeros(3)

    def loss_for(tau, scene):
        c = synth.Camera(W, H, cam.fx, cam.fy, cam.cx, cam.cy, torch.linalg.inv(synth.se3_exp(tau) @ torch.linalg.inv(cam.T_wc)))
        view, full, proj, _ = c.matrices()
        cr = CpuRaster3()
        r = cr.forward(scene, (view, full, proj), W, H, c.tanfovx, c.tanfovy, bg, want_views=False)
        L = float((r["color"].double() * wc).sum() + (r["depth"].double() * wd).sum())
        return L, cr, r

    L0, cr, r = loss_for(torch.zeros(6), sc)
    gr = cr.backward(r["radii"], wc, wd)
    ana = gr["tau"].double().sum(0)
    eps = 2e-3
    num = torch.zeros(6, dtype=torch.float64)
    for k in range(6):
        d = torch.zeros(6)
        d[k] = eps
        num[k] = (loss_for(d, sc)[0] - loss_for(-d, sc)[0]) / (2 * eps)
    cos = float((ana * num).sum() / (ana.norm() * num.norm()))
    assert cos > 0.98, (cos, ana, num)
    # dL/dcolors is exact up to fp32 (the loss is linear in the colours)
    dirc = torch.randn(P, 3, generator=g)
    sc2 = dict(sc)
    sc2["colors_precomp"] = sc["colors_precomp"] + 1e-2 * dirc
    fd = (loss_for(torch.zeros(6), sc2)[0] - L0) / 1e-2
    an = float((gr["colors"].double() * dirc).sum())
    assert abs(fd - an) <= 2e-3 * max(1.0, abs(an)), (fd, an)
    # dL/dmeans3D and dL/dscales along random directions.  Opacities in [0.2, 0.5]: with fainter splats the
    # 1/255 alpha cut-off runs through the bright part of the footprint and its jump dominates the
    # finite difference (the analytic gradient, like the reference's, ignores that discontinuity).
    sc3 = dict(sc)
    sc3["opacities"] = (sc["opacities"] - 0.02) / 0.15 * 0.3 + 0.2
    L3, cr3, r3 = loss_for(torch.zeros(6), sc3)
    gr3 = cr3.backward(r3["radii"], wc, wd)
    h = 2e-3
    for name, scale_by in (("means3D", None), ("scales", "scales")):
        d = torch.randn(sc3[name].shape, generator=g)
        if scale_by:
            d = d * sc3[scale_by]
        scp, scm = dict(sc3), dict(sc3)
        scp[name] = sc3[name] + h * d
        scm[name] = sc3[name] - h * d
        fd = (loss_for(torch.zeros(6), scp)[0] - loss_for(torch.zeros(6), scm)[0]) / (2 * h)
        an = float((gr3[name].double() * d).sum())
        scale = float((gr3[name].double().abs() * d.abs()).sum())   # size of the terms that are summed
        assert abs(fd - an) <= 0.05 * abs(an) + 0.01 * scale, (name, fd, an, scale)
