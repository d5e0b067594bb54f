"""CPU checks of the fp64 surfel arbiter (oracle/cpu_rasters64.c) itself.

The arbiter decides GPU parity rows that the noisy reference backward cannot (tests/test_rasters_gpu.py), so
its calculus is checked here against central finite differences of its OWN fp64 forward.  The reference's
backward is the true gradient of its forward wherever the forward is smooth and no quirk applies, i.e. with
the two blend discontinuities (alpha < 1/255 skip, T < 1e-4 stop) switched off and the non-surface config
(no x10 normal term q3, no unweighted per-pixel-depth term q4, no z-scale rule q2/q5); with a centred
principal point the pose Jacobian (RS backward.cu:487-548) is exact as well.
"""
import pytest
import torch

from pings_b200 import synth


def _tiny_scene(P=40, W=64, H=48, seed=4):
    cam = synth.hd_camera(seed, W, H)   # centred principal point
    cam = synth.Camera(W, H, 60.0, 60.0, W / 2.0, H / 2.0, cam.T_wc)
    sc = synth.gaussians_in_view(cam, P, seed, surfel=False, scale_lo=0.2, scale_hi=1.5)
    sc["opacities"] = sc["opacities"] * 0.85          # stay below the 0.99 alpha clamp
    g = torch.Generator().manual_seed(seed + 9)
    ct = dict(color=torch.randn(3, H, W, generator=g), normal=torch.randn(3, H, W, generator=g),
              depth=torch.randn(1, H, W, generator=g) * 0.1, opac=torch.randn(1, H, W, generator=g))
    return cam, sc, ct


def _lists(cam, sc, W, H):
    """Every tile gets every in-front Gaussian, sorted by view depth: the lists are INPUTS of the arbiter."""
    view = cam.matrices()[0].double()
    z = (torch.cat([sc["means3D"].double(), torch.ones(sc["means3D"].shape[0], 1, dtype=torch.float64)], 1) @ view)[:, 2]
    vis = z > 0.5
    order = torch.argsort(z[vis])
    ids = torch.nonzero(vis).flatten()[order].to(torch.int32)
    T = ((W + 15) // 16) * ((H + 15) // 16)
    n = ids.numel()
    pl = ids.repeat(T)
    rg = torch.stack([torch.arange(T) * n, (torch.arange(T) + 1) * n], 1).to(torch.int32)
    return vis.to(torch.int32), pl, rg


def _loss(out, ct):
    return sum(float((out[k].double() * ct[k].double()).sum()) for k in ("color", "normal", "depth", "opac"))


@pytest.mark.parametrize("cfg", [[0, 1, 0, 1, 0], [0, 0, 0, 1, 0]])
def test_fp64_backward_is_the_gradient_of_the_fp64_forward(cfg):
    from oracle import cpu_rasters64 as O
    W, H = 64, 48
    cam, sc, ct = _tiny_scene(W=W, H=H)
    radii, pl, rg = _lists(cam, sc, W, H)
    config = torch.tensor(cfg, dtype=torch.float32)
    bg = torch.tensor([0.3, 0.6, 0.1])
    mats = cam.matrices()
    L = O.lib()
    L.rs64_set_thresholds(0.0, 0.0)
    try:
        def fwd(scene, m=mats):
            return O.blend(scene, m, W, H, cam.tanfovx, cam.tanfovy, bg, config, radii, pl, rg, ct)
        base = fwd(sc)
        g = O.chain(sc, mats, W, H, cam.tanfovx, cam.tanfovy, config, radii, base["g_means2D"], base["g_conic"],
                    base["g_normal"], base["g_depths"])
        analytic = dict(means3D=g["means3D"], scales=g["scales"], rotations=g["rotations"],
                        opacities=base["g_opacity"], colors_precomp=base["g_colors"])
        ids = torch.nonzero(radii).flatten()[:6].tolist()
        for name, rel_h in (("means3D", 2e-4), ("scales", 1e-3), ("rotations", 1e-3), ("opacities", 1e-3),
                            ("colors_precomp", 1e-3)):
            A = analytic[name]
            scale = float(A.abs().max())
            for i in ids:
                for k in range(sc[name].shape[1]):
                    x0 = float(sc[name][i, k])
                    h = rel_h * max(1.0, abs(x0))
                    plus, minus = {**sc, name: sc[name].clone()}, {**sc, name: sc[name].clone()}
                    plus[name][i, k] = x0 + h
                    minus[name][i, k] = x0 - h
                    step = float(plus[name][i, k].double() - minus[name][i, k].double())   # the step actually taken in fp32
                    fd = (_loss(fwd(plus), ct) - _loss(fwd(minus), ct)) / step
                    assert abs(fd - float(A[i, k])) <= 2e-4 * scale + 2e-3 * abs(fd), (name, i, k, fd, float(A[i, k]))
        # pose: T_cw <- SE3_exp(tau) T_cw (utils/campose_utils.py:79-98); analytic = sum over surfels of dL_dtau
        tau_sum = g["tau"].sum(0)
        T_cw = torch.linalg.inv(cam.T_wc)
        for k in range(6):
            h = 2e-4
            vals = []
            for s in (+1, -1):
                tau = torch.zeros(6, dtype=torch.float64)
                tau[k] = s * h
                cam2 = synth.Camera(W, H, cam.fx, cam.fy, cam.cx, cam.cy, torch.linalg.inv(synth.se3_exp(tau) @ T_cw))
                vals.append(_loss(fwd(sc, cam2.matrices()), ct))
            fd = (vals[0] - vals[1]) / (2 * h)
            assert abs(fd - float(tau_sum[k])) <= 2e-3 * float(tau_sum.abs().max()) + 5e-3 * abs(fd), (k, fd, float(tau_sum[k]))
    finally:
        L.rs64_set_thresholds(1.0 / 255.0, 0.0001)


def test_surface_quirks_are_kept():
    """q2/q5: dL_dscale.z == 0 for surfaces; q3: the normal blend sum carries the x10; q4: the per-pixel-depth
    term only changes the depth image."""
    from oracle import cpu_rasters64 as O
    W, H = 64, 48
    cam, sc, ct = _tiny_scene(W=W, H=H, seed=6)
    sc = synth.surfels_in_view(cam, 40, 6, scale_lo=0.2, scale_hi=1.5)
    radii, pl, rg = _lists(cam, sc, W, H)
    bg = torch.zeros(3)
    mats = cam.matrices()
    on = O.blend(sc, mats, W, H, cam.tanfovx, cam.tanfovy, bg, torch.tensor([1., 1, 1, 1, 0]), radii, pl, rg, ct)
    off = O.blend(sc, mats, W, H, cam.tanfovx, cam.tanfovy, bg, torch.tensor([1., 1, 0, 1, 0]), radii, pl, rg, ct)
    g = O.chain(sc, mats, W, H, cam.tanfovx, cam.tanfovy, torch.tensor([1., 1, 1, 1, 0]), radii, on["g_means2D"],
                on["g_conic"], on["g_normal"], on["g_depths"])
    assert float(g["scales"][:, 2].abs().max()) == 0.0
    assert float(g["scales"][:, :2].abs().max()) > 0.0
    # x10: with only the normal cotangent, dL_dnormal = 10 * sum_pix alpha T dL_dpixN; the colour channel has the
    # same weights, so swapping the roles of the cotangents reproduces the sum without the factor
    zc, zd = torch.zeros_like(ct["color"]), torch.zeros_like(ct["depth"])
    ct_n = dict(color=zc, normal=ct["normal"], depth=zd, opac=zd)
    ct_c = dict(color=ct["normal"], normal=zc, depth=zd, opac=zd)
    cfg = torch.tensor([1., 1, 0, 1, 0])
    a = O.blend(sc, mats, W, H, cam.tanfovx, cam.tanfovy, bg, cfg, radii, pl, rg, ct_n)
    b = O.blend(sc, mats, W, H, cam.tanfovx, cam.tanfovy, bg, cfg, radii, pl, rg, ct_c)
    torch.testing.assert_close(a["g_normal"], 10.0 * b["g_colors"], rtol=1e-12, atol=1e-15)
    torch.testing.assert_close(on["color"], off["color"], rtol=0, atol=0)
    assert float((on["depth"] - off["depth"]).abs().max()) > 0
