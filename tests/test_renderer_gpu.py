"""GPU tests of the render front-end mirror (pings_b200/gaussian_renderer.py).

spawn_gaussians is compared with tests/golden/spawn_gs.npz, produced by the reference's own
spawn_gaussians + Decoder on CPU (tests/golden/make_spawn_golden.py).  Tolerances: fp32 MLPs with a
different summation order (ours: fused SIMT kernel, reference: oneDNN sgemm) -> rtol 1e-4 / atol 1e-5;
the alpha > 0 compaction must select exactly the same Gaussians.
render() is exercised end to end (both gs types) for its contract: keys, shapes, None rules, gradients.
"""
import math
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from pings_b200 import synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "spawn_gs.npz")
RTOL, ATOL = 1e-4, 1e-5


def _decoders(z, dev="cuda"):
    from pings_b200.decoder import Decoder
    from pings_b200.neural_points import default_config
    N, Fg, Fc, Hd, Lv, K = [int(x) for x in z["N"]]
    cfg = default_config(device=dev)
    decs = {"gauss_xyz": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_rot": Decoder(cfg, Fg, Hd, Lv, 4, K, 0),
            "gauss_scale": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_alpha": Decoder(cfg, Fg, Hd, Lv, 1, K, 0),
            "gauss_color": Decoder(cfg, Fc, Hd, Lv, 3, K, 3)}
    for name, d in decs.items():
        pre = f"dec.{name}."
        d.load_state_dict({k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)})
        d.to(dev)
    return decs


def _data(z, dev="cuda"):
    d = {k[5:]: torch.from_numpy(z[k]).to(dev) for k in z.files if k.startswith("data.")}
    d["resolution"] = 0.25
    d["geo_feature"].requires_grad_(True)
    d["color_feature"].requires_grad_(True)
    return d


def test_spawn_gaussians_matches_reference_golden():
    from pings_b200.gaussian_renderer import spawn_gaussians
    z = np.load(GOLD)
    decs, data = _decoders(z), _data(z)
    visible = torch.from_numpy(z["visible"]).cuda()
    out = spawn_gaussians(data, decs, visible, torch.from_numpy(z["cam_origin"]).cuda(), False, True,
                          learn_color_residual=True, gs_type="gaussian_surfel")
    assert out["local_view_gaussian_count"] == z["out.gaussian_xyz"].shape[0]
    assert torch.equal(out["gaussian_free_mask"].cpu(), torch.from_numpy(z["out.gaussian_free_mask"]))
    for k in ("gaussian_xyz", "gaussian_scale", "gaussian_rot", "gaussian_alpha", "gaussian_color", "alpha_all"):
        torch.testing.assert_close(out[k].cpu(), torch.from_numpy(z["out." + k]), rtol=RTOL, atol=ATOL)
    loss = sum((out[k] * torch.from_numpy(z["w." + k]).cuda()).sum()
               for k in ("gaussian_xyz", "gaussian_scale", "gaussian_rot", "gaussian_alpha", "gaussian_color"))
    loss.backward()
    for k in ("geo_feature", "color_feature"):
        ref = torch.from_numpy(z["grad." + k])
        torch.testing.assert_close(data[k].grad.cpu(), ref, rtol=1e-3, atol=1e-4 * float(ref.abs().max()))
    for name, d in decs.items():
        ref = torch.from_numpy(z[f"grad.{name}.lout.weight"])
        torch.testing.assert_close(d.lout.weight.grad.cpu(), ref, rtol=1e-3, atol=1e-4 * float(ref.abs().max()))
    # fewer than 10 spawning points -> None (GR:571-572)
    few = torch.zeros_like(visible)
    few[:5] = True
    assert spawn_gaussians(data, decs, few, None) is None


class _Cam(SimpleNamespace):
    def full_patch(self, down_rate=0):
        s = 2 ** down_rate
        return torch.tensor([0, 0, int(self.image_height / s) - 1, int(self.image_width / s) - 1], dtype=torch.float32,
                            device=self.device)


def _camera(W, H, seed, dev="cuda"):
    cam = synth.kitti_camera(seed, W, H)
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    return _Cam(device=torch.device(dev), uid=0, image_width=W, image_height=H, FoVx=2 * math.atan(cam.tanfovx),
                FoVy=2 * math.atan(cam.tanfovy), zfar=100.0, fx=cam.fx, fy=cam.fy,
                prcppoint=torch.tensor([cam.cx / W, cam.cy / H], device=dev), projection_matrix=proj,
                world_view_transform=view, full_proj_transform=full, camera_center=center,
                cam_rot_delta=torch.zeros(3, device=dev, requires_grad=True),
                cam_trans_delta=torch.zeros(3, device=dev, requires_grad=True),
                exposure_mat=(torch.eye(3, device=dev) * 1.1).requires_grad_(True),
                exposure_offset=torch.full((1, 3), 0.01, device=dev).requires_grad_(True),
                exposure_a=torch.zeros(1, device=dev), exposure_b=torch.zeros(1, device=dev)), cam


@pytest.mark.parametrize("gs_type", ["gaussian_surfel", "3d_gs"])
def test_render_contract(gs_type):
    from pings_b200.gaussian_renderer import render
    z = np.load(GOLD)
    decs, data = _decoders(z), _data(z)
    W, H = 320, 200
    vc, cam = _camera(W, H, 3)
    # put the neural points in front of the camera
    g = torch.Generator().manual_seed(1)
    N = data["position"].shape[0]
    pc = torch.stack([(torch.rand(N, generator=g) - 0.5) * 8, (torch.rand(N, generator=g) - 0.5) * 3,
                      torch.rand(N, generator=g) * 10 + 2, torch.ones(N)], 1).double()
    data["position"] = (cam.T_wc @ pc.T).T[:, :3].float().cuda()
    bg = torch.zeros(3, device="cuda")
    res = render(vc, None, data, decs, None, bg, view_concat_on=True, learn_color_residual=True, gs_type=gs_type,
                 d2n_on=True, max_scale_ratio=2.0, unit_scale_ratio=0.5)
    assert res is not None
    assert res["render"].shape == (3, H, W) and res["surf_depth"].shape == (1, H, W) and res["rend_alpha"].shape == (1, H, W)
    assert res["surf_normal"].shape == (3, H, W)
    G = res["gaussian_xyz"].shape[0]
    assert res["radii"].shape == (G,) and res["viewspace_points"].shape == (G, 3)
    assert bool(res["visibility_filter"].any()) and 0.0 < res["visible_neural_point_ratio"] <= 1.0
    if gs_type == "gaussian_surfel":
        assert res["rend_normal"].shape == (3, H, W) and res["contributions"].shape == (G,)
    else:
        assert res["rend_normal"] is None
        assert float(res["surf_depth"][res["rend_alpha"] <= 1e-3].abs().sum()) == 0.0   # GR:437
    assert float(res["rend_alpha"].max()) > 0.05
    loss = res["render"].mean() + 0.1 * res["surf_depth"].mean() + res["rend_alpha"].mean()
    loss.backward()
    for t in (data["geo_feature"], data["color_feature"], vc.cam_rot_delta, vc.cam_trans_delta, vc.exposure_mat,
              decs["gauss_xyz"].lout.weight, decs["gauss_color"].lout.weight):
        assert t.grad is not None and torch.isfinite(t.grad).all() and float(t.grad.abs().sum()) > 0
    assert res["viewspace_points"].grad is not None
    # None rules: inputs missing (GR:264-265), nothing visible (GR:221-224), cam_pose path recomputes the matrices
    assert render(vc, None, None, decs, None, bg, gs_type=gs_type) is None
    far = dict(data)
    far["position"] = 2 * vc.camera_center - data["position"]   # mirrored through the camera centre: all behind it
    assert render(vc, None, far, decs, None, bg, view_concat_on=True, gs_type=gs_type) is None
    with torch.no_grad():
        res2 = render(vc, cam.T_wc.float().cuda(), data, decs, None, bg, view_concat_on=True, learn_color_residual=True,
                      gs_type=gs_type, max_scale_ratio=2.0, unit_scale_ratio=0.5)
    # the fp32 inverse of the pose differs in the last bits from the precomputed matrices: an alpha cut-off can
    # flip on isolated pixels, everything else must agree
    diff = (res2["render"] - res["render"].detach()).abs()
    assert float((diff > 1e-3).float().mean()) < 1e-3 and float(diff.max()) < 5e-2


# ---- numeric parity of everything render() does AROUND the rasterizer ---------------------------------------
# tests/golden/render_post.npz was produced by the REFERENCE render() (GR:27-466) over the recording stand-in
# of tests/fake_raster.py (tests/golden/make_render_golden.py); the mirror runs over the same stand-in.
RGOLD = os.path.join(os.path.dirname(__file__), "golden", "render_post.npz")


@pytest.mark.parametrize("case", ["surfel", "surfel_down", "gs3d"])
def test_render_postprocessing_matches_reference_golden(case, monkeypatch):
    import math as m
    import sys
    sys.path.insert(0, os.path.dirname(__file__))
    import fake_raster as FR
    from pings_b200 import gaussian_renderer as GRM
    from pings_b200.decoder import Decoder
    from pings_b200.neural_points import default_config
    z = np.load(RGOLD)
    p = case + "."
    N, Fg, Fc, Hd, Lv, K, down, d2n, affine, surround, seed = [int(x) for x in z[p + "meta"]]
    dev = "cuda"
    cfg = default_config(device=dev)
    decs = {"gauss_xyz": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_rot": Decoder(cfg, Fg, Hd, Lv, 4, K, 0),
            "gauss_scale": Decoder(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_alpha": Decoder(cfg, Fg, Hd, Lv, 1, K, 0),
            "gauss_color": Decoder(cfg, Fc, Hd, Lv, 3, K, 3)}
    for name, d in decs.items():
        pre = f"{p}dec.{name}."
        d.load_state_dict({k[len(pre):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(pre)})
        d.to(dev)
    data = {k[len(p) + 5:]: torch.from_numpy(z[k]).to(dev) for k in z.files if k.startswith(p + "data.")}
    data["resolution"] = 0.25
    gaussians = None
    if surround:
        gaussians = {k[len(p) + 4:]: torch.from_numpy(z[k]).to(dev) for k in z.files if k.startswith(p + "sur.")}
    W, H, fx = 96, 64, 80.0
    T_wc = torch.eye(4)
    T_wc[:3, 3] = torch.tensor([0.3, -0.2, 0.1])
    wvt = torch.linalg.inv(T_wc).T.contiguous().to(dev)
    proj = torch.eye(4, device=dev)
    cam = _Cam(device=torch.device(dev), uid=0, image_width=W, image_height=H, FoVx=2 * m.atan(W / (2 * fx)),
               FoVy=2 * m.atan(H / (2 * fx)), zfar=100.0, fx=fx, fy=fx,
               prcppoint=torch.from_numpy(z[p + "cam.prcppoint"]).to(dev), projection_matrix=proj,
               world_view_transform=wvt, full_proj_transform=wvt @ proj, camera_center=T_wc[:3, 3].to(dev),
               cam_rot_delta=torch.zeros(3, device=dev), cam_trans_delta=torch.zeros(3, device=dev),
               exposure_mat=torch.from_numpy(z[p + "cam.exposure_mat"]).to(dev),
               exposure_offset=torch.from_numpy(z[p + "cam.exposure_offset"]).to(dev),
               exposure_a=torch.from_numpy(z[p + "cam.exposure_a"]).to(dev),
               exposure_b=torch.from_numpy(z[p + "cam.exposure_b"]).to(dev))
    surfel = case != "gs3d"
    log = {}
    fake = FR.make_fake(surfel, torch.from_numpy(z[p + "visible"]), seed, log)
    if surfel:
        import pings_b200.diff_gaussian_surfel_rasterization as pkg
    else:
        import pings_b200.diff_gaussian_rasterization as pkg
    monkeypatch.setattr(pkg, "GaussianRasterizer", fake)
    res = GRM.render(cam, None, data, decs, gaussians, torch.tensor([0.1, 0.2, 0.3], device=dev), down_rate=down,
                     view_concat_on=True, correct_exposure=True, correct_exposure_affine=bool(affine),
                     learn_color_residual=True, front_only_on=(case != "surfel_down"), d2n_on=bool(d2n),
                     gs_type="gaussian_surfel" if surfel else "3d_gs")
    # what reached the rasterizer: settings (q20: full-resolution tanfov, truncated H/W), config, bbox, inputs
    s = log["settings"]
    hw = z[p + "settings.hw_tan"]
    assert (s.image_height, s.image_width) == (int(hw[0]), int(hw[1]))
    assert abs(s.tanfovx - hw[2]) < 1e-12 and abs(s.tanfovy - hw[3]) < 1e-12 and s.scale_modifier == hw[4]
    if surfel:
        assert np.array_equal(s.config.cpu().numpy(), z[p + "settings.config"])
        assert np.array_equal(s.patch_bbox.cpu().numpy(), z[p + "settings.patch_bbox"])
    assert log["inputs"]["means3D"].shape[0] == int(z[p + "in.P"][0])      # same alpha > 0 selection + concat
    torch.testing.assert_close(log["inputs"]["means3D"].detach().cpu(), torch.from_numpy(z[p + "in.means3D"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(log["inputs"]["opacities"].detach().cpu(), torch.from_numpy(z[p + "in.opacities"]), rtol=1e-4, atol=1e-5)
    # the result dict
    assert sorted(res.keys()) == list(z[p + "res.keys"])
    assert sorted(k for k, v in res.items() if v is None) == sorted(z[p + "res.none_keys"].tolist())
    assert abs(res["visible_neural_point_ratio"] - float(z[p + "res.visible_ratio"][0])) < 1e-12
    for k in ("render", "surf_depth", "rend_alpha", "rend_normal", "surf_normal", "contributions"):
        key = p + "res." + k
        if key in z.files:
            # surf_normal: a normalised sum of cross products of neighbour differences (fp32 evaluation order differs
            # from torch's: the kernel is pinned tighter against an eager restatement in tests/test_losses_gpu.py)
            tol = dict(rtol=1e-3, atol=2e-5) if k == "surf_normal" else dict(rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(res[k].detach().cpu(), torch.from_numpy(z[key]), **tol)
    for k in ("visibility_filter", "radii"):
        assert np.array_equal(res[k].cpu().numpy(), z[p + "res." + k])
    # backward of the post-processing: gradient of the fixed loss w.r.t. the rasterizer outputs
    loss = FR.render_loss(res, seed)
    assert abs(float(loss) - float(z[p + "loss"][0])) <= 1e-5 * abs(float(z[p + "loss"][0])) + 1e-4
    loss.backward()
    for k, leaf in log["leaves"].items():
        ref = torch.from_numpy(z[p + "grad." + k])
        got = leaf.grad.cpu()
        atol = 1e-6 * float(ref.abs().max()) + 1e-7
        if k == "depth" and d2n and (p + "res.surf_normal") in z.files:
            # depth2normal backward at a pixel whose own normal is exactly zero next to the image border (replicated
            # neighbour + masked neighbours: a near-degenerate cross-product sum): the eager reference adds -X and +X
            # (|X| ~ 1e4..1e12) into that pixel's depth gradient around the neighbours' contributions, which quantises
            # them to ulp(X) - the golden holds that artefact (0.3 % at one pixel of `surfel_down`).  Our kernel
            # combines the pair in registers (deterministic, unquantised), so such pixels get the looser bar.
            zero_n = torch.from_numpy(z[p + "res.surf_normal"]).abs().amax(0, keepdim=True) == 0
            loose = ~torch.isclose(got, ref, rtol=1e-4, atol=atol) & zero_n
            assert int(loose.sum()) <= 2 and float((got - ref)[loose].abs().max() if int(loose.sum()) else 0.0) <= 5e-3 * float(ref.abs().max())
            got = torch.where(loose, ref, got)
        torch.testing.assert_close(got, ref, rtol=1e-4, atol=atol)
