"""Generates tests/golden/render_post.npz by RUNNING THE REFERENCE's render() (GR:27-466) on CPU with the
rasterizer packages replaced by the recording stand-in of tests/fake_raster.py (the CUDA rasterizers cannot run
in the build container; they are pinned separately against the reference CUDA build).  What this fixture pins
is everything render() does AROUND the rasterizer: settings passed down (q20 down_rate handling, config
tensor, patch bbox), visibility / spawn / concat with surrounding Gaussians, depth-by-alpha normalisation
(3DGS), depth2normal (point_utils.py:83-148), exposure correction, and the gradient of a fixed loss w.r.t.
the rasterizer outputs.  Build container only:

    python tests/golden/make_render_golden.py
"""
import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_python as rp  # noqa: E402
import fake_raster as FR  # noqa: E402

CASES = {
    # name: (gs_type, down_rate, d2n_on, affine exposure, with surrounding gaussians)
    "surfel": ("gaussian_surfel", 0, True, True, True),
    "surfel_down": ("gaussian_surfel", 1, True, False, False),
    "gs3d": ("3d_gs", 0, True, True, False),
}


class Cam(types.SimpleNamespace):
    def full_patch(self, img_down_rate: int = 0):
        s = 2 ** img_down_rate
        return torch.tensor([0, 0, int(self.image_height / s) - 1, int(self.image_width / s) - 1]).to(
            dtype=torch.float32, device=self.device)


def make_camera(g, W=96, H=64, device="cpu"):
    fx = fy = 80.0
    T_wc = torch.eye(4)
    T_wc[:3, 3] = torch.tensor([0.3, -0.2, 0.1])
    T_cw = torch.linalg.inv(T_wc)
    wvt = T_cw.T.contiguous()
    proj = torch.eye(4)
    return Cam(device=device, uid=0, image_width=W, image_height=H, FoVx=2 * math.atan(W / (2 * fx)),
               FoVy=2 * math.atan(H / (2 * fy)), zfar=100.0, fx=fx, fy=fy,
               prcppoint=torch.tensor([0.48, 0.53]), projection_matrix=proj, world_view_transform=wvt,
               full_proj_transform=wvt @ proj, camera_center=T_wc[:3, 3].clone(),
               cam_rot_delta=torch.zeros(3), cam_trans_delta=torch.zeros(3),
               exposure_mat=(torch.eye(3) + 0.1 * torch.randn(3, 3, generator=g)),
               exposure_offset=0.05 * torch.randn(1, 3, generator=g),
               exposure_a=torch.tensor([0.2]), exposure_b=torch.tensor([-0.03]))


def main():
    rp._stub()
    cfg = rp.load_config("run_kitti_gs.yaml")
    Dec = rp.decoder_cls()
    out = {}
    for ci, (name, (gs_type, down, d2n, affine, surround)) in enumerate(CASES.items()):
        torch.manual_seed(20 + ci)
        g = torch.Generator().manual_seed(20 + ci)
        Fg, Fc, Hd, Lv, K, N = 32, 16, 128, 1, 8, 300
        decs = {"gauss_xyz": Dec(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_rot": Dec(cfg, Fg, Hd, Lv, 4, K, 0),
                "gauss_scale": Dec(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_alpha": Dec(cfg, Fg, Hd, Lv, 1, K, 0),
                "gauss_color": Dec(cfg, Fc, Hd, Lv, 3, K, 3)}
        data = {"position": torch.randn(N, 3, generator=g) * torch.tensor([10.0, 10.0, 1.0]),
                "orientation": torch.nn.functional.normalize(torch.randn(N, 4, generator=g), dim=1),
                "color": torch.rand(N, 3, generator=g),
                "geo_feature": 0.3 * torch.randn(N + 1, Fg, generator=g),
                "color_feature": 0.3 * torch.randn(N + 1, Fc, generator=g), "resolution": 0.25}
        visible = torch.rand(N, generator=g) < 0.7
        cam = make_camera(g)
        gaussians = None
        if surround:
            M = 40
            gaussians = {"gaussian_xyz": torch.randn(M, 3, generator=g), "gaussian_scale": torch.rand(M, 3, generator=g),
                         "gaussian_rot": torch.nn.functional.normalize(torch.randn(M, 4, generator=g), dim=1),
                         "gaussian_alpha": torch.rand(M, 1, generator=g), "gaussian_color": torch.rand(M, 3, generator=g)}
        log = {}
        fake = types.ModuleType("fake")
        surfel = gs_type == "gaussian_surfel"
        fake.GaussianRasterizer = FR.make_fake(surfel, visible, 100 + ci, log)
        fake.GaussianRasterizationSettings = FR.SettingsS if surfel else FR.Settings3
        sys.modules["diff_gaussian_surfel_rasterization" if surfel else "diff_gaussian_rasterization"] = fake
        from gaussian_splatting.gaussian_renderer import render  # type: ignore
        res = render(cam, None, data, decs, gaussians, torch.tensor([0.1, 0.2, 0.3]), down_rate=down,
                     view_concat_on=True, correct_exposure=True, correct_exposure_affine=affine,
                     learn_color_residual=True, front_only_on=(ci != 1), d2n_on=d2n, gs_type=gs_type)
        loss = FR.render_loss(res, 100 + ci)
        loss.backward()
        p = name + "."
        out[p + "meta"] = np.array([N, Fg, Fc, Hd, Lv, K, down, int(d2n), int(affine), int(surround), 100 + ci])
        out[p + "visible"] = visible.numpy()
        for k, v in data.items():
            if isinstance(v, torch.Tensor):
                out[p + "data." + k] = v.numpy()
        for dn, d in decs.items():
            for pk, pv in d.state_dict().items():
                out[f"{p}dec.{dn}.{pk}"] = pv.numpy()
        if gaussians is not None:
            for k, v in gaussians.items():
                out[p + "sur." + k] = v.numpy()
        for k in ("exposure_mat", "exposure_offset", "exposure_a", "exposure_b", "prcppoint"):
            out[p + "cam." + k] = getattr(cam, k).numpy()
        s = log["settings"]
        out[p + "settings.hw_tan"] = np.array([s.image_height, s.image_width, s.tanfovx, s.tanfovy, s.scale_modifier])
        if surfel:
            out[p + "settings.config"] = s.config.numpy()
            out[p + "settings.patch_bbox"] = s.patch_bbox.numpy()
        out[p + "in.P"] = np.array([log["inputs"]["means3D"].shape[0]])
        out[p + "in.means3D"] = log["inputs"]["means3D"].detach().numpy()
        out[p + "in.opacities"] = log["inputs"]["opacities"].detach().numpy()
        for k, v in res.items():
            if isinstance(v, torch.Tensor) and k in ("render", "surf_depth", "rend_alpha", "rend_normal", "surf_normal",
                                                     "visibility_filter", "radii", "contributions"):
                out[p + "res." + k] = v.detach().numpy()
        out[p + "res.none_keys"] = np.array([k for k, v in res.items() if v is None])
        out[p + "res.keys"] = np.array(sorted(res.keys()))
        out[p + "res.visible_ratio"] = np.array([res["visible_neural_point_ratio"]])
        out[p + "loss"] = np.array([float(loss)])
        for k, v in log["leaves"].items():
            out[p + "grad." + k] = v.grad.numpy() if v.grad is not None else np.zeros(0)
        print(name, "P", int(out[p + "in.P"][0]), "loss", float(loss), "keys", len(res))
    path = os.path.join(HERE, "render_post.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
