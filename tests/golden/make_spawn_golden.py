"""Generates tests/golden/spawn_gs.npz by RUNNING THE REFERENCE's spawn_gaussians (GR:469-778) and Decoder
(model/decoder.py) on CPU, imported through oracle/ref_python.py.  Build container only:

    python tests/golden/make_spawn_golden.py

The fixture holds the neural-point data, the five decoders' weights and the reference outputs (and the
gradient of a fixed scalar w.r.t. the geo/colour features) for the KITTI-GS configuration
(feature_dim 32, color_feature_dim 16, hidden 128, 8 Gaussians per point, view_concat, colour residual).
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_python as rp  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    torch.manual_seed(5)
    g = torch.Generator().manual_seed(5)
    cfg = rp.load_config("run_kitti_gs.yaml")
    rp._stub()
    sys.modules.setdefault("diff_gaussian_rasterization", __import__("unittest.mock").mock.MagicMock())
    from gaussian_splatting.gaussian_renderer import spawn_gaussians  # type: ignore
    Dec = rp.decoder_cls()
    Fg, Fc, Hd, Lv, K = 32, 16, 128, 1, 8
    decs = {"gauss_xyz": Dec(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_rot": Dec(cfg, Fg, Hd, Lv, 4, K, 0),
            "gauss_scale": Dec(cfg, Fg, Hd, Lv, 3, K, 0), "gauss_alpha": Dec(cfg, Fg, Hd, Lv, 1, K, 0),
            "gauss_color": Dec(cfg, Fc, Hd, Lv, 3, K, 3)}
    N = 700
    data = {
        "position": torch.randn(N, 3, generator=g) * torch.tensor([15.0, 15.0, 1.0]),
        "orientation": torch.nn.functional.normalize(torch.randn(N, 4, generator=g), dim=1),
        "color": torch.rand(N, 3, generator=g),
        "geo_feature": (0.3 * torch.randn(N + 1, Fg, generator=g)).requires_grad_(True),
        "color_feature": (0.3 * torch.randn(N + 1, Fc, generator=g)).requires_grad_(True),
        "resolution": 0.25,
        "free_mask": torch.rand(N, generator=g) < 0.1,
        "valid_mask": torch.rand(N, generator=g) < 0.9,
    }
    visible = torch.rand(N, generator=g) < 0.7
    cam_origin = torch.tensor([1.0, -2.0, 0.5])
    out = spawn_gaussians(data, decs, visible, cam_origin, False, True, learn_color_residual=True,
                          gs_type="gaussian_surfel", displacement_range_ratio=1.0, max_scale_ratio=1.0,
                          unit_scale_ratio=0.2)
    w = {k: torch.randn(out[k].shape, generator=g) for k in ("gaussian_xyz", "gaussian_scale", "gaussian_rot",
                                                            "gaussian_alpha", "gaussian_color")}
    loss = sum((out[k] * w[k]).sum() for k in w)
    loss.backward()
    z = {"N": np.array([N, Fg, Fc, Hd, Lv, K]), "visible": visible.numpy(), "cam_origin": cam_origin.numpy()}
    for k, v in data.items():
        if isinstance(v, torch.Tensor):
            z["data." + k] = v.detach().numpy()
    z["grad.geo_feature"] = data["geo_feature"].grad.numpy()
    z["grad.color_feature"] = data["color_feature"].grad.numpy()
    for name, d in decs.items():
        for pk, pv in d.state_dict().items():
            z[f"dec.{name}.{pk}"] = pv.numpy()
        z[f"grad.{name}.lout.weight"] = d.lout.weight.grad.numpy()
    for k, v in out.items():
        if isinstance(v, torch.Tensor):
            z["out." + k] = v.detach().numpy()
    for k, v in w.items():
        z["w." + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, "spawn_gs.npz"), **z)
    print("wrote spawn_gs.npz", {k: tuple(v.shape) for k, v in out.items() if isinstance(v, torch.Tensor)})


if __name__ == "__main__":
    main()
