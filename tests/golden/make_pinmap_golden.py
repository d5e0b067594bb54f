"""Generates tests/golden/pin_map_small.pth + pin_map_small.npz by RUNNING THE REFERENCE on CPU (build container only):

    python tests/golden/make_pinmap_golden.py

* pin_map_small.pth: written by the reference's own utils/tools.py:469-491 save_implicit_map - a pickle of the live
  `model.neural_gaussians.NeuralPoints` module plus the decoders' state_dicts (the format inspect_pings.py:184-200 reads).
* pin_map_small.npz: the same map's tensors as plain arrays, grid queries, and the outputs of the reference's
  Mesher.query_points (utils/mesher.py:40-166) on them (sdf, colour, marching-cubes mask).
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
from oracle import ref_python as rp  # noqa: E402
from make_neural_golden import synth_scan  # noqa: E402


def main():
    torch.manual_seed(31)
    g = torch.Generator().manual_seed(31)
    cfg = rp.load_config(buffer_size=int(5e4), feature_dim=8, color_feature_dim=4, feature_std=0.1, color_on=True)
    cfg.silence = True
    cfg.color_channel = 3
    NP, Dec = rp.neural_points_cls(), rp.decoder_cls()
    from utils.mesher import Mesher  # type: ignore
    from utils.tools import save_implicit_map  # type: ignore
    npm = NP(cfg)
    npm.travel_dist = torch.tensor([0.0, 2.0])
    for ts, shift in enumerate([(0, 0, 0), (2.0, 0, 0)]):
        pts = synth_scan(g, 3000, extent=(16.0, 16.0, 3.0), shift=shift)
        cols = torch.rand(pts.shape[0], 3, generator=g)
        npm.update(pts, cols, None, torch.tensor(shift, dtype=torch.float32), None, ts)
    n = npm.neural_points.shape[0]
    npm.valid_gs_mask = torch.rand(n, generator=g) < 0.9
    sdf_mlp = Dec(cfg, 8, 64, 1, 1)
    color_mlp = Dec(cfg, 4, 64, 1, 3)
    decoders = {"sdf": sdf_mlp, "semantic": None, "color": color_mlp}
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "model"))
        save_implicit_map(d, npm, decoders)
        os.replace(os.path.join(d, "model", "pin_map.pth"), os.path.join(HERE, "pin_map_small.pth"))
    mesher = Mesher(cfg, npm, decoders)
    xs = torch.arange(-6.0, 6.0, 0.45)
    zs = torch.arange(-2.4, 1.0, 0.45)
    X, Y, Z = torch.meshgrid(xs, xs, zs, indexing="ij")
    coord = torch.stack((X.flatten(), Y.flatten(), Z.flatten()), 1).contiguous()
    sdf, _, color, mask = mesher.query_points(coord, 3000, query_sdf=True, query_sem=False, query_color=True,
                                              query_mask=True, query_locally=False, mask_min_nn_count=4, out_torch=True)
    out = {"coord": coord.numpy(), "sdf": sdf.numpy(), "color": color.numpy(), "mask": mask.numpy(),
           "neural_points": npm.neural_points.numpy(), "geo_features": npm.geo_features.detach().numpy(),
           "color_features": npm.color_features.detach().numpy(), "valid_gs_mask": npm.valid_gs_mask.numpy(),
           "point_ts_create": npm.point_ts_create.numpy(),
           "hash_slots": torch.nonzero(npm.buffer_pt_index >= 0).flatten().numpy(),
           "meta": np.array([cfg.voxel_size_m, cfg.buffer_size, cfg.query_nn_k, cfg.num_nei_cells, cfg.search_alpha,
                             int(cfg.weighted_first), sdf_mlp.sdf_scale])}
    out["hash_values"] = npm.buffer_pt_index[torch.from_numpy(out["hash_slots"])].numpy()
    for name, dec in (("sdf", sdf_mlp), ("color", color_mlp)):
        for k, v in dec.state_dict().items():
            out[f"dec.{name}.{k}"] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "pin_map_small.npz"), **out)
    print("points", n, "queries", coord.shape[0], "mask true", int(mask.sum()),
          os.path.getsize(os.path.join(HERE, "pin_map_small.pth")) // 1024, "KiB pth")


if __name__ == "__main__":
    main()
