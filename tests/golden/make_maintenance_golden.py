"""Generates tests/golden/maintenance.npz by RUNNING THE REFERENCE's NeuralPoints map maintenance on CPU
(model/neural_gaussians.py: update :214-375, reset_local_map :378-478, assign_local_to_global :482-494,
recreate_hash :939-1024, query_certainty :1117-1133 with the neighbourhood switch of utils/mapper.py:462-476),
imported through oracle/ref_python.py.  Build container only:

    python tests/golden/make_maintenance_golden.py

Three coloured scans along a short trajectory are inserted (temporal local map on, some invalid colours, the
third frame far enough along the travelled distance to re-insert occupied voxels); the fixture stores the scans
and the complete map state after every stage.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_python as rp  # noqa: E402
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_neural_golden import synth_scan  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))

STATE_KEYS = ["neural_points", "point_orientations", "point_ts_create", "point_ts_update", "point_certainties",
              "point_colors", "valid_color_mask", "valid_gs_mask", "free_gs_mask", "local_neural_points",
              "local_point_certainties", "local_point_ts_update", "local_point_colors", "local_valid_color_mask",
              "local_valid_gs_mask", "local_free_gs_mask", "local_mask", "sorrounding_mask", "global2local"]


def snapshot(npm, tag, out):
    for k in STATE_KEYS:
        v = getattr(npm, k, None)
        if isinstance(v, torch.Tensor):
            out[f"{tag}.{k}"] = v.detach().numpy().copy()
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    out[f"{tag}.hash_slots"] = occ.numpy()
    out[f"{tag}.hash_values"] = npm.buffer_pt_index[occ].numpy()
    out[f"{tag}.geo_features_shape"] = np.array(npm.geo_features.shape)


def main():
    torch.manual_seed(11)
    g = torch.Generator().manual_seed(11)
    cfg = rp.load_config(buffer_size=int(2e5), feature_dim=8, color_feature_dim=4, feature_std=0.0, color_on=True,
                         local_map_radius=20.0, sorrounding_map_radius=28.0, local_map_travel_dist_ratio=2.0)
    NP = rp.neural_points_cls()
    npm = NP(cfg)
    npm.travel_dist = torch.tensor([0.0, 5.0, 80.0])
    out = {"meta": np.array([cfg.voxel_size_m, cfg.buffer_size, cfg.local_map_radius, npm.sorrounding_map_radius,
                             npm.diff_travel_dist_local, int(npm.temporal_local_map_on), int(cfg.range_filter_2d)],
                            dtype=np.float64),
           "travel_dist": npm.travel_dist.numpy()}
    ratios = []
    for ts, shift in enumerate([(0, 0, 0), (3.0, 0, 0), (4.0, 1.0, 0)]):
        pts = synth_scan(g, 5000, shift=shift)
        cols = torch.rand(pts.shape[0], 3, generator=g)
        cols[torch.rand(pts.shape[0], generator=g) < 0.2] = -1.0     # invalid colours (utils/tools.py convention)
        out[f"scan{ts}.points"], out[f"scan{ts}.colors"] = pts.numpy(), cols.numpy()
        out[f"scan{ts}.sensor"] = np.array(shift, dtype=np.float32)
        r = npm.update(pts, cols, None, torch.tensor(shift, dtype=torch.float32), None, ts, is_reliable=(ts != 1))
        ratios.append(r)
        snapshot(npm, f"after{ts}", out)
    out["new_point_ratio"] = np.array(ratios)
    # local training edits, written back (assign_local_to_global)
    n_local = npm.local_neural_points.shape[0]
    npm.local_point_certainties = torch.rand(n_local, generator=g)
    npm.local_point_ts_update = torch.randint(0, 3, (n_local,), generator=g, dtype=torch.int32)
    npm.local_geo_features.data = torch.randn(npm.local_geo_features.shape, generator=g)
    npm.local_color_features.data = torch.randn(npm.local_color_features.shape, generator=g)
    out["edit.local_point_certainties"] = npm.local_point_certainties.numpy().copy()
    out["edit.local_point_ts_update"] = npm.local_point_ts_update.numpy().copy()
    out["edit.local_geo_features"] = npm.local_geo_features.data.numpy().copy()
    out["edit.local_color_features"] = npm.local_color_features.data.numpy().copy()
    npm.assign_local_to_global()
    snapshot(npm, "assigned", out)
    out["assigned.geo_features"] = npm.geo_features.numpy().copy()
    out["assigned.color_features"] = npm.color_features.numpy().copy()
    # query_certainty with the single-cell neighbourhood the mapper switches to (utils/mapper.py:462-476)
    q_src = npm.neural_points[torch.randint(0, npm.neural_points.shape[0], (800,), generator=g)]
    q = (q_src + 0.2 * torch.randn(800, 3, generator=g)).contiguous()
    q[:20] += 50.0
    npm.set_search_neighborhood(num_nei_cells=1, search_alpha=0.0)
    out["certainty.q"] = q.numpy()
    out["certainty.K"] = np.array([npm.neighbor_K])
    out["certainty.value"] = npm.query_certainty(q).numpy()
    npm.set_search_neighborhood(num_nei_cells=cfg.num_nei_cells, search_alpha=cfg.search_alpha)
    out["certainty.value_default_nbh"] = npm.query_certainty(q).numpy()
    # recreate_hash: kept_points (index rebuild only), then the merging variant
    npm.recreate_hash(torch.tensor([4.0, 1.0, 0.0]), None, kept_points=True, with_ts=True, cur_ts=2)
    snapshot(npm, "rehash_kept", out)
    npm.recreate_hash(torch.tensor([4.0, 1.0, 0.0]), None, kept_points=False, with_ts=False, cur_ts=2)
    snapshot(npm, "rehash_merged", out)
    out["rehash_merged.geo_features"] = npm.geo_features.numpy().copy()
    path = os.path.join(OUT, "maintenance.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB; points per stage:",
          [out[f"after{t}.neural_points"].shape[0] for t in range(3)], "merged", npm.neural_points.shape[0],
          "ratios", ratios)


if __name__ == "__main__":
    main()
