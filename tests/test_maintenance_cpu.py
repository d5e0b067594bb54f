"""Neural-point map maintenance mirrors (pings_b200/neural_points.py: update, reset_local_map,
assign_local_to_global, recreate_hash) against tests/golden/maintenance.npz, which was produced by running the
reference's own NeuralPoints (model/neural_gaussians.py:214-494,939-1024) on the same three scans
(tests/golden/make_maintenance_golden.py).  Index state must match exactly; these are plain torch tensor ops
(SURVEY.md §8f row 3), so the test runs on the CPU."""
import os

import numpy as np
import torch

from pings_b200.neural_points import NeuralPoints, default_config

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "maintenance.npz"))

KEYS = ["neural_points", "point_orientations", "point_ts_create", "point_ts_update", "point_certainties",
        "point_colors", "valid_color_mask", "valid_gs_mask", "free_gs_mask", "local_neural_points",
        "local_point_certainties", "local_point_ts_update", "local_point_colors", "local_valid_color_mask",
        "local_valid_gs_mask", "local_free_gs_mask", "local_mask", "sorrounding_mask", "global2local"]


def _check(npm, tag):
    for k in KEYS:
        key = f"{tag}.{k}"
        if key not in G.files:
            continue
        got = getattr(npm, k).detach().cpu().numpy()
        assert got.shape == G[key].shape, (key, got.shape, G[key].shape)
        assert np.array_equal(got, G[key]), key
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    assert np.array_equal(occ.numpy(), G[f"{tag}.hash_slots"]), tag
    assert np.array_equal(npm.buffer_pt_index[occ].numpy(), G[f"{tag}.hash_values"]), tag
    assert tuple(npm.geo_features.shape) == tuple(G[f"{tag}.geo_features_shape"]), tag


def build_map(device="cpu"):
    voxel, buf, radius, ring, diff_td, temporal, filter_2d = G["meta"]
    cfg = default_config(device=device, buffer_size=int(buf), voxel_size_m=float(voxel), feature_dim=8,
                         color_feature_dim=4, color_on=True, local_map_radius=float(radius),
                         sorrounding_map_radius=float(ring), local_map_travel_dist_ratio=float(diff_td / radius),
                         range_filter_2d=bool(filter_2d))
    npm = NeuralPoints(cfg)
    assert npm.temporal_local_map_on == bool(temporal)
    npm.travel_dist = torch.from_numpy(G["travel_dist"]).to(device)
    return npm


def insert_scans(npm, check=None):
    dev = npm.device
    for ts in range(3):
        pts, cols = torch.from_numpy(G[f"scan{ts}.points"]).to(dev), torch.from_numpy(G[f"scan{ts}.colors"]).to(dev)
        r = npm.update(pts, cols, None, torch.from_numpy(G[f"scan{ts}.sensor"]).to(dev), None, ts, is_reliable=(ts != 1))
        assert abs(r - float(G["new_point_ratio"][ts])) < 1e-12
        if check:
            check(npm, f"after{ts}")


def apply_edits(npm):
    dev = npm.device
    npm.local_point_certainties = torch.from_numpy(G["edit.local_point_certainties"]).to(dev)
    npm.local_point_ts_update = torch.from_numpy(G["edit.local_point_ts_update"]).to(dev)
    npm.local_geo_features.data = torch.from_numpy(G["edit.local_geo_features"]).to(dev)
    npm.local_color_features.data = torch.from_numpy(G["edit.local_color_features"]).to(dev)
    npm.assign_local_to_global()


def test_update_reset_assign_rehash_match_reference_state():
    npm = build_map()
    insert_scans(npm, _check)
    apply_edits(npm)
    _check(npm, "assigned")
    assert np.array_equal(npm.geo_features.numpy(), G["assigned.geo_features"])
    assert np.array_equal(npm.color_features.numpy(), G["assigned.color_features"])
    npm.recreate_hash(torch.tensor([4.0, 1.0, 0.0]), None, kept_points=True, with_ts=True, cur_ts=2)
    _check(npm, "rehash_kept")
    npm.recreate_hash(torch.tensor([4.0, 1.0, 0.0]), None, kept_points=False, with_ts=False, cur_ts=2)
    _check(npm, "rehash_merged")
    assert np.array_equal(npm.geo_features.numpy(), G["rehash_merged.geo_features"])
