"""Generates tests/golden/neural_*.npz by RUNNING THE REFERENCE ITSELF on CPU
(/root/reference/model/neural_gaussians.py + model/decoder.py, imported through oracle/ref_python.py).
Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_neural_golden.py

Each fixture holds the full map state the kernels read, the seeded queries, and the reference's
outputs of radius_neighborhood_search / query_feature / Decoder.sdf and their autograd gradients.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_python as rp  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def synth_scan(g, n, extent=(30.0, 30.0, 3.0), shift=(0.0, 0.0, 0.0)):
    """Points on a ground plane + two walls, so that voxels are locally dense like a LiDAR scan."""
    a = torch.rand(n, 3, generator=g)
    ground = torch.stack([(a[:, 0] - 0.5) * extent[0], (a[:, 1] - 0.5) * extent[1], -1.7 + 0.02 * torch.randn(n, generator=g)], 1)
    wall = torch.stack([(a[:, 0] - 0.5) * extent[0], torch.full((n,), extent[1] * 0.3), -1.7 + a[:, 2] * extent[2]], 1)
    pts = torch.cat([ground[: n // 2], wall[n // 2:]], 0)
    return pts + torch.tensor(shift)


def make(name, seed, cfg_over, F_geo, F_color, hidden, level, after_pgo, color_query, flags):
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed)
    cfg = rp.load_config(buffer_size=int(2e5), feature_dim=F_geo, color_feature_dim=F_color, feature_std=0.1,
                         color_on=F_color > 0, **cfg_over)
    NP, Dec = rp.neural_points_cls(), rp.decoder_cls()
    npm = NP(cfg)
    npm.travel_dist = torch.tensor([0.0, 30.0, 500.0])
    # three frames; the third is far along the trajectory so the travel-distance window excludes frame 0/1 points
    for ts, shift in enumerate([(0, 0, 0), (6.0, 0, 0), (12.0, 0, 0)]):
        pts = synth_scan(g, 6000, shift=shift)
        cols = torch.rand(pts.shape[0], 3, generator=g) if F_color > 0 else None
        npm.update(pts, cols, None, torch.tensor(shift, dtype=torch.float32), None, ts)
    n = npm.neural_points.shape[0]
    # exercise the masks / orientations / certainties
    npm.free_gs_mask = torch.rand(n, generator=g) < 0.1
    npm.valid_gs_mask = torch.rand(n, generator=g) < 0.9
    npm.point_certainties = torch.rand(n, generator=g)
    npm.point_orientations = torch.nn.functional.normalize(torch.randn(n, 4, generator=g), dim=1)
    npm.after_pgo = after_pgo
    cur_ts = flags.pop("cur_ts", 1)
    npm.reset_local_map(torch.tensor([6.0, 0.0, 0.0]), None, cur_ts)
    dec = Dec(cfg, F_geo, hidden, level, 1)
    q_src = npm.neural_points[torch.randint(0, n, (1500,), generator=g)]
    q = (q_src + 0.15 * torch.randn(1500, 3, generator=g)).contiguous()
    q[:40] += 40.0  # queries with no neighbours at all
    q_ts = torch.randint(0, 3, (1500,), generator=g, dtype=torch.int32)

    dist2, idx = npm.radius_neighborhood_search(q, time_filtering=npm.temporal_local_map_on)
    cert_before = npm.local_point_certainties.clone()
    ts_before = npm.local_point_ts_update.clone()
    qg = q.clone().requires_grad_(True)
    geo_vec, color_vec, weight, nn_counts, certainty = npm.query_feature(
        qg, q_ts, accumulate_stability=True, query_color_feature=color_query, **flags)
    sdf = dec.sdf(geo_vec)
    if not cfg.weighted_first:
        sdf = torch.sum(sdf * weight, dim=1).squeeze(1)
    r1 = torch.randn(sdf.shape, generator=g)
    r2 = torch.randn(weight.shape, generator=g)
    loss = (sdf * r1).sum() + (weight * r2).sum()
    r3 = None
    if color_vec is not None:
        r3 = torch.randn(color_vec.shape, generator=g)
        loss = loss + (color_vec * r3).sum()
    loss.backward()
    st = {
        "neural_points": npm.neural_points, "point_orientations": npm.point_orientations,
        "geo_features": npm.geo_features, "color_features": npm.color_features,
        "point_ts_create": npm.point_ts_create, "point_ts_update": npm.point_ts_update,
        "point_certainties": npm.point_certainties, "valid_gs_mask": npm.valid_gs_mask, "free_gs_mask": npm.free_gs_mask,
        "local_neural_points": npm.local_neural_points, "local_point_orientations": npm.local_point_orientations,
        "local_geo_features": npm.local_geo_features.data, "local_point_certainties": cert_before,
        "local_point_ts_update": ts_before, "global2local": npm.global2local, "local_mask": npm.local_mask,
        "travel_dist": npm.travel_dist,
    }
    if F_color > 0:
        st["local_color_features"] = npm.local_color_features.data
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    out = {f"state.{k}": v.detach().numpy() for k, v in st.items() if v is not None}
    out["state.hash_slots"] = occ.numpy()
    out["state.hash_values"] = npm.buffer_pt_index[occ].numpy()
    out.update({f"dec.{k}": v.detach().numpy() for k, v in dec.state_dict().items()})
    out.update(dict(
        meta=np.array([cfg.voxel_size_m, cfg.num_nei_cells, cfg.search_alpha, cfg.query_nn_k, cfg.buffer_size, F_geo, F_color,
                       hidden, level, int(cfg.weighted_first), int(after_pgo), cur_ts, npm.diff_travel_dist_local,
                       int(npm.temporal_local_map_on), int(color_query), int(flags.get("use_only_valid_points", False)),
                       int(flags.get("use_only_measured_points", True)), dec.sdf_scale], dtype=np.float64),
        q=q.numpy(), q_ts=q_ts.numpy(), dist2=dist2.numpy(), idx=idx.numpy().astype(np.int32),
        geo_vec=geo_vec.detach().numpy(), weight=weight.detach().numpy(), nn_counts=nn_counts.numpy(),
        certainty=certainty.numpy(), sdf=sdf.detach().numpy(), r1=r1.numpy(), r2=r2.numpy(),
        cert_after=npm.local_point_certainties.numpy(), ts_after=npm.local_point_ts_update.numpy(),
        d_q=qg.grad.numpy(), d_geo=npm.local_geo_features.grad.numpy(),
    ))
    for k, p in dec.named_parameters():
        out[f"dgrad.{k}"] = p.grad.numpy()
    if color_vec is not None:
        out.update(color_vec=color_vec.detach().numpy(), r3=r3.numpy(), d_color=npm.local_color_features.grad.numpy())
    path = os.path.join(OUT, f"neural_{name}.npz")
    np.savez_compressed(path, **out)
    print(name, "points", n, "local", npm.local_neural_points.shape[0], "K", npm.neighbor_K, "mean nn", float(nn_counts.float().mean()),
          "zero-nn", int((nn_counts == 0).sum()), os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    # A: PIN-SLAM configuration (config/run_pin_slam.yaml): K=81, k=6, F=8, weighted_first, MLP 11->64->1
    make("pin", 1, dict(), 8, 0, 64, 1, False, False, dict(cur_ts=1))
    # B: KITTI-GS-like (config/run_kitti_gs.yaml:21-30): F=32/16, !weighted_first, colour features, rotated
    #    neighbour vectors (after_pgo), valid-mask filtering, two hidden levels of 128
    make("gs", 2, dict(weighted_first=False, voxel_size_m=0.25, search_alpha=0.8), 32, 16, 128, 2, True, True,
         dict(cur_ts=1, use_only_valid_points=True))
    # C: travel-distance window active (cur_ts=2 is 500 m along: only frame-2 points pass the filter), small K
    make("window", 3, dict(num_nei_cells=1, search_alpha=0.2, local_map_travel_dist_ratio=1.0), 8, 0, 32, 1, False, False,
         dict(cur_ts=2))
