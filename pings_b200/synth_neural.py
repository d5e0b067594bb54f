"""Seeded synthetic inputs for the SDF-query configuration C1 (BASELINE.json configs[0]):
one 64-beam x 1800-azimuth LiDAR scan of an analytic scene (ground plane, four walls, 20 cylinders),
the neural-point map built from it with the PIN-SLAM settings (config/run_pin_slam.yaml: voxel 0.4 m,
81-cell neighbourhood, k = 6, F = 8) and a decoder 11 -> 64 -> 1.  Generated on the CPU with a
torch.Generator, then moved to the target device."""
from __future__ import annotations

import math

import torch

from .decoder import Decoder
from .neural_points import NeuralPoints, default_config


def lidar_scan(seed: int = 42, n_az: int = 1800, n_el: int = 64) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    az = torch.arange(n_az, dtype=torch.float64) * (2 * math.pi / n_az)
    el = torch.deg2rad(torch.linspace(-24.8, 2.0, n_el, dtype=torch.float64))
    A, E = torch.meshgrid(az, el, indexing="ij")
    d = torch.stack([torch.cos(E) * torch.cos(A), torch.cos(E) * torch.sin(A), torch.sin(E)], -1).reshape(-1, 3)
    inf = torch.full((d.shape[0],), 1e9, dtype=torch.float64)
    t = torch.where(d[:, 2] < -1e-9, -1.73 / d[:, 2], inf)                       # ground z = -1.73
    t = torch.minimum(t, 25.0 / d[:, 0].abs().clamp_min(1e-9))                    # walls x = +-25
    t = torch.minimum(t, 25.0 / d[:, 1].abs().clamp_min(1e-9))                    # walls y = +-25
    cyl = (torch.rand(20, 2, generator=g, dtype=torch.float64) - 0.5) * 40.0      # 20 cylinders r = 0.5
    dxy = d[:, :2]
    a = (dxy ** 2).sum(-1).clamp_min(1e-12)
    for c in cyl:
        b = -(dxy @ c)
        disc = b * b - a * ((c ** 2).sum() - 0.25)
        tc = (-b - torch.sqrt(disc.clamp_min(0))) / a
        t = torch.where((disc > 0) & (tc > 0), torch.minimum(t, tc), t)
    t = t + 0.02 * torch.randn(t.shape, generator=g, dtype=torch.float64)
    t = t.clamp(2.5, 80.0)  # range gate (config.py:58, run_pin_slam.yaml:6); keeps all 115,200 returns
    return (d * t[:, None]).float().contiguous()


def c1_map_and_queries(device="cuda", seed: int = 42, n_az: int = 1800, buffer_size: int = int(1e8)):
    """-> (NeuralPoints, Decoder, queries [N,3]) on `device`."""
    pts = lidar_scan(seed, n_az)
    g = torch.Generator().manual_seed(seed + 1)
    cfg = default_config(device=device, buffer_size=buffer_size, track_on=False, temporal_local_map_off=True)
    npm = NeuralPoints(cfg)
    npm.travel_dist = torch.zeros(1, device=device)
    npm.update(pts.to(device), None, None, torch.zeros(3, device=device), None, 0)
    n = npm.count()
    feats = 0.1 * torch.randn(n + 1, cfg.feature_dim, generator=g)
    npm.geo_features = feats.to(device)
    npm.reset_local_map(torch.zeros(3, device=device), None, 0)
    torch.manual_seed(seed)
    dec = Decoder(cfg, cfg.feature_dim, 64, 1, 1)
    q = (pts + 0.05 * torch.randn(pts.shape, generator=g)).contiguous().to(device)
    return npm, dec, q


def oracle_state(npm: NeuralPoints) -> dict:
    """CPU copies of the tensors the query path reads, keyed like the reference attributes."""
    keys = ["buffer_pt_index", "neural_points", "point_orientations", "geo_features", "point_ts_create",
            "point_certainties", "valid_gs_mask", "free_gs_mask", "local_neural_points", "local_point_orientations",
            "local_point_certainties", "local_point_ts_update", "global2local", "neighbor_dx", "travel_dist"]
    S = {k: getattr(npm, k).detach().cpu() for k in keys if getattr(npm, k, None) is not None}
    S["local_geo_features"] = npm.local_geo_features.detach().cpu()
    S.update(resolution=float(npm.resolution), buffer_size=int(npm.buffer_size), max_valid_dist2=npm.max_valid_dist2,
             cur_ts=int(npm.cur_ts), diff_travel_dist_local=float(npm.diff_travel_dist_local))
    return S


def large_map_and_queries(device="cuda", seed: int = 7, nx: int = 1000, ny: int = 500, n_queries: int = 115_200,
                          buffer_size: int = int(1e8)):
    """A map-sized variant of C1: ~5e5 neural points - one per 0.4 m voxel of a 400 m x 200 m
    ground sheet with gentle relief, as a multi-frame trajectory would leave it - inserted through update(), PIN-SLAM
    settings otherwise, and scan-sized query batches drawn uniformly over the sheet, so that the 81 slot-table probes of
    a query land in DRAM (at C1's 7,033 points the touched slots sit in L2)."""
    g = torch.Generator().manual_seed(seed)
    cfg = default_config(device=device, buffer_size=buffer_size, track_on=False, temporal_local_map_off=True,
                         local_map_radius=1e4, sorrounding_map_radius=1.4e4)
    res = cfg.voxel_size_m
    ix, iy = torch.meshgrid(torch.arange(nx, dtype=torch.float32), torch.arange(ny, dtype=torch.float32), indexing="ij")
    x = (ix.reshape(-1) + 0.5) * res - nx * res / 2
    y = (iy.reshape(-1) + 0.5) * res - ny * res / 2
    z = 0.15 * torch.sin(x / 7.0) * torch.cos(y / 9.0) - 1.73
    pts = torch.stack([x, y, z], 1) + (torch.rand(nx * ny, 3, generator=g) - 0.5) * torch.tensor([0.3, 0.3, 0.02])
    npm = NeuralPoints(cfg)
    npm.travel_dist = torch.zeros(1, device=device)
    npm.update(pts.to(device), None, None, torch.zeros(3, device=device), None, 0)
    n = npm.count()
    npm.geo_features = (0.1 * torch.randn(n + 1, cfg.feature_dim, generator=g)).to(device)
    npm.reset_local_map(torch.zeros(3, device=device), None, 0)
    torch.manual_seed(seed)
    dec = Decoder(cfg, cfg.feature_dim, 64, 1, 1)
    sel = torch.randint(0, pts.shape[0], (n_queries,), generator=g)
    q = (pts[sel] + 0.05 * torch.randn(n_queries, 3, generator=g)).contiguous().to(device)
    return npm, dec, q
