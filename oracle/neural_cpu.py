"""TEST INFRASTRUCTURE — CPU restatement (torch CPU tensor ops) of the reference's neural-point
query path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this; the product never does.

Follows, step by step:
  radius_search   <- model/neural_gaussians.py:1061-1115 (radius_neighborhood_search)
  query_feature   <- model/neural_gaussians.py:506-725
  mlp / sdf       <- model/decoder.py:62-104, utils/mapper.py:2273-2289 (Mapper.sdf)
Pinned against tests/golden/neural_*.npz, which were produced by running the reference itself
(tests/golden/make_neural_golden.py) — see tests/test_oracle_neural.py.
"""
from __future__ import annotations

import torch

PRIMES = torch.tensor([73856093, 19349669, 83492791], dtype=torch.int64)


def neighbor_offsets(num_nei_cells: int, search_alpha: float) -> torch.Tensor:
    """NP:1026-1043."""
    d = torch.arange(-num_nei_cells, num_nei_cells + 1, dtype=torch.int64)
    dx = torch.stack(torch.meshgrid(d, d, d, indexing="ij"), dim=-1).reshape(-1, 3)
    return dx[(dx ** 2).sum(-1) < (num_nei_cells + search_alpha) ** 2]


def radius_search(S: dict, q: torch.Tensor, time_filtering: bool):
    """S: map state dict (tensors named like the reference attributes). NP:1061-1115."""
    res, B = S["resolution"], int(S["buffer_size"])
    cells = (q / res).floor().to(torch.int64)[:, None, :] + S["neighbor_dx"]        # NP:1069-1073
    h = torch.fmod((cells * PRIMES).sum(-1), B)                                      # NP:1078-1080
    idx = S["buffer_pt_index"][h]                                                    # NP:1084 (negative wraps)
    if time_filtering:                                                               # NP:1088-1094
        td = S["travel_dist"]
        diff = torch.abs(td[S["cur_ts"]] - td[S["point_ts_create"][idx].long()])
        idx[~(diff < S["diff_travel_dist_local"])] = -1
    sub = S["neural_points"][idx] - q.view(-1, 1, 3)                                 # NP:1098-1099
    d2 = torch.sum(sub ** 2, dim=-1)
    d2[idx == -1] = S["max_valid_dist2"]                                             # NP:1102
    idx[d2 > S["max_valid_dist2"]] = -1                                              # NP:1105
    return d2, idx


def rotate_passive(quat: torch.Tensor, p: torch.Tensor) -> torch.Tensor:
    """utils/tools.py:743-751."""
    w = quat[..., :1]
    v = -quat[..., 1:]
    t = 2 * torch.linalg.cross(v, p)
    return p + w * t + torch.linalg.cross(v, t)


def query_feature(S: dict, q: torch.Tensor, nn_k: int, weighted_first: bool, query_locally=True, time_filtering=True,
                  use_only_measured=True, use_only_valid=False, after_pgo=False, geo=None, color=None):
    """NP:506-725 without the in-place stability side effects (returned separately).
    geo / color: feature tables (with the padding row) or None. Returns dict."""
    d2, idx = radius_search(S, q, time_filtering)
    if use_only_measured:
        idx[S["free_gs_mask"][idx]] = -1                                             # NP:544-546
    if use_only_valid:
        idx[~S["valid_gs_mask"][idx]] = -1                                           # NP:548-550
    if query_locally:
        idx = S["global2local"][idx]                                                 # NP:553-554
    nn_counts = (idx >= 0).sum(-1)                                                   # NP:557
    d2[idx == -1] = 9e3                                                              # NP:562
    sd, order = torch.sort(d2, dim=1)                                                # NP:565
    sidx = idx.gather(1, order)
    d2, idx = sd[:, :nn_k], sidx[:, :nn_k]
    valid = idx >= 0
    pts = S["local_neural_points"] if query_locally else S["neural_points"]
    quat = S["local_point_orientations"] if query_locally else S["point_orientations"]
    cert = S["local_point_certainties"] if query_locally else S["point_certainties"]
    nv = q.view(-1, 1, 3) - pts[idx]                                                 # NP:612-625
    if after_pgo:
        nv = rotate_passive(quat[idx], nv)
    nv = torch.where(valid[..., None], nv, torch.zeros_like(nv))
    w = 1.0 / (d2 + 1e-15)                                                           # NP:642-662
    w = torch.where(valid, w, torch.zeros_like(w))
    w = torch.where((nn_counts == 0)[:, None], torch.full_like(w, 1e-15), w)
    w = w / w.sum(1, keepdim=True)
    w = torch.where(valid, w, torch.zeros_like(w))
    c = torch.where(valid, cert[idx], torch.zeros_like(w))
    out = dict(weight=w.unsqueeze(-1), nn_counts=nn_counts, certainty=(c * w).sum(1), knn_idx=idx, knn_dist2=d2)
    for name, table in (("geo", geo), ("color", color)):
        if table is None:
            continue
        f = torch.where(valid[..., None], table[idx], torch.zeros(1, dtype=table.dtype))
        vec = torch.cat((f, nv), dim=2)
        out[name] = (vec * w.unsqueeze(-1)).sum(1) if weighted_first else vec        # NP:701-710
    return out


def stability_side_effects(knn_idx, weight, query_ts, cert, ts_update):
    """NP:664-689: returns the updated (certainties, ts_update)."""
    valid = knn_idx >= 0
    idx = torch.where(valid, knn_idx, torch.zeros_like(knn_idx))
    cert = cert.clone().scatter_add_(0, idx.flatten(), (weight.squeeze(-1) * valid).flatten())
    if query_ts is not None:
        ts = torch.where(valid, query_ts.view(-1, 1).expand_as(idx), torch.zeros_like(idx).int())
        ts_update = ts_update.clone().scatter_reduce_(0, idx.flatten(), ts.flatten().int(), reduce="amax", include_self=True)
    return cert, ts_update


def mlp(x: torch.Tensor, weights: list, biases: list, leaky: bool = False) -> torch.Tensor:
    """DEC:62-82."""
    h = x
    for i, (W, b) in enumerate(zip(weights, biases)):
        h = torch.nn.functional.linear(h, W, b)
        if i < len(weights) - 1:
            h = torch.nn.functional.leaky_relu(h) if leaky else torch.relu(h)
    return h


def sdf_query(S, q, nn_k, weighted_first, weights, biases, sdf_scale, leaky=False, **kw):
    """Mapper.sdf (utils/mapper.py:2273-2289) -> (sdf [N], std [N] | None, nn_counts)."""
    r = query_feature(S, q, nn_k, weighted_first, geo=S["local_geo_features"] if kw.get("query_locally", True) else S["geo_features"], **kw)
    pred = mlp(r["geo"], weights, biases, leaky).squeeze(1) * sdf_scale if weighted_first else \
        mlp(r["geo"], weights, biases, leaky) * sdf_scale
    std = None
    if not weighted_first:
        mean = torch.sum(pred * r["weight"], dim=1)
        std = torch.sqrt(torch.sum(r["weight"] * (pred - mean.unsqueeze(-1)) ** 2, dim=1)).squeeze(1)
        pred = mean.squeeze(1)
    return pred, std, r["nn_counts"]


def state_from_golden(z, prefix="state.") -> dict:
    """Rebuild the state dict (incl. the dense slot table) from a tests/golden/neural_*.npz file."""
    S = {k[len(prefix):]: torch.from_numpy(z[k]) for k in z.files if k.startswith(prefix)}
    meta = z["meta"]
    S["resolution"] = float(meta[0])
    S["buffer_size"] = int(meta[4])
    S["neighbor_dx"] = neighbor_offsets(int(meta[1]), float(meta[2]))
    S["max_valid_dist2"] = 3 * ((int(meta[1]) + 1) * float(meta[0])) ** 2
    S["cur_ts"] = int(meta[11])
    S["diff_travel_dist_local"] = float(meta[12])
    table = torch.full((S["buffer_size"],), -1, dtype=torch.int64)
    table[S["hash_slots"]] = S["hash_values"]
    S["buffer_pt_index"] = table
    return S
