"""Generates tests/golden/tracker_*.npz by RUNNING THE REFERENCE's Tracker.registration_step / implicit_reg /
implicit_color_reg (utils/tracker.py:353-737) on CPU, imported through oracle/ref_python.py.  Build container only:

    python tests/golden/make_tracker_golden.py

The map and the decoder are built exactly like tests/golden/make_neural_golden.py builds them (same seeds), the
source points are map points moved by a small rigid motion + noise; the gradient-norm window is set from the
quantiles of the reference's own gradient norms so that the validity mask really selects.
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


def build_map(seed, cfg_over, F_geo, hidden, level):
    torch.manual_seed(seed)
    g = torch.Generator().manual_seed(seed)
    cfg = rp.load_config(buffer_size=int(2e5), feature_dim=F_geo, color_feature_dim=0, feature_std=0.1, color_on=False,
                         **cfg_over)
    NP, Dec = rp.neural_points_cls(), rp.decoder_cls()
    npm = NP(cfg)
    npm.travel_dist = torch.tensor([0.0, 30.0, 500.0])
    for ts, shift in enumerate([(0, 0, 0), (6.0, 0, 0), (12.0, 0, 0)]):
        npm.update(synth_scan(g, 6000, shift=shift), None, None, torch.tensor(shift, dtype=torch.float32), None, ts)
    n = npm.neural_points.shape[0]
    npm.valid_gs_mask = torch.rand(n, generator=g) < 0.9
    npm.point_certainties = torch.rand(n, generator=g)
    npm.reset_local_map(torch.tensor([6.0, 0.0, 0.0]), None, 1)
    dec = Dec(cfg, F_geo, hidden, level, 1)
    return cfg, npm, dec, g


def state_dict(npm, dec, cfg, F_geo, hidden, level):
    st = {
        "neural_points": npm.neural_points, "point_orientations": npm.point_orientations, "geo_features": npm.geo_features,
        "point_ts_create": npm.point_ts_create, "point_ts_update": npm.point_ts_update,
        "point_certainties": npm.point_certainties, "valid_gs_mask": npm.valid_gs_mask, "free_gs_mask": npm.free_gs_mask,
        "local_neural_points": npm.local_neural_points, "local_point_orientations": npm.local_point_orientations,
        "local_geo_features": npm.local_geo_features.data, "local_point_certainties": npm.local_point_certainties,
        "local_point_ts_update": npm.local_point_ts_update, "global2local": npm.global2local, "local_mask": npm.local_mask,
        "travel_dist": npm.travel_dist,
    }
    occ = torch.nonzero(npm.buffer_pt_index >= 0).flatten()
    out = {f"state.{k}": v.detach().numpy() for k, v in st.items() if v is not None}
    out["state.hash_slots"] = occ.numpy()
    out["state.hash_values"] = npm.buffer_pt_index[occ].numpy()
    out.update({f"dec.{k}": v.detach().numpy() for k, v in dec.state_dict().items()})
    out["meta"] = np.array([cfg.voxel_size_m, cfg.num_nei_cells, cfg.search_alpha, cfg.query_nn_k, cfg.buffer_size, F_geo, 0,
                            hidden, level, int(cfg.weighted_first), 0, 1, npm.diff_travel_dist_local,
                            int(npm.temporal_local_map_on), 0, 1, 1, dec.sdf_scale], dtype=np.float64)
    return out


def make(name, seed, cfg_over, F_geo, hidden, level):
    rp._stub()
    sys.path.insert(0, rp.REF)
    from utils import tracker as RT  # type: ignore
    cfg, npm, dec, g = build_map(seed, cfg_over, F_geo, hidden, level)
    trk = RT.Tracker(cfg, npm, {"sdf": dec, "semantic": None, "color": None})
    trk.silence = True
    n = npm.local_neural_points.shape[0]
    src = npm.local_neural_points[torch.randint(0, n, (4000,), generator=g)]
    ang = torch.tensor(0.02)
    R = torch.tensor([[torch.cos(ang), -torch.sin(ang), 0.0], [torch.sin(ang), torch.cos(ang), 0.0], [0.0, 0.0, 1.0]])
    pts = (src @ R.T + torch.tensor([0.05, -0.03, 0.02]) + 0.05 * torch.randn(4000, 3, generator=g)).contiguous()
    pts[:60] += 60.0                                   # no neighbours: masked out
    normals = torch.nn.functional.normalize(torch.randn(4000, 3, generator=g), dim=1)
    labels = 0.02 * torch.randn(4000, generator=g)
    out = state_dict(npm, dec, cfg, F_geo, hidden, level)
    # the reference's own query results (for the per-stage comparison) and its gradient-norm quantiles
    q = trk.query_source_points(pts.clone(), cfg.infer_bs, True, True, False, False, query_locally=True,
                                mask_min_nn_count=cfg.track_mask_query_nn_k)
    sdf_pred, sdf_grad, _, _, _, mask, certainty, sdf_std = q
    gn = sdf_grad.norm(dim=-1)[mask]
    lo, hi = float(gn.quantile(0.15)), float(gn.quantile(0.9))
    out.update(points=pts.numpy(), normals=normals.numpy(), labels=labels.numpy(), q_sdf=sdf_pred.numpy(),
               q_grad=sdf_grad.numpy(), q_mask=mask.numpy(), q_std=sdf_std.numpy(), q_cert=certainty.numpy(),
               cfgv=np.array([cfg.infer_bs, cfg.track_mask_query_nn_k, cfg.surface_sample_range_m, cfg.max_sdf_std_ratio,
                              int(cfg.reg_dist_div_grad_norm), lo, hi], dtype=np.float64))
    calls = {"plain": dict(normals=None, GM_dist=None, GM_grad=None, lm=0.0, vis=False),
             "robust": dict(normals=normals, GM_dist=0.3, GM_grad=0.1, lm=1e-4, vis=False),
             "cov": dict(normals=None, GM_dist=0.3, GM_grad=None, lm=0.0, vis=True)}
    for cname, c in calls.items():
        T, cov, eig, _, valid_points, res_cm, _ = trk.registration_step(pts.clone(), c["normals"], labels, None, lo, hi,
                                                                        c["GM_dist"], c["GM_grad"], c["lm"], c["vis"])
        out[f"{cname}.T"] = T.numpy()
        out[f"{cname}.res_cm"] = np.array([res_cm])
        out[f"{cname}.valid_count"] = np.array([valid_points.shape[0]])
        out[f"{cname}.valid_points"] = valid_points.numpy()
        if cov is not None:
            out[f"{cname}.cov"] = cov.numpy()
            out[f"{cname}.eig"] = eig.numpy()
        print(name, cname, "valid", valid_points.shape[0], "res_cm", round(res_cm, 4), "t", T[:3, 3].numpy().round(5))
    # too few valid points: identity
    T, _, _, _, vp, res_cm, _ = trk.registration_step(pts[:70].clone(), None, labels[:70], None, lo, hi, None, None, 0.0, False)
    out["few.T"], out["few.valid_count"] = T.numpy(), np.array([vp.shape[0]])
    # the two solver functions on plain tensors (weight = tensor and the float 1.0 case, covariance / eigenvalues)
    P = torch.randn(3000, 3, generator=g) * 10
    G = torch.nn.functional.normalize(torch.randn(3000, 3, generator=g), dim=1) * (1 + 0.1 * torch.randn(3000, 1, generator=g))
    r = 0.05 * torch.randn(3000, generator=g)
    w = torch.rand(3000, 1, generator=g)
    T1, cov1, eig1 = RT.implicit_reg(P, G, r, w, lm_lambda=1e-3, require_cov=True, require_eigen=True)
    T2, _, _ = RT.implicit_reg(P, G, r, 1.0)
    col = torch.rand(3000, 1, generator=g)
    cgrad = 0.3 * torch.randn(3000, 1, 3, generator=g)
    cres = 0.1 * torch.randn(3000, 1, generator=g)
    T3 = RT.implicit_color_reg(P, G, r, col, cgrad, cres, w, w_photo_loss=0.1, lm_lambda=1e-4)
    out.update({"fn.P": P.numpy(), "fn.G": G.numpy(), "fn.r": r.numpy(), "fn.w": w.numpy(), "fn.T1": T1.numpy(),
                "fn.cov1": cov1.numpy(), "fn.eig1": eig1.numpy(), "fn.T2": T2.numpy(), "fn.col": col.numpy(),
                "fn.cgrad": cgrad.numpy(), "fn.cres": cres.numpy(), "fn.T3": T3.numpy()})
    path = os.path.join(OUT, f"tracker_{name}.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    make("pin", 1, dict(), 8, 64, 1)                                                        # weighted_first (sdf_std = 0)
    make("gs", 2, dict(weighted_first=False, voxel_size_m=0.25, search_alpha=0.8), 32, 128, 2)   # per-neighbour SDF + std
