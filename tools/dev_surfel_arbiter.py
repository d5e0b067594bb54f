"""Dev probe (GPU box): for the small surfel parity cases, distance of OUR backward and of the REFERENCE CUDA
backward (several runs) from the fp64 arbiter (oracle/cpu_rasters64.c), per tensor and split into blend sums
/ per-surfel chain / conditioning class.  usage: python tools/dev_surfel_arbiter.py [out.json]"""
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cpu_rasters64 as O64   # noqa: E402
from oracle.ref_api import RefRasterS     # noqa: E402
from pings_b200 import _raster_c as RC    # noqa: E402

_spec = importlib.util.spec_from_file_location("t_rasters", os.path.join(ROOT, "tests", "test_rasters_gpu.py"))
_t = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_t)
CONFIGS, _scene = _t.CONFIGS, _t._scene

CASES = [(3000, 320, 200, 3, 0), (40000, 640, 360, 1, 0), (20000, 333, 217, 5, 1), (20000, 320, 200, 6, 2),
         (10000, 320, 200, 8, 3)]


def relmax(a, b):
    a, b = a.double().cpu().reshape(b.shape[0], -1), b.double().cpu().reshape(b.shape[0], -1)
    return float((a - b).abs().max()) / (float(b.abs().max()) + 1e-300)


def main():
    rows = []
    for (P, W, H, seed, cfg) in CASES:
        cam, sc, mats, ct, ex = _scene(P, W, H, seed)
        view, full, proj, center = mats
        config = torch.tensor(CONFIGS[cfg], dtype=torch.float32, device="cuda")
        bg = torch.tensor([1.0, 0.5, 0.25], device="cuda")
        e = torch.empty(0)
        res = RC.rasterize_gaussians_surfel(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"],
                                            sc["rotations"], 1.0, e, view, full, proj, ex["prcppoint"], ex["patch_bbox"],
                                            cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, True, config)
        R, color, normal, depth, opac, radii, contributions, geom, binning, img = res
        dv = RC.debug_views(1, P, W, H, R, geom, binning, img)
        out = RC.rasterize_gaussians_surfel_backward(
            bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
            ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"],
            ct["opac"], e, 0, center, geom, R, binning, img, True, config, full_tau=True)
        names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations", "tau"]
        ours = dict(zip(names, out))
        bs = {k: v.clone() for k, v in RC.debug_blend_sums(1, P, geom).items()}
        ref = RefRasterS()
        r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"],
                        config, campos=center)
        runs = [ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"],
                             config, r["radii"], ct["color"], ct["normal"], ct["depth"], ct["opac"], campos=center)
                for _ in range(4)]
        a = O64.blend(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, config, radii, dv["point_list"],
                      dv["ranges"], ct)
        g64 = O64.chain(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, config, radii, a["g_means2D"],
                        a["g_conic"], a["g_normal"], a["g_depths"])
        row = dict(case=[P, W, H, seed, CONFIGS[cfg]], visible=int((radii > 0).sum()), R=int(R))
        row["img_ours_vs_64"] = {k: relmax(v, a[k].reshape(v.shape)) for k, v in
                                 dict(color=color, normal=normal, depth=depth, opac=opac).items()}
        row["img_ref_vs_64"] = {k: relmax(r[k], a[k].reshape(r[k].shape)) for k in ("color", "normal", "depth", "opac")}
        row["ncontrib_diff_pixels"] = int((dv["n_contrib"].cpu() != a["n_contrib"]).sum())
        blend = {}
        ours_b = dict(means2D=ours["means2D"][:, :2], conic=bs["conic"], opacity=ours["opacity"], colors=ours["colors"],
                      normal=bs["normal"], depths=bs["depths"])
        for k in ours_b:
            refk = [(x[k].reshape(P, -1)[:, [0, 1, 3]] if k == "conic" else x[k].reshape(P, -1)[:, :a["g_" + k].shape[1]])
                    for x in runs]
            blend[k] = dict(ours=relmax(ours_b[k], a["g_" + k]), ref=[relmax(x, a["g_" + k]) for x in refk])
        row["blend_vs_64"] = blend
        full_ = {}
        for k in ("means3D", "cov3D", "scales", "rotations", "tau"):
            full_[k] = dict(ours=relmax(ours[k], g64[k]), ref=[relmax(x[k], g64[k]) for x in runs],
                            ref_vs_ref=relmax(runs[1][k], runs[0][k].double().cpu().reshape(P, -1)))
        row["full_vs_64"] = full_
        g_own = O64.chain(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, config, radii, ours["means2D"],
                          bs["conic"], bs["normal"], bs["depths"])
        x0 = runs[0]
        g_ref_own = O64.chain(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, config, radii, x0["means2D"],
                              x0["conic"].reshape(P, 4)[:, [0, 1, 3]], x0["normal"], x0["depths"])
        row["chain_only"] = {k: dict(ours=relmax(ours[k], g_own[k]), ref=relmax(x0[k], g_ref_own[k]))
                             for k in ("means3D", "cov3D", "scales", "rotations", "tau")}
        # conditioning class of a row: kappa = a c / (a c - b^2) of the projected covariance (from the conic record)
        co = dv["records"][:, 1, :3].double().cpu()
        kappa = (co[:, 0] * co[:, 2]) / (co[:, 0] * co[:, 2] - co[:, 1] ** 2).clamp_min(1e-300)
        vis = radii.cpu() > 0
        cls = {}
        for k in ("means3D", "cov3D", "scales", "rotations", "tau"):
            o = ours[k].double().cpu().reshape(P, -1)
            mx = float(g64[k].abs().max())
            den = g64[k].abs().max(1).values + 1e-4 * mx
            rel = (o - g64[k]).abs().max(1).values / den
            rr = (runs[0][k].double().cpu().reshape(P, -1) - g64[k]).abs().max(1).values / den
            cls[k] = {}
            for lo, hi in ((0, 10), (10, 100), (100, 1000), (1000, 1e30)):
                m = vis & (kappa >= lo) & (kappa < hi)
                if int(m.sum()):
                    cls[k][f"k<{hi:g}"] = dict(n=int(m.sum()), ours=float(rel[m].max()), ref=float(rr[m].max()),
                                               ours_q99=float(rel[m].quantile(0.99)))
        row["by_kappa"] = cls
        tau_o = ours["tau"].double().sum(0).cpu()
        tau64 = g64["tau"].sum(0)
        row["tau_sum"] = dict(ours=float((tau_o - tau64).abs().max() / tau64.abs().max()),
                              ref=[float((x["tau"].double().sum(0).cpu() - tau64).abs().max() / tau64.abs().max()) for x in runs])
        rows.append(row)
        print(json.dumps(row), flush=True)
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as f:
            json.dump(rows, f, indent=1)


if __name__ == "__main__":
    main()
