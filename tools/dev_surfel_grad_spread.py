"""Dev probe: run-to-run spread of the per-surfel gradient chain, ours vs the reference build, on the
test case that sits closest to its bar (tests/test_rasters_gpu.py [20000-320-200-6-2])."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
sys.path.insert(0, "tests")
import test_rasters_gpu as T
from oracle.ref_api import RefRasterS
from pings_b200 import _raster_c as RC

def main(P=20000, W=320, H=200, seed=6, cfg=2, K=6):
    cam, sc, mats, ct, ex = T._scene(P, W, H, seed)
    view, full, proj, center = mats
    config = torch.tensor(T.CONFIGS[cfg], dtype=torch.float32, device="cuda")
    bg = torch.tensor([1.0, 0.5, 0.25], device="cuda")
    e = torch.empty(0)
    ref = RefRasterS()
    r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"], config, campos=center)
    fw = RC.rasterize_gaussians_surfel(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                                       view, full, proj, ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, H, W, e, 0,
                                       center, False, True, config)
    R, color, normal, depth, opac, radii, contributions, geom, binning, img = fw
    names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations", "tau"]
    ours, refs = [], []
    for _ in range(K):
        out = RC.rasterize_gaussians_surfel_backward(
            bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
            ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"], ct["opac"],
            e, 0, center, geom, R, binning, img, True, config, full_tau=True)
        ours.append({n: t.double().reshape(P, -1) for n, t in zip(names, out) if n != "sh"})
        g = ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"], config,
                         r["radii"], ct["color"], ct["normal"], ct["depth"], ct["opac"], campos=center)
        refs.append({n: g[n].double().reshape(P, -1) for n in names if n != "sh"})
    for n in ["means3D", "cov3D", "scales", "rotations", "tau"]:
        mx = refs[0][n].abs().max()
        rr = max(float((refs[i][n] - refs[j][n]).abs().max() / mx) for i in range(K) for j in range(i))
        oo = max(float((ours[i][n] - ours[j][n]).abs().max() / mx) for i in range(K) for j in range(i))
        orr = [float((ours[i][n] - refs[j][n]).abs().max() / mx) for i in range(K) for j in range(K)]
        print(f"{n:10s} ref-ref {rr:.4f}  ours-ours {oo:.4f}  ours-ref max {max(orr):.4f} median {sorted(orr)[len(orr)//2]:.4f}")

if __name__ == "__main__":
    main()
