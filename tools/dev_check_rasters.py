"""Dev probe: surfel bwd error vs the reference and the reference's own run-to-run noise."""
import sys, torch
sys.path.insert(0, ".")
from pings_b200 import synth, _raster_c as RC
from oracle.ref_api import RefRasterS
sys.path.insert(0, "tests")
from test_rasters_gpu import _scene, CONFIGS
P, W, H, seed, cfg = [int(x) for x in sys.argv[1:6]]
cam, sc, mats, ct, ex = _scene(P, W, H, seed)
view, full, proj, center = mats
config = torch.tensor(CONFIGS[cfg], dtype=torch.float32, device="cuda")
bg = torch.tensor([1.0, 0.5, 0.25], device="cuda")
e = torch.empty(0)
R, color, normal, depth, opac, radii, contributions, geom, binning, img = RC.rasterize_gaussians_surfel(
    bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
    ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, True, config)
ref = RefRasterS()
r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"], config, campos=center)
out = RC.rasterize_gaussians_surfel_backward(
    bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
    ex["prcppoint"], ex["patch_bbox"], cam.tanfovx, cam.tanfovy, ct["color"], ct["normal"], ct["depth"], ct["opac"],
    e, 0, center, geom, R, binning, img, True, config, full_tau=True)
gs = [ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, ex["prcppoint"], ex["patch_bbox"], config,
                   r["radii"], ct["color"], ct["normal"], ct["depth"], ct["opac"], campos=center) for _ in range(2)]
names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations", "tau"]
for n, t in zip(names, out):
    if n == "sh": continue
    a, b, c = t.double(), gs[0][n].double(), gs[1][n].double()
    mx = b.abs().max()
    d = (a - b).abs()
    i = int(d.flatten().argmax())
    row = i // a[0].numel()
    print(f"{n:10s} ours-ref {float(d.max()/mx):.2e}  ref-ref {float((b-c).abs().max()/mx):.2e}  worst row {row} ours {a.flatten()[i]:.6e} ref {b.flatten()[i]:.6e} max {float(mx):.3e}")
    if n == "cov3D":
        print("  conic grad ref:", gs[0]["conic"][row].flatten().tolist(), "radii", int(radii[row]), "scales", sc["scales"][row].tolist())
        print("  ours row", a[row].tolist()); print("  ref  row", b[row].tolist()); print("  ref2 row", c[row].tolist())
