"""Dev diagnostic: run pings_b200 3DGS fwd+bwd and the reference build side by side, print stats."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pings_b200 import synth, _raster_c as RC, _lib as L
from oracle.ref_api import RefRaster3

def stats(name, a, b):
    a = a.double().flatten(); b = b.double().flatten()
    d = (a - b).abs()
    denom = b.abs().max().item() + 1e-30
    print(f"  {name:14s} maxabs {d.max().item():.3e}  rel-to-max {d.max().item()/denom:.3e}  mean {d.mean().item():.3e}  ref|max| {denom:.3e}")

def run(P, W, H, seed=0):
    cam = synth.kitti_camera(seed, W, H)
    sc = {k: v.cuda() for k, v in synth.gaussians_in_view(cam, P, seed).items()}
    view, full, proj, center = [m.cuda() for m in cam.matrices()]
    bg = torch.ones(3, device="cuda")
    ct = {k: v.cuda() for k, v in synth.cotangents(cam, seed).items()}
    e = torch.empty(0)
    t0 = time.time()
    R, color, radii, geom, binning, img, depth, opac, n_touched = RC.rasterize_gaussians(
        bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
        cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, True)
    torch.cuda.synchronize()
    print(f"P={P} {W}x{H}: mine R={R} ({time.time()-t0:.3f}s)")
    ref = RefRaster3()
    r = ref.forward(sc, (view, full, proj), W, H, cam.tanfovx, cam.tanfovy, bg, campos=center)
    print(f"  ref R={r['num_rendered']}")
    dv = RC.debug_views(0, P, W, H, R, geom, binning, img)
    print("  radii equal:", torch.equal(radii, r["radii"]), " mismatches:", (radii != r["radii"]).sum().item())
    vis = radii > 0
    print("  depth bits equal:", torch.equal(dv["records"][vis][:, 2, 0].view(torch.int32), r["depths"][vis].view(torch.int32)))
    print("  means2D bits equal:", torch.equal(dv["records"][vis][:, 0, :2].contiguous().view(torch.int32), r["means2D"][vis].contiguous().view(torch.int32)))
    print("  conic bits equal:", torch.equal(dv["records"][vis][:, 1, :].contiguous().view(torch.int32), r["conic_opacity"][vis].contiguous().view(torch.int32)))
    if R == r["num_rendered"]:
        print("  point_list equal:", torch.equal(dv["point_list"], r["point_list"]))
        print("  ranges equal:", torch.equal(dv["ranges"], r["ranges"]))
    print("  n_contrib equal:", torch.equal(dv["n_contrib"], r["n_contrib"]), (dv["n_contrib"] != r["n_contrib"]).sum().item())
    print("  n_touched equal:", torch.equal(n_touched, r["n_touched"]), (n_touched != r["n_touched"]).sum().item())
    stats("color", color, r["color"]); stats("depth", depth, r["depth"]); stats("opacity", opac, r["opacity"])
    stats("final_T", dv["final_T"], r["final_T"])
    # backward
    out = RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e,
                                          view, full, proj, cam.tanfovx, cam.tanfovy, ct["color"], ct["depth"], e, 0, center,
                                          geom, R, binning, img, True, full_tau=True)
    torch.cuda.synchronize()
    g = ref.backward(sc, (view, full, proj), cam.tanfovx, cam.tanfovy, bg, r["radii"], ct["color"], ct["depth"], campos=center)
    names = ["means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations", "tau"]
    for n, t in zip(names, out):
        if n == "sh": continue
        stats("d" + n, t, g[n])
    out2 = RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e,
                                           view, full, proj, cam.tanfovx, cam.tanfovy, ct["color"], ct["depth"], e, 0, center,
                                           geom, R, binning, img, False)
    stats("tau_sum", out2[-1].flatten(), g["tau"].double().sum(0).float())
    print("   tau mine", out2[-1].flatten().tolist())
    print("   tau ref ", g["tau"].double().sum(0).tolist())

if __name__ == "__main__":
    run(2000, 320, 200, 3)
    run(50000, 640, 360, 1)
    run(500000, 1226, 370, 0)
