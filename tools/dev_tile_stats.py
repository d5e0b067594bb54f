"""Dev diagnostic: per-tile list length / max n_contrib distribution of the C2 bench scene."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pings_b200 import synth, _raster_c as RC

def run(P=500000, W=1226, H=370, seed=0):
    cam = synth.kitti_camera(seed, W, H)
    sc = {k: v.cuda() for k, v in synth.gaussians_in_view(cam, P, seed).items()}
    view, full, proj, center = [m.cuda() for m in cam.matrices()]
    bg = torch.ones(3, device="cuda")
    e = torch.empty(0)
    R, color, radii, geom, binning, img, depth, opac, n_touched = RC.rasterize_gaussians(
        bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full, proj,
        cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, True)
    dv = RC.debug_views(0, P, W, H, R, geom, binning, img)
    rg = dv["ranges"].view(-1, 2).long()
    ln = (rg[:, 1] - rg[:, 0]).float()
    gx, gy = (W + 15) // 16, (H + 15) // 16
    nc = dv["n_contrib"].view(H, W).float()
    pad = torch.zeros(gy * 16, gx * 16, device="cuda"); pad[:H, :W] = nc
    ncmax = pad.view(gy, 16, gx, 16).amax((1, 3)).flatten()
    ncmean = pad.view(gy, 16, gx, 16).mean((1, 3)).flatten()
    for name, t in (("len", ln), ("ncmax", ncmax), ("ncmean", ncmean)):
        q = torch.quantile(t, torch.tensor([0.0, 0.25, 0.5, 0.75, 0.9, 0.99, 1.0], device="cuda"))
        print(name, "sum", t.sum().item(), "mean", t.mean().item(), "q", [round(x, 1) for x in q.tolist()])
    top = torch.topk(ncmax, 10)
    print("top ncmax", top.values.tolist(), "len there", ln[top.indices].tolist())
    print("sum ncmax / 148 SMs / 8 CTAs:", ncmax.sum().item() / 148 / 8, " max:", ncmax.max().item())

if __name__ == "__main__":
    run()
