"""Dev diagnostic (GPU box): per-tile walk length of the C5 views (5 M surfels, 1080p): is rs_render_fwd bound by the
longest tile (critical path) or by the sum (throughput)?  usage: python tools/dev_c5_tiles.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pings_b200 import _raster_c as RC, synth   # noqa: E402

P, W, H = 5_000_000, 1920, 1080
dev = torch.device("cuda")
sc = {k: v.to(dev) for k, v in synth.large_map(P, 3).items()}
cams = synth.loop_cameras(64, W, H)
cfg = torch.tensor([1, 1, 1, 1, 1], dtype=torch.float32, device=dev)
bg = torch.zeros(3, device=dev)
e = torch.empty(0)
gx, gy = (W + 15) // 16, (H + 15) // 16
for i in (0, 9, 23, 40):
    cam = cams[i]
    view, full, proj, center = [m.to(dev) for m in cam.matrices()]
    out = RC.rasterize_gaussians_surfel(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
                                        1.0, e, view, full, proj, torch.tensor([cam.cx / W, cam.cy / H], device=dev),
                                        torch.tensor([0.0, 0.0, H - 1.0, W - 1.0], device=dev), cam.tanfovx, cam.tanfovy, H, W,
                                        e, 0, center, False, True, cfg)
    R, geom, binning, img = out[0], out[7], out[8], out[9]
    dv = RC.debug_views(1, P, W, H, R, geom, binning, img)
    rg = dv["ranges"].view(-1, 2).long()
    ln = (rg[:, 1] - rg[:, 0]).float()
    nc = dv["n_contrib"].view(H, W).float()
    pad = torch.zeros(gy * 16, gx * 16, device=dev)
    pad[:H, :W] = nc
    ncmax = pad.view(gy, 16, gx, 16).amax((1, 3)).flatten()
    q = torch.quantile(ncmax, torch.tensor([0.5, 0.9, 0.99, 0.999, 1.0], device=dev))
    ql = torch.quantile(ln, torch.tensor([0.5, 0.9, 0.99, 0.999, 1.0], device=dev))
    slots = 148 * 14
    print(f"view {i}: R={R} tiles={gx * gy} walk(ncmax) sum={ncmax.sum().item():.0f} q50/90/99/99.9/max={[round(x) for x in q.tolist()]} "
          f"len q={[round(x) for x in ql.tolist()]} | sum/slots={ncmax.sum().item() / slots:.0f} vs max={ncmax.max().item():.0f}")
    rows = ncmax.view(gy, gx).sum(1)
    top = torch.topk(rows, 5)
    print("   busiest tile rows:", top.indices.tolist(), [round(x) for x in top.values.tolist()], "of total", round(rows.sum().item()))
