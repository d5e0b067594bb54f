"""Dev probe: GPU busy/idle timeline of the autograd-module step (no H2D), via torch.profiler."""
import sys, time
import torch
from torch.profiler import profile, ProfilerActivity
sys.path.insert(0, ".")
import bench
from pings_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from pings_b200.losses import photo_depth_loss

dev = torch.device("cuda", 0)
cam, sc_cpu, ct_cpu, target_cpu = bench.make_scene(0)
P, W, H = bench.C2["P"], bench.C2["W"], bench.C2["H"]
sc = {k: v.to(dev) for k, v in sc_cpu.items()}
view, full, proj, center = [m.to(dev) for m in cam.matrices()]
bg = torch.ones(3, device=dev)
target = target_cpu.to(dev)
tdepth = torch.rand(1, H, W, device=dev) * 60 + 1
rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, full, proj, 0, center, False, False)
rast = GaussianRasterizer(rs)

def step():
    leaves = {k: sc[k].detach().requires_grad_(True) for k in sc}
    theta = torch.zeros(1, 3, device=dev, requires_grad=True)
    rho = torch.zeros(1, 3, device=dev, requires_grad=True)
    m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
    color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                   colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                   rotations=leaves["rotations"], theta=theta, rho=rho)
    loss, _, _ = photo_depth_loss(color, target, depth, tdepth, opacity, 0.01, 0.5, 80.0, 0.5)
    loss.backward()

for _ in range(5): step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(6): step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start; t1 = max(e.time_range.end for e in ev)
busy = 0; cur_end = t0; gaps = []
for e in ev:
    s, en = e.time_range.start, e.time_range.end
    if s > cur_end:
        gaps.append((s - cur_end, e.name[:50]))
    busy += max(0, en - max(s, cur_end)); cur_end = max(cur_end, en)
print("span us", t1 - t0, "busy us", busy, "per step", (t1 - t0) / 6, "idle per step", (t1 - t0 - busy) / 6)
gaps.sort(reverse=True)
for g in gaps[:14]: print(round(g[0], 1), g[1])
from collections import Counter
c = Counter()
for e in ev: c[e.name[:45]] += (e.time_range.end - e.time_range.start) / 6
for k, v in c.most_common(22): print(round(v, 1), k)
