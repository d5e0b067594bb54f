"""Dev probe: where the e2e step (bench.py) spends its time beyond the resident kernels."""
import sys, time
import torch
sys.path.insert(0, ".")
import bench
from pings_b200 import _raster_c as RC
from pings_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

dev = torch.device("cuda", 0)
cam, sc_cpu, ct_cpu, target_cpu = bench.make_scene(0)
P, W, H = bench.C2["P"], bench.C2["W"], bench.C2["H"]
sc = {k: v.to(dev) for k, v in sc_cpu.items()}
ct = {k: v.to(dev) for k, v in ct_cpu.items()}
view, full, proj, center = [m.to(dev) for m in cam.matrices()]
bg = torch.ones(3, device=dev)
e = torch.empty(0)
target = target_cpu.to(dev)
rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, full, proj, 0, center, False, False)
rast = GaussianRasterizer(rs)


def timeit(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


def raw():
    R, color, radii, geom, binning, img, depth, opac, n_touched = RC.rasterize_gaussians(
        bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"], 1.0, e, view, full,
        proj, cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, False)
    RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"], 1.0, e,
                                    view, full, proj, cam.tanfovx, cam.tanfovy, ct["color"], ct["depth"], e, 0, center,
                                    geom, R, binning, img, False)


def raw_sync():
    raw()
    torch.cuda.synchronize()


def module(loss_kind):
    leaves = {k: sc[k].detach().requires_grad_(True) for k in sc}
    theta = torch.zeros(1, 3, device=dev, requires_grad=True)
    rho = torch.zeros(1, 3, device=dev, requires_grad=True)
    m2d = torch.zeros(P, 3, device=dev, requires_grad=True)
    color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                   colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                   rotations=leaves["rotations"], theta=theta, rho=rho)
    if loss_kind == "sum":
        torch.autograd.backward([color, depth], [ct["color"], ct["depth"]])
    else:
        loss = (color - target).abs().mean() + 0.01 * depth.mean()
        loss.backward()
    return rho


def module_sync(kind):
    r = module(kind)
    float(r.grad.flatten()[0])


print("raw C-ABI fwd+bwd, async          : %.3f ms" % timeit(raw))
print("raw C-ABI fwd+bwd, sync per step  : %.3f ms" % timeit(raw_sync))
print("autograd module, given cotangents : %.3f ms" % timeit(lambda: module("sum")))
print("autograd module, L1+depth loss    : %.3f ms" % timeit(lambda: module("l1")))
print("  ... + D2H read per step         : %.3f ms" % timeit(lambda: module_sync("l1")))
