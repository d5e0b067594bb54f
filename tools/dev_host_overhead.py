"""Dev probe (GPU box): host-side cost per call of the Python wrappers, measured with a scene so small that the GPU
is never the bottleneck (P = 256, 64x48 image).  usage: python tools/dev_host_overhead.py"""
import cProfile
import os
import pstats
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pings_b200 import _raster_c as RC, synth   # noqa: E402
from pings_b200.diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer   # noqa: E402
from pings_b200.losses import photo_depth_loss   # noqa: E402

W, H, P = 64, 48, 256
cam = synth.hd_camera(0, W, H)
sc = {k: v.cuda() for k, v in synth.gaussians_in_view(cam, P, 0).items()}
view, full, proj, center = [m.cuda() for m in cam.matrices()]
bg = torch.zeros(3, device="cuda")
e = torch.empty(0)
ct = {k: v.cuda() for k, v in synth.cotangents(cam, 0).items()}
N = 2000


def t(fn, n=N):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e6


def fwd():
    return RC.rasterize_gaussians(bg, sc["means3D"], sc["colors_precomp"], sc["opacities"], sc["scales"], sc["rotations"],
                                  1.0, e, view, full, proj, cam.tanfovx, cam.tanfovy, H, W, e, 0, center, False, False)


res = fwd()


def bwd():
    R, color, radii, geom, binning, img, depth, opac, n_touched = res
    return RC.rasterize_gaussians_backward(bg, sc["means3D"], radii, sc["colors_precomp"], sc["scales"], sc["rotations"],
                                           1.0, e, view, full, proj, cam.tanfovx, cam.tanfovy, ct["color"], ct["depth"],
                                           e, 0, center, geom, R, binning, img, False)


rs = GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, bg, 1.0, view, full, proj, 0, center, False, False)
rast = GaussianRasterizer(rs)
target = torch.rand(3, H, W, device="cuda")
target_depth = torch.rand(1, H, W, device="cuda") * 60 + 1


def step():
    leaves = {k: v.detach().requires_grad_(True) for k, v in sc.items()}
    theta = torch.zeros(1, 3, device="cuda", requires_grad=True)
    rho = torch.zeros(1, 3, device="cuda", requires_grad=True)
    m2d = torch.zeros(P, 3, device="cuda", requires_grad=True)
    color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                   colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                   rotations=leaves["rotations"], theta=theta, rho=rho)
    loss, _, _ = photo_depth_loss(color, target, depth, target_depth, opacity, 0.01, 0.5, 80.0, 0.5)
    loss.backward()
    return loss


def step_noloss():
    leaves = {k: v.detach().requires_grad_(True) for k, v in sc.items()}
    theta = torch.zeros(1, 3, device="cuda", requires_grad=True)
    rho = torch.zeros(1, 3, device="cuda", requires_grad=True)
    m2d = torch.zeros(P, 3, device="cuda", requires_grad=True)
    color, radii, depth, opacity, n_touched = rast(leaves["means3D"], m2d, leaves["opacities"],
                                                   colors_precomp=leaves["colors_precomp"], scales=leaves["scales"],
                                                   rotations=leaves["rotations"], theta=theta, rho=rho)
    torch.autograd.backward([color, depth], [ct["color"], ct["depth"]])


print("us per call: RC forward %.1f, RC backward %.1f, module fwd+bwd %.1f, + fused loss %.1f" %
      (t(fwd), t(bwd), t(step_noloss), t(step)))
os.environ["X"] = "1"
pr = cProfile.Profile()
pr.enable()
for _ in range(500):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
