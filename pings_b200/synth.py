"""Seeded synthetic inputs for the BASELINE.json configs.

Everything is generated on the CPU with a torch.Generator so that tests, bench.py and the
golden-fixture scripts see identical data; nothing here touches the reference checkout.
Camera conventions follow gaussian_splatting/utils/cameras.py:57-75,208-216 and
graphics_utils.py:51-73 of the reference (projection_matrix stored transposed,
world_view_transform = T_cw.T, full_proj_transform = world_view_transform @ projection_matrix).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch


def se3_exp(tau: torch.Tensor) -> torch.Tensor:
    """SE(3) exponential of tau = (rho, theta) (utils/campose_utils.py:28-61)."""
    rho, theta = tau[:3].double(), tau[3:].double()
    W = torch.tensor([[0, -theta[2], theta[1]], [theta[2], 0, -theta[0]], [-theta[1], theta[0], 0]], dtype=torch.float64)
    W2 = W @ W
    a = float(theta.norm())
    I = torch.eye(3, dtype=torch.float64)
    if a < 1e-5:
        R = I + W + 0.5 * W2
        V = I + 0.5 * W + W2 / 6.0
    else:
        R = I + math.sin(a) / a * W + (1 - math.cos(a)) / (a * a) * W2
        V = I + (1 - math.cos(a)) / (a * a) * W + (a - math.sin(a)) / (a ** 3) * W2
    T = torch.eye(4, dtype=torch.float64)
    T[:3, :3] = R
    T[:3, 3] = V @ rho
    return T


def projection_matrix(znear, zfar, fx, fy, cx, cy, W, H) -> torch.Tensor:
    """getProjectionMatrix (graphics_utils.py:51-73), returned un-transposed."""
    top = znear * cy / fy
    bottom = -znear * (H - cy) / fy
    right = znear * (W - cx) / fx
    left = -znear * cx / fx
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = -(right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


@dataclass
class Camera:
    W: int
    H: int
    fx: float
    fy: float
    cx: float
    cy: float
    T_wc: torch.Tensor  # [4,4] camera-to-world, float64
    znear: float = 0.1
    zfar: float = 100.0

    @property
    def tanfovx(self) -> float:
        return self.W / (2.0 * self.fx)

    @property
    def tanfovy(self) -> float:
        return self.H / (2.0 * self.fy)

    def matrices(self):
        """(world_view_transform, full_proj_transform, projection_matrix, camera_center), fp32."""
        T_cw = torch.linalg.inv(self.T_wc).float()
        wvt = T_cw.T.contiguous()
        proj = projection_matrix(self.znear, self.zfar, self.fx, self.fy, self.cx, self.cy, self.W, self.H).T.contiguous()
        full = (wvt @ proj).contiguous()
        center = torch.linalg.inv(wvt)[3, :3].contiguous()
        return wvt, full, proj, center


def kitti_camera(seed: int = 0, W: int = 1226, H: int = 370) -> Camera:
    g = torch.Generator().manual_seed(seed)
    tau = torch.randn(6, generator=g) * 0.05
    return Camera(W=W, H=H, fx=718.856, fy=718.856, cx=607.19 * W / 1226.0, cy=185.22 * H / 370.0, T_wc=se3_exp(tau))


def hd_camera(seed: int = 1, W: int = 1920, H: int = 1080) -> Camera:
    g = torch.Generator().manual_seed(seed)
    tau = torch.randn(6, generator=g) * 0.05
    return Camera(W=W, H=H, fx=1000.0 * W / 1920.0, fy=1000.0 * W / 1920.0, cx=W / 2.0, cy=H / 2.0, T_wc=se3_exp(tau))


def gaussians_in_view(cam: Camera, P: int, seed: int = 0, surfel: bool = False,
                      scale_lo: float = 0.01, scale_hi: float = 0.3) -> dict:
    """Config C2/C3 scene: 90 % of the Gaussians inside (slightly beyond) the frustum with
    log-uniform depth in [2, 80] m, 10 % uniform in a 160 m cube (exercises culling)."""
    g = torch.Generator().manual_seed(seed + 1000)
    n_in = int(P * 0.9)
    u = torch.rand(n_in, 2, generator=g) * 2.2 - 1.1
    depth = torch.exp(torch.rand(n_in, generator=g) * (math.log(80.0) - math.log(2.0)) + math.log(2.0))
    # NDC -> camera frame: x_ndc = (2 fx x/z + (W - 2cx)) / W  (from the projection matrix)
    xc = (u[:, 0] * cam.W - (cam.W - 2 * cam.cx)) / (2 * cam.fx) * depth
    yc = (u[:, 1] * cam.H - (cam.H - 2 * cam.cy)) / (2 * cam.fy) * depth
    pc = torch.stack([xc, yc, depth, torch.ones_like(depth)], dim=1).double()
    pw = (cam.T_wc @ pc.T).T[:, :3].float()
    cube = (torch.rand(P - n_in, 3, generator=g) - 0.5) * 160.0
    means = torch.cat([pw, cube], dim=0)
    perm = torch.randperm(P, generator=g)
    means = means[perm].contiguous()
    scales = torch.exp(torch.rand(P, 3, generator=g) * (math.log(scale_hi) - math.log(scale_lo)) + math.log(scale_lo))
    rots = torch.nn.functional.normalize(torch.randn(P, 4, generator=g), dim=1)
    if surfel:
        scales[:, 2] = 1e-7
    opac = torch.rand(P, 1, generator=g) * 0.95 + 0.05
    colors = torch.rand(P, 3, generator=g)
    return dict(means3D=means, scales=scales.contiguous(), rotations=rots.contiguous(), opacities=opac,
                colors_precomp=colors)


def cotangents(cam: Camera, seed: int = 0, normal: bool = False) -> dict:
    g = torch.Generator().manual_seed(seed + 2000)
    hw = cam.W * cam.H
    out = dict(color=torch.randn(3, cam.H, cam.W, generator=g) / hw, depth=torch.randn(1, cam.H, cam.W, generator=g) / hw)
    if normal:
        out["normal"] = torch.randn(3, cam.H, cam.W, generator=g) / hw
        out["opac"] = torch.randn(1, cam.H, cam.W, generator=g) / hw
    return out


def surfels_in_view(cam: Camera, P: int, seed: int = 1, facing: float = 0.7, **kw) -> dict:
    """Config C3 scene: `gaussians_in_view(surfel=True)` with the quaternions re-oriented
    so that about `facing` of the surfels face the camera (normal = third column of R, as in
    RS forward.cu:274-276; a surfel is front-facing when dot(p - campos, n) < 0)."""
    sc = gaussians_in_view(cam, P, seed, surfel=True, **kw)
    g = torch.Generator().manual_seed(seed + 3000)
    q = sc["rotations"]
    w, x, y, z = q.unbind(1)
    n = torch.stack([2 * (x * z + w * y), 2 * (y * z - w * x), 1 - 2 * (x * x + y * y)], dim=1)
    campos = cam.T_wc[:3, 3].float()
    back = ((sc["means3D"] - campos) * n).sum(1) > 0
    frac_back = float(back.float().mean())
    p_flip = 0.0 if frac_back <= 0 else max(0.0, min(1.0, (frac_back - (1.0 - facing)) / frac_back))
    flip = back & (torch.rand(P, generator=g) < p_flip)
    qf = torch.stack([-x, w, z, -y], dim=1)  # q * (0,1,0,0): half turn about the local x axis flips the normal
    sc["rotations"] = torch.where(flip[:, None], qf, q).contiguous()
    return sc


# ---- config C5 (BASELINE.json configs[4]): inspect_pings-style large-map render ----------------
def large_map(P: int = 5_000_000, seed: int = 3) -> dict:
    """P surfels in a 400 m x 400 m x 20 m slab, street-like: 60 % within 2 m of the ground (normals up, random
    yaw, slight tilt), 40 % anywhere in the slab with random orientation.  Same dictionary as surfels_in_view."""
    g = torch.Generator().manual_seed(seed + 5000)
    n_ground = int(P * 0.6)
    xy = (torch.rand(P, 2, generator=g) - 0.5) * 400.0
    z = torch.cat([torch.rand(n_ground, generator=g) * 2.0, torch.rand(P - n_ground, generator=g) * 20.0])
    means = torch.cat([xy, z[:, None]], 1).contiguous()
    yaw = torch.rand(n_ground, generator=g) * (2 * math.pi)
    qg = torch.stack([torch.cos(yaw / 2), 0.05 * torch.randn(n_ground, generator=g),
                      0.05 * torch.randn(n_ground, generator=g), torch.sin(yaw / 2)], 1)
    qr = torch.randn(P - n_ground, 4, generator=g)
    rots = torch.nn.functional.normalize(torch.cat([qg, qr], 0), dim=1).contiguous()
    scales = torch.exp(torch.rand(P, 3, generator=g) * (math.log(0.4) - math.log(0.05)) + math.log(0.05))
    scales[:, 2] = 1e-7
    opac = torch.rand(P, 1, generator=g) * 0.95 + 0.05
    colors = torch.rand(P, 3, generator=g)
    perm = torch.randperm(P, generator=g)
    return dict(means3D=means[perm].contiguous(), scales=scales[perm].contiguous(), rotations=rots[perm].contiguous(),
                opacities=opac[perm].contiguous(), colors_precomp=colors[perm].contiguous())


def loop_cameras(n: int = 64, W: int = 1920, H: int = 1080, radius: float = 100.0, height: float = 1.7) -> list:
    """n poses on a closed loop (circle of `radius` m, `height` m above the ground), looking along the tangent;
    camera axes: z forward, x right, y down (cameras.py convention)."""
    cams = []
    for i in range(n):
        phi = 2 * math.pi * i / n
        c = torch.tensor([radius * math.cos(phi), radius * math.sin(phi), height], dtype=torch.float64)
        zc = torch.tensor([-math.sin(phi), math.cos(phi), 0.0], dtype=torch.float64)
        yc = torch.tensor([0.0, 0.0, -1.0], dtype=torch.float64)
        xc = torch.linalg.cross(yc, zc)
        T = torch.eye(4, dtype=torch.float64)
        T[:3, 0], T[:3, 1], T[:3, 2], T[:3, 3] = xc, yc, zc, c
        cams.append(Camera(W=W, H=H, fx=1000.0 * W / 1920.0, fy=1000.0 * W / 1920.0, cx=W / 2.0, cy=H / 2.0, T_wc=T))
    return cams


def pings_like_gaussians(cam: Camera, P: int = 300_000, seed: int = 7) -> dict:
    """A second 3DGS scene with PINGS-like occupancy (R/P about 5-15, tile lists that do not saturate
    after a few percent): spawned-Gaussian sized splats (0.02-0.15 m) on depth shells with low opacity, all inside
    the frustum - the shape a PINGS mapping frame has (GR:661-685: scales <= voxel size, alpha = tanh(.))."""
    sc = gaussians_in_view(cam, P, seed, scale_lo=0.02, scale_hi=0.15)
    g = torch.Generator().manual_seed(seed + 4000)
    sc["opacities"] = (torch.rand(P, 1, generator=g) * 0.35 + 0.05).contiguous()
    return sc
