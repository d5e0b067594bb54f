"""Consumer side of the pose gradient (reference utils/campose_utils.py:15-98).

The rasterizers return dL/dtau for a LEFT-multiplicative perturbation of the world-to-camera pose,
tau = [rho (translation), theta (rotation)]: T_w2c <- SE3_exp(tau) @ T_w2c.  These helpers apply that
update.  Unlike the reference (which branches on `angle < 1e-5` in Python, i.e. one device sync per
call), the exponential maps here are branch-free: the small-angle series and the closed form are
blended with torch.where, so they work on batches [..., 3] / [..., 6] and never leave the device.
"""
from __future__ import annotations

import torch


def skew_sym_mat(x: torch.Tensor) -> torch.Tensor:
    """[..., 3] -> [..., 3, 3] cross-product matrix."""
    z = torch.zeros_like(x[..., 0])
    rows = [torch.stack([z, -x[..., 2], x[..., 1]], -1), torch.stack([x[..., 2], z, -x[..., 0]], -1),
            torch.stack([-x[..., 1], x[..., 0], z], -1)]
    return torch.stack(rows, -2)


def _coeffs(theta: torch.Tensor):
    """A = sin a / a, B = (1 - cos a) / a^2, C = (a - sin a) / a^3 with the reference's series below 1e-5."""
    a = torch.linalg.norm(theta, dim=-1)
    small = a < 1e-5
    s = torch.where(small, torch.ones_like(a), a)
    A = torch.where(small, torch.ones_like(a), torch.sin(s) / s)
    B = torch.where(small, torch.full_like(a, 0.5), (1.0 - torch.cos(s)) / (s * s))
    C = torch.where(small, torch.full_like(a, 1.0 / 6.0), (s - torch.sin(s)) / (s * s * s))
    return A[..., None, None], B[..., None, None], C[..., None, None]


def SO3_exp(theta: torch.Tensor) -> torch.Tensor:
    W = skew_sym_mat(theta)
    A, B, _ = _coeffs(theta)
    eye = torch.eye(3, device=theta.device, dtype=theta.dtype)
    return eye + A * W + B * (W @ W)


def V(theta: torch.Tensor) -> torch.Tensor:
    W = skew_sym_mat(theta)
    _, B, C = _coeffs(theta)
    eye = torch.eye(3, device=theta.device, dtype=theta.dtype)
    return eye + B * W + C * (W @ W)


def SE3_exp(tau: torch.Tensor) -> torch.Tensor:
    """[..., 6] (rho, theta) -> [..., 4, 4]."""
    rho, theta = tau[..., :3], tau[..., 3:]
    R = SO3_exp(theta)
    t = (V(theta) @ rho[..., None])[..., 0]
    T = torch.zeros(*tau.shape[:-1], 4, 4, device=tau.device, dtype=tau.dtype)
    T[..., :3, :3] = R
    T[..., :3, 3] = t
    T[..., 3, 3] = 1.0
    return T


def update_pose(camera, converged_threshold: float = 1e-4) -> bool:
    """utils/campose_utils.py:79-98: T_w2c <- SE3_exp([cam_trans_delta, cam_rot_delta]) @ T_w2c, deltas zeroed.
    Returns True when the step is below the threshold (or when pose optimisation is off: zero step)."""
    tau = torch.cat([camera.cam_trans_delta, camera.cam_rot_delta], dim=0)
    tau_norm = float(tau.norm())          # the one host read the convergence test needs
    if tau_norm == 0.0:
        return True
    T_w2c = torch.eye(4, device=tau.device, dtype=tau.dtype)
    T_w2c[:3, :3] = camera.R
    T_w2c[:3, 3] = camera.T
    new_w2c = SE3_exp(tau) @ T_w2c
    camera.set_pose(torch.linalg.inv(new_w2c))   # set_pose takes camera-to-world
    camera.cam_rot_delta.data.fill_(0)
    camera.cam_trans_delta.data.fill_(0)
    return tau_norm < converged_threshold
