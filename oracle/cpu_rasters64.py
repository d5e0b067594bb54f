"""TEST INFRASTRUCTURE — ctypes wrapper of oracle/cpu_rasters64.c, the fp64 restatement of the reference
surfel rasterizer's blend (forward + backward) and per-surfel backward chain (RS cuda_rasterizer/
forward.cu:375-624, backward.cu:160-914).  Imported only by tests/ and tools/ (the arbiter between the CUDA
implementation and the reference CUDA build, whose float-atomic backward is noisy on ill-conditioned rows).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import torch

from .cpu_raster import _cpu_tag

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "cpu_rasters64.c")
_lib = None
vp = C.c_void_p


def build() -> str:
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    target = os.path.join(out_dir, f"libcpu_rasters64_{_cpu_tag()}.so")
    if not os.path.exists(target) or os.path.getmtime(target) < os.path.getmtime(_SRC):
        # no -ffast-math and no FMA contraction: this is the exact-arithmetic arbiter
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared", "-o", target, _SRC, "-lm"])
    return target


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.rs64_blend.restype = None
        L.rs64_blend.argtypes = ([C.c_int] * 3 + [vp] * 5 + [C.c_float] + [vp] * 3 + [C.c_float] * 2 + [vp] * 8
                                 + [vp] * 5 + [vp] * 6)
        L.rs64_chain.restype = None
        L.rs64_chain.argtypes = ([C.c_int] * 3 + [vp] * 2 + [C.c_float] + [vp] * 4 + [C.c_float] * 2 + [vp] * 2
                                 + [vp] * 4 + [vp] * 5)
        L.rs64_set_thresholds.restype = None
        L.rs64_set_thresholds.argtypes = [C.c_double, C.c_double]
        _lib = L
    return _lib


def _f(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to("cpu", torch.float32).contiguous()


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def blend(scene: dict, cam_mats, W, H, tanx, tany, bg, config, radii, point_list, ranges, ct: dict, mod=1.0,
          images=True) -> dict:
    """fp64 blend forward + backward on the given tile lists.  `ct` holds the four image cotangents
    (color [3,H,W], normal [3,H,W], depth [1,H,W], opac [1,H,W]).  Returns double CPU tensors."""
    L = lib()
    view, proj = _f(cam_mats[0]), _f(cam_mats[1])
    means, colors, opac = _f(scene["means3D"]), _f(scene["colors_precomp"]), _f(scene["opacities"])
    scales, rots = _f(scene["scales"]), _f(scene["rotations"])
    P = means.shape[0]
    bgc, cfg = _f(bg), _f(config)
    rad = radii.detach().to("cpu", torch.int32).contiguous()
    pl = point_list.detach().to("cpu", torch.int32).contiguous()
    rg = ranges.detach().to("cpu", torch.int32).contiguous()
    dC, dN, dD, dO = _f(ct["color"]), _f(ct["normal"]), _f(ct["depth"]), _f(ct["opac"])
    z = lambda *s: torch.zeros(*s, dtype=torch.float64)
    out = dict(color=z(3, H, W), normal=z(3, H, W), depth=z(1, H, W), opac=z(1, H, W)) if images else {}
    nc = torch.zeros(H * W, dtype=torch.int32)
    g = dict(means2D=z(P, 2), conic=z(P, 3), opacity=z(P, 1), colors=z(P, 3), normal=z(P, 3), depths=z(P, 1))
    L.rs64_blend(P, W, H, _p(bgc), _p(means), _p(colors), _p(opac), _p(scales), float(mod), _p(rots), _p(view), _p(proj),
                 float(tanx), float(tany), _p(cfg), _p(rad), _p(pl), _p(rg), _p(dC), _p(dN), _p(dD), _p(dO),
                 _p(out.get("color")), _p(out.get("normal")), _p(out.get("depth")), _p(out.get("opac")), _p(nc),
                 _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]), _p(g["normal"]), _p(g["depths"]))
    out["n_contrib"] = nc
    out.update({f"g_{k}": v for k, v in g.items()})
    return out


def chain(scene: dict, cam_mats, W, H, tanx, tany, config, radii, means2D, conic, normal, depths, mod=1.0) -> dict:
    """fp64 per-surfel backward chain from blend sums (any float dtype; converted to double):
    means2D [P,>=2], conic [P,3] as (x, y, w), normal [P,3], depths [P] or [P,1]."""
    L = lib()
    view, proj, proj_raw = _f(cam_mats[0]), _f(cam_mats[1]), _f(cam_mats[2])
    means, scales, rots = _f(scene["means3D"]), _f(scene["scales"]), _f(scene["rotations"])
    P = means.shape[0]
    cfg = _f(config)
    rad = radii.detach().to("cpu", torch.int32).contiguous()
    d = lambda t, n: t.detach().to("cpu", torch.float64).reshape(P, -1)[:, :n].contiguous()
    m2, cn, nm, dp = d(means2D, 2), d(conic, 3), d(normal, 3), d(depths, 1)
    z = lambda *s: torch.zeros(*s, dtype=torch.float64)
    g = dict(means3D=z(P, 3), cov3D=z(P, 6), scales=z(P, 3), rotations=z(P, 4), tau=z(P, 6))
    L.rs64_chain(P, W, H, _p(means), _p(scales), float(mod), _p(rots), _p(view), _p(proj), _p(proj_raw), float(tanx),
                 float(tany), _p(cfg), _p(rad), _p(m2), _p(cn), _p(nm), _p(dp), _p(g["means3D"]), _p(g["cov3D"]),
                 _p(g["scales"]), _p(g["rotations"]), _p(g["tau"]))
    return g
