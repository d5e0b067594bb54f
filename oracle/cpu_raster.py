"""TEST INFRASTRUCTURE — ctypes wrapper of oracle/cpu_raster.c (plain-C restatement of the reference
3DGS rasterizer, R3).  Imported only by tests/ and bench.py's `cpu_baseline` leg.

The C file is compiled on first use with the host's own gcc (`-O3 -march=native -fopenmp`) into
oracle/_build/ (git-ignored; keyed by the CPU model so a library built in one container is never run
on another micro-architecture).
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import platform
import subprocess
import time

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "cpu_raster.c")
_lib = None
vp = C.c_void_p


def _cpu_tag() -> str:
    model = platform.processor()
    try:
        with open("/proc/cpuinfo") as f:
            for ln in f:
                if ln.startswith("model name") or ln.startswith("flags"):
                    model += ln
                    if ln.startswith("flags"):
                        break
    except OSError:
        pass
    return hashlib.sha1(model.encode()).hexdigest()[:10]


def build() -> str:
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    target = os.path.join(out_dir, f"libcpu_raster_{_cpu_tag()}.so")
    if not os.path.exists(target) or os.path.getmtime(target) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O3", "-march=native", "-fopenmp", "-fPIC", "-shared", "-o", target, _SRC, "-lm"])
    return target


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.cr_forward.restype = vp
        L.cr_forward.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, vp, C.c_float, vp, vp, vp, C.c_float,
                                 C.c_float, vp, vp, vp, vp, vp]
        L.cr_backward.restype = None
        L.cr_backward.argtypes = [vp, vp, vp, vp, vp, C.c_float, vp, vp, vp, vp, C.c_float, C.c_float, vp, vp, vp] + [vp] * 10
        L.cr_free.argtypes = [vp]
        L.cr_num_rendered.restype = C.c_int64
        L.cr_num_rendered.argtypes = [vp]
        for n in ("cr_point_list", "cr_ranges", "cr_n_contrib"):
            getattr(L, n).restype = vp
            getattr(L, n).argtypes = [vp]
        _lib = L
    return _lib


def _c(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to("cpu", torch.float32).contiguous()


class CpuRaster3:
    """forward()/backward() on CPU tensors; same dictionaries as oracle.ref_api.RefRaster3."""

    def __init__(self):
        self.L = lib()
        self.h = None

    def __del__(self):
        if getattr(self, "h", None):
            self.L.cr_free(self.h)

    def forward(self, scene: dict, cam_mats, W, H, tanx, tany, bg, mod=1.0, want_views=True):
        view, proj, proj_raw = [_c(m) for m in cam_mats]
        sc = {k: _c(v) for k, v in scene.items()}
        bg = _c(bg)
        P = sc["means3D"].shape[0]
        out_color, out_depth, out_opacity = torch.zeros(3, H, W), torch.zeros(1, H, W), torch.zeros(1, H, W)
        radii = torch.zeros(P, dtype=torch.int32)
        n_touched = torch.zeros(P, dtype=torch.int32)
        if self.h:
            self.L.cr_free(self.h)
        self.h = self.L.cr_forward(P, W, H, bg.data_ptr(), sc["means3D"].data_ptr(), sc["colors_precomp"].data_ptr(),
                                   sc["opacities"].data_ptr(), sc["scales"].data_ptr(), mod, sc["rotations"].data_ptr(),
                                   view.data_ptr(), proj.data_ptr(), tanx, tany, out_color.data_ptr(),
                                   out_depth.data_ptr(), out_opacity.data_ptr(), radii.data_ptr(), n_touched.data_ptr())
        self._keep = (sc, view, proj, proj_raw, bg, W, H, tanx, tany, mod)
        R = int(self.L.cr_num_rendered(self.h))
        res = dict(num_rendered=R, color=out_color, depth=out_depth, opacity=out_opacity, radii=radii,
                   n_touched=n_touched)
        if want_views:
            T = ((W + 15) // 16) * ((H + 15) // 16)

            def arr(p, n):
                if n == 0:
                    return torch.zeros(0, dtype=torch.int32)
                buf = (C.c_int32 * n).from_address(p)
                return torch.frombuffer(buf, dtype=torch.int32).clone()
            res["point_list"] = arr(self.L.cr_point_list(self.h), R)
            res["ranges"] = arr(self.L.cr_ranges(self.h), 2 * T).view(T, 2)
            res["n_contrib"] = arr(self.L.cr_n_contrib(self.h), W * H)
        return res

    def backward(self, radii, dL_dcolor, dL_ddepth):
        sc, view, proj, proj_raw, bg, W, H, tanx, tany, mod = self._keep
        P = sc["means3D"].shape[0]
        z = lambda *s: torch.zeros(*s)
        g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), depths=z(P, 1), means3D=z(P, 3),
                 cov3D=z(P, 6), scales=z(P, 3), rotations=z(P, 4), tau=z(P, 6))
        dc, dd, radii = _c(dL_dcolor), _c(dL_ddepth), radii.to("cpu", torch.int32).contiguous()
        self.L.cr_backward(self.h, bg.data_ptr(), sc["means3D"].data_ptr(), sc["colors_precomp"].data_ptr(),
                           sc["scales"].data_ptr(), mod, sc["rotations"].data_ptr(), view.data_ptr(), proj.data_ptr(),
                           proj_raw.data_ptr(), tanx, tany, radii.data_ptr(), dc.data_ptr(), dd.data_ptr(),
                           g["means2D"].data_ptr(), g["conic"].data_ptr(), g["opacity"].data_ptr(),
                           g["colors"].data_ptr(), g["depths"].data_ptr(), g["means3D"].data_ptr(),
                           g["cov3D"].data_ptr(), g["scales"].data_ptr(), g["rotations"].data_ptr(), g["tau"].data_ptr())
        return g


def bench_sample(cfg: dict, seconds_budget: float = 20.0) -> dict:
    """Timed CPU arm of bench.py: fwd+bwd of the C2 view on all host cores.  The scene is the first
    `n` Gaussians of the C2 scene, n chosen so that one pass fits the budget (the full 500k-Gaussian
    view is used when it does)."""
    from pings_b200 import synth
    cam = synth.kitti_camera(cfg["seed"], cfg["W"], cfg["H"])
    full = synth.gaussians_in_view(cam, cfg["P"], cfg["seed"])
    ct = synth.cotangents(cam, cfg["seed"])
    view, fullp, proj, center = cam.matrices()
    bg = torch.ones(3)
    W, H = cfg["W"], cfg["H"]

    def run(n):
        sc = {k: v[:n].contiguous() for k, v in full.items()}
        cr = CpuRaster3()
        t0 = time.perf_counter()
        r = cr.forward(sc, (view, fullp, proj), W, H, cam.tanfovx, cam.tanfovy, bg, want_views=False)
        cr.backward(r["radii"], ct["color"], ct["depth"])
        return time.perf_counter() - t0, r["num_rendered"]

    n = min(cfg["P"], 50_000)
    t, R = run(n)
    while n < cfg["P"] and t * (min(cfg["P"], n * 4) / n) < seconds_budget / 2:
        n = min(cfg["P"], n * 4)
        t, R = run(n)
    times = [t]
    t_end = time.perf_counter() + max(0.0, seconds_budget - t)
    while len(times) < 3 and time.perf_counter() < t_end:
        times.append(run(n)[0])
    best = min(times)
    return {"value": round(W * H / best / 1e6, 4), "unit": "Mpix/s", "cores": os.cpu_count() or 1, "kind": "port",
            "sample": f"C2 view {W}x{H}, first {n} of {cfg['P']} Gaussians (num_rendered {R}), fwd+bwd with pose "
                      f"grads, oracle/cpu_raster.c (gcc -O3 -march=native -fopenmp), best of {len(times)} runs, "
                      f"{best:.2f} s per pass"}
