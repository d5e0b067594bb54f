"""TEST INFRASTRUCTURE — builds the reference's own CUDA code into oracle/_ref/ (git-ignored).

The reference sources are compiled *where they lie* under /root/reference (never copied);
only the small C-ABI shims oracle/ref_*_capi.cu come from this repository.  The resulting
shared objects travel to the GPU box with the gpurun snapshot.  On a machine without
/root/reference (the GPU box) this script is a no-op and the prebuilt files are used.

Flags: sm_100a, `-include cstdint` (rasterizer_impl.h:24,40 use std::uintptr_t / uint32_t
without <cstdint> under gcc 13), GLM from the submodule's third_party/ (SURVEY.md §8c).
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("PINGS_REFERENCE", "/root/reference")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target: str, sources: list[str]) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def _build_raster(name: str, subdir: str, shim: str) -> str | None:
    root = os.path.join(REF, "submodules", subdir)
    if not os.path.isdir(root):
        return None
    srcs = [os.path.join(root, "cuda_rasterizer", f) for f in ("forward.cu", "backward.cu", "rasterizer_impl.cu")]
    shim_path = os.path.join(HERE, shim)
    target = os.path.join(OUT, f"lib{name}.so")
    if _newer(target, srcs + [shim_path]):
        return target
    cmd = [NVCC, "-std=c++17", "-O3", *ARCH, "-include", "cstdint", "-I", os.path.join(root, "third_party", "glm"),
           "-I", root, "-shared", "-Xcompiler", "-fPIC", "-w", "-o", target, *srcs, shim_path]
    subprocess.check_call(cmd)
    return target


def _build_ssim() -> str | None:
    """The reference fused-ssim extension (ssim.cu + ext.cpp, torch C++ API) compiled in place as a
    torch extension module `fused_ssim_cuda_ref` with the reference's own flags (-O3 --use_fast_math,
    FS/setup.py:13-23) for sm_100a.  ~4 min once; cached by mtime."""
    root = os.path.join(REF, "submodules", "fused-ssim")
    srcs = [os.path.join(root, "ssim.cu"), os.path.join(root, "ext.cpp")]
    if not all(os.path.exists(s) for s in srcs):
        return None
    bdir = os.path.join(OUT, "fused_ssim_build")
    target = os.path.join(OUT, "fused_ssim_cuda_ref.so")
    if _newer(target, srcs):
        return target
    os.makedirs(bdir, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    from torch.utils import cpp_extension
    cpp_extension.load(name="fused_ssim_cuda_ref", sources=srcs, build_directory=bdir, is_python_module=False,
                       extra_cflags=["-O3"], extra_cuda_cflags=["-O3", "--use_fast_math", *ARCH], verbose=False)
    import shutil
    shutil.copy(os.path.join(bdir, "fused_ssim_cuda_ref.so"), target)
    return target


PY_STAGE = ["model/__init__.py", "model/neural_gaussians.py", "model/decoder.py", "utils/__init__.py", "utils/config.py",
            "utils/tools.py", "utils/campose_utils.py", "config/run_pin_slam.yaml", "config/run_kitti_gs.yaml"]


def _stage_python() -> str | None:
    """Stage the reference's own PyTorch neural-point / decoder modules into oracle/_ref/py/ (git-ignored, travels
    with the gpurun snapshot like the built .so files) so that bench.py can time the REFERENCE's CPU SDF path on the
    GPU box's host cores (`cpu_baseline.kind = "reference"`, BASELINE.json north_star).  Files are copied verbatim;
    nothing under oracle/_ref/ enters the repository history."""
    import shutil
    dst_root = os.path.join(OUT, "py")
    n = 0
    for rel in PY_STAGE:
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            continue
        dst = os.path.join(dst_root, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copy2(src, dst)
        n += 1
    for pkg in ("model", "utils"):
        init = os.path.join(dst_root, pkg, "__init__.py")
        if os.path.isdir(os.path.dirname(init)) and not os.path.exists(init):
            open(init, "w").close()
    return dst_root if n else None


def build(verbose: bool = True) -> dict[str, str | None]:
    if not os.path.isdir(REF):
        if verbose:
            print(f"[oracle] {REF} absent: using prebuilt oracle/_ref as is")
        return {}
    os.makedirs(OUT, exist_ok=True)
    res = {
        "ref_r3": _build_raster("ref_r3", "diff-gaussian-rasterization-w-pose", "ref_r3_capi.cu"),
    }
    if os.path.exists(os.path.join(HERE, "ref_rs_capi.cu")):
        res["ref_rs"] = _build_raster("ref_rs", "diff-gaussian-surfel-rasterization-w-pose", "ref_rs_capi.cu")
    res["ref_ssim"] = _build_ssim()
    res["ref_python"] = _stage_python()
    if verbose:
        for k, v in res.items():
            print(f"[oracle] {k}: {v}")
    return res


if __name__ == "__main__":
    build()
    sys.exit(0)
