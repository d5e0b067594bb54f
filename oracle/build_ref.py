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
    shim_path = os.path.join(HERE, "ref_ssim_capi.cu")
    src = os.path.join(REF, "submodules", "fused-ssim", "ssim.cu")
    if not (os.path.exists(src) and os.path.exists(shim_path)):
        return None
    # ssim.cu includes torch/extension.h: build it as a torch extension-free object is not
    # possible, so the SSIM reference is compiled through the shim which includes the kernel
    # source text with the torch host part compiled out (see ref_ssim_capi.cu).
    target = os.path.join(OUT, "libref_ssim.so")
    if _newer(target, [src, shim_path]):
        return target
    cmd = [NVCC, "-std=c++17", "-O3", "--use_fast_math", *ARCH, "-I", os.path.dirname(src), "-shared", "-Xcompiler", "-fPIC", "-w",
           "-o", target, shim_path]
    subprocess.check_call(cmd)
    return target


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
    if verbose:
        for k, v in res.items():
            print(f"[oracle] {k}: {v}")
    return res


if __name__ == "__main__":
    build()
    sys.exit(0)
