"""Builds the native parts in-tree (no network, no JIT cache):

  python/vulkan_forge/libvfcuda.so      C-ABI library: hand-written sm_100a CUDA kernels (nvcc)
  python/vulkan_forge/_vulkan_forge*.so pybind11 host module above the C ABI (g++)

nvcc cross-compiles for sm_100a without a GPU.  Flags that are part of the arithmetic contract
(DESIGN.md): --fmad=false (every FMA in the kernels is an explicit fmaf) and -ffp-contract=off
for host code.  -lineinfo keeps ncu's source page usable.
"""
from __future__ import annotations

import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
PKG = os.path.join(HERE, "python", "vulkan_forge")
NVCC = os.environ.get("VF_NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("VF_CXX", "/usr/bin/g++")

CUDA_SRCS = ["vf_context.cu", "vf_terrain.cu", "vf_dem.cu", "vf_misc.cu", "vf_png_dev.cu", "vf_png.cpp"]
CUDA_DEPS = CUDA_SRCS + ["vf_device.cuh", "vf_terrain_kernels.cuh", "vf_geom_march.cuh", "vf_raster_lean.cuh", "vf_internal.h", "vf_png_dev.h", "vf_png_kernels.cuh", "vf_png_encode.cuh", "vf_png_encode_tail.inc", "vf_png_encode_emit.inc", "vf_png_pack.cuh", "vf_lut_data.inc",
                         "../../include/vf_cuda.h"]
HOST_DEPS = ["host/module.cpp", "host/module_renderer.inc", "host/vf_camera.hpp", "../../include/vf_cuda.h"]


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in deps)


def lib_path() -> str:
    return os.path.join(PKG, "libvfcuda.so")


def module_path() -> str:
    return os.path.join(PKG, "_vulkan_forge" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    out = lib_path()
    if force or _stale(out, CUDA_DEPS):
        cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "--fmad=false",
               "-std=c++17", "-ccbin", CXX, "-Xcompiler", "-fPIC,-ffp-contract=off",
               "-shared", "-o", out] + [os.path.join(CSRC, s) for s in CUDA_SRCS] + ["-lz"]
        cmd[1:1] = os.environ.get("VF_NVCC_EXTRA", "").split()      # e.g. -DVF_GM_WARPS=6 -DVF_GM_CTAS=3 for tuning runs
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        subprocess.check_call(cmd)
    return out


def build_module(force: bool = False) -> str:
    import pybind11
    out = module_path()
    if force or _stale(out, HOST_DEPS) or os.path.getmtime(lib_path()) > os.path.getmtime(out):
        cmd = [CXX, "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fvisibility=hidden",
               "-I", pybind11.get_include(), "-I", sysconfig.get_paths()["include"],
               os.path.join(CSRC, "host", "module.cpp"), "-o", out,
               "-L", PKG, "-l:libvfcuda.so", "-Wl,-rpath,$ORIGIN"]
        subprocess.check_call(cmd)
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_cuda(force, verbose), build_module(force)


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
