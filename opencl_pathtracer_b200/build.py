#!/usr/bin/env python3
"""Builds the native libraries IN-TREE (opencl_pathtracer_b200/_lib/):

  libpathtracer_cuda.so   the product: sm_100a kernels + the C ABI of include/pathtracer_cuda.h
  libpathtracer_scene.so  scene side of the host mirror of the reference interface (importers, BVH builder) + pth_scene_*;
                          does not depend on the CUDA library
  libpathtracer_host.so   controller side (PathTracer_Main, CUDA_* shims, headless dialog) + pth_main / pth_run_default
  _lib_checked/           the same three libraries with the device code compiled -DPTC_CHECKED (bounds checks, ptc_counters.faults)
  pathtracer_alone        headless stand-alone front-end (counterpart of the reference's Alone/main.cpp)

nvcc cross-compiles for sm_100a without a GPU. Bit-exact arithmetic: -fmad=false on the device,
-ffp-contract=off on the host.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "_lib")
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
INC = os.path.join(ROOT, "include")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC,-fopenmp,-ffp-contract=off,-fvisibility=hidden", "-I", INC, "-I", CSRC]
CXX_FLAGS = ["-std=c++17", "-O2", "-fPIC", "-fopenmp", "-ffp-contract=off", "-fno-fast-math", "-Wall", "-Wno-unknown-pragmas",
             "-I", INC, "-I", HOST]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def _build_cuda(out_dir, extra, verbose, force):
    os.makedirs(out_dir, exist_ok=True)
    cuda_so = os.path.join(out_dir, "libpathtracer_cuda.so")
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + \
              [os.path.join(INC, f) for f in os.listdir(INC)]
    cu = [os.path.join(CSRC, "pt_kernels.cu"), os.path.join(CSRC, "pt_wide_build.cu"), os.path.join(CSRC, "pt_lbvh.cu"),
          os.path.join(CSRC, "pathtracer_cuda.cu"), os.path.join(CSRC, "pt_multi.cu"), os.path.join(CSRC, "pt_static.cu")]
    if force or _newer(cuda_so, cu + headers + [__file__]):
        objs = []
        for src in cu:
            obj = os.path.join(out_dir, os.path.basename(src) + ".o")
            _run(["nvcc"] + NVCC_FLAGS + extra + ["-c", src, "-o", obj], verbose)
            objs.append(obj)
        _run(["nvcc", "-shared", "-o", cuda_so] + objs + ["-Xcompiler", "-fopenmp", "-lgomp", "-ldl"], verbose)
        for o in objs:
            os.remove(o)
    return cuda_so, headers


def build(verbose=False, force=False):
    cuda_so, headers = _build_cuda(LIB, [], verbose, force)
    host_hdr = [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".h")] + headers
    # scene side of the host mirror (importers, BVH builder): no dependency on the CUDA library — bench.py's reference arm
    # generates its scenes through it without mapping any device code
    scene_so = os.path.join(LIB, "libpathtracer_scene.so")
    scene_src = [os.path.join(HOST, f) for f in ("PathTracer_BVH.cpp", "PathTracer_ProceduralImporter.cpp", "PathTracer_FileImporter.cpp",
                                                 "pth_scene_capi.cpp")]
    if force or _newer(scene_so, scene_src + host_hdr + [__file__]):
        _run(["g++"] + CXX_FLAGS + ["-shared", "-o", scene_so] + scene_src, verbose)
    # controller side: PathTracer_Main, the CUDA_* shims, the headless dialog
    host_so = os.path.join(LIB, "libpathtracer_host.so")
    host_src = [os.path.join(HOST, f) for f in ("PathTracer_CUDA.cpp", "PathTracer_Dialog.cpp", "PathTracer.cpp", "pth_capi.cpp")]
    if force or _newer(host_so, host_src + host_hdr + [cuda_so, scene_so, __file__]):
        _run(["g++"] + CXX_FLAGS + ["-shared", "-o", host_so] + host_src +
             ["-L", LIB, "-lpathtracer_scene", "-lpathtracer_cuda", "-Wl,-rpath,$ORIGIN"], verbose)
    exe = os.path.join(LIB, "pathtracer_alone")
    if force or _newer(exe, [os.path.join(HOST, "main.cpp"), host_so]):
        _run(["g++"] + CXX_FLAGS + ["-o", exe, os.path.join(HOST, "main.cpp"), "-L", LIB, "-lpathtracer_host", "-lpathtracer_scene",
                                    "-lpathtracer_cuda", "-Wl,-rpath,$ORIGIN"], verbose)
    # the CHECKED variant of the device library (bounds checks on every index the kernels form, reported through
    # ptc_counters.faults): tests/test_checked_build.py runs the kernels under it, in the place of compute-sanitizer's memcheck,
    # which is closed on the GPU pool. Selected with PTC_LIB_DIR=<repo>/opencl_pathtracer_b200/_lib_checked.
    checked = os.path.join(HERE, "_lib_checked")
    checked_so, _ = _build_cuda(checked, ["-DPTC_CHECKED"], verbose, force)
    for so in (scene_so, host_so):
        dst = os.path.join(checked, os.path.basename(so))
        if force or _newer(dst, [so]):
            shutil.copy2(so, dst)
    return cuda_so, host_so


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
