# SPDX-License-Identifier: WTFPL
"""Builds the product in-tree with nvcc / g++ (no JIT cache: the .so travels with the repo snapshot).

    python -m vulkancomputeraytracing_b200.build [--force]

Outputs (all under vulkancomputeraytracing_b200/lib/, git-ignored):
    libvcrt.so                 the C ABI of include/vcrt.h + include/SceneGenerator.hpp (CUDA, sm_100a)
    VulkanComputeRayTracing    the headless host program (Begin / Draw / End over the C ABI)
    SceneGenerator             prints the scene as GLSL text like the reference's tool
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
LIB = os.path.join(HERE, "lib")
INCLUDE = os.path.join(ROOT, "include")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
GXX = "/usr/bin/g++"   # the image's $CXX (/opt/gcc) lacks libgomp.spec; pin the distro compiler

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++20",
              "-Xcompiler", "-fPIC,-ffp-contract=off", "-ccbin", GXX] + os.environ.get("VCRT_EXTRA_NVCC", "").split()
GXX_FLAGS = ["-std=c++20", "-O2", "-ffp-contract=off", "-Wall", "-I", INCLUDE]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def _listing(folder, suffixes):
    return [os.path.join(folder, f) for f in sorted(os.listdir(folder)) if f.endswith(suffixes)]


def build(force=False):
    os.makedirs(LIB, exist_ok=True)
    headers = _listing(INCLUDE, (".h", ".hpp"))
    target = os.path.join(LIB, "libvcrt.so")
    scene = os.path.join(HOST, "SceneGenerator.cpp")
    if force or _newer(target, _listing(CSRC, (".cu", ".cuh")) + headers + [scene]):
        _run([NVCC] + NVCC_FLAGS + ["-shared", "-o", target, os.path.join(CSRC, "vcrt_cabi.cu"), scene, "-ldl"])
    exe = os.path.join(LIB, "VulkanComputeRayTracing")
    main_sources = [os.path.join(HOST, f) for f in ("VulkanComputeRayTracing.cpp", "Renderer.cpp", "Frontend.cpp", "Platform.cpp")]
    if force or _newer(exe, main_sources + headers + [target]):
        _run([GXX] + GXX_FLAGS + ["-o", exe] + main_sources + ["-L", LIB, "-lvcrt", "-Wl,-rpath,$ORIGIN"])
    gen_exe = os.path.join(LIB, "SceneGenerator")
    gen = os.path.join(HOST, "SceneGeneratorMain.cpp")
    if force or _newer(gen_exe, [gen, scene] + headers):
        _run([GXX] + GXX_FLAGS + ["-o", gen_exe, gen, scene])
    return target


if __name__ == "__main__":
    build(force="--force" in sys.argv)
