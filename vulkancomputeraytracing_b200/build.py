# SPDX-License-Identifier: WTFPL
"""Builds lib/libvcrt.so (CUDA, sm_100a) and the C++ host programs in-tree with nvcc / g++.

    python -m vulkancomputeraytracing_b200.build [--force]
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
LIB = os.path.join(HERE, "lib")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
GXX = "/usr/bin/g++"   # the image's $CXX (/opt/gcc) lacks libgomp.spec; pin the distro compiler

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-ffp-contract=off", "-ccbin", GXX]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def _run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def build(force=False):
    os.makedirs(LIB, exist_ok=True)
    cu_sources = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(ROOT, "include", "vcrt.h")]
    target = os.path.join(LIB, "libvcrt.so")
    if force or _newer(target, cu_sources):
        _run([NVCC] + NVCC_FLAGS + ["-shared", "-o", target, os.path.join(CSRC, "vcrt_cabi.cu"), "-ldl"])
    if os.path.isdir(HOST):
        host_sources = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST)) if f.endswith(".cpp")]
        headers = [os.path.join(ROOT, "include", f) for f in sorted(os.listdir(os.path.join(ROOT, "include")))]
        exe = os.path.join(LIB, "VulkanComputeRayTracing")
        main_sources = [s for s in host_sources if not s.endswith("SceneGeneratorMain.cpp")]
        if main_sources and (force or _newer(exe, main_sources + headers + [target])):
            _run([GXX, "-std=c++20", "-O2", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"), "-o", exe] + main_sources +
                 ["-L", LIB, "-lvcrt", "-Wl,-rpath,$ORIGIN"])
        gen = os.path.join(HOST, "SceneGeneratorMain.cpp")
        gen_exe = os.path.join(LIB, "SceneGenerator")
        if os.path.exists(gen) and (force or _newer(gen_exe, [gen, os.path.join(HOST, "SceneGenerator.cpp")] + headers)):
            _run([GXX, "-std=c++20", "-O2", "-ffp-contract=off", "-I", os.path.join(ROOT, "include"), "-o", gen_exe, gen,
                  os.path.join(HOST, "SceneGenerator.cpp")])
    return target


if __name__ == "__main__":
    build(force="--force" in sys.argv)
