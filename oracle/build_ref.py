#!/usr/bin/env python3
# SPDX-License-Identifier: WTFPL
"""TEST INFRASTRUCTURE (oracle/): build the REFERENCE's own code for the hot path on the CPU.

Nothing from /root/reference is copied into the repository. At build time this script
reads the reference's shader sources where they lie, applies the purely mechanical
GLSL-450 -> C++20 rewrite listed below, writes the result to oracle/_ref/gen/ (git-ignored)
and compiles it with oracle/ref_shim/{glsl_compat.hpp,ref_driver.cpp} into
oracle/_ref/libvcrt_ref.so.  It also compiles the reference's SceneGenerator.cpp unmodified
into oracle/_ref/SceneGenerator.

Rewrite rules (each one is syntax only; no arithmetic is altered):
  R1  drop `#version`, `#extension`, `layout(...)` declarations and `precision ...;`
  R2  float literals get an `f` suffix (GLSL literals are binary32)
  R3  `inout T x` / `out T x` -> `T& x`; `const in T x` / `in T x` -> `const T& x`
  R4  read swizzles `.xy .xz .yz .rgb` -> accessor calls; `c.rgb += e` / `c.rgb /= e`
      -> `c.set_rgb(c.rgb() + e)` / `c.set_rgb(c.rgb() / e)`
  R5  `world.length()` -> element count; `ivec2(gl_GlobalInvocationID.xy)` -> helper
  R6  the four compile-time #defines SAMPLES_PER_PIXEL / MAX_RECURSION_LEVEL / IMAGE_WIDTH /
      IMAGE_HEIGHT become run-time ints (integer division semantics kept, bug B1)
  R7  `void main()` -> `void shader_main()`; #include targets get `.hpp` appended
  R8  bug B7 (ray_color falls off its end): `return vec3(0.0f);` appended, plus two
      counters (rays traced, depth exhausted) that do not touch the arithmetic
"""
import os
import re
import subprocess
import sys

REF = os.environ.get("VCRT_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")

SOURCES = ["shaders/shader.comp", "shaders/include/functions.glsl", "shaders/include/textures.glsl",
           "shaders/include/structures.glsl", "shaders/include/globals.glsl"]

FLOAT_LIT = re.compile(r"(?<![\w.])((?:\d+\.\d*|\.\d+)(?:[eE][+-]?\d+)?|\d+[eE][+-]?\d+)(?![\w.])")


def rewrite(text: str, name: str) -> str:
    out = []
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("#version") or s.startswith("#extension") or s.startswith("precision "):
            continue  # R1
        if s.startswith("layout"):
            continue  # R1
        if s.startswith("#include"):
            line = re.sub(r'#include\s+"(?:include/)?([^"]+)"', r'#include "\1.hpp"', line)  # R7
            out.append(line)
            continue
        m = re.match(r"#define\s+(SAMPLES_PER_PIXEL|MAX_RECURSION_LEVEL|IMAGE_WIDTH|IMAGE_HEIGHT)\s+\d+", s)
        if m:  # R6
            var = {"SAMPLES_PER_PIXEL": "ref_samples_per_pixel", "MAX_RECURSION_LEVEL": "ref_max_recursion_level",
                   "IMAGE_WIDTH": "ref_image_width", "IMAGE_HEIGHT": "ref_image_height"}[m.group(1)]
            out.append(f"#undef {m.group(1)}\n#define {m.group(1)} {var}")
            continue
        if not s.startswith("#"):
            line = FLOAT_LIT.sub(lambda k: k.group(1) + "f", line)  # R2
        line = re.sub(r"\bconst\s+in\s+(\w+)\s+(\w+)", r"const \1& \2", line)  # R3
        line = re.sub(r"\binout\s+(\w+)\s+(\w+)", r"\1& \2", line)
        line = re.sub(r"\bout\s+(\w+)\s+(\w+)", r"\1& \2", line)
        line = re.sub(r"(\w+)\.rgb\s*\+=\s*(.+);", r"\1.set_rgb(\1.rgb() + (\2));", line)  # R4
        line = re.sub(r"(\w+)\.rgb\s*/=\s*(.+);", r"\1.set_rgb(\1.rgb() / (\2));", line)
        line = line.replace("ivec2(gl_GlobalInvocationID.xy)", "make_ivec2(gl_GlobalInvocationID.xy())")  # R5
        line = re.sub(r"\.(xy|xz|yz|rgb)\b(?!\()", r".\1()", line)  # R4
        line = line.replace("world.length()", "int(sizeof(world)/sizeof(world[0]))")  # R5
        line = re.sub(r"\bvoid\s+main\s*\(\s*\)", "void shader_main()", line)  # R7
        out.append(line)
    text = "\n".join(out) + "\n"
    if name == "functions.glsl":  # R8
        text = text.replace("for(int pass=0;pass<MAX_RECURSION_LEVEL;pass++) {",
                            "for(int pass=0;pass<MAX_RECURSION_LEVEL;pass++) {\n        ref_rays_traced++;", 1)
        idx = text.rstrip().rfind("}")
        text = text[:idx] + "    ref_depth_exhausted++;\n    return vec3(0.0f);\n}\n"
    return "#pragma once\n" + text if name != "shader.comp" else text


def run(cmd):
    print("+", " ".join(cmd), flush=True)
    subprocess.check_call(cmd)


def main() -> int:
    if not os.path.isdir(REF):
        print(f"reference not present at {REF}; keeping any prebuilt oracle/_ref", file=sys.stderr)
        return 0
    os.makedirs(GEN, exist_ok=True)
    for rel in SOURCES:
        name = os.path.basename(rel)
        with open(os.path.join(REF, rel), encoding="utf-8-sig") as f:
            src = f.read()
        with open(os.path.join(GEN, name + ".hpp"), "w") as f:
            f.write(rewrite(src, name))
    shim = os.path.join(HERE, "ref_shim")
    # -O2 -ffp-contract=off: bit-stable binary32 arithmetic, no FMA contraction.
    run(["g++", "-std=c++20", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fopenmp", "-fPIC", "-shared",
         "-I", shim, "-I", OUT, "-o", os.path.join(OUT, "libvcrt_ref.so"), os.path.join(shim, "ref_driver.cpp")])
    # Same sources, -O3 -march=x86-64-v3 (AVX2+FMA; portable to the GPU box's host, unlike
    # -march=native): the timed CPU baseline (bench.py --impl reference).
    run(["g++", "-std=c++20", "-O3", "-march=x86-64-v3", "-fopenmp", "-fPIC", "-shared",
         "-I", shim, "-I", OUT, "-o", os.path.join(OUT, "libvcrt_ref_fast.so"), os.path.join(shim, "ref_driver.cpp")])
    run(["g++", "-O2", "-o", os.path.join(OUT, "SceneGenerator"), os.path.join(REF, "SceneGenerator.cpp")])
    return 0


if __name__ == "__main__":
    sys.exit(main())
