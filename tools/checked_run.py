# SPDX-License-Identifier: WTFPL
"""Bounds checks of our own, because compute-sanitizer is closed on the GPU pool (profiles/r02_sanitizer_closed.txt).

    python tools/checked_run.py --build     # here (no GPU needed): nvcc -DVCRT_CHECKED -> build/checked/libvcrt.so
    python tools/checked_run.py             # on the GPU box: every kernel, every mode, adversarial scenes; prints the counters

With -DVCRT_CHECKED every hand-computed index (candidate lists and ray state in shared memory, the traversal stack in local
memory, node / leaf / sphere indices, the sum planes and the fixed-point sums in global memory, the constant-bank table) is
compared with its bound before the access, and every sample added to the fixed-point sums is checked to lie in [0, 1].
Violations are counted in VcrtRenderStats.checkFailures; the images are still compared with the exhaustive in-order kernel.
Covers what memcheck would (out-of-bounds) - not races: the only cross-thread traffic in the kernels is the mbarrier-guarded
TMA copy of the scene blob (one producer, everybody waits) and integer atomics."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
OUT = os.path.join(ROOT, "build", "checked", "libvcrt.so")

if "--build" in sys.argv:
    from vulkancomputeraytracing_b200 import build as B
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [B.NVCC] + B.NVCC_FLAGS + ["-DVCRT_CHECKED", "-shared", "-o", OUT, os.path.join(B.CSRC, "vcrt_cabi.cu"), os.path.join(B.HOST, "SceneGenerator.cpp"), "-ldl"]
    print("+", " ".join(cmd)); subprocess.check_call(cmd)
    sys.exit(0)

import numpy as np
from vulkancomputeraytracing_b200 import cabi
cabi.LIB_PATH = OUT
from vulkancomputeraytracing_b200 import host

rng = np.random.default_rng(7)
def random_scene(count):
    s = np.zeros((count, 10), np.float32)
    s[:, 0:3] = rng.uniform(-6, 6, (count, 3)); s[:, 1] = np.abs(s[:, 1]) * 0.3
    s[:, 3] = rng.uniform(0.05, 0.6, count); s[:, 4:7] = rng.uniform(0.1, 1, (count, 3))
    s[:, 7] = rng.integers(1, 4, count); s[:, 8] = np.where(s[:, 7] == 3, 1.5, rng.uniform(0, 1, count))
    return s
line = np.zeros((200, 10), np.float32); line[:, 0] = np.linspace(12, -12, 200); line[:, 1] = 0.3 * np.linspace(1.8, -1.8, 200); line[:, 2] = 0.23 * np.linspace(2.7, -2.7, 200)
line[:, 3] = 0.4; line[:, 4:7] = 0.9; line[:, 7] = 3; line[:, 8] = 1.5           # 200 glass spheres on the view axis: candidate-list overflow
scenes = {"rtiow": host.generate_scene(), "tiny": host.generate_scene(1), "stress": host.stress_scene(), "10k": host.generate_scene(50),
          "1100": random_scene(1100), "2200": random_scene(2200), "line": line, "one": random_scene(1), "empty": np.zeros((0, 10), np.float32)}
total, renders, worst = 0, 0, 0.0
for name, scene in scenes.items():
    host.upload_scene(scene)
    n = scene.shape[0]
    for trav in (1, 2, 3):
        if trav == 1 and n > 2400:
            continue
        for sem in (0, 1):
            for chunks in (0, 1, 3):
                for (w, h, spp, depth) in ((72, 40, 3, 12), (33, 1, 2, 50), (130, 70, 1, 3)):
                    img, st = host.render(w, h, spp, depth=depth, traversal=trav, semantics=sem, sample_chunks=chunks, flags=cabi.VCRT_FLAG_COUNT_STATS if chunks == 3 else 0)
                    total += st.checkFailures; renders += 1
                    if trav != 3 and sem == 0:
                        ref, _ = host.render(w, h, spp, depth=depth, traversal=3, semantics=sem)
                        worst = max(worst, float(np.abs(img - ref).max()))
    host.render(72, 40, 2, traversal=2); _, st = host.render(72, 40, 1, traversal=2, sample_offset=2, flags=cabi.VCRT_FLAG_ACCUMULATE); total += st.checkFailures
    print(f"{name:7s} {n:6d} spheres: check failures so far {total}, renders {renders}, max |diff| to the in-order kernel {worst:.2e}", flush=True)
host.upload_scene(scenes["rtiow"])
_, st = host.render(1920, 1080, 16, traversal=1); total += st.checkFailures
_, st = host.render(1920, 1080, 16, traversal=2); total += st.checkFailures
host.destroy()
print(f"CHECKED_RUN renders {renders + 2} check_failures {total} max_abs_diff_vs_in_order_kernel {worst:.3e} -> {'PASS' if total == 0 and worst <= 2.0 ** -20 else 'FAIL'}")
sys.exit(0 if total == 0 and worst <= 2.0 ** -20 else 1)
