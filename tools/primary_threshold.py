# dev: from how many spheres on does the primary-visibility pass pay?  Run twice: VCRT_NO_PRIMARY=0 / 1 (read once per process).
import sys; sys.path.insert(0, '.')
import numpy as np
from vulkancomputeraytracing_b200 import host
for g in (1, 3, 5, 6, 7, 8, 9, 11):
    sc = host.generate_scene(g); host.upload_scene(sc)
    best = 1e9
    for i in range(3):
        _, st = host.render(1920, 1080, 64, traversal=1, download=False); best = min(best, st.renderKernelMs)
    print(f"{sc.shape[0]:4d} spheres  brute {best:7.2f} ms  (primary kernel {st.primaryKernelMs:.2f} ms)")
host.destroy()
