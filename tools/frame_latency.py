# dev: per-frame latency at interactive sizes (the reference renders 1 spp per DrawNextFrame): kernel time, cudaRender total, wall clock per call with and without the download
import sys, time
sys.path.insert(0, '.')
import numpy as np
from vulkancomputeraytracing_b200 import host, cabi
scene = host.generate_scene(); host.upload_scene(scene)
for trav in (0, 1):
    for (w, h, spp) in ((1280, 720, 1), (1280, 720, 4), (400, 225, 16)):
        img = np.empty((h, w, 4), np.float32)
        for i in range(5): host.render(w, h, spp, traversal=trav, out=img)
        t = time.perf_counter(); n = 50
        for i in range(n): _, st = host.render(w, h, spp, traversal=trav, download=False)
        t1 = (time.perf_counter() - t) / n
        t = time.perf_counter()
        for i in range(n): _, st = host.render(w, h, spp, traversal=trav, out=img)
        t2 = (time.perf_counter() - t) / n
        print(f"trav {trav} {w}x{h}x{spp}: kernel {st.renderKernelMs:.3f} ms, cudaRender total (events) {st.totalMs:.3f} ms, wall per call {t1*1e3:.3f} ms, with download {t2*1e3:.3f} ms, launches {st.kernelLaunches}")
host.destroy()
