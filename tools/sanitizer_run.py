# dev: the smallest run that touches every kernel (for compute-sanitizer). All traversals, both semantics, both accumulators.
import sys
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi, host
import numpy as np
big = np.tile(host.generate_scene(), (5, 1))[:2200]; big[:, 0] += np.repeat(np.arange(5), 485)[:2200] * 0.01   # ~2200 spheres: first-generation scan kernel
mid = big[:1100]                                                                                                # ~1100 spheres: scanu 512 x 4
for scene in (host.generate_scene(), host.generate_scene(3), host.stress_scene(), mid, big):
    host.upload_scene(scene)
    for trav in (1, 2, 3):
        for sem in (0, 1):
            for chunks in (0, 1):
                img, st = host.render(72, 40, 2, depth=12, traversal=trav, semantics=sem, sample_chunks=chunks)
    host.render(72, 40, 2, traversal=2)
    host.render(72, 40, 1, traversal=2, sample_offset=2, flags=cabi.VCRT_FLAG_ACCUMULATE)   # progressive frame on the fixed-point sums
    host.render(72, 40, 3, traversal=1, flags=cabi.VCRT_FLAG_COUNT_STATS)
cos = np.linspace(-15, 2, 4096, dtype=np.float32); out = np.empty_like(cos)
cabi.check(cabi.load().cudaEvaluateSchlick(cos.ctypes.data, np.full_like(cos, 1.5).ctypes.data, cos.size, out.ctypes.data))
host.destroy()
print("sanitizer run done")
