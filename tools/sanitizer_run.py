# dev: the smallest run that touches every kernel (for compute-sanitizer). All traversals, both semantics, both accumulators.
import sys
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi, host
for scene in (host.generate_scene(), host.generate_scene(3), host.stress_scene()):
    host.upload_scene(scene)
    for trav in (1, 2, 3):
        for sem in (0, 1):
            for chunks in (0, 1):
                img, st = host.render(72, 40, 2, depth=12, traversal=trav, semantics=sem, sample_chunks=chunks)
    img, st = host.render(72, 40, 2, traversal=1, shape=1)      # first-generation scan kernel
    host.render(72, 40, 1, traversal=2, sample_offset=2, flags=cabi.VCRT_FLAG_ACCUMULATE)
host.destroy()
print("sanitizer run done")
