# dev: wall-clock of cudaUploadScene and of the first render after it (which builds the LBVH). usage: upload_timing.py [grid]
import sys, time
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import host
grid = int(sys.argv[1]) if len(sys.argv) > 1 else 158
scene = host.generate_scene(grid)
host.upload_scene(scene); host.render(64, 36, 1, traversal=2)
for i in range(4):
    t0 = time.perf_counter(); host.upload_scene(scene); t1 = time.perf_counter()
    _, st = host.render(64, 36, 1, traversal=2, download=False); t2 = time.perf_counter()
    _, st2 = host.render(64, 36, 1, traversal=2, download=False); t3 = time.perf_counter()
    print(f"n={scene.shape[0]}: upload {1e3*(t1-t0):.2f} ms, first render (tree build + 64x36) {1e3*(t2-t1):.2f} ms, second render {1e3*(t3-t2):.2f} ms", flush=True)
host.destroy()
