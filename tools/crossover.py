# dev: where VCRT_TRAVERSAL_AUTO should switch from the brute-force scan to the LBVH (kBruteForceLimit in vcrt_cabi.cu)
import sys; sys.path.insert(0,'.')
from vulkancomputeraytracing_b200 import host
import numpy as np
full = host.generate_scene()
scenes = [np.ascontiguousarray(np.concatenate([full[-k:-1], full[-1:]])) for k in (2, 3, 4, 5, 7)] + [host.generate_scene(g) for g in (1, 2, 3, 4, 5)]
for sc in scenes:
    host.upload_scene(sc); g = 0
    out=[]
    for trav in (1,2):
        best=1e9
        for i in range(3):
            _, st = host.render(1920,1080,64,traversal=trav,download=False); best=min(best, st.renderKernelMs)
        out.append(best)
    print(f"grid {g}: {sc.shape[0]} spheres  brute {out[0]:.2f} ms  lbvh {out[1]:.2f} ms  lbvh/brute speed {out[0]/out[1]:.2f}")
host.destroy()
