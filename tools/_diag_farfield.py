import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from vulkancomputeraytracing_b200 import cabi, host
from test_gpu_parity import far_field_scene
for case in ["slab_3000", "slab_3000_ground", "cube_800", "far_camera", "tiny_radii"]:
    s, cam = far_field_scene(case)
    host.upload_scene(s, cam)
    for depth in (1, 12):
        simple, ss = host.render(480, 270, 2, depth=depth, traversal=3, sample_chunks=1)
        tree, st = host.render(480, 270, 2, depth=depth, traversal=2, sample_chunks=1)
        nested, sn = host.render(480, 270, 2, depth=depth, traversal=2, sample_chunks=1, shape=31)
        print(case, "depth", depth, "hits", ss.rays - 480*270*2, "rays", ss.rays, st.rays, sn.rays, "flat!=simple px", int((np.abs(simple-tree).max(-1)>0).sum()), "nested!=simple px", int((np.abs(simple-nested).max(-1)>0).sum()),
              "nodes", st.nodeVisits, sn.nodeVisits, "hit px", int((simple[...,:3].min(-1) < 0.49).sum()), flush=True)
    host.destroy()
