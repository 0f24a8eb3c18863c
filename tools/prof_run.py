# dev script: one render through the C ABI (for ncu). usage: prof_run.py W H spp traversal [chunks]
import ctypes, sys, numpy as np
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi
L = cabi.load(); vp = ctypes.c_void_p
O = ctypes.CDLL('oracle/_build/liboracle.so')
O.oracle_generate_scene.argtypes=[ctypes.c_int, ctypes.c_double, ctypes.c_double, vp, ctypes.c_int]
g = np.zeros((600,10), np.float32); n = O.oracle_generate_scene(11, 0.8, 0.95, g.ctypes.data_as(vp), 600); w = np.ascontiguousarray(g[:n])
c = cabi.VcrtCamera(); L.vcrtDefaultCamera(ctypes.byref(c))
cabi.check(L.cudaUploadScene(w.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), n, ctypes.byref(c)))
W,H,spp,trav = map(int, sys.argv[1:5]); chunks = int(sys.argv[5]) if len(sys.argv) > 5 else 0
p = cabi.VcrtRenderParams(); L.vcrtDefaultRenderParams(ctypes.byref(p))
p.width, p.height, p.samplesPerPixel, p.traversal, p.sampleChunks = W,H,spp,trav,chunks
for i in range(2):
    cabi.check(L.cudaRender(ctypes.byref(p)))
    st = cabi.VcrtRenderStats(); L.cudaGetRenderStats(ctypes.byref(st))
    print(f'kernel {st.renderKernelMs:.2f} ms  {st.rays*n/st.renderKernelMs/1e9:.3f} Ttests/s  chunks {st.sampleChunksUsed}')
