# dev: time per sample vs spp / chunks / depth
import ctypes, sys, numpy as np
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi
L = cabi.load(); vp = ctypes.c_void_p
O = ctypes.CDLL('oracle/_build/liboracle.so')
O.oracle_generate_scene.argtypes=[ctypes.c_int, ctypes.c_double, ctypes.c_double, vp, ctypes.c_int]
g = np.zeros((600,10), np.float32); n = O.oracle_generate_scene(11, 0.8, 0.95, g.ctypes.data_as(vp), 600); w = np.ascontiguousarray(g[:n])
c = cabi.VcrtCamera(); L.vcrtDefaultCamera(ctypes.byref(c))
cabi.check(L.cudaUploadScene(w.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), n, ctypes.byref(c)))
def run(W,H,spp,chunks,depth=50,spi=0):
    p = cabi.VcrtRenderParams(); L.vcrtDefaultRenderParams(ctypes.byref(p))
    p.width, p.height, p.samplesPerPixel, p.traversal, p.sampleChunks, p.maxDepth = W,H,spp,1,chunks,depth; p.reserved[0] = spi
    cabi.check(L.cudaRender(ctypes.byref(p))); st = cabi.VcrtRenderStats(); L.cudaGetRenderStats(ctypes.byref(st))
    print(f'{W}x{H} spp={spp} chunks={st.sampleChunksUsed} spi={spi} depth={depth}: {st.renderKernelMs:.1f} ms  {st.rays*n/st.renderKernelMs/1e9:.3f} Ttests/s  rays/sample {st.rays/st.samples:.3f} cand/ray {st.candidates/st.rays:.2f}', flush=True)
run(1920,1080,16,1)
for spp,spi in [(128,1),(128,2),(128,4),(128,8),(500,1),(500,2),(500,4),(500,8),(500,16)]: run(1920,1080,spp,0,spi=spi)
run(400,225,16,0,spi=1); run(400,225,16,0,spi=4); run(3840,2160,125,0,spi=4)
