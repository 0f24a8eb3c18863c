# dev script: GPU vs oracle quick check + timing
import ctypes, sys, time, numpy as np
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi
L = cabi.load(); vp = ctypes.c_void_p
O = ctypes.CDLL('oracle/_build/liboracle.so')
O.oracle_generate_scene.argtypes=[ctypes.c_int, ctypes.c_double, ctypes.c_double, vp, ctypes.c_int]
O.oracle_render.argtypes=[vp, ctypes.c_int, vp] + [ctypes.c_int]*6 + [ctypes.c_uint64, ctypes.c_int, ctypes.c_int, vp, vp]
g = np.zeros((600,10), np.float32); n = O.oracle_generate_scene(11, 0.8, 0.95, g.ctypes.data_as(vp), 600); w = np.ascontiguousarray(g[:n])
cam = np.array([13,2,3, 0,0,0, 0,1,0, 20], np.float32)
def orender(W,H,spp,seed=0,off=0,depth=50):
    img = np.zeros((H,W,4), np.float32); st = np.zeros(14, np.uint64)
    O.oracle_render(w.ctypes.data_as(vp), n, cam.ctypes.data_as(vp), W,H,spp,off,depth,1,seed,1,0, img.ctypes.data_as(vp), st.ctypes.data_as(vp)); return img, st
c = cabi.VcrtCamera(); L.vcrtDefaultCamera(ctypes.byref(c))
cabi.check(L.cudaUploadScene(w.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), n, ctypes.byref(c)))
def grender(W,H,spp,trav,seed=0,off=0,depth=50,chunks=1):
    p = cabi.VcrtRenderParams(); L.vcrtDefaultRenderParams(ctypes.byref(p))
    p.width, p.height, p.samplesPerPixel, p.maxDepth, p.seed, p.sampleOffset, p.traversal, p.sampleChunks = W,H,spp,depth,seed,off,trav,chunks
    cabi.check(L.cudaRender(ctypes.byref(p)))
    img = np.zeros((H,W,4), np.float32); cabi.check(L.cudaDownloadImage(img.ctypes.data_as(vp), img.nbytes))
    st = cabi.VcrtRenderStats(); L.cudaGetRenderStats(ctypes.byref(st)); return img, st
def cmp(a,b,name):
    eq = np.array_equal(a,b); d = np.abs(a-b); print(f'{name}: bit-equal={eq} maxdiff={d.max():.3e} frac_diff_pixels={(d.max(-1)>0).mean():.5f}', flush=True)
for (W,H,spp) in [(64,36,2),(400,225,4)]:
    o, ost = orender(W,H,spp)
    s, sst = grender(W,H,spp,3); cmp(s,o,f'simple vs oracle {W}x{H}x{spp}'); print(' rays', sst.rays, int(ost[0]), 'exh', sst.depthExhausted, int(ost[1]))
    b, bst = grender(W,H,spp,1); cmp(b,o,f'scan   vs oracle {W}x{H}x{spp}'); cmp(b,s,'scan vs simple')
    f_, fst = grender(W,H,spp,1,chunks=0); d=np.abs(f_-b); print(' fixed-point mode vs strict: maxdiff %.3e'%d.max(), 'rays equal', fst.rays==bst.rays)
    print(' rays', bst.rays, 'cand/ray %.2f'%(bst.candidates/max(1,bst.rays)), 'overflow', bst.listOverflows, 'ms %.2f'%bst.renderKernelMs)
tf = ctypes.c_double(); mhz = ctypes.c_double(); cabi.check(L.cudaMeasureFp32Peak(ctypes.byref(tf), ctypes.byref(mhz))); print('fp32 peak TF', tf.value, 'MHz', mhz.value)
for (W,H,spp,ch) in [(400,225,16,1),(1920,1080,16,1),(1920,1080,64,0),(1920,1080,500,0)]:
    b, st = grender(W,H,spp,1,chunks=ch)
    tests = st.rays*n
    print(f'{W}x{H}x{spp} chunks={st.sampleChunksUsed}: kernel {st.renderKernelMs:.1f} ms total {st.totalMs:.1f} ms  {st.samples/st.renderKernelMs/1e3:.1f} Msamples/s  {st.rays/st.renderKernelMs/1e3:.1f} Mrays/s  {tests/st.renderKernelMs/1e9:.2f} Ttests/s  alg {tests*17/st.renderKernelMs/1e9:.1f} TF  rays/sample {st.rays/st.samples:.4f} cand/ray {st.candidates/st.rays:.2f} ovf {st.listOverflows} mean {b[...,:3].reshape(-1,3).mean(0)}', flush=True)
s, st = grender(1920,1080,4,3); print('simple 1080p x4: %.1f ms'%st.renderKernelMs)
