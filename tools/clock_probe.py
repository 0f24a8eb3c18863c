# dev: watch SM clock / power while a long render and a long FFMA burst run
import ctypes, subprocess, sys, time, numpy as np
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import cabi
L = cabi.load(); vp = ctypes.c_void_p
O = ctypes.CDLL('oracle/_build/liboracle.so')
O.oracle_generate_scene.argtypes=[ctypes.c_int, ctypes.c_double, ctypes.c_double, vp, ctypes.c_int]
g = np.zeros((600,10), np.float32); n = O.oracle_generate_scene(11, 0.8, 0.95, g.ctypes.data_as(vp), 600); w = np.ascontiguousarray(g[:n])
c = cabi.VcrtCamera(); L.vcrtDefaultCamera(ctypes.byref(c))
cabi.check(L.cudaUploadScene(w.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), n, ctypes.byref(c)))
mon = subprocess.Popen(['nvidia-smi','--query-gpu=clocks.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,temperature.gpu','--format=csv,noheader','-lms','100'], stdout=open('gpurun_out/clocks_probe.csv','w'))
time.sleep(0.5)
p = cabi.VcrtRenderParams(); L.vcrtDefaultRenderParams(ctypes.byref(p))
p.width, p.height, p.samplesPerPixel, p.traversal = 1920,1080,500,1
for i in range(3):
    t=time.time(); cabi.check(L.cudaRender(ctypes.byref(p))); st = cabi.VcrtRenderStats(); L.cudaGetRenderStats(ctypes.byref(st))
    print(f'render {i}: kernel {st.renderKernelMs:.1f} ms, {st.rays*n/st.renderKernelMs/1e9:.3f} Ttests/s', flush=True)
print('MARK peak', flush=True); time.sleep(1.0)
tf = ctypes.c_double(); mhz = ctypes.c_double()
for i in range(20):
    cabi.check(L.cudaMeasureFp32Peak(ctypes.byref(tf), ctypes.byref(mhz)))
print('fp32 peak after 20 back-to-back:', tf.value)
time.sleep(0.3); mon.terminate()
rows = [l.strip() for l in open('gpurun_out/clocks_probe.csv')]
print(len(rows)); print('\n'.join(rows[::3][:60]))
