# dev script: renders through the C ABI (for ncu / quick timing). usage: prof_run.py W H spp traversal [chunks] [scene] [reps] [shape] [semantics]
import sys
sys.path.insert(0, '.')
from vulkancomputeraytracing_b200 import host
W, H, spp, trav = map(int, sys.argv[1:5]); chunks = int(sys.argv[5]) if len(sys.argv) > 5 else 0
kind = sys.argv[6] if len(sys.argv) > 6 else "reference"; reps = int(sys.argv[7]) if len(sys.argv) > 7 else 2; shape = int(sys.argv[8]) if len(sys.argv) > 8 else 0; semantics = int(sys.argv[9]) if len(sys.argv) > 9 else 0
scene = host.stress_scene() if kind == "stress" else host.generate_scene(11 if kind == "reference" else int(kind))
n = scene.shape[0]
host.upload_scene(scene)
for i in range(reps):
    _, st = host.render(W, H, spp, traversal=trav, sample_chunks=chunks, download=False, shape=shape, semantics=semantics)
    print(f'{W}x{H}x{spp} trav={st.traversalUsed} n={n}: kernel {st.renderKernelMs:.2f} ms total {st.totalMs:.2f} ms  {st.samples/st.renderKernelMs/1e3:.1f} Msamples/s  '
          f'{st.rays/st.renderKernelMs/1e3:.1f} Mrays/s  {st.rays*n/st.renderKernelMs/1e9:.3f} Ttests/s  alg {st.rays*n*17/st.renderKernelMs/1e9:.1f} TF  '
          f'primary {st.primaryKernelMs:.2f} ms  rays/sample {st.rays/st.samples:.4f} cand/ray {st.candidates/max(1,st.rays):.2f} ovf {st.listOverflows} nodes/ray {st.nodeVisits/max(1,st.rays):.1f}', flush=True)
tf, mhz = host.fp32_peak(); print(f'fp32 peak {tf:.1f} TF at {mhz:.0f} MHz')
host.destroy()
