#!/usr/bin/env python3
# SPDX-License-Identifier: WTFPL
"""Headline benchmark: samples/s (and Mrays/s) of the per-pixel path-tracing loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c1|c2|c3|c4a|c4b|c5]
                    [--traversal auto|brute|lbvh|simple] [--no-cpu] [--no-extras]

A "step" is one full render of the workload's frame (every pixel, every sample, depth 50).
  N = 1  -> BASELINE.json configs[1]: RTIOW final scene (485 spheres), 1920x1080, 500 spp, depth 50
  N > 1  -> BASELINE.json configs[2]: the same scene at 3840x2160, 1000 spp in total, samples-per-pixel sharded
            over the N ranks, one NCCL reduce of the per-GPU accumulation buffers per step (inside cudaRender)
`value` = samples/s of the whole job with the scene resident on the device (max over ranks of the CUDA-event
time); `e2e` = the same through the C ABI from HOST buffers: cudaUploadScene (host AoS -> device) + cudaRender +
cudaDownloadImage (device -> host image) inside the timed region.  One JSON line on stdout (rank 0).

The headline traversal is north_star's linear scan (VCRT_TRAVERSAL_BRUTE: every ray tests every sphere - the FP32
roofline case); the same workload through VCRT_TRAVERSAL_AUTO (what a default caller gets: the LBVH) rides on the same
line as `auto_same_workload`, with its own e2e and clocks.  At N = 1 the line also carries every other BASELINE.json
config as a short leg (`configs`: c1 on GPU and CPU, c3 on one GPU, c4a / c4b LBVH, c5 brute + LBVH) and `parity`: the
image that was timed, checked against the other traversal (bit for bit) and the in-order exhaustive kernel.  At N > 1
`parity` says whether the sharded + reduced image equals the image of one GPU rendering all samples (bit for bit), and
`strong_scaling` gives configs[2] on ONE GPU of the same box, so that the scaling factor can be read at equal config.
The process exits non-zero when a parity check fails.

--impl reference times the reference's own shader code for this path compiled as C++ (oracle/_ref, built by
oracle/build_ref.py; else the oracle port) on the host cores, on a bounded sample of the same workload."""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
import zlib

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (width, height, total spp, depth, scene, description)
    "c1": (400, 225, 16, 50, "reference", "configs[0]: RTIOW final scene (485 spheres), 400x225, 16 spp, depth 50"),
    "c2": (1920, 1080, 500, 50, "reference", "configs[1]: RTIOW final scene (485 spheres), 1920x1080, 500 spp, depth 50"),
    "c3": (3840, 2160, 1000, 50, "reference", "configs[2]: RTIOW final scene (485 spheres), 3840x2160, 1000 spp, depth 50, spp-sharded"),
    "c4a": (1920, 1080, 256, 50, "grid50", "configs[3]: 10k random spheres (grid half-extent 50), 1920x1080, 256 spp, LBVH"),
    "c4b": (1920, 1080, 256, 50, "grid158", "configs[3]: 100k random spheres (grid half-extent 158), 1920x1080, 256 spp, LBVH"),
    "c5": (1920, 1080, 256, 50, "stress", "configs[4]: all metal/dielectric divergence stress scene, 1920x1080, 256 spp, depth 50"),
}
FLOP_PER_TEST = 17      # one ray-sphere test of the reference = 17 flop (11 FP32-pipe instructions)
FLOP_PER_FILTER = 14    # what the path kernel executes per (ray, sphere): 7 FMAs of the conservative filter
FLOP_PER_PRIMARY_FILTER = 6   # what the primary-visibility kernel executes per (camera ray, sphere): 1 multiply + 2 FMAs, each counted as an FMA-pipe slot of 2 flop
FLOP_PER_BOX = 24       # LBVH: one box = 6 FMAs + 6 min/max, counted as 2 x 6 + 6 x 2
TRAVERSALS = {"auto": 0, "brute": 1, "lbvh": 2, "simple": 3}
KERNEL_NAMES = {1: "render_scanu_kernel", 2: "render_lbvh_flat_kernel", 3: "render_simple_kernel"}
PRIMARY_KERNEL = "primary_scan_kernel"
CPU_BUILD = "g++ -O3 -march=x86-64-v3 -fopenmp, FMA contraction on (not the bit-stable -O2 -ffp-contract=off build the goldens come from)"


def clock_sampler(stop, rows, index):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    try:
        proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                stdout=subprocess.PIPE, text=True)
    except OSError:
        return
    def pump():
        for line in proc.stdout:
            rows.append(line.strip())
    t = threading.Thread(target=pump, daemon=True); t.start()
    stop.wait()
    proc.terminate()
    t.join(timeout=2)


def summarise_clocks(rows):
    sm, mx, reasons = [], [], set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in rows:
        f = [x.strip() for x in r.split(",")]
        if len(f) < 7:
            continue
        try:
            sm.append(float(f[0])); mx.append(float(f[1]))
        except ValueError:
            continue
        for name, v in zip(names, f[3:7]):
            if v.lower().startswith("active"):
                reasons.add(name)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    sm.sort()
    return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def build_scene(host, kind):
    if kind == "reference":
        return host.generate_scene()
    if kind == "stress":
        return host.stress_scene()
    return host.generate_scene(int(kind[4:]))


# --------------------------------------------------------------------------------------- the CPU side (reference arm, cpu_baseline)
def cpu_runner(scene_kind, depth):
    """run(w, h, spp) -> stats on all host cores: the reference's own shader code compiled as C++ where it was built and the
    scene is the baked one, else the oracle port (same arithmetic, run-time scene)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_harness as oh
    cores = os.cpu_count() or 1
    ref = oh.Reference.open(fast=True) if scene_kind == "reference" else None
    if ref is not None:
        return (lambda w, h, s: ref.render(w, h, s, depth, threads=cores)[1]), "reference", cores
    orc = oh.Oracle(fast=True)
    scene = orc.scene() if scene_kind == "reference" else (orc.scene(11, 0.0, 0.5) if scene_kind == "stress" else orc.scene(int(scene_kind[4:])))
    return (lambda w, h, s: orc.render(scene, w, h, s, depth=depth, rng=oh.RNG_HASH, threads=cores)[1]), "port", cores


def calibrate(run, w, h, spp, budget_s):
    cw, ch = max(1, w // 8), max(1, h // 8)
    run(cw, ch, 1)   # thread pool start-up stays out of the calibration
    t = time.perf_counter(); run(cw, ch, 2); per_sample = (time.perf_counter() - t) / (cw * ch * 2)
    return max(1, min(spp, int(budget_s / (per_sample * w * h))))


def run_reference(args):
    """The reference's CPU implementation of the path on the host cores: a bounded sample of the workload per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    name = args.workload or ("c2" if args.gpus == 1 else "c3")
    w, h, spp, depth, scene_kind, desc = WORKLOADS[name]
    run, kind, cores = cpu_runner(scene_kind, depth)
    step_spp = calibrate(run, w, h, spp, 6.0)   # the full frame at reduced spp, ~6 s per step
    for _ in range(1 if args.warmup else 0):
        run(w, h, step_spp)
    total_t, rays = 0.0, 0
    for _ in range(args.steps):
        t = time.perf_counter(); st = run(w, h, step_spp); total_t += time.perf_counter() - t
        rays += st["rays"]
    value = w * h * step_spp * args.steps / total_t
    sample = f"full {w}x{h} frame, depth {depth}, {step_spp} of {spp} spp per step, {args.steps} steps, hash RNG (reference B2)"
    base = {"value": value, "unit": "samples/s", "cores": cores, "kind": kind, "sample": sample, "spp_per_step": step_spp, "spp_of_config": spp, "build": CPU_BUILD}
    line = {
        "impl": "reference", "metric": "samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_t / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "sample": sample, "spp_per_step": step_spp, "spp_of_config": spp, "build": CPU_BUILD,
                   "note": "samples/s does not depend on spp (every sample is an independent path); the frame, scene and depth are the config's"},
        "mrays_per_sec": rays / total_t * 1e-6, "cpu_baseline": base,
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline(w, h, spp, depth, scene_kind, budget_s=12.0, exact=False):
    """cpu_baseline leg of the GPU arm: the same CPU code on rank 0's host cores; `exact` = the whole config, not a sample."""
    run, kind, cores = cpu_runner(scene_kind, depth)
    s = spp if exact else calibrate(run, w, h, spp, budget_s)
    if exact:
        run(w, h, 1)
    t = time.perf_counter(); st = run(w, h, s); dt = time.perf_counter() - t
    return {"value": w * h * s / dt, "unit": "samples/s", "cores": cores, "kind": kind, "build": CPU_BUILD,
            "sample": (f"the whole config: {w}x{h}, {s} spp, depth {depth}" if exact else f"full {w}x{h} frame, depth {depth}, {s} of {spp} spp") + f", hash RNG (reference B2), {dt:.2f} s",
            "seconds": dt, "mrays_per_sec": st["rays"] / dt * 1e-6}


# --------------------------------------------------------------------------------------- B200 arm
class Bench:
    def __init__(self, args):
        import numpy as np
        import torch
        from vulkancomputeraytracing_b200 import cabi, host, sharding
        self.np, self.torch, self.cabi, self.host, self.sharding, self.args = np, torch, cabi, host, sharding, args
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1")); self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            if self.world == 1 and args.gpus > 1:
                raise SystemExit("launch with torchrun: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
            args.gpus = self.world
        self.lib = cabi.load()
        cabi.check(self.lib.cudaSelectDevice(self.local))
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            sharding.init_communicator(dist, device=torch.device("cuda", self.local))
            self.dist = dist
        self.cam = host.default_camera()
        self.pinned = {}

    # ---- plumbing
    def barrier(self):
        if self.world > 1:
            self.cabi.check(self.lib.cudaCommBarrier())

    def max_over_ranks(self, v):
        out = ctypes.c_float(); self.cabi.check(self.lib.cudaCommMaxFloat(ctypes.c_float(v), ctypes.byref(out))); return out.value

    def sum_over_ranks(self, v):
        if self.world == 1:
            return v
        t = self.torch.tensor([float(v)], dtype=self.torch.float64, device=f"cuda:{self.local}"); self.dist.all_reduce(t); return float(t.item())

    def host_image(self, w, h):
        """pinned host memory for the read-back of the e2e legs"""
        if (w, h) not in self.pinned:
            self.pinned[(w, h)] = self.torch.empty((h, w, 4), dtype=self.torch.float32).pin_memory().numpy()
        return self.pinned[(w, h)]

    def timed(self, step, steps, clocks_rows=None, collective=True):
        """K steps, each bracketed by CUDA events on the layer's stream (the stream every kernel of cudaRender is launched on);
        L2 flushed between steps.  Returns device ms (sum over steps), wall s and the summed per-step stats of this rank."""
        lib, cabi = self.lib, self.cabi
        stop = threading.Event(); sampler = None
        if clocks_rows is not None and self.rank == 0:
            sampler = threading.Thread(target=clock_sampler, args=(stop, clocks_rows, self.local), daemon=True); sampler.start()
            time.sleep(0.3)
        if collective:
            self.barrier()   # after the sampler start-up: a late rank would otherwise make the others wait inside their first reduce
        tot = {"dev_ms": 0.0, "kernel_ms": 0.0, "rays": 0, "launches": 0, "cand": 0, "nodes": 0, "tests": 0, "primary_ms": 0.0, "primary_rays": 0}
        t0 = time.perf_counter()
        for _ in range(steps):
            cabi.check(lib.cudaFlushL2())
            cabi.check(lib.cudaStreamMark(0))
            step()
            cabi.check(lib.cudaStreamMark(1))
            ms = ctypes.c_float(); cabi.check(lib.cudaStreamElapsed(0, 1, ctypes.byref(ms)))
            st = self.host.stats()
            tot["dev_ms"] += ms.value; tot["kernel_ms"] += st.renderKernelMs; tot["rays"] += st.rays; tot["launches"] += st.kernelLaunches
            tot["cand"] += st.candidates; tot["nodes"] += st.nodeVisits; tot["tests"] += st.sphereTests
            tot["primary_ms"] += st.primaryKernelMs; tot["primary_rays"] += st.primaryRays
        tot["wall_s"] = time.perf_counter() - t0
        if collective:
            self.barrier()
        if sampler is not None:
            time.sleep(0.15); stop.set(); sampler.join(timeout=3)
        return tot

    def leg(self, name, traversal, steps, warmup, scene=None, sharded=False, e2e=True, clocks=True):
        """One measured workload: `warmup` + `steps` resident steps, then (optionally) as many through host buffers.
        Returns the summary dict (rank 0; every rank must call it when `sharded`)."""
        host, cabi, lib = self.host, self.cabi, self.lib
        w, h, spp, depth, scene_kind, desc = WORKLOADS[name]
        if scene is None:
            scene = build_scene(host, scene_kind)
        count = scene.shape[0]
        world = self.world if sharded else 1
        offset, my_spp = self.sharding.shard_samples(spp, world, self.rank) if sharded else (0, spp)
        flags = 0 if sharded else cabi.VCRT_FLAG_NO_REDUCE
        host.upload_scene(scene)
        params = host.render_params(w, h, my_spp, depth=depth, seed=self.args.seed, sample_offset=offset, traversal=traversal, total_samples=spp, flags=flags)
        scene_ptr = scene.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere))
        image = self.host_image(w, h)

        def step_resident():
            cabi.check(lib.cudaRender(ctypes.byref(params)))

        def step_e2e():
            cabi.check(lib.cudaUploadScene(scene_ptr, count, ctypes.byref(self.cam)))
            cabi.check(lib.cudaRender(ctypes.byref(params)))
            if self.rank == 0 or not sharded:
                cabi.check(lib.cudaDownloadImage(image.ctypes.data_as(ctypes.c_void_p), image.nbytes))

        for _ in range(warmup):
            step_resident()
        rows = [] if clocks else None
        r = self.timed(step_resident, steps, rows, collective=sharded)
        used = host.stats().traversalUsed
        dev_ms = self.max_over_ranks(r["dev_ms"]) if sharded else r["dev_ms"]
        rays = self.sum_over_ranks(r["rays"]) if sharded else r["rays"]
        samples = w * h * spp * steps
        out = {"workload": desc, "traversal": {1: "brute", 2: "lbvh", 3: "simple"}.get(used, "?"), "kernel": KERNEL_NAMES.get(used, "?"), "spheres": count,
               "value": samples / (dev_ms * 1e-3), "unit": "samples/s", "ms_per_step": dev_ms / steps, "steps": steps, "warmup": warmup,
               "mrays_per_sec": rays / (dev_ms * 1e-3) * 1e-6, "rays_per_sample": rays / samples,
               "kernel_ms_per_step": r["kernel_ms"] / steps, "gpu_launches": int(r["launches"])}
        if used == 2:
            out["node_visits_per_ray"] = r["nodes"] / max(1, r["rays"]); out["sphere_tests_per_ray"] = r["tests"] / max(1, r["rays"])
        if used == 1:
            out["filter_candidates_per_ray"] = r["cand"] / max(1, r["rays"])
        if clocks:
            out["clocks"] = summarise_clocks(rows)
        out["_raw"] = dict(r, rays_all=rays, dev_ms_max=dev_ms, count=count, my_spp=my_spp, used=used, scene_bytes=int(scene.nbytes), image_bytes=int(image.nbytes), params=params, image=image, scene=scene)
        if e2e:
            step_e2e()
            e = self.timed(step_e2e, steps, None, collective=sharded)
            e_wall = self.max_over_ranks(e["wall_s"]) if sharded else e["wall_s"]
            out["e2e"] = {"value": samples / e_wall, "unit": "samples/s", "h2d_bytes_per_step": int(scene.nbytes + ctypes.sizeof(self.cam)),
                          "d2h_bytes_per_step": int(image.nbytes),
                          "timing": "wall clock around cudaUploadScene + cudaRender + cudaDownloadImage from / to host memory, max over ranks"}
        return out

    def crc(self, image):
        return zlib.crc32(self.np.ascontiguousarray(image).tobytes())


def roofline(b, leg, name):
    """roofline object of a leg's path-tracing kernel: FP32 FMA pipe, denominator = nominal peak at the SM clock sampled under load."""
    raw = leg["_raw"]
    sms, max_khz, gpu_name = b.host.device_info()
    clocks = leg.get("clocks") or {}
    mhz = clocks.get("sm_mhz") or max_khz / 1e3
    peak = sms * 128 * 2 * mhz * 1e6 * 1e-12
    if not hasattr(b, "ffma"):
        b.ffma = b.host.fp32_peak()
    ffma_tf, ffma_mhz = b.ffma
    kernel_s = raw["kernel_ms"] * 1e-3
    out = {"bound": "fp32", "unit": "TFLOP/s", "peak": peak, "kernel": leg["kernel"],
           "peak_source": f"nominal {sms} SMs x 128 FP32 lanes x 2 flop x {mhz:.0f} MHz (median SM clock sampled by nvidia-smi during the timed steps); "
                          f"MEASURED_PEAKS.json has no FP32 figure; an FFMA-only kernel (cudaMeasureFp32Peak, 512 FFMAs per loop trip) reached "
                          f"{ffma_tf:.2f} TFLOP/s = {ffma_tf / peak:.3f} of it in this run",
           "ffma_only_kernel_tflops": ffma_tf, "kernel_ms_per_step": leg["kernel_ms_per_step"],
           "kernel_share_of_step": raw["kernel_ms"] / raw["dev_ms"] if raw["dev_ms"] else None, "traffic": None}
    if raw["used"] == 1:
        # Two kernels share a brute-force step: primary_scan_kernel (camera rays + first scatter event) and render_scanu_kernel (every
        # other ray).  The roofline object is that of the DOMINANT one, the path kernel, from its own CUDA-event time; the primary
        # kernel and the step as a whole ride beside it.
        n = raw["count"]
        p_rays, p_s = raw["primary_rays"], raw["primary_ms"] * 1e-3
        s_rays, s_s = raw["rays"] - p_rays, kernel_s - p_s
        out["kernel_ms_per_step"] = s_s * 1e3 / leg["steps"]
        out["kernel_share_of_step"] = s_s * 1e3 / raw["dev_ms"] if raw["dev_ms"] else None
        out["achieved"] = s_rays * n * FLOP_PER_TEST / s_s * 1e-12
        out["algorithmic"] = (f"rays of this kernel x {n} spheres x {FLOP_PER_TEST} flop (SURVEY.md 8d: the reference's test with a and r^2 hoisted); "
                              "rays from the in-kernel counters, this rank; time = CUDA events around this kernel's launches")
        # candidates are counted over both kernels; split them by the rays each kernel traced
        cand_s = raw["cand"] * s_rays / max(1, raw["rays"])
        executed = (s_rays * n * FLOP_PER_FILTER + cand_s * FLOP_PER_TEST) / s_s * 1e-12
        out["achieved_executed"] = executed
        out["frac_executed"] = executed / peak
        out["executed"] = f"what the kernel issues: {FLOP_PER_FILTER} flop (7 FMAs of the conservative filter) per (ray, sphere) + {FLOP_PER_TEST} per exact test of a survivor; shading not counted"
        out["filter_candidates_per_ray"] = leg.get("filter_candidates_per_ray")
        if p_rays and p_s > 0:
            out["primary_kernel"] = {"kernel": PRIMARY_KERNEL, "kernel_ms_per_step": p_s * 1e3 / leg["steps"], "share_of_step": p_s * 1e3 / raw["dev_ms"],
                                     "rays_per_step": p_rays / leg["steps"], "achieved": p_rays * n * FLOP_PER_TEST / p_s * 1e-12,
                                     "frac": p_rays * n * FLOP_PER_TEST / p_s * 1e-12 / peak,
                                     "frac_executed": p_rays * n * FLOP_PER_PRIMARY_FILTER / p_s * 1e-12 / peak,
                                     "executed": f"{FLOP_PER_PRIMARY_FILTER} flop per (camera ray, sphere): the camera-relative filter"}
        try:   # the primary kernel's DRAM traffic and pipe figures from the same committed capture
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                t = json.load(f).get(PRIMARY_KERNEL, {}).get(name)
            if t and "primary_kernel" in out:
                out["primary_kernel"].update({"traffic": t["dram_bytes_per_launch"], "ncu_fma_pipe_active_pct": t.get("fma_pipe_active_pct"),
                                              "ncu_issue_active_pct": t.get("issue_active_pct"), "ncu_threads_per_instruction": t.get("threads_per_instruction")})
        except (OSError, ValueError):
            pass
        out["step"] = {"kernels": [PRIMARY_KERNEL, leg["kernel"]] if p_rays else [leg["kernel"]], "kernel_ms_per_step": kernel_s * 1e3 / leg["steps"],
                       "achieved": raw["rays"] * n * FLOP_PER_TEST / kernel_s * 1e-12, "frac": raw["rays"] * n * FLOP_PER_TEST / kernel_s * 1e-12 / peak,
                       "note": "all rays of the step x spheres x 17 flop over the time of both kernels: the algorithmic rate of the brute-force path as a whole"}
    else:
        flop = raw["nodes"] * 2 * FLOP_PER_BOX + raw["tests"] * FLOP_PER_TEST
        out["achieved"] = flop / kernel_s * 1e-12
        out["bound"] = "issue slots / divergence (tree traversal); FP32 flop shown for scale"
        out["algorithmic"] = f"node visits x 2 boxes x {FLOP_PER_BOX} flop + leaf sphere tests x {FLOP_PER_TEST} flop, from the in-kernel counters"
        out["node_visits_per_ray"] = leg.get("node_visits_per_ray"); out["sphere_tests_per_ray"] = leg.get("sphere_tests_per_ray")
    out["frac"] = out["achieved"] / peak
    # traffic + pipe utilisation: from the committed ncu --set full capture of this kernel on this workload (profiles/traffic.json)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(leg["kernel"], {}).get(name)
        if t:
            out["traffic"] = t["dram_bytes_per_launch"]; out["traffic_source"] = t["source"]
            for key in ("fma_pipe_active_pct", "issue_active_pct", "threads_per_instruction"):
                if key in t:
                    out["ncu_" + key] = t[key]
    except (OSError, ValueError):
        pass
    return out, gpu_name


def public(leg):
    return {k: v for k, v in leg.items() if k != "_raw"}


def run_b200(args):
    b = Bench(args)
    np, host, cabi, lib = b.np, b.host, b.cabi, b.lib
    rank, world = b.rank, b.world
    name = args.workload or ("c2" if world == 1 else "c3")
    w, h, spp, depth, scene_kind, desc = WORKLOADS[name]
    # headline traversal: north_star's linear scan (every ray tests every sphere; FP32-FMA roofline) for the
    # 485-sphere workloads, the GPU-built LBVH for the 10k / 100k scenes that cannot be scanned
    default_traversal = "lbvh" if name.startswith("c4") else "brute"
    traversal = TRAVERSALS[args.traversal or default_traversal]
    failures = []

    head = b.leg(name, traversal, args.steps, args.warmup, sharded=world > 1)
    raw = head["_raw"]
    per_rank_ms = [raw["dev_ms"]]
    if world > 1:
        gathered = [b.torch.zeros(1, dtype=b.torch.float64, device=f"cuda:{b.local}") for _ in range(world)]
        b.dist.all_gather(gathered, b.torch.tensor([raw["dev_ms"]], dtype=b.torch.float64, device=f"cuda:{b.local}"))
        per_rank_ms = [float(g.item()) for g in gathered]
    head_image = raw["image"].copy() if rank == 0 else None   # the frame of the last e2e step = the frame that was timed

    # ---- the same workload the way a default caller runs it (VCRT_TRAVERSAL_AUTO -> the LBVH), as a second headline
    auto = None
    if raw["used"] == 1 and not args.no_extras:
        auto = b.leg(name, TRAVERSALS["auto"], args.steps, 3, scene=raw["scene"], sharded=world > 1)
        auto["speedup_vs_headline"] = auto["value"] / head["value"]

    parity = {}
    if world == 1:
        # ---- what was timed is what is checked: the e2e frame of the headline against (a) the same frame through the other
        # traversal - two different algorithms, same Philox draws, order-free fixed-point sums: the bits must agree - and (b) the
        # in-order exhaustive kernel (reference statement order, float sums) on a 2-spp replica of the frame: |diff| <= 2^-22
        parity["image_checked"] = True
        parity["image_crc32"] = b.crc(head_image)
        parity["range_ok"] = bool((head_image[..., 3] == 1).all() and head_image[..., :3].min() >= 0 and head_image[..., :3].max() <= 1)
        if auto is not None:
            parity["brute_equals_lbvh_bits"] = bool(np.array_equal(head_image, auto["_raw"]["image"]))
            if not parity["brute_equals_lbvh_bits"]:
                failures.append("the brute-force and LBVH images of the timed frame differ")
        host.upload_scene(raw["scene"])
        fast, sf = host.render(w, h, 2, depth=depth, seed=args.seed, traversal=traversal)
        slow, ss = host.render(w, h, 2, depth=depth, seed=args.seed, traversal=TRAVERSALS["simple"])
        parity["vs_exhaustive_in_order_kernel_2spp"] = {"max_abs_diff": float(np.abs(fast - slow).max()), "bound": 2.0 ** -22, "rays_equal": bool(sf.rays == ss.rays)}
        if not (np.abs(fast - slow).max() <= 2.0 ** -22 and sf.rays == ss.rays and parity["range_ok"]):
            failures.append("the timed configuration differs from the exhaustive in-order kernel")
    else:
        # ---- N-GPU image == 1-GPU image (the invariant is one store per pixel, shader.comp:56-57, split over
        # ranks): a reduced frame of the workload (same size, 8 spp in total and 37 = unequal shards) sharded through cudaRender
        # with the reduce, against rank 0 rendering all samples alone.  int64 sums: bit for bit.
        parity = {"ranks": world, "traversals": ["brute", "lbvh"], "frame": f"{w}x{h}", "cases": []}
        worst, same = 0.0, True
        host.upload_scene(raw["scene"])
        for trav in ("brute", "lbvh"):
            for total, flags in ((8, 0), (37, 0), (world // 2 or 1, cabi.VCRT_FLAG_ALL_REDUCE)):
                off, cnt = b.sharding.shard_samples(total, world, rank)
                img, st = host.render(w, h, cnt, depth=depth, seed=args.seed, sample_offset=off, traversal=TRAVERSALS[trav], total_samples=total, flags=flags,
                                      download=(rank == 0 or flags != 0))
                rays = b.sum_over_ranks(st.rays)
                if rank == 0:
                    full, sf = host.render(w, h, total, depth=depth, seed=args.seed, traversal=TRAVERSALS[trav], flags=cabi.VCRT_FLAG_NO_REDUCE)
                    d = float(np.abs(img - full).max()); ok = bool(np.array_equal(img, full) and int(rays) == sf.rays)
                    worst = max(worst, d); same &= ok
                    parity["cases"].append({"traversal": trav, "spp_total": total, "all_reduce": bool(flags), "equal": ok, "max_abs_diff": d})
                b.barrier()
        parity["sharded_equals_single"] = bool(same); parity["max_abs_diff"] = worst
        if rank == 0 and not same:
            failures.append("the sharded + reduced image differs from the single-GPU image")

    # ---- the other configs of BASELINE.json as short legs (N = 1), configs[2] on ONE GPU (any N: the same-config denominator)
    configs = {}
    strong = None
    if not args.no_extras:
        if world == 1 and name == "c2":
            for key, trav, steps in (("c1", "brute", 20), ("c1_auto", "auto", 20), ("c4a", "lbvh", 5), ("c4b", "lbvh", 5), ("c5", "brute", 3), ("c5_auto", "auto", 5)):
                leg = b.leg(key.split("_")[0], TRAVERSALS[trav], steps, 3)
                if key.startswith("c4") or key.startswith("c5"):
                    leg["roofline"], _ = roofline(b, leg, key.split("_")[0])
                configs[key] = public(leg)
            leg = b.leg("c3", TRAVERSALS["brute"], 2, 3, e2e=False)
            leg["note"] = "configs[2] rendered by ONE GPU (no sharding, no reduce): the same-config denominator of the N-GPU scaling factor"
            configs["c3_single_gpu"] = public(leg)
            leg = b.leg("c3", TRAVERSALS["auto"], 3, 3, e2e=False)
            configs["c3_single_gpu_auto"] = public(leg)
        elif world > 1 and name == "c3":
            # rank 0 alone renders the whole of configs[2] (the other ranks wait at the barrier): 3 warm-up + 2 timed steps
            if rank == 0:
                leg = b.leg("c3", traversal, 2, 3, scene=raw["scene"], e2e=False, clocks=False)
                strong = {"single_gpu": public(leg), "speedup": head["value"] / leg["value"], "efficiency": head["value"] / leg["value"] / world,
                          "note": "configs[2] on ONE GPU of this box (rank 0, VCRT_FLAG_NO_REDUCE) against the N-GPU value of this line: strong scaling at equal config"}
            b.barrier()

    if rank == 0:
        roof, gpu_name = roofline(b, head, name)
        line = {
            "metric": "samples_per_sec", "value": head["value"], "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": desc, "spheres": raw["count"], "spp_per_gpu": raw["my_spp"],
                       "parallelism": f"spp-shard x{world} + 1 NCCL reduce (int64 fixed-point sums)" if world > 1 else "single GPU",
                       "traversal": head["traversal"], "rng": "philox4x32-10", "seed": args.seed, "l2": "flushed between timed steps (256 MB write)",
                       "n1_vs_nN": "N=1 runs configs[1], N>1 runs configs[2] (total work fixed for N >= 2: strong scaling); samples/s is comparable across the two "
                                   "(same scene, framing and per-sample cost) and configs[2] on ONE GPU is measured too: `configs.c3_single_gpu` at N=1, `strong_scaling` at N>1"},
            "mrays_per_sec": head["mrays_per_sec"], "rays_per_sample": head["rays_per_sample"],
            "sphere_tests_per_sec": raw["rays_all"] * raw["count"] / (raw["dev_ms_max"] * 1e-3),
            "e2e": head["e2e"], "gpu_launches": head["gpu_launches"], "roofline": roof, "clocks": head["clocks"], "gpu": gpu_name,
            "parity": parity, "per_rank_ms_per_step": [m / args.steps for m in per_rank_ms],
        }
        if auto is not None:
            auto["roofline"], _ = roofline(b, auto, name)
            line["auto_same_workload"] = public(auto)
        if configs:
            line["configs"] = configs
        if strong is not None:
            line["strong_scaling"] = strong
        if not args.no_cpu and world == 1:   # rank 0 at N=1 only
            line["cpu_baseline"] = cpu_baseline(w, h, spp, depth, scene_kind)
            if configs and name == "c2":
                c1 = cpu_baseline(*WORKLOADS["c1"][:5], exact=True)     # C1 timed exactly
                configs["c1"]["cpu"] = c1
                configs["c1"]["gpu_over_cpu_e2e"] = configs["c1"]["e2e"]["value"] / c1["value"]
                w3, h3, s3 = WORKLOADS["c3"][:3]
                configs["c3_cpu_extrapolated"] = {"seconds": w3 * h3 * s3 / line["cpu_baseline"]["value"], "label": "EXTRAPOLATED, not run",
                                                  "from": "configs[2] samples (3840 x 2160 x 1000) / the measured CPU samples/s of cpu_baseline (same scene, depth and per-sample cost)",
                                                  "cores": line["cpu_baseline"]["cores"]}
        if failures:
            line["parity"]["failures"] = failures
        print(json.dumps(line), flush=True)
    b.barrier()
    host.destroy()
    if b.dist is not None:
        b.dist.destroy_process_group()
    return 1 if failures else 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--traversal", default=None, choices=sorted(TRAVERSALS))
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-extras", action="store_true", help="headline only: no auto / other-config / single-GPU legs")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3   # timing hygiene: at least 3 warm-up steps
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
