#!/usr/bin/env python3
# SPDX-License-Identifier: WTFPL
"""Headline benchmark: samples/s (and Mrays/s) of the per-pixel path-tracing loop.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload c1|c2|c3|c4a|c4b|c5]

A "step" is one full render of the workload's frame (every pixel, every sample, depth 50).
  N = 1  -> BASELINE.json configs[1]: RTIOW final scene (485 spheres), 1920x1080, 500 spp, depth 50
  N > 1  -> BASELINE.json configs[2]: the same scene at 3840x2160, 1000 spp in total, samples-per-pixel sharded
            over the N ranks, one NCCL reduce of the per-GPU sum buffers per step (inside cudaRender)
`value` = samples/s of the whole job with the scene resident on the device (max over ranks of the CUDA-event
time); `e2e` = the same through the C ABI from HOST buffers: cudaUploadScene (host AoS -> device) + cudaRender +
cudaDownloadImage (device -> host image) inside the timed region.  One JSON line on stdout (rank 0).

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
FLOP_PER_TEST = 17   # one ray-sphere test = 17 flop (11 FP32-pipe instructions)


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


# --------------------------------------------------------------------------------------- reference arm (CPU)
def run_reference(args):
    """The reference's CPU implementation of the path on the host cores: a bounded sample of the workload per step."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import oracle_harness as oh
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    name = args.workload or ("c2" if args.gpus == 1 else "c3")
    w, h, spp, depth, scene_kind, desc = WORKLOADS[name]
    cores = os.cpu_count() or 1
    ref = oh.Reference.open(fast=True) if scene_kind == "reference" else None
    kind = "reference" if ref is not None else "port"
    orc = None
    if ref is None:
        orc = oh.Oracle(fast=True)
        scene = orc.scene() if scene_kind == "reference" else (orc.scene(11, 0.0, 0.5) if scene_kind == "stress" else orc.scene(int(scene_kind[4:])))
    # bounded sample: the full frame at reduced spp, sized from a quick calibration to ~6 s per step
    def render(s):
        t = time.perf_counter()
        if ref is not None:
            _, st = ref.render(w, h, s, depth, threads=cores)
        else:
            _, st = orc.render(scene, w, h, s, depth=depth, rng=oh.RNG_HASH, threads=cores)
        return time.perf_counter() - t, st
    cw, ch = max(1, w // 8), max(1, h // 8)
    render(1) if w * h <= 100000 else None   # thread pool start-up stays out of the calibration
    t = time.perf_counter()
    if ref is not None:
        ref.render(cw, ch, 2, depth, threads=cores)
    else:
        orc.render(scene, cw, ch, 2, depth=depth, rng=oh.RNG_HASH, threads=cores)
    per_sample = (time.perf_counter() - t) / (cw * ch * 2)
    step_spp = max(1, min(spp, int(6.0 / (per_sample * w * h))))
    for _ in range(args.warmup if args.warmup < 2 else 1):
        render(step_spp)
    total_t, rays = 0.0, 0
    for _ in range(args.steps):
        dt, st = render(step_spp)
        total_t += dt; rays += st["rays"]
    samples = w * h * step_spp * args.steps
    value = samples / total_t
    sample = f"full {w}x{h} frame, depth {depth}, {step_spp} of {spp} spp per step, {args.steps} steps, hash RNG (reference B2)"
    line = {
        "impl": "reference", "metric": "samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": total_t / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": desc, "sample": sample},
        "mrays_per_sec": rays / total_t * 1e-6,
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline(w, h, spp, depth, scene_kind, budget_s=12.0):
    """cpu_baseline leg of the GPU arm: the same CPU code, ~10-15 s on rank 0's host cores."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_harness as oh
    cores = os.cpu_count() or 1
    ref = oh.Reference.open(fast=True) if scene_kind == "reference" else None
    if ref is None:
        orc = oh.Oracle(fast=True)
        scene = orc.scene() if scene_kind == "reference" else (orc.scene(11, 0.0, 0.5) if scene_kind == "stress" else orc.scene(int(scene_kind[4:])))
        run = lambda ww, hh, s: orc.render(scene, ww, hh, s, depth=depth, rng=oh.RNG_HASH, threads=cores)
    else:
        run = lambda ww, hh, s: ref.render(ww, hh, s, depth, threads=cores)
    cw, ch = max(1, w // 8), max(1, h // 8)
    run(cw, ch, 1)   # thread pool start-up stays out of the calibration
    t = time.perf_counter(); run(cw, ch, 2); per_sample = (time.perf_counter() - t) / (cw * ch * 2)
    s = max(1, min(spp, int(budget_s / (per_sample * w * h))))
    t = time.perf_counter(); _, st = run(w, h, s); dt = time.perf_counter() - t
    return {"value": w * h * s / dt, "unit": "samples/s", "cores": cores, "kind": "reference" if ref is not None else "port",
            "sample": f"full {w}x{h} frame, depth {depth}, {s} of {spp} spp, hash RNG (reference B2), {dt:.1f} s",
            "mrays_per_sec": st["rays"] / dt * 1e-6}


# --------------------------------------------------------------------------------------- B200 arm
def run_b200(args):
    import numpy as np
    import torch
    from vulkancomputeraytracing_b200 import cabi, host, sharding
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with torchrun: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N")
        args.gpus = world
    lib = cabi.load()
    cabi.check(lib.cudaSelectDevice(local))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        sharding.init_communicator(dist, device=torch.device("cuda", local))

    name = args.workload or ("c2" if world == 1 else "c3")
    w, h, spp, depth, scene_kind, desc = WORKLOADS[name]
    scene = build_scene(host, scene_kind)
    count = scene.shape[0]
    offset, my_spp = sharding.shard_samples(spp, world, rank)
    # headline traversal: north_star's linear scan (every ray tests every sphere; FP32-FMA roofline) for the
    # 485-sphere workloads, the GPU-built LBVH for the 10k / 100k scenes that cannot be scanned
    default_traversal = "lbvh" if name.startswith("c4") else "brute"
    traversal = {"auto": 0, "brute": 1, "lbvh": 2, "simple": 3}[args.traversal or default_traversal]
    host.upload_scene(scene)
    params = host.render_params(w, h, my_spp, depth=depth, seed=args.seed, sample_offset=offset, traversal=traversal, total_samples=spp)
    image = np.empty((h, w, 4), np.float32)
    # pinned host image for the e2e leg
    from torch import from_numpy
    pinned = torch.empty((h, w, 4), dtype=torch.float32).pin_memory() if torch.cuda.is_available() else None
    host_image = pinned.numpy() if pinned is not None else image
    scene_ptr = scene.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)); cam = host.default_camera()

    def barrier():
        if world > 1:
            cabi.check(lib.cudaCommBarrier())

    def step_resident():
        cabi.check(lib.cudaRender(ctypes.byref(params)))

    def step_e2e():
        cabi.check(lib.cudaUploadScene(scene_ptr, count, ctypes.byref(cam)))
        cabi.check(lib.cudaRender(ctypes.byref(params)))
        if rank == 0:
            cabi.check(lib.cudaDownloadImage(host_image.ctypes.data_as(ctypes.c_void_p), host_image.nbytes))

    def timed(step, steps, clocks_rows=None):
        """K steps, each bracketed by CUDA events on the layer's stream; L2 flushed between steps. Returns per-rank
        device ms (sum), wall s, and the per-step stats."""
        stop = threading.Event(); sampler = None
        if clocks_rows is not None and rank == 0:
            sampler = threading.Thread(target=clock_sampler, args=(stop, clocks_rows, local), daemon=True); sampler.start()
            time.sleep(0.3)
        barrier()   # after the sampler start-up: a late rank would otherwise make the others wait inside their first reduce
        dev_ms, kernel_ms, rays, launches, cand = 0.0, 0.0, 0, 0, 0
        t0 = time.perf_counter()
        for _ in range(steps):
            cabi.check(lib.cudaFlushL2())
            cabi.check(lib.cudaStreamMark(0))
            step()
            cabi.check(lib.cudaStreamMark(1))
            ms = ctypes.c_float(); cabi.check(lib.cudaStreamElapsed(0, 1, ctypes.byref(ms)))
            st = host.stats()
            dev_ms += ms.value; kernel_ms += st.renderKernelMs; rays += st.rays; launches += st.kernelLaunches; cand += st.candidates
        wall = time.perf_counter() - t0
        barrier()
        if sampler is not None:
            stop.set(); sampler.join(timeout=3)
        return dev_ms, wall, kernel_ms, rays, launches, cand

    def max_over_ranks(v):
        out = ctypes.c_float(); cabi.check(lib.cudaCommMaxFloat(ctypes.c_float(v), ctypes.byref(out))); return out.value

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([float(v)], dtype=torch.float64, device=f"cuda:{local}"); dist.all_reduce(t); return float(t.item())

    for _ in range(args.warmup):
        step_resident()
    clocks_rows = []
    dev_ms, wall, kernel_ms, rays, launches, cand = timed(step_resident, args.steps, clocks_rows)
    dev_ms_max = max_over_ranks(dev_ms)
    per_rank_ms = [dev_ms]
    if world > 1:
        gathered = [torch.zeros(1, dtype=torch.float64, device=f"cuda:{local}") for _ in range(world)]
        dist.all_gather(gathered, torch.tensor([dev_ms], dtype=torch.float64, device=f"cuda:{local}"))
        per_rank_ms = [float(g.item()) for g in gathered]
    total_rays = sum_over_ranks(rays)
    kernel_ms_max = max_over_ranks(kernel_ms)
    for _ in range(1):
        step_e2e()
    e_ms, e_wall, *_ = timed(step_e2e, args.steps)
    # the same workload through the LBVH (optional in north_star, but faster even at 485 spheres): reported beside the headline
    lbvh = None
    if traversal == 1:
        params.traversal = 2
        for _ in range(2):
            step_resident()
        l_ms, _, l_kernel, l_rays, _, _ = timed(step_resident, args.steps)
        l_ms_max = max_over_ranks(l_ms); l_rays_total = sum_over_ranks(l_rays)
        st_l = host.stats()
        lbvh = {"value": w * h * spp * args.steps / (l_ms_max * 1e-3), "unit": "samples/s", "ms_per_step": l_ms_max / args.steps,
                "mrays_per_sec": l_rays_total / (l_ms_max * 1e-3) * 1e-6, "node_visits_per_ray": st_l.nodeVisits / max(1, st_l.rays),
                "sphere_tests_per_ray": st_l.sphereTests / max(1, st_l.rays)}
        params.traversal = traversal
        step_resident()   # leave the layer's stats / image on the headline traversal
    # e2e is wall-clock around the public calls (uploads, render, read-back), max over ranks
    e_wall_max = max_over_ranks(e_wall)

    samples_per_step = w * h * spp
    value = samples_per_step * args.steps / (dev_ms_max * 1e-3)
    e2e_value = samples_per_step * args.steps / e_wall_max
    if rank == 0:
        peak_tf, peak_mhz = host.fp32_peak()
        sms, max_khz, gpu_name = host.device_info()
        # roofline of the dominant kernel (the persistent path-tracing kernel), live CUDA-event time of that kernel alone
        my_tests = rays * count
        achieved = my_tests * FLOP_PER_TEST / (kernel_ms * 1e-3) * 1e-12 if kernel_ms > 0 else 0.0
        nominal_peak = sms * 128 * 2 * max_khz * 1e3 * 1e-12
        clocks = summarise_clocks(clocks_rows)
        line = {
            "metric": "samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms_max / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": desc, "spheres": count, "spp_per_gpu": my_spp, "parallelism": f"spp-shard x{world} + 1 NCCL reduce" if world > 1 else "single GPU",
                       "traversal": {1: "brute", 2: "lbvh", 3: "simple"}.get(host.stats().traversalUsed, "?"), "rng": "philox4x32-10", "seed": args.seed,
                       "l2": "flushed between timed steps (256 MB write)",
                       "n1_vs_nN": "N=1 runs configs[1], N>1 runs configs[2]; samples/s is comparable (same scene and framing)"},
            "mrays_per_sec": total_rays / (dev_ms_max * 1e-3) * 1e-6,
            "rays_per_sample": total_rays / (samples_per_step * args.steps),
            "sphere_tests_per_sec": total_rays * count / (dev_ms_max * 1e-3),
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": int(scene.nbytes + ctypes.sizeof(cam)),
                    "d2h_bytes_per_step": int(host_image.nbytes), "timing": "wall clock around cudaUploadScene + cudaRender + cudaDownloadImage, max over ranks"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf if peak_tf else None,
                         "traffic": None, "kernel": "render_scanu_kernel" if host.stats().traversalUsed == 1 else "render_lbvh_kernel",
                         "kernel_ms_per_step": kernel_ms / args.steps, "kernel_share_of_step": kernel_ms / dev_ms if dev_ms else None,
                         "algorithmic": f"rays x {count} spheres x {FLOP_PER_TEST} flop (SURVEY.md 8d); rays from the in-kernel counter",
                         "peak_source": f"FFMA-only kernel measured live on this GPU (cudaMeasureFp32Peak, {peak_mhz:.0f} MHz effective); "
                                        f"MEASURED_PEAKS.json has no FP32 figure; nominal {sms} SMs x 128 x 2 x {max_khz / 1e3:.0f} MHz = {nominal_peak:.1f}",
                         "filter_candidates_per_ray": cand / rays if rays else None},
            "clocks": clocks, "gpu": gpu_name,
        }
        # roofline.traffic: DRAM bytes of one launch of this kernel on this workload, from a committed ncu --set full capture
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                t = json.load(f).get(line["roofline"]["kernel"], {}).get(name)
            if t:
                line["roofline"]["traffic"] = t["dram_bytes_per_launch"]
                line["roofline"]["traffic_source"] = t["source"]
        except (OSError, ValueError):
            pass
        if lbvh is not None:
            lbvh["speedup_vs_headline"] = lbvh["value"] / value
            line["lbvh_same_workload"] = lbvh
        if host.stats().traversalUsed == 2:
            st = host.stats()
            line["roofline"].update({"bound": "latency (divergent tree traversal; FP32 flop shown for scale)",
                                     "algorithmic": "node visits x 2 boxes x 24 flop + leaf sphere tests x 17 flop, from the in-kernel counters",
                                     "achieved": (st.nodeVisits * 48 + st.sphereTests * 17) / (st.renderKernelMs * 1e-3) * 1e-12,
                                     "node_visits_per_ray": st.nodeVisits / max(1, st.rays), "sphere_tests_per_ray": st.sphereTests / max(1, st.rays)})
            line["roofline"]["frac"] = line["roofline"]["achieved"] / peak_tf
        line["per_rank_ms_per_step"] = [m / args.steps for m in per_rank_ms]
        if not args.no_cpu and world == 1:   # rank 0 at N=1 only
            line["cpu_baseline"] = cpu_baseline(w, h, spp, depth, scene_kind)
        print(json.dumps(line), flush=True)
    barrier()
    host.destroy()
    if dist is not None:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS))
    ap.add_argument("--traversal", default=None, choices=["auto", "brute", "lbvh", "simple"])
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return run_reference(args)
    if args.warmup < 3:
        args.warmup = 3   # timing hygiene: at least 3 warm-up steps
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
