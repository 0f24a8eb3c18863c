# SPDX-License-Identifier: WTFPL
"""Parity of the CUDA path (through the C ABI, lib/libvcrt.so) against the CPU oracle.

Bar: BIT-EXACT.  With the same Philox draws the kernels perform the reference's binary32
operations in the reference's order, so image, ray count and depth-exhaustion count must equal the
oracle's exactly when samples are summed in index order (sampleChunks = 1).  The default
accumulation (32.32 fixed point, order-free) is bounded against that.  Against the reference's OWN
random numbers (the sin-hash, bug B2, not reproducible on a GPU) parity is statistical, with the
tolerances of BASELINE.json: channel means < 0.5 %, RMSE < 1e-2 at >= 128 spp."""
import ctypes
import json
import os
import subprocess
import sys
import zlib

import numpy as np
import pytest

import oracle_harness as oh
from vulkancomputeraytracing_b200 import cabi, host

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
BRUTE, SIMPLE, LBVH = cabi.VCRT_TRAVERSAL_BRUTE, cabi.VCRT_TRAVERSAL_SIMPLE, cabi.VCRT_TRAVERSAL_LBVH


def assert_same(gpu, st, cpu, cst):
    assert st.rays == cst["rays"], (st.rays, cst["rays"])
    assert st.depthExhausted == cst["exhausted"]
    diff = np.abs(gpu - cpu)
    assert np.array_equal(gpu, cpu), f"max |diff| {diff.max():.3e} on {(diff.max(-1) > 0).mean():.4%} of the pixels"


@pytest.mark.parametrize("traversal", [BRUTE, SIMPLE, LBVH])
@pytest.mark.parametrize("shape", [(64, 36, 2, 50), (203, 113, 3, 50), (400, 225, 1, 7)])
def test_bit_exact_vs_oracle_reference_scene(oracle, gpu_scene, traversal, shape):
    w, h, spp, depth = shape
    cpu, cst = oracle.render(gpu_scene, w, h, spp, depth=depth)
    gpu, st = host.render(w, h, spp, depth=depth, traversal=traversal, sample_chunks=1)
    assert st.traversalUsed == traversal and st.kernelLaunches >= 2
    assert_same(gpu, st, cpu, cst)


def test_config0_bit_exact_and_statistical_vs_reference_golden(oracle, gpu_scene):
    # BASELINE.json configs[0]: 400x225, 16 spp, depth 50
    cpu, cst = oracle.render(gpu_scene, 400, 225, 16)
    gpu, st = host.render(400, 225, 16, traversal=BRUTE, sample_chunks=1)
    assert_same(gpu, st, cpu, cst)
    with open(os.path.join(GOLDEN, "ref_kat.json")) as f:
        want = json.load(f)["render_400x225_s16"]          # the reference's shader code, its own hash "RNG"
    mean = gpu[..., :3].reshape(-1, 3).astype(np.float64).mean(0)
    assert np.all(np.abs(mean / np.array(want["mean_rgb"]) - 1) < 5e-3)
    assert abs(st.rays / want["rays"] - 1) < 5e-3
    blocks = gpu[:224, :, :3].astype(np.float64).reshape(28, 8, 50, 8, 3).mean((1, 3))
    assert np.sqrt(np.mean((blocks - np.load(os.path.join(GOLDEN, "ref_hash_400x225_s16_blocks8.npy"))) ** 2)) < 1e-2
    assert (gpu[..., 3] == 1).all() and gpu[..., :3].min() >= 0 and gpu[..., :3].max() <= 1


def test_default_fixed_point_accumulation_is_order_free_and_close(oracle, gpu_scene):
    w, h, spp = 160, 90, 12
    cpu, cst = oracle.render(gpu_scene, w, h, spp)
    a, sa = host.render(w, h, spp, traversal=BRUTE)                 # sampleChunks = 0: 32.32 fixed-point sums
    b, sb = host.render(w, h, spp, traversal=BRUTE)
    assert np.array_equal(a, b) and sa.rays == sb.rays == cst["rays"]   # any schedule, same bits
    assert np.abs(a - cpu).max() <= 2.0 ** -22                      # vs float sums in sample order: rounding of the sum only
    # two half renders accumulated == one render, to the bit (integer sums)
    host.render(w, h, spp // 2, traversal=BRUTE, download=False)
    c, sc = host.render(w, h, spp // 2, traversal=BRUTE, sample_offset=spp // 2, flags=cabi.VCRT_FLAG_ACCUMULATE)
    assert np.array_equal(a, c)


@pytest.mark.parametrize("traversal", [BRUTE, SIMPLE, LBVH])
def test_seed_offset_and_camera_are_honoured(oracle, vcrt, traversal):
    scene = host.generate_scene(4)
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(-6, 3.5, 9), (ctypes.c_float * 3)(0.5, 0.4, -1), (ctypes.c_float * 3)(0.1, 1, 0), 35.0)
    host.upload_scene(scene, cam)
    try:
        cpu, cst = oracle.render(scene, 120, 120, 2, seed=0x1234567890ABCDEF, sample_offset=77, camera=host.camera_array(cam))
        gpu, st = host.render(120, 120, 2, seed=0x1234567890ABCDEF, sample_offset=77, traversal=traversal, sample_chunks=1)
        assert_same(gpu, st, cpu, cst)
        other, _ = host.render(120, 120, 2, seed=1, traversal=traversal, sample_chunks=1)
        assert not np.array_equal(other, gpu)
    finally:
        host.destroy()


def random_scene(rng, count, big=0):
    s = np.zeros((count, 10), np.float32)
    s[:, 0:3] = rng.uniform(-6, 6, (count, 3))
    s[:, 3] = rng.uniform(0.05, 1.2, count)
    s[:big, 3] = rng.uniform(60, 400, big)                       # huge spheres (the "always tested" list)
    s[:big, 0:3] = rng.uniform(-500, 500, (big, 3))
    s[:, 4:7] = rng.uniform(0.2, 1, (count, 3))
    s[:, 7] = rng.integers(1, 4, count)
    s[:, 8] = np.where(s[:, 7] == 3, 1.5, rng.uniform(0, 1, count))
    return s


@pytest.mark.parametrize("traversal", [BRUTE, LBVH])
@pytest.mark.parametrize("case", ["overlapping", "inside", "huge", "stress", "single", "empty", "dense_line"])
def test_bit_exact_on_adversarial_scenes(oracle, vcrt, traversal, case):
    rng = np.random.default_rng(zlib.crc32(case.encode()))
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(0, 1, 14), (ctypes.c_float * 3)(0, 0, 0), (ctypes.c_float * 3)(0, 1, 0), 50.0)
    if case == "overlapping":
        scene = random_scene(rng, 300)                           # mutually intersecting spheres of all materials
    elif case == "inside":
        scene = random_scene(rng, 60)
        scene[0] = [0, 1, 14, 5.0, .8, .8, .8, 3, 1.5, 0]        # the camera sits inside a glass sphere
        scene[1] = [0, 1, 14, 30.0, .9, .7, .5, 2, 0.3, 0]       # ... which sits inside a metal one
    elif case == "huge":
        scene = random_scene(rng, 120, big=5)
    elif case == "stress":
        scene = host.stress_scene()
        cam = host.default_camera()
    elif case == "single":
        scene = np.array([[0, 0, 0, 2.0, .5, .6, .7, 1, 0.9, 0]], np.float32)
    elif case == "empty":
        scene = np.zeros((0, 10), np.float32)
    else:                                                        # 200 spheres strung along the view axis: long candidate lists (overflow path)
        scene = random_scene(rng, 200)
        scene[:, 0] = rng.normal(0, 0.02, 200); scene[:, 1] = 1 + rng.normal(0, 0.02, 200)
        scene[:, 2] = np.linspace(-40, 10, 200); scene[:, 3] = 0.3; scene[:, 7] = 3; scene[:, 8] = 1.5
    host.upload_scene(scene, cam)
    try:
        cpu, cst = oracle.render(scene, 96, 64, 2, depth=20, camera=host.camera_array(cam))
        gpu, st = host.render(96, 64, 2, depth=20, traversal=traversal, sample_chunks=1)
        assert_same(gpu, st, cpu, cst)
        if case == "dense_line" and traversal == BRUTE:
            assert st.listOverflows > 0                          # the exhaustive fallback really ran
        # the default accumulation (fixed-point sums; brute force: primary pass + compact hand-over list, whose "not pre-traced"
        # entries the overflowing lists of dense_line produce): the same paths, sums that differ by their rounding only
        fixed, sf = host.render(96, 64, 2, depth=20, traversal=traversal)
        assert sf.rays == st.rays and sf.depthExhausted == st.depthExhausted
        assert np.abs(fixed - gpu).max() <= 2.0 ** -22 * max(1.0, float(np.abs(gpu[..., :3]).max()))
    finally:
        host.destroy()


@pytest.mark.parametrize("traversal", [BRUTE, SIMPLE, LBVH])
@pytest.mark.parametrize("shape", [(1, 1, 3), (7, 3, 2), (9, 5, 1), (33, 1, 2), (1, 40, 2)])
def test_tiny_and_ragged_images(oracle, gpu_scene, traversal, shape):
    w, h, spp = shape                                            # bug B9: every pixel is rendered, whatever the tile size
    cpu, cst = oracle.render(gpu_scene, w, h, spp)
    gpu, st = host.render(w, h, spp, traversal=traversal, sample_chunks=1)
    assert_same(gpu, st, cpu, cst)


@pytest.mark.parametrize("traversal", [BRUTE, SIMPLE, LBVH])
def test_depth_limits(oracle, gpu_scene, traversal):
    img, st = host.render(32, 18, 2, depth=0, traversal=traversal)     # no pass of functions.glsl:73 runs: black (bug B7)
    assert (img[..., :3] == 0).all() and (img[..., 3] == 1).all() and st.rays == 0 and st.depthExhausted == st.samples
    for depth in (1, 2):
        cpu, cst = oracle.render(gpu_scene, 64, 36, 2, depth=depth)
        gpu, st = host.render(64, 36, 2, depth=depth, traversal=traversal, sample_chunks=1)
        assert_same(gpu, st, cpu, cst)
        assert st.depthExhausted > 0


def test_unknown_material_is_reported(vcrt):
    scene = host.generate_scene(2)
    scene[:, 7] = 7.0                                            # bug B13: the reference silently loops to the depth limit
    host.upload_scene(scene)
    try:
        p = host.render_params(32, 18, 1, traversal=BRUTE)
        assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_SCENE
        assert host.stats().badMaterial > 0
    finally:
        host.destroy()


def test_call_order_errors(vcrt):
    host.destroy()
    p = host.render_params(16, 16, 1)
    assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_NOT_READY           # render before upload
    host.upload_scene(host.generate_scene(1))
    buf = np.zeros((16, 16, 4), np.float32)
    assert vcrt.cudaDownloadImage(buf.ctypes.data_as(ctypes.c_void_p), buf.nbytes) == cabi.VCRT_ERROR_NOT_READY
    assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_SUCCESS
    assert vcrt.cudaDownloadImage(buf.ctypes.data_as(ctypes.c_void_p), 16) == cabi.VCRT_ERROR_INVALID_ARGUMENT
    p.width = 0
    assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_INVALID_ARGUMENT
    p.width = 16; p.samplesPerPixel = 0                                              # 0 samples: only for a rank that joins a reduce
    assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_INVALID_ARGUMENT
    p.samplesPerPixel = 1
    for slot in (0, 1):                                                              # reserved fields are reserved
        p.reserved[slot] = 1
        assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_INVALID_ARGUMENT
        p.reserved[slot] = 0
    assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_SUCCESS
    host.destroy()
    assert vcrt.cudaDestroyRenderer() == cabi.VCRT_SUCCESS


def test_full_size_properties_config1(gpu_scene):
    # BASELINE.json configs[1] at full resolution (1920x1080; 128 spp keeps the test short): size-independent properties
    w, h, spp = 1920, 1080, 128
    full, st = host.render(w, h, spp, traversal=BRUTE)
    assert st.samples == w * h * spp and st.sphereTests == st.rays * 485
    assert abs(st.rays / st.samples / 2.2975 - 1) < 5e-3                            # SURVEY.md 3.4 probe
    assert (full[..., 3] == 1).all() and np.isfinite(full).all() and full[..., :3].min() >= 0 and full[..., :3].max() <= 1
    with open(os.path.join(GOLDEN, "ref_kat.json")) as f:
        want = np.array(json.load(f)["render_400x225_s16"]["mean_rgb"])
    mean = full[..., :3].reshape(-1, 3).astype(np.float64).mean(0)
    assert np.all(np.abs(mean / want - 1) < 5e-3)                                    # same framing at any 16:9 size (bug B1)
    # linearity in the sample axis: mean(first half) + mean(second half) == 2 * mean(all), bit-exact integer sums
    a, sa = host.render(w, h, spp // 2, traversal=BRUTE)
    b, sb = host.render(w, h, spp // 2, traversal=BRUTE, sample_offset=spp // 2)
    assert sa.rays + sb.rays == st.rays
    assert np.abs((a.astype(np.float64) + b) / 2 - full).max() < 1e-6
    rmse = np.sqrt(np.mean((a[..., :3].astype(np.float64) - b[..., :3]) ** 2))       # two independent 64-spp estimates
    assert rmse < 0.0805 * np.sqrt(2 / 64) * 1.25
    # 8x8 block means against the 400x225 reference-golden blocks are not comparable pixel for pixel at another size;
    # compare column / row profiles instead (image content is resolution independent)
    prof = full[..., :3].astype(np.float64).mean((1, 2))
    assert prof[0] > prof[-1] * 0.5 and np.isfinite(prof).all()


def test_host_program_end_to_end(tmp_path, oracle):
    # the headless program (Begin -> loop -> End over the C ABI) writes the same image the ABI returns
    exe = os.path.join(os.path.dirname(cabi.LIB_PATH), "VulkanComputeRayTracing")
    prefix = str(tmp_path / "frame")
    out = subprocess.run([exe, "--width", "96", "--height", "54", "--spp", "2", "--frames", "2", "--progressive", "--out", prefix],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "Hello CUDA." in out.stdout and "Bye CUDA." in out.stdout and "frame 1: 4 spp" in out.stdout
    with open(prefix + ".pfm", "rb") as f:
        assert f.readline() == b"PF\n" and f.readline() == b"96 54\n" and f.readline() == b"-1.0\n"
        pfm = np.frombuffer(f.read(), "<f4").reshape(54, 96, 3)[::-1]
    cpu, _ = oracle.render(oracle.scene(), 96, 54, 4)
    assert np.abs(pfm - cpu[..., :3]).max() <= 2.0 ** -22       # 2 progressive frames of 2 spp == 4 spp
    with open(prefix + ".ppm", "rb") as f:
        assert f.readline() == b"P6\n" and f.readline() == b"96 54\n" and f.readline() == b"255\n"
        ppm = f.read()
    with open(prefix + ".png", "rb") as f:                      # the same 8 bits in a PNG (stored deflate blocks)
        png = f.read()
    assert png[:8] == b"\x89PNG\r\n\x1a\n"
    idat = png.index(b"IDAT")
    rows = np.frombuffer(zlib.decompress(png[idat + 4:-16]), np.uint8).reshape(54, 1 + 96 * 3)
    assert rows[:, 1:].tobytes() == ppm


@pytest.mark.parametrize("grid", [50, 158])
def test_lbvh_bit_exact_on_scaled_scenes(oracle, vcrt, grid):
    # BASELINE.json configs[3]: 10 k (grid 50) and 100 k (grid 158) spheres through the GPU-built LBVH
    scene = host.generate_scene(grid)
    assert scene.shape[0] == {50: 10003, 158: 99857}[grid]
    host.upload_scene(scene)
    try:
        w, h, spp = (96, 54, 2) if grid == 50 else (64, 36, 1)
        cpu, cst = oracle.render(scene, w, h, spp)
        gpu, st = host.render(w, h, spp, traversal=LBVH, sample_chunks=1)
        assert_same(gpu, st, cpu, cst)
        assert st.nodeVisits > 0 and st.sphereTests < st.rays * scene.shape[0] // 50      # the tree really prunes
        auto, sa = host.render(w, h, spp, sample_chunks=1)                                 # AUTO picks the LBVH at this size
        assert sa.traversalUsed == LBVH and np.array_equal(auto, gpu)
        # a larger frame against the on-device exhaustive kernel (every sphere tested exactly for every ray)
        simple, ss = host.render(320, 180, 1, traversal=SIMPLE, sample_chunks=1)
        tree, stt = host.render(320, 180, 1, traversal=LBVH, sample_chunks=1)
        assert ss.rays == stt.rays and np.array_equal(simple, tree)
        fixed, _ = host.render(320, 180, 1, traversal=LBVH)                                # default fixed-point sums
        assert np.abs(fixed - tree).max() <= 2.0 ** -22
    finally:
        host.destroy()


def far_field_scene(case):
    rng = np.random.default_rng(zlib.crc32(case.encode()))
    n = 20000 if case in ("slab_3000", "slab_3000_ground", "tiny_radii") else 3000
    s = np.zeros((n, 10), np.float32)
    s[:, 4:7] = rng.uniform(0.3, 1, (n, 3)); s[:, 7] = rng.integers(1, 4, n)
    s[:, 8] = np.where(s[:, 7] == 3, 1.5, rng.uniform(0, 1, n))
    eye, at, fov = (0, 2, 0), (100, 1.5, 30), 60.0
    if case.startswith("slab_3000"):     # L up to 4000: 19 eps L^2 = 18 against r^2 = 1 .. 36
        s[:, 0] = rng.uniform(-3000, 3000, n); s[:, 1] = rng.uniform(0, 12, n); s[:, 2] = rng.uniform(-3000, 3000, n); s[:, 3] = rng.uniform(1, 6, n)
        if case.endswith("ground"): s[0] = [0, -1000, 0, 1000, .5, .5, .5, 1, 1.0, 0]
    elif case == "cube_800":
        s[:, 0:3] = rng.uniform(-800, 800, (n, 3)); s[:, 3] = rng.uniform(0.5, 8, n)
        eye, at = (10, 20, -30), (400, 100, 300)
    elif case == "far_camera":           # the whole scene is far from the origin of every primary ray
        s[:, 0:3] = rng.uniform(-40, 40, (n, 3)); s[:, 3] = rng.uniform(0.2, 1.0, n)
        eye, at, fov = (2500, 300, -1800), (0, 0, 0), 3.0
    else:                                # radii far below the error of the discriminant: what is "hit" is decided by rounding
        s[:, 0] = rng.uniform(-400, 400, n); s[:, 1] = rng.uniform(0, 3, n); s[:, 2] = rng.uniform(-400, 400, n); s[:, 3] = rng.uniform(0.02, 0.1, n)
        eye, at, fov = (-390, 1.5, -390), (400, 1.5, 400), 40.0
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(*eye), (ctypes.c_float * 3)(*at), (ctypes.c_float * 3)(0, 1, 0), fov)
    return s, cam


@pytest.mark.parametrize("case", ["slab_3000", "slab_3000_ground", "cube_800", "far_camera", "tiny_radii"])
def test_lbvh_widening_is_conservative_in_the_far_field(vcrt, case):
    # At a distance L the reference's binary32 discriminant is off by ~19 eps L^2, so far spheres are "hit" by rays that miss them
    # (L = 3000: 19 eps L^2 = 10).  The LBVH must still reach every sphere the exact test accepts: its boxes are widened per ray by
    # a bound on that error (vcrt_kernel_lbvh.cuh, distance-aware widening).  The exhaustive in-order kernel tests every sphere
    # exactly for every ray, so a sphere the tree misses shows up as a different bit.
    s, cam = far_field_scene(case)
    host.upload_scene(s, cam)
    try:
        for depth in (1, 12):
            simple, ss = host.render(480, 270, 2, depth=depth, traversal=SIMPLE, sample_chunks=1)
            tree, st = host.render(480, 270, 2, depth=depth, traversal=LBVH, sample_chunks=1)
            assert ss.rays == st.rays and np.array_equal(simple, tree)
        assert ss.rays - 480 * 270 * 2 > 1000                     # enough hits for the comparison to mean something
    finally:
        host.destroy()


@pytest.mark.parametrize("case", ["slab_3000", "cube_800", "far_camera"])
def test_brute_force_filter_is_conservative_in_the_far_field(vcrt, case):
    # the same scenes, thinned to what fits shared memory, through the candidate filter of the brute-force scan: its margin is
    # relative (kappa (|c|^2 + |o|^2 + r^2)), so it has to hold where the reference's discriminant is dominated by rounding too
    s, cam = far_field_scene(case)
    s = s[:: max(1, s.shape[0] // 1500)][:1500]
    host.upload_scene(s, cam)
    try:
        simple, ss = host.render(480, 270, 2, depth=12, traversal=SIMPLE, sample_chunks=1)
        brute, sb = host.render(480, 270, 2, depth=12, traversal=BRUTE, sample_chunks=1)
        assert sb.traversalUsed == BRUTE and sb.rays == ss.rays and np.array_equal(brute, simple)
        assert sb.listOverflows < sb.rays // 2 and ss.rays - 480 * 270 * 2 > 300   # the filter ran for most rays (far_camera: its margin, 1.5e-5 |o|^2 = 145, passes so much that 16 % of the lists overflow), and something was hit
    finally:
        host.destroy()


@pytest.mark.parametrize("case", ["forty_large", "equal_large", "few_spheres", "all_large_but_one"])
def test_lbvh_outlier_spheres_outside_the_tree(oracle, vcrt, case):
    # Spheres much larger than the median radius are kept out of the tree (at most 16, the largest first) and tested exactly
    # for every ray; the image must not depend on which side of that policy a sphere falls.
    rng = np.random.default_rng(zlib.crc32(case.encode()))
    n, large = {"forty_large": (240, 40), "equal_large": (240, 40), "few_spheres": (20, 3), "all_large_but_one": (60, 0)}[case]
    s = random_scene(rng, n)
    s[:, 3] = rng.uniform(0.08, 0.15, n)
    pick = rng.choice(n, large, replace=False)                     # scattered through the caller's order
    s[pick, 3] = 1.5 if case == "equal_large" else rng.uniform(0.8, 2.5, large)
    if case == "all_large_but_one":
        s[:, 3] = rng.uniform(1.0, 2.0, n); s[7, 3] = 0.01         # the median is large: nothing counts as an outlier
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(0, 1, 14), (ctypes.c_float * 3)(0, 0, 0), (ctypes.c_float * 3)(0, 1, 0), 50.0)
    host.upload_scene(s, cam)
    try:
        cpu, cst = oracle.render(s, 120, 80, 2, depth=16, camera=host.camera_array(cam))
        gpu, st = host.render(120, 80, 2, depth=16, traversal=LBVH, sample_chunks=1)
        assert_same(gpu, st, cpu, cst)
    finally:
        host.destroy()


def test_lbvh_duplicates_and_coincident_centres(oracle, vcrt):
    # identical spheres (exact ties: the lower index must win, as in the reference's in-order scan), concentric shells
    rng = np.random.default_rng(5)
    base = random_scene(rng, 40)
    scene = np.concatenate([base, base[::-1], base[:10]])             # every sphere at least twice, different materials order
    scene[80:, 7] = 2; scene[80:, 8] = 0.5                            # the third copies are metal: a wrong tie-break changes the image
    scene = np.concatenate([scene, [[0, 0, 0, r, .9, .9, .9, 3, 1.5, 0] for r in (0.5, 1.0, 1.5, 2.0)]]).astype(np.float32)
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(0, 1, 14), (ctypes.c_float * 3)(0, 0, 0), (ctypes.c_float * 3)(0, 1, 0), 50.0)
    host.upload_scene(scene, cam)
    try:
        cpu, cst = oracle.render(scene, 96, 64, 2, depth=12, camera=host.camera_array(cam))
        for traversal in (LBVH, BRUTE, SIMPLE):
            gpu, st = host.render(96, 64, 2, depth=12, traversal=traversal, sample_chunks=1)
            assert_same(gpu, st, cpu, cst)
    finally:
        host.destroy()


@pytest.mark.parametrize("count", [650, 1100, 2200])
def test_brute_force_kernel_ladder(oracle, vcrt, count):
    # VCRT_TRAVERSAL_BRUTE picks its kernel by what fits shared memory beside the scene blob: render_scanu<1024,1> up to ~700
    # spheres, render_scanu<512,2> up to ~1200, the first-generation render_scan<256,2> up to ~2400.  Same bits from all.
    rng = np.random.default_rng(count)
    scene = random_scene(rng, count, big=1)
    host.upload_scene(scene)
    try:
        cpu, cst = oracle.render(scene, 120, 68, 2, depth=9)
        gpu, st = host.render(120, 68, 2, depth=9, traversal=BRUTE, sample_chunks=1)
        assert st.traversalUsed == BRUTE
        assert_same(gpu, st, cpu, cst)
        fixed, sf = host.render(120, 68, 2, depth=9, traversal=BRUTE)
        assert sf.rays == cst["rays"] and np.abs(fixed - cpu).max() <= 2.0 ** -22
    finally:
        host.destroy()


@pytest.mark.parametrize("traversal", [BRUTE, SIMPLE, LBVH])
def test_count_stats_match_oracle(oracle, gpu_scene, traversal):
    # VCRT_FLAG_COUNT_STATS: hits by material and the path-length histogram (SURVEY.md 8c acceptance test) equal the oracle's
    cpu, cst = oracle.render(gpu_scene, 160, 90, 4)
    gpu, st = host.render(160, 90, 4, traversal=traversal, sample_chunks=1, flags=cabi.VCRT_FLAG_COUNT_STATS)
    assert_same(gpu, st, cpu, cst)
    assert list(st.hitsByMaterial) == cst["hits"] and list(st.pathLength) == cst["path_hist"]
    assert sum(st.pathLength) + st.depthExhausted == st.samples
    frac = np.array(st.hitsByMaterial[1:], float) / sum(st.hitsByMaterial)
    assert np.allclose(frac, [0.669, 0.199, 0.133], atol=0.02)              # SURVEY.md 3.4 probe
    plain, sp = host.render(160, 90, 4, traversal=traversal, sample_chunks=1)
    assert sum(sp.hitsByMaterial) == 0 and np.array_equal(plain, gpu)       # off by default, and never changes the image


def test_acceptance_path_lengths_and_materials_vs_the_references_own_rng(oracle, gpu_scene):
    # SURVEY.md 8c acceptance test, last clause: with Philox instead of the reference's hash "RNG" (bug B2) the path-length
    # histogram stays within 2 % relative in bins 1-6 and rays / sample within 0.5 %.  The HASH oracle is the restatement that
    # is bit-exact to the reference's shader code (tests/test_oracle_golden.py).
    w, h, spp = 400, 225, 64
    _, ref = oracle.render(gpu_scene, w, h, spp, rng=oh.RNG_HASH)
    _, st = host.render(w, h, spp, traversal=LBVH, flags=cabi.VCRT_FLAG_COUNT_STATS)
    assert abs(st.rays / ref["rays"] - 1) < 5e-3
    got, want = np.array(st.pathLength[:6], float), np.array(ref["path_hist"][:6], float)
    assert np.all(np.abs(got / want - 1) < 2e-2), (got, want)
    hits, rhits = np.array(st.hitsByMaterial, float), np.array(ref["hits"], float)
    assert np.all(np.abs(hits[1:] / rhits[1:] - 1) < 2e-2) and hits[0] == rhits[0] == 0


def test_more_than_2_to_the_32_pixel_samples_in_one_call(gpu_scene):
    # 3840x2160x520 spp = 4.31e9 pixel-samples: the 32-bit work-item numbering has to group samples (it used to wrap silently)
    w, h, spp = 3840, 2160, 520
    img, st = host.render(w, h, spp, traversal=LBVH)
    assert st.samples == w * h * spp > 2 ** 32
    assert abs(st.rays / st.samples / 2.2975 - 1) < 5e-3
    with open(os.path.join(GOLDEN, "ref_kat.json")) as f:
        want = json.load(f)["render_400x225_s16"]["mean_rgb"]          # same framing at any 16:9 size (bug B1)
    mean = img[..., :3].reshape(-1, 3).astype(np.float64).mean(0)
    assert np.all(np.abs(mean / np.array(want) - 1) < 5e-3)
    host.render(16, 16, 1, traversal=LBVH, download=False)              # give the 4K buffers back to smaller ones


def test_item_numbering_corner_cases(oracle, gpu_scene):
    # one tile and very many samples (runs per pixel are capped), an image of exactly one band and one pixel more, strict chunks
    img, st = host.render(8, 4, 5_000_000, depth=2, traversal=BRUTE)
    assert st.samples == 32 * 5_000_000 and 0 < st.rays <= 2 * st.samples and np.isfinite(img).all()
    for w, h in ((512, 512), (513, 512), (1024, 257)):                  # 2^18 padded pixels, and just above
        cpu, cst = oracle.render(gpu_scene, w, h, 1, depth=3)
        for trav in (BRUTE, LBVH):
            gpu, sg = host.render(w, h, 1, depth=3, traversal=trav, sample_chunks=1)
            assert_same(gpu, sg, cpu, cst)
            fixed, sf = host.render(w, h, 1, depth=3, traversal=trav)
            assert sf.rays == cst["rays"] and np.abs(fixed - cpu).max() <= 2.0 ** -22
    cpu, cst = oracle.render(gpu_scene, 96, 54, 7)
    for chunks in (2, 3, 7):                                              # float sums per chunk, chunks combined in order
        gpu, sg = host.render(96, 54, 7, traversal=LBVH, sample_chunks=chunks)
        assert sg.rays == cst["rays"] and np.abs(gpu - cpu).max() <= 2.0 ** -20


@pytest.mark.parametrize("shape", [(65535, 1), (1, 65535), (16381, 5), (3, 20011), (8191, 9)])
def test_extreme_aspect_ratios(oracle, gpu_scene, shape):
    # the kernels turn tile and pixel indices into coordinates with reciprocal estimates instead of integer divisions
    # (vcrt_device.cuh: udiv_small, tile_to_xy, pixel_to_xy): the widest and the tallest images the ABI accepts, odd sizes between
    w, h = shape
    cpu, cst = oracle.render(gpu_scene, w, h, 2, depth=4)
    for trav in (BRUTE, LBVH):
        gpu, sg = host.render(w, h, 2, depth=4, traversal=trav, sample_chunks=1)
        assert_same(gpu, sg, cpu, cst)
        fixed, sf = host.render(w, h, 2, depth=4, traversal=trav)          # brute force: primary pass + compact hand-over list
        assert sf.rays == cst["rays"] and np.abs(fixed - cpu).max() <= 2.0 ** -22


@pytest.mark.parametrize("traversal", [BRUTE, LBVH])
def test_chunk_counts_that_do_not_divide_the_samples(oracle, gpu_scene, traversal):
    # 7 samples in 5 or 6 chunks are 4 runs of 2 (ceil(7 / ceil(7 / 5)) = 4), 9 samples in 4 chunks are 3 runs of 3: only the
    # planes a kernel wrote may be summed.  A render with MORE planes goes first, so that a stale plane would show.
    host.render(96, 54, 16, traversal=traversal, sample_chunks=16, download=False)
    for spp, chunk_list in ((7, (5, 6, 4, 7)), (9, (4, 5, 8)), (5, (3, 4))):
        cpu, cst = oracle.render(gpu_scene, 96, 54, spp)
        for chunks in chunk_list:
            gpu, sg = host.render(96, 54, spp, traversal=traversal, sample_chunks=chunks)
            per_run = -(-spp // chunks)
            assert sg.sampleChunksUsed == -(-spp // per_run) <= chunks
            assert sg.rays == cst["rays"] and np.abs(gpu - cpu).max() <= 2.0 ** -20, (spp, chunks)
            host.render(96, 54, 16, traversal=traversal, sample_chunks=16, download=False)


def test_attenuation_outside_0_1_takes_float_sums(oracle, vcrt):
    # The default accumulator (32.32 fixed point) needs radiance in [0, 1] per sample.  A scene with an albedo of 40 or a
    # lambertian ratio of 3 (radiance far above 1, overflowing to inf after a few bounces) is summed in binary32 in sample order
    # instead, whatever sampleChunks says: inf and NaN then propagate exactly as in the reference arithmetic.
    scene = host.generate_scene(2)
    scene[3, 4:7] = (40.0, 0.5, 2.0)
    scene[5, 8] = 3.0 if int(scene[5, 7]) == 1 else scene[5, 8]
    scene[0, 7] = 1.0; scene[0, 8] = 1e30; scene[0, 4:7] = 1.0      # a lambertian sphere that multiplies by 1e30 per hit
    host.upload_scene(scene)
    try:
        cpu, cst = oracle.render(scene, 96, 54, 3, depth=12)
        assert cpu[..., :3].max() > 1.5
        for traversal in (BRUTE, LBVH):
            for chunks in (0, 1):
                gpu, st = host.render(96, 54, 3, depth=12, traversal=traversal, sample_chunks=chunks)
                assert st.rays == cst["rays"] and np.array_equal(gpu, cpu, equal_nan=True), (traversal, chunks)
    finally:
        host.destroy()


def test_uploading_the_same_scene_again_keeps_the_tree_and_the_bits(vcrt):
    # cudaUploadScene with byte-identical spheres (a caller that uploads every frame) copies them again but keeps the LBVH;
    # any change - one sphere, or just the order - rebuilds it.
    scene = host.generate_scene(30)
    host.upload_scene(scene)
    try:
        first, s1 = host.render(160, 90, 2, traversal=LBVH)
        host.upload_scene(scene.copy())
        again, s2 = host.render(160, 90, 2, traversal=LBVH)
        assert np.array_equal(first, again) and s1.nodeVisits == s2.nodeVisits
        moved = scene.copy(); moved[:-1, 1] += 0.07                # every sphere but the ground, 7 cm up
        host.upload_scene(moved)
        third, s3 = host.render(160, 90, 2, traversal=LBVH)
        simple, s4 = host.render(160, 90, 2, traversal=SIMPLE)
        assert s3.rays == s4.rays and np.abs(third - simple).max() <= 2.0 ** -22 and not np.array_equal(third, first)
        cam = cabi.VcrtCamera((ctypes.c_float * 3)(-6, 3.5, 9), (ctypes.c_float * 3)(0.5, 0.4, -1), (ctypes.c_float * 3)(0.1, 1, 0), 35.0)
        host.upload_scene(moved, cam)                                   # same spheres, another camera: tree kept, new view
        other, s5 = host.render(160, 90, 2, traversal=LBVH)
        simple, s6 = host.render(160, 90, 2, traversal=SIMPLE)
        assert s5.rays == s6.rays and np.abs(other - simple).max() <= 2.0 ** -22
    finally:
        host.destroy()


def test_primary_visibility_pass_changes_no_bit(gpu_scene):
    # VCRT_TRAVERSAL_BRUTE runs the camera rays and their first scatter event in a kernel of their own (vcrt_kernel_primary.cuh) and
    # hands the scattered rays to the path kernel.  VCRT_NO_PRIMARY=1 (read once per process) switches that off: a second process
    # renders the old way, and image bits, ray counts and the statistics counters must agree in both accumulation modes.
    code = (
        "import sys, zlib, json; sys.path.insert(0, '.');"
        "from vulkancomputeraytracing_b200 import cabi, host;"
        "host.upload_scene(host.generate_scene()); out = [];\n"
        "for chunks, flags in ((0, 0), (1, 0), (3, cabi.VCRT_FLAG_COUNT_STATS)):\n"
        "    img, st = host.render(203, 113, 5, depth=9, traversal=1, sample_chunks=chunks, flags=flags)\n"
        "    out.append([zlib.crc32(img.tobytes()), st.rays, st.depthExhausted, list(st.hitsByMaterial), list(st.pathLength), st.primaryRays])\n"
        "print(json.dumps(out))")
    env = dict(os.environ, VCRT_NO_PRIMARY="1")
    other = json.loads(subprocess.run([sys.executable, "-c", code], cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))), env=env,
                                      capture_output=True, text=True, check=True).stdout.strip().splitlines()[-1])
    for (chunks, flags), want in zip(((0, 0), (1, 0), (3, cabi.VCRT_FLAG_COUNT_STATS)), other):
        img, st = host.render(203, 113, 5, depth=9, traversal=BRUTE, sample_chunks=chunks, flags=flags)
        got = [zlib.crc32(img.tobytes()), st.rays, st.depthExhausted, list(st.hitsByMaterial), list(st.pathLength)]
        assert got == want[:5], (chunks, flags)
        assert st.primaryRays == 203 * 113 * 5 and want[5] == 0      # every camera ray went through the primary kernel here, none there
        assert st.primaryKernelMs > 0 and st.renderPasses == 1


def test_schlick_dense(vcrt):
    # functions.glsl:58-62 on 3 x 2^20 inputs: the device routine (pow(x, 5.) restated from glibc's powf, csrc/vcrt_powf.cuh)
    # against what the reference build returned (golden from oracle/_ref; live too where that build travelled to this machine).
    # Bar: 0 mismatches.
    with open(os.path.join(GOLDEN, "ref_schlick_dense.json")) as f:
        want = json.load(f)
    cos, ior = oh.schlick_inputs()
    out = np.empty_like(cos)
    cabi.check(vcrt.cudaEvaluateSchlick(cos.ctypes.data, ior.ctypes.data, cos.size, out.ctypes.data))
    host.destroy()
    got = oh.block_crcs(out, want["block"])
    differing = sum(a != b for a, b in zip(got, want["crc32"]))
    ref = oh.Reference.open()
    mismatches = None
    if ref is not None:
        live = ref.schlick_many(cos, ior)
        mismatches = int((out.view(np.uint32) != live.view(np.uint32)).sum())
    print(f"schlick: {cos.size} inputs, {differing} of {len(got)} golden blocks differ, live mismatches: {mismatches}")
    assert differing == 0 and not mismatches


@pytest.mark.parametrize("case", ["zero_and_negative_radius", "tiny_and_huge", "nan_and_inf_spheres", "degenerate_camera", "coincident_centres",
                                  "nan_spheres_after_a_huge_first_sphere", "beyond_the_tame_range", "edge_of_the_tame_range"])
def test_pathological_inputs_follow_the_reference_arithmetic(oracle, vcrt, case):
    # inputs the reference never guards against: whatever binary32 arithmetic makes of them, all traversals make the same
    rng = np.random.default_rng(zlib.crc32(case.encode()))
    scene = random_scene(rng, 40)
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(0, 1, 14), (ctypes.c_float * 3)(0, 0, 0), (ctypes.c_float * 3)(0, 1, 0), 50.0)
    if case == "zero_and_negative_radius":
        scene[0:5, 3] = 0.0; scene[5:10, 3] = -scene[5:10, 3]          # r = 0 never hits; r < 0 hits with an inward "outward" normal
    elif case == "tiny_and_huge":
        scene[0:4, 3] = 1e-20; scene[4:6, 0:3] *= 1e17; scene[6, 3] = 3e18; scene[7, 0:3] = [1e-30, 1e-30, 1e-30]
    elif case == "nan_and_inf_spheres":
        scene[0, 0] = np.nan; scene[1, 3] = np.inf; scene[2, 1] = -np.inf; scene[3, 3] = np.nan; scene[4, 4:7] = np.nan
    elif case == "degenerate_camera":
        cam = cabi.VcrtCamera((ctypes.c_float * 3)(1, 2, 3), (ctypes.c_float * 3)(1, 2, 3), (ctypes.c_float * 3)(0, 1, 0), 50.0)   # lookfrom == lookat
    elif case == "nan_spheres_after_a_huge_first_sphere":
        # a NaN sphere is "hit" at t = NaN and from then on the LAST sphere of the caller's order with a root > t_min wins: the layer
        # keeps spheres of radius > 50 at the end of its own arrays, which must not change that order
        scene[0] = [0, -1000, 0, 1000, .5, .5, .5, 1, 1.0, 0]; scene[3] = [0, 0, -300, 200, .9, .2, .2, 2, 0.1, 0]
        scene[7, 0] = np.nan; scene[20, 3] = np.nan
    elif case == "beyond_the_tame_range":
        # finite spheres whose squares overflow: disc = inf - inf = NaN, the sphere is "hit" at t = NaN like a NaN sphere.  The filter and
        # the tree are only conservative against rounding, so such a scene goes to the in-order kernel whatever was asked for
        scene[0] = [0, -1000, 0, 1000, .5, .5, .5, 2, 0.0, 0]; scene[1] = [0, 0, -300, 200, .9, .2, .2, 2, 0.1, 0]
        scene[4:6, 0:3] *= 1e17; scene[6, 3] = 3e18; scene[8:12, 0:3] = [[3e38, 0, 0], [0, 3e38, 1], [1e30, 1e30, 1e30], [-3e38, 1, 1]]
        scene[8:12, 7] = 2
    elif case == "edge_of_the_tame_range":
        # a cluster around a point with |p|^2 = 1e18 (t_max = 1e5 keeps anything farther than 1.35e6 from a ray's origin invisible): hit
        # points land on both sides of |o|^2 = 1e18, where rays switch to the exhaustive scan inside the kernels
        base = np.array([7.0710678e8, 7.0710678e8, 0.0])
        scene[:, 0:3] = base + rng.uniform(-8e5, 8e5, (40, 3)); scene[:, 3] = rng.uniform(5e4, 3e5, 40)
        eye = base + [0, 1e5, 1.2e6]
        cam = cabi.VcrtCamera((ctypes.c_float * 3)(*eye), (ctypes.c_float * 3)(*base), (ctypes.c_float * 3)(0, 1, 0), 50.0)
    else:
        scene[0:20, 0:3] = scene[0, 0:3]; scene[0:20, 3] = np.linspace(0.5, 1.5, 20)   # 20 concentric spheres of three materials
    host.upload_scene(scene, cam)
    try:
        cpu, cst = oracle.render(scene, 64, 48, 2, depth=12, camera=host.camera_array(cam))
        for traversal in (SIMPLE, BRUTE, LBVH):
            gpu, st = host.render(64, 48, 2, depth=12, traversal=traversal, sample_chunks=1)
            assert st.traversalUsed == (SIMPLE if case in ("tiny_and_huge", "nan_and_inf_spheres", "nan_spheres_after_a_huge_first_sphere", "beyond_the_tame_range") else traversal)
            assert st.rays == cst["rays"] and st.depthExhausted == cst["exhausted"], (traversal, st.rays, cst["rays"])
            assert np.array_equal(gpu, cpu, equal_nan=True), traversal
            again, _ = host.render(64, 48, 2, depth=12, traversal=traversal)          # fixed-point accumulation must not trap on NaN either
            assert again.shape == gpu.shape
    finally:
        host.destroy()


@pytest.mark.parametrize("seed", range(40))
def test_randomised_scenes_cameras_and_parameters(oracle, vcrt, seed):
    # 40 seeded draws over everything a caller controls: sphere count, sizes (incl. the always-tested > 50 class), materials with
    # fuzz > 1 and ior <= 1 or 0, colours outside [0,1], camera, field of view, image shape, depth, spp, seed, sample offset
    rng = np.random.default_rng(1000 + seed)
    count = int(rng.choice([1, 2, 3, 7, 33, 64, 65, 150, 400]))
    scene = random_scene(rng, count, big=int(rng.integers(0, min(4, count))))
    scene[:, 0:3] *= rng.choice([0.01, 1.0, 30.0])
    scene[:, 3] *= rng.choice([0.1, 1.0, 3.0])
    scene[:, 4:7] = rng.uniform(-0.5, 2.0, (count, 3))
    scene[:, 8] = np.where(scene[:, 7] == 3, rng.choice([0.0, 0.5, 1.0, 1.5, 2.4], count), rng.uniform(0, 3, count))
    eye = rng.uniform(-15, 15, 3); at = rng.uniform(-2, 2, 3); up = rng.normal(0, 1, 3)
    cam = cabi.VcrtCamera((ctypes.c_float * 3)(*eye), (ctypes.c_float * 3)(*at), (ctypes.c_float * 3)(*up), float(rng.uniform(1, 170)))
    w, h = int(rng.integers(1, 130)), int(rng.integers(1, 90))
    spp, depth = int(rng.integers(1, 6)), int(rng.choice([1, 2, 5, 50]))
    rs, off = int(rng.integers(0, 2 ** 63)), int(rng.integers(0, 10 ** 6))
    host.upload_scene(scene, cam)
    try:
        cpu, cst = oracle.render(scene, w, h, spp, depth=depth, seed=rs, sample_offset=off, camera=host.camera_array(cam))
        for traversal in (SIMPLE, BRUTE, LBVH):
            gpu, st = host.render(w, h, spp, depth=depth, seed=rs, sample_offset=off, traversal=traversal, sample_chunks=1)
            assert st.rays == cst["rays"] and st.depthExhausted == cst["exhausted"], (traversal, count, w, h, spp, depth)
            assert np.array_equal(gpu, cpu, equal_nan=True), (traversal, count, w, h, spp, depth)
        if seed % 2 == 0:                                                # the same draw in BOOK semantics
            book, cb = oracle.render_book(scene, w, h, spp, depth=depth, seed=rs, sample_offset=off, camera=host.camera_array(cam))
            for traversal in (SIMPLE, BRUTE, LBVH):
                gb, sb = host.render(w, h, spp, depth=depth, seed=rs, sample_offset=off, traversal=traversal, sample_chunks=1,
                                     semantics=cabi.VCRT_SEMANTICS_BOOK)
                assert sb.rays == cb["rays"] and sb.depthExhausted == cb["exhausted"], (traversal, count, w, h, spp, depth)
                assert np.array_equal(gb, book, equal_nan=True), (traversal, count, w, h, spp, depth)
    finally:
        host.destroy()
