# SPDX-License-Identifier: WTFPL
"""The acceptance test of SURVEY.md 8c / BASELINE.json north_star against the REFERENCE's own image:

    per-pixel RMSE < 1e-2 wherever both sides have >= 128 spp, per-channel image mean within 0.5 %, alpha == 1

GPU side: the product as a caller gets it - Philox draws, REFERENCE semantics, the default accumulator (32.32 fixed point),
VCRT_TRAVERSAL_AUTO (the LBVH at 485 spheres) and, beside it, VCRT_TRAVERSAL_BRUTE.  Reference side: the reference's shader
code compiled as C++ with its own hash "RNG" (bug B2) - as a committed golden made by tests/golden/make_golden.py from
oracle/_ref at 400x225x256, and live (oracle/_ref where it travelled, else the HASH oracle that is bit-exact to it) at the
benched size 1920x1080x128.  The two sides use different random numbers, so the floor of the RMSE is the Monte-Carlo noise of
both: sigma sqrt(2 / spp) with sigma = 0.0805 (the per-sample standard deviation of this scene)."""
import os

import numpy as np
import pytest

import oracle_harness as oh
from vulkancomputeraytracing_b200 import cabi, host

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SIGMA = 0.0805


def check(gpu, ref_rgb, spp, label):
    assert (gpu[..., 3] == 1).all() and gpu[..., :3].min() >= 0 and gpu[..., :3].max() <= 1
    diff = gpu[..., :3].astype(np.float64) - ref_rgb.astype(np.float64)
    rmse = float(np.sqrt(np.mean(diff ** 2)))
    floor = SIGMA * np.sqrt(2.0 / spp)
    means = gpu[..., :3].reshape(-1, 3).astype(np.float64).mean(0) / ref_rgb.reshape(-1, 3).astype(np.float64).mean(0) - 1
    print(f"{label}: RMSE {rmse:.5f} (Monte-Carlo floor {floor:.5f}, tolerance 1e-2), channel means off by {np.abs(means).max():.3%}")
    assert rmse < 1e-2                                  # north_star's tolerance
    assert rmse < 1.6 * floor                           # and no more than the noise of two independent estimates allows
    assert np.all(np.abs(means) < 5e-3)                 # per-channel image mean within 0.5 %
    return rmse


@pytest.mark.parametrize("traversal", [cabi.VCRT_TRAVERSAL_AUTO, cabi.VCRT_TRAVERSAL_BRUTE])
def test_rmse_vs_reference_image_400x225_256spp(gpu_scene, traversal):
    z = np.load(os.path.join(GOLDEN, "ref_hash_400x225_s256_u16.npz"))
    ref_rgb = z["rgb"].astype(np.float64) / 65535.0      # quantised at 2^-16: 7.7e-6 against a tolerance of 1e-2
    gpu, st = host.render(400, 225, 256, traversal=traversal)
    assert st.traversalUsed == (cabi.VCRT_TRAVERSAL_LBVH if traversal == cabi.VCRT_TRAVERSAL_AUTO else traversal) and st.sampleChunksUsed == 1
    check(gpu, ref_rgb, 256, f"400x225x256 traversal {traversal}")
    assert abs(st.rays / float(z["rays"]) - 1) < 5e-3   # rays per sample within 0.5 %
    # 16 spp is not enough for 1e-2 (floor 0.028): the tolerance SURVEY.md 8c gives for configs[0] is 2.5e-2 there... at equal spp;
    # against the 256-spp golden the 16-spp render is bounded by its own noise
    low, _ = host.render(400, 225, 16, traversal=traversal)
    rm = float(np.sqrt(np.mean((low[..., :3].astype(np.float64) - ref_rgb) ** 2)))
    assert rm < 2.5e-2 and rm > 1e-2


def test_rmse_vs_reference_image_at_the_benched_size(gpu_scene):
    # 1920x1080 (BASELINE.json configs[1]) at 128 spp: ~1 minute of the reference's code on the host cores
    ref = oh.Reference.open(fast=False)
    if ref is not None:
        img, st = ref.render(1920, 1080, 128, 50)
    else:
        img, st = oh.Oracle().render(gpu_scene, 1920, 1080, 128, rng=oh.RNG_HASH)
    for traversal in (cabi.VCRT_TRAVERSAL_AUTO, cabi.VCRT_TRAVERSAL_BRUTE):
        gpu, sg = host.render(1920, 1080, 128, traversal=traversal)
        check(gpu, img[..., :3], 128, f"1920x1080x128 traversal {traversal}")
        assert abs(sg.rays / st["rays"] - 1) < 5e-3
    # and per pixel, the two GPU traversals are the same estimator with the same draws: fixed-point sums, identical bits
    a, _ = host.render(1920, 1080, 8, traversal=cabi.VCRT_TRAVERSAL_LBVH)
    b, _ = host.render(1920, 1080, 8, traversal=cabi.VCRT_TRAVERSAL_BRUTE)
    assert np.array_equal(a, b)


def test_benched_configuration_per_pixel_against_the_oracle(oracle, gpu_scene):
    # What bench.py times (1920x1080, default accumulator, BRUTE and AUTO) against the ORACLE with the same Philox draws, per
    # pixel: the oracle sums in binary32 in sample order, the product in 32.32 fixed point, so |diff| <= 2^-22 per channel.
    # 4 spp keeps the CPU side at ~2 s; the kernels, the CTA shape and the work-item plan are those of the 500-spp bench frame.
    cpu, cst = oracle.render(gpu_scene, 1920, 1080, 4)
    for traversal in (cabi.VCRT_TRAVERSAL_BRUTE, cabi.VCRT_TRAVERSAL_AUTO):
        gpu, st = host.render(1920, 1080, 4, traversal=traversal)
        assert st.rays == cst["rays"] and st.depthExhausted == cst["exhausted"]
        assert np.abs(gpu - cpu).max() <= 2.0 ** -22
    strict, ss = host.render(1920, 1080, 4, traversal=cabi.VCRT_TRAVERSAL_AUTO, sample_chunks=1)
    assert np.array_equal(strict, cpu)
