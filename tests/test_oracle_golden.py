# SPDX-License-Identifier: WTFPL
"""Pins the CPU oracle (oracle/vcrt_oracle.cpp) to the reference: against golden vectors that
tests/golden/make_golden.py produced from the reference's own shader code compiled as C++, and,
where oracle/_ref exists on this machine, against that code run live.  CPU only."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_harness as oh

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="module")
def kat():
    with open(os.path.join(GOLDEN, "ref_kat.json")) as f:
        return json.load(f)


def test_scene_table_equals_baked_world(oracle):
    # SceneGenerator.cpp:23-56 restated (MT19937 + generate_canonical by hand) vs globals.glsl:31-517
    world = golden("ref_world.npy")
    assert world.shape == (485, 10)
    np.testing.assert_array_equal(oracle.scene(), world)
    kinds = world[:, 7].astype(int)
    assert (np.bincount(kinds, minlength=4)[1:] == [373, 75, 37]).all()     # SURVEY.md 8c KAT-scene
    assert world[484, 3] == 1000.0 and world[481, 7] == 3 and world[482, 7] == 1 and world[483, 7] == 2


def test_scene_text_sha_is_the_surveyed_one():
    with open(os.path.join(GOLDEN, "ref_scene_text.sha256")) as f:
        assert f.read().strip() == "98a54e3a9ff2ae74c90eef5f296b69d3190751d867737f271ef1aa348e3079c7"


def test_hash_image_bit_exact_vs_reference_golden(oracle, kat):
    img, st = oracle.render(golden("ref_world.npy"), 64, 36, 4, depth=50, rng=oh.RNG_HASH)
    np.testing.assert_array_equal(img, golden("ref_hash_64x36_s4.npy"))
    assert st["rays"] == kat["render_64x36_s4"]["rays"] and st["exhausted"] == kat["render_64x36_s4"]["exhausted"]


def test_depth_exhaustion_matches_reference_golden(oracle, kat):
    # depth 3: bug B7 path (black) is taken by ~12 % of the samples
    img, st = oracle.render(golden("ref_world.npy"), 96, 54, 2, depth=3, rng=oh.RNG_HASH)
    np.testing.assert_array_equal(img, golden("ref_hash_96x54_s2_d3.npy"))
    assert st["exhausted"] == kat["render_96x54_s2_d3"]["exhausted"] > 1000


def test_function_known_answers(oracle, kat):
    for x, y, want in kat["rand"]:
        assert oracle.lib.oracle_hash_rand(x, y) == np.float32(want)
    for c, i, want in kat["schlick"]:
        got = oracle.lib.oracle_schlick(c, i)
        assert got == np.float32(want) or (np.isnan(got) and np.isnan(want))
    for case in kat["refract"]:
        ok, out = oracle.refract(case["v"], case["n"], case["eta"])
        assert int(ok) == case["ok"]
        np.testing.assert_array_equal(out, np.array(case["out"], np.float32))
    # SURVEY.md 8c KAT-functions (derived from the reference source at survey time)
    assert oracle.lib.oracle_hash_rand(1, 1) == 0.6875 and abs(oracle.lib.oracle_hash_rand(2, 2) - 0.4677734) < 1e-7
    assert abs(oracle.lib.oracle_schlick(0.5, 1.5) - 0.07) < 1e-7 and abs(oracle.lib.oracle_schlick(1.0, 1.5) - 0.04) < 1e-7


def test_camera_known_answers(oracle):
    # SURVEY.md 8c KAT-camera (bug B1: both deltas / HEIGHT, vw = vh * (W / H as integers))
    f = oracle.camera(400, 225)
    np.testing.assert_allclose(f["du"], [0.004754592, 0, -0.02060323], rtol=2e-6, atol=1e-9)
    np.testing.assert_allclose(f["dv"], [0.003054426, -0.02091107, 0.0007048675], rtol=2e-6)
    np.testing.assert_allclose(f["pixel00"], [-0.8746099, 2.34204, 2.228616], rtol=2e-6)
    f = oracle.camera(1280, 720)
    np.testing.assert_allclose(f["du"], [0.00148581, 0, -0.006438509], rtol=2e-6, atol=1e-9)
    np.testing.assert_allclose(f["pixel00"], [-0.8772943, 2.349228, 2.235456], rtol=2e-6)


def test_first_hits_match_reference_golden(oracle, kat):
    world = golden("ref_world.npy")
    for h in kat["first_hit"]:
        idx, t = oracle.first_hit(world, h["o"], h["d"])
        assert idx == h["index"], h
        if idx >= 0:
            assert np.float32(t) == np.float32(h["t"]), h
    # SURVEY.md 8c KAT-first-hit
    by_px = {(h["w"], h["px"], h["py"]): h for h in kat["first_hit"]}
    assert by_px[(400, 0, 0)]["index"] == -1
    assert by_px[(400, 112, 112)]["index"] == 483 and abs(by_px[(400, 112, 112)]["t"] - 0.6923077) < 1e-6
    assert by_px[(400, 399, 224)]["index"] == 484 and abs(by_px[(400, 399, 224)]["t"] - 0.4676228) < 1e-6
    assert by_px[(1280, 640, 360)]["index"] == 483 and abs(by_px[(1280, 640, 360)]["t"] - 0.6391953) < 1e-6


def test_philox_known_answers(oracle):
    # Random123 kat_vectors for philox4x32-10
    np.testing.assert_array_equal(oracle.philox([0, 0, 0, 0], [0, 0]), np.array([0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8], np.uint32))
    np.testing.assert_array_equal(oracle.philox([0xffffffff] * 4, [0xffffffff] * 2),
                                  np.array([0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd], np.uint32))
    np.testing.assert_array_equal(oracle.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]),
                                  np.array([0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1], np.uint32))


def test_config0_statistics_vs_reference_golden(oracle, kat):
    # BASELINE.json configs[0]: 400x225, 16 spp, depth 50.  HASH mode equals the reference's statistics exactly;
    # PHILOX mode (what the GPU runs) agrees statistically: channel means < 0.5 %, rays/sample < 0.5 %.
    world = golden("ref_world.npy")
    want = kat["render_400x225_s16"]
    img, st = oracle.render(world, 400, 225, 16, rng=oh.RNG_HASH)
    assert st["rays"] == want["rays"] and st["exhausted"] == want["exhausted"]
    np.testing.assert_allclose(img[..., :3].reshape(-1, 3).astype(np.float64).mean(0), want["mean_rgb"], rtol=1e-12)
    np.testing.assert_allclose(want["mean_rgb"], [0.317935, 0.373134, 0.452783], rtol=2e-4)   # SURVEY.md 8c KAT-image
    img_p, st_p = oracle.render(world, 400, 225, 16, rng=oh.RNG_PHILOX)
    mean_p = img_p[..., :3].reshape(-1, 3).astype(np.float64).mean(0)
    assert np.all(np.abs(mean_p / np.array(want["mean_rgb"]) - 1) < 5e-3)
    assert abs(st_p["rays"] / want["rays"] - 1) < 5e-3
    assert (img_p[..., 3] == 1).all() and img_p[..., :3].min() >= 0 and img_p[..., :3].max() <= 1
    rmse = float(np.sqrt(np.mean((img_p[..., :3].astype(np.float64) - img[..., :3]) ** 2)))
    assert rmse < 4e-2   # two independent 16-spp estimates: floor 0.0805 * sqrt(2/16) = 0.0285
    blocks = img_p[:224, :, :3].astype(np.float64).reshape(28, 8, 50, 8, 3).mean((1, 3))
    assert np.sqrt(np.mean((blocks - golden("ref_hash_400x225_s16_blocks8.npy")) ** 2)) < 1e-2


def test_sharded_sums_add_up(oracle):
    # sample_offset splits a render: sums over [0,4) == sums over [0,2) + [2,4)  (spp sharding)
    world = golden("ref_world.npy")
    full, sf = oracle.render(world, 48, 27, 4, normalise=False)
    a, sa = oracle.render(world, 48, 27, 2, normalise=False)
    b, sb = oracle.render(world, 48, 27, 2, sample_offset=2, normalise=False)
    assert sa["rays"] + sb["rays"] == sf["rays"]
    np.testing.assert_allclose(a[..., :3] + b[..., :3], full[..., :3], rtol=1e-6, atol=1e-7)


def test_oracle_vs_live_reference_build(oracle, reference):
    # skipped where oracle/_ref is absent.  A size / spp the golden files do not cover.
    np.testing.assert_array_equal(oracle.scene(), reference.world())
    img, st = oracle.render(reference.world(), 80, 45, 3, depth=50, rng=oh.RNG_HASH)
    rimg, rst = reference.render(80, 45, 3, 50)
    np.testing.assert_array_equal(img, rimg)
    assert st["rays"] == rst["rays"] and st["exhausted"] == rst["exhausted"]
    rng = np.random.default_rng(1)
    f = oracle.camera(80, 45)
    for _ in range(200):
        d = (f["pixel00"] + np.float32(rng.uniform(0, 80)) * f["du"] + np.float32(rng.uniform(0, 45)) * f["dv"] - f["center"]).astype(np.float32)
        assert oracle.first_hit(reference.world(), f["center"], d) == reference.first_hit(f["center"], d)
