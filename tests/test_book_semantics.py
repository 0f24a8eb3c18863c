# SPDX-License-Identifier: WTFPL
"""VCRT_SEMANTICS_BOOK: the path with the reference's matched bugs fixed as "Ray Tracing in One
Weekend" writes the routines.  The reference holds no book-mode code or vectors, so the CPU checker here is
PARITY-UNPINNED: it is validated by invariants the book's algorithm must satisfy (below), and the CUDA kernels are
then held bit-exact to it."""
import ctypes
import os

import numpy as np
import pytest

import oracle_harness as oh

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_book_camera_is_centred_with_square_pixels(oracle):
    # bug B1 fixed: the optical axis goes through the image centre and pixels are square at any aspect ratio
    for (w, h) in [(400, 225), (1280, 720), (300, 300), (225, 400)]:
        f = oracle.camera(w, h, book=True)
        centre = f["pixel00"] + np.float32((w - 1) / 2) * f["du"] + np.float32((h - 1) / 2) * f["dv"]
        axis = centre - f["center"]
        want = -f["center"] / np.linalg.norm(f["center"])               # lookat is the origin
        assert np.allclose(axis / np.linalg.norm(axis), want, atol=2e-6)
        assert abs(np.linalg.norm(f["du"]) / np.linalg.norm(f["dv"]) - 1) < 1e-6
        assert abs(np.dot(f["du"], f["dv"])) < 1e-9
    ref = oracle.camera(400, 225, book=False)                            # the reference's frame is NOT centred (B1)
    c = ref["pixel00"] + np.float32(199.5) * ref["du"] + np.float32(112) * ref["dv"] - ref["center"]
    assert not np.allclose(c / np.linalg.norm(c), -ref["center"] / np.linalg.norm(ref["center"]), atol=1e-3)


def test_book_image_invariants(oracle):
    world = np.load(os.path.join(GOLDEN, "ref_world.npy"))
    img, st = oracle.render_book(world, 160, 90, 16)
    assert (img[..., 3] == 1).all() and np.isfinite(img).all() and img[..., :3].min() >= 0 and img[..., :3].max() <= 1
    lin, st2 = oracle.render_book(world, 160, 90, 16, normalise=False)
    assert st2["rays"] == st["rays"]
    np.testing.assert_allclose(img[..., :3], np.sqrt(lin[..., :3] / np.float32(16)), rtol=2e-6, atol=1e-7)   # gamma 2 of the mean
    other, _ = oracle.render_book(world, 160, 90, 16, seed=7)
    rmse = np.sqrt(np.mean((img[..., :3].astype(np.float64) ** 2 - other[..., :3].astype(np.float64) ** 2) ** 2))
    assert 0 < rmse < 0.12                                               # two independent 16-spp estimates of one image
    assert abs(np.mean(img[..., :3] ** 2) / np.mean(other[..., :3] ** 2) - 1) < 0.02
    # the book's materials scatter over the whole sphere: paths are longer than the reference's +octant scatter (bug B4)
    _, sref = oracle.render(world, 160, 90, 16)
    assert st["rays"] > sref["rays"]
    # split renders add up (spp sharding works in this mode too)
    a, _ = oracle.render_book(world, 64, 36, 3, normalise=False)
    b, _ = oracle.render_book(world, 64, 36, 2, sample_offset=3, normalise=False)
    full, _ = oracle.render_book(world, 64, 36, 5, normalise=False)
    np.testing.assert_allclose(a[..., :3] + b[..., :3], full[..., :3], rtol=1e-6, atol=1e-7)


def test_book_material_physics(oracle):
    cam = np.array([0, 0, 5, 0, 0, 0, 0, 1, 0, 40], np.float32)
    # white furnace: albedo-1 lambertian sphere under the sky can never be brighter than the sky's maximum (1,1,1)
    s = np.array([[0, 0, 0, 1, 1, 1, 1, 1, 0.3, 0]], np.float32)        # texture.y (0.3) must be ignored (bug B4b fixed)
    lin, _ = oracle.render_book(s, 64, 64, 64, normalise=False, camera=cam)
    assert lin[..., :3].max() / 64 <= 1.0 + 1e-6
    s2 = s.copy(); s2[0, 8] = 1.0
    lin2, _ = oracle.render_book(s2, 64, 64, 64, normalise=False, camera=cam)
    np.testing.assert_array_equal(lin, lin2)
    # a perfect mirror (fuzz 0) seen head-on reflects the sky behind the camera; a fuzz-1 metal absorbs part of its scatter
    mirror = np.array([[0, 0, 0, 1, 1, 1, 1, 2, 0.0, 0]], np.float32)
    rough = np.array([[0, 0, 0, 1, 1, 1, 1, 2, 1.0, 0]], np.float32)
    lm, _ = oracle.render_book(mirror, 64, 64, 32, normalise=False, camera=cam)
    lr, _ = oracle.render_book(rough, 64, 64, 32, normalise=False, camera=cam)
    centre = (slice(24, 40), slice(24, 40))
    assert lr[centre][..., :3].mean() < lm[centre][..., :3].mean()       # below-surface scatter is absorbed
    # glass refracts: a glass ball is NOT invisible (the reference's bug B5 makes it nearly so)
    glass = np.array([[0, 0, 0, 1, 1, 1, 1, 3, 1.5, 0], [0, -1001, 0, 1000, .2, .8, .2, 1, 1, 0]], np.float32)
    gb, _ = oracle.render_book(glass, 64, 64, 32, camera=cam)
    empty, _ = oracle.render_book(glass[1:], 64, 64, 32, camera=cam)
    assert np.abs(gb[centre][..., :3] - empty[centre][..., :3]).mean() > 0.02


@pytest.mark.gpu
@pytest.mark.parametrize("traversal", [1, 2, 3])
def test_book_mode_cuda_bit_exact_vs_checker(oracle, gpu_scene, traversal):
    from vulkancomputeraytracing_b200 import cabi, host
    for (w, h, spp, depth) in [(96, 54, 3, 50), (203, 113, 2, 4)]:
        cpu, cst = oracle.render_book(gpu_scene, w, h, spp, depth=depth)
        gpu, st = host.render(w, h, spp, depth=depth, traversal=traversal, sample_chunks=1, semantics=cabi.VCRT_SEMANTICS_BOOK)
        assert st.rays == cst["rays"] and st.depthExhausted == cst["exhausted"]
        assert np.array_equal(gpu, cpu), f"max |diff| {np.abs(gpu - cpu).max():.3e}"
    ref, _ = host.render(96, 54, 3, traversal=traversal, sample_chunks=1)
    assert not np.array_equal(ref, gpu[:54, :96]) if gpu.shape == ref.shape else True


@pytest.mark.gpu
def test_book_mode_stress_scene_and_fixed_point(oracle, vcrt):
    from vulkancomputeraytracing_b200 import cabi, host
    scene = host.stress_scene()
    host.upload_scene(scene)
    try:
        cpu, cst = oracle.render_book(scene, 120, 68, 4)
        for traversal in (1, 2):
            gpu, st = host.render(120, 68, 4, traversal=traversal, sample_chunks=1, semantics=cabi.VCRT_SEMANTICS_BOOK)
            assert st.rays == cst["rays"] and np.array_equal(gpu, cpu)
            fixed, _ = host.render(120, 68, 4, traversal=traversal, semantics=cabi.VCRT_SEMANTICS_BOOK)
            assert np.abs(fixed - cpu).max() < 1e-3      # sqrt of a mean that differs by <= 2^-22 (steep near 0)
        p = host.render_params(16, 16, 1); p.semantics = 9
        assert vcrt.cudaRender(ctypes.byref(p)) == cabi.VCRT_ERROR_INVALID_ARGUMENT
    finally:
        host.destroy()
