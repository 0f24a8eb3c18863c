# SPDX-License-Identifier: WTFPL
"""cudaRenderNoAlloc (include/vcrt.h): the render call that allocates nothing, its scratch in a static __device__ array.

Collected LAST on purpose (the file name).  The test stands on its own in a fresh process
(the new entry point needs no cudaUploadScene and no state of the other entry points) and after every other test.

Bar: the same BITS as the existing entry point (cudaRender + cudaDownloadImage) with the same parameters - the bar of
tests/test_gpu_parity.py - on two calls in a row that share the scratch array (different image size, scene, seed and run
count, so that anything left over from the first call would show in the second), then the first call again.
Inputs and outputs are device memory of the caller: torch tensors here, allocated before the calls."""
import ctypes

import numpy as np
import pytest

from vulkancomputeraytracing_b200 import cabi, host

pytestmark = pytest.mark.gpu
SIMPLE, BRUTE, LBVH = cabi.VCRT_TRAVERSAL_SIMPLE, cabi.VCRT_TRAVERSAL_BRUTE, cabi.VCRT_TRAVERSAL_LBVH


def no_alloc(lib, torch, spheres_dev, count, image_dev, cam, stream=None, **kw):
    p = host.render_params(**kw)
    s = stream if stream is not None else torch.cuda.current_stream()
    rc = lib.cudaRenderNoAlloc(spheres_dev.data_ptr(), count, ctypes.byref(cam), ctypes.byref(p), image_dev.data_ptr(),
                               image_dev.numel() * 4, s.cuda_stream)
    return rc


def existing(scene, cam, width, height, spp, **kw):
    host.upload_scene(scene, cam)
    img, st = host.render(width, height, spp, **kw)
    return img, st


def test_no_alloc_render_equals_the_existing_entry_point_on_two_calls_in_a_row(vcrt):
    import torch
    torch.cuda.set_device(0)
    big = host.generate_scene()        # the reference scene, 485 spheres
    small = host.generate_scene(3)     # a 6 x 6 grid
    cam_a = host.default_camera()
    cam_b = cabi.VcrtCamera((ctypes.c_float * 3)(-6, 3.5, 9), (ctypes.c_float * 3)(0.5, 0.4, -1), (ctypes.c_float * 3)(0.1, 1, 0), 35.0)
    # everything the calls need is allocated here, before them; the images are poisoned so that an unwritten pixel shows
    dev_big, dev_small = torch.from_numpy(big).cuda(), torch.from_numpy(small).cuda()
    img_a = torch.full((90, 160, 4), float("nan"), device="cuda")
    img_b = torch.full((57, 101, 4), float("nan"), device="cuda")
    img_c = torch.full((90, 160, 4), float("nan"), device="cuda")
    torch.cuda.synchronize()
    a_kw = dict(width=160, height=90, spp=3, depth=50, seed=7)
    b_kw = dict(width=101, height=57, spp=7, depth=12, seed=0x1234567890ABCDEF, sample_offset=5, sample_chunks=5, semantics=1)
    try:
        # the first call on a device loads the module (the static array comes with it) and looks the array's address up
        warm = torch.zeros((8, 8, 4), device="cuda")
        assert no_alloc(vcrt, torch, dev_small, len(small), warm, cam_a, width=8, height=8, spp=1, depth=2) == 0, vcrt.cudaGetLastErrorString()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        free_before = torch.cuda.mem_get_info()[0]
        # two calls in a row on one stream, no synchronisation in between: the second overwrites the first one's scratch
        assert no_alloc(vcrt, torch, dev_big, len(big), img_a, cam_a, traversal=SIMPLE, **a_kw) == 0, vcrt.cudaGetLastErrorString()
        assert no_alloc(vcrt, torch, dev_small, len(small), img_b, cam_b, **b_kw) == 0, vcrt.cudaGetLastErrorString()
        assert no_alloc(vcrt, torch, dev_big, len(big), img_c, cam_a, sample_chunks=1, **a_kw) == 0, vcrt.cudaGetLastErrorString()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before          # torch's allocator was not asked for anything
        assert torch.cuda.mem_get_info()[0] >= free_before      # and the device's free memory did not shrink (no cudaMalloc)
        got_a, got_b, got_c = img_a.cpu().numpy(), img_b.cpu().numpy(), img_c.cpu().numpy()

        want_a, st = existing(big, cam_a, 160, 90, 3, depth=50, seed=7, traversal=SIMPLE)
        assert st.traversalUsed == SIMPLE
        assert np.array_equal(got_a, want_a), f"call 1: {(got_a != want_a).any(-1).mean():.4%} of the pixels differ"
        assert np.array_equal(got_c, want_a), f"call 3: {(got_c != want_a).any(-1).mean():.4%} of the pixels differ"
        for traversal in (BRUTE, LBVH):   # 7 samples in 5 chunks = 4 runs of 2: the float modes' split, BOOK semantics
            want_b, st = existing(small, cam_b, 101, 57, 7, depth=12, seed=0x1234567890ABCDEF, sample_offset=5, sample_chunks=5, semantics=1, traversal=traversal)
            assert st.sampleChunksUsed == 4
            assert np.array_equal(got_b, want_b), f"call 2 vs traversal {traversal}: {(got_b != want_b).any(-1).mean():.4%} of the pixels differ"
        assert (got_a[..., 3] == 1).all() and (got_b[..., 3] == 1).all() and len(np.unique(got_a)) > 1000
    finally:
        host.destroy()


def test_no_alloc_render_rejects_what_its_static_array_cannot_hold(vcrt):
    import torch
    torch.cuda.set_device(0)
    scene = host.generate_scene(2)
    dev = torch.from_numpy(scene).cuda()
    img = torch.zeros((1080, 1920, 4), device="cuda")
    cam = host.default_camera()
    invalid, unsupported = cabi.VCRT_ERROR_INVALID_ARGUMENT, cabi.VCRT_ERROR_UNSUPPORTED
    # 1920 x 1080 in one run fits (2 073 600 <= 2^21 plane pixels) ...
    assert no_alloc(vcrt, torch, dev, len(scene), img, cam, width=1920, height=1080, spp=1, depth=3) == 0, vcrt.cudaGetLastErrorString()
    torch.cuda.synchronize()
    assert bool((img[..., 3] == 1).all())
    # ... in two it does not, nor does a larger image, nor more spheres than the array has room for
    assert no_alloc(vcrt, torch, dev, len(scene), img, cam, width=1920, height=1080, spp=2, depth=3, sample_chunks=2) == invalid
    assert b"VCRT_NOALLOC_MAX_PLANE_PIXELS" in vcrt.cudaGetLastErrorString()
    assert no_alloc(vcrt, torch, dev, len(scene), img, cam, width=1921, height=1092, spp=1, depth=3) == invalid
    assert no_alloc(vcrt, torch, dev, cabi.VCRT_NOALLOC_MAX_SPHERES + 1, img, cam, width=64, height=64, spp=1) == invalid
    assert b"VCRT_NOALLOC_MAX_SPHERES" in vcrt.cudaGetLastErrorString()
    # what the entry point does not do: other traversals, flags, an image buffer that is too small, host pointers
    assert no_alloc(vcrt, torch, dev, len(scene), img, cam, width=64, height=64, spp=1, traversal=BRUTE) == unsupported
    assert no_alloc(vcrt, torch, dev, len(scene), img, cam, width=64, height=64, spp=1, flags=cabi.VCRT_FLAG_ACCUMULATE) == invalid
    assert no_alloc(vcrt, torch, dev, len(scene), img[:1], cam, width=1920, height=1080, spp=1) == invalid
    host_img = torch.zeros((64, 64, 4))
    assert no_alloc(vcrt, torch, dev, len(scene), host_img, cam, stream=torch.cuda.current_stream(), width=64, height=64, spp=1) == invalid
    assert no_alloc(vcrt, torch, torch.from_numpy(scene), len(scene), img, cam, width=64, height=64, spp=1) == invalid
    torch.cuda.synchronize()
