# SPDX-License-Identifier: WTFPL
"""numpy-facing helpers over the C ABI (cabi.py) for the parity tests and bench.py.

This is harness convenience, not a second implementation: every function is one or two calls into
lib/libvcrt.so.  There is no CPU path here; without the library import fails, without a GPU the
calls raise VcrtError(VCRT_ERROR_NO_DEVICE).
"""
import ctypes

import numpy as np

from . import cabi

_VP = ctypes.c_void_p


def generate_scene(grid_half=11, lambert_below=0.8, metal_below=0.95, variant=cabi.VCRT_SCENE_REFERENCE):
    """(count, 10) float32 sphere table from the product's scene producer (host/SceneGenerator.cpp)."""
    lib = cabi.load()
    recipe = cabi.VcrtSceneRecipe()
    lib.vcrtDefaultSceneRecipe(ctypes.byref(recipe))
    recipe.gridHalf, recipe.lambertBelow, recipe.metalBelow, recipe.variant = grid_half, lambert_below, metal_below, variant
    capacity = lib.vcrtSceneCapacity(ctypes.byref(recipe))
    table = np.zeros((capacity, 10), np.float32)
    count = lib.vcrtGenerateScene(ctypes.byref(recipe), table.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), capacity)
    if count < 0:
        raise cabi.VcrtError(count, "vcrtGenerateScene")
    return np.ascontiguousarray(table[:count])


def stress_scene(grid_half=11):
    """BASELINE.json configs[4]: 0 % lambertian / 50 % metal / 50 % glass, no lambertian big sphere or ground."""
    return generate_scene(grid_half, 0.0, 0.5, cabi.VCRT_SCENE_STRESS)


def default_camera():
    cam = cabi.VcrtCamera()
    cabi.load().vcrtDefaultCamera(ctypes.byref(cam))
    return cam


def camera_array(cam=None):
    cam = cam or default_camera()
    return np.array(list(cam.lookfrom) + list(cam.lookat) + list(cam.vup) + [cam.vfov], np.float32)


def upload_scene(spheres, camera=None):
    spheres = np.ascontiguousarray(spheres, np.float32).reshape(-1, 10)
    cam = camera or default_camera()
    lib = cabi.load()
    cabi.check(lib.cudaUploadScene(spheres.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), spheres.shape[0], ctypes.byref(cam)))


def render_params(width, height, spp, depth=50, seed=0, sample_offset=0, traversal=cabi.VCRT_TRAVERSAL_AUTO, flags=0,
                  sample_chunks=0, total_samples=0, shape=0, semantics=0):
    p = cabi.VcrtRenderParams()
    cabi.load().vcrtDefaultRenderParams(ctypes.byref(p))
    p.width, p.height, p.samplesPerPixel, p.maxDepth, p.seed = width, height, spp, depth, seed
    p.sampleOffset, p.traversal, p.flags, p.sampleChunks, p.totalSamples = sample_offset, traversal, flags, sample_chunks, total_samples
    p.reserved[1] = shape   # experiments build only (tools/experiments/README.md); the product rejects non-zero reserved fields
    p.semantics = semantics
    return p


def stats():
    st = cabi.VcrtRenderStats()
    cabi.check(cabi.load().cudaGetRenderStats(ctypes.byref(st)))
    return st


def render(width, height, spp, download=True, out=None, **kw):
    """cudaRender (+ cudaDownloadImage).  Returns (image HxWx4 float32 or None, VcrtRenderStats)."""
    lib = cabi.load()
    p = render_params(width, height, spp, **kw)
    cabi.check(lib.cudaRender(ctypes.byref(p)))
    img = None
    if download:
        img = out if out is not None else np.empty((height, width, 4), np.float32)
        cabi.check(lib.cudaDownloadImage(img.ctypes.data_as(_VP), img.nbytes))
    return img, stats()


def download_sums(width, height):
    img = np.empty((height, width, 4), np.float32)
    cabi.check(cabi.load().cudaDownloadSums(img.ctypes.data_as(_VP), img.nbytes))
    return img


def fp32_peak():
    tf, mhz = ctypes.c_double(), ctypes.c_double()
    cabi.check(cabi.load().cudaMeasureFp32Peak(ctypes.byref(tf), ctypes.byref(mhz)))
    return tf.value, mhz.value


def device_info():
    sms, khz = ctypes.c_int(), ctypes.c_int()
    name = ctypes.create_string_buffer(256)
    cabi.check(cabi.load().cudaGetDeviceInfo(ctypes.byref(sms), ctypes.byref(khz), name, 256))
    return sms.value, khz.value, name.value.decode()


def destroy():
    cabi.load().cudaDestroyRenderer()
