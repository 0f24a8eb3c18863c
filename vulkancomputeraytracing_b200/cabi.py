# SPDX-License-Identifier: WTFPL
"""ctypes binding of include/vcrt.h (the C ABI in lib/libvcrt.so).

This is the same stub a maintainer of the reference would write for any FFI: structs and
prototypes only.  There is no fallback: if the CUDA library is missing, importing this
module raises; if no GPU is present, every compute call returns VCRT_ERROR_NO_DEVICE and
`check()` raises VcrtError.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvcrt.so")

VCRT_SUCCESS = 0
VCRT_ERROR_INVALID_ARGUMENT = -1
VCRT_ERROR_NO_DEVICE = -2
VCRT_ERROR_NOT_READY = -3
VCRT_ERROR_OUT_OF_MEMORY = -4
VCRT_ERROR_CUDA = -5
VCRT_ERROR_NCCL = -6
VCRT_ERROR_UNSUPPORTED = -7
VCRT_ERROR_SCENE = -8

VCRT_TRAVERSAL_AUTO, VCRT_TRAVERSAL_BRUTE, VCRT_TRAVERSAL_LBVH, VCRT_TRAVERSAL_SIMPLE = 0, 1, 2, 3
VCRT_FLAG_ACCUMULATE, VCRT_FLAG_NO_REDUCE, VCRT_FLAG_ALL_REDUCE, VCRT_FLAG_COUNT_STATS = 1, 2, 4, 8
VCRT_NOALLOC_MAX_SPHERES, VCRT_NOALLOC_MAX_PLANE_PIXELS = 4096, 1 << 21   # cudaRenderNoAlloc: what its static scratch array holds


class VcrtSphere(ctypes.Structure):
    _fields_ = [("center", ctypes.c_float * 3), ("radius", ctypes.c_float),
                ("colour", ctypes.c_float * 3), ("texture", ctypes.c_float * 3)]


class VcrtCamera(ctypes.Structure):
    _fields_ = [("lookfrom", ctypes.c_float * 3), ("lookat", ctypes.c_float * 3),
                ("vup", ctypes.c_float * 3), ("vfov", ctypes.c_float)]


class VcrtRenderParams(ctypes.Structure):
    _fields_ = [("width", ctypes.c_uint32), ("height", ctypes.c_uint32), ("samplesPerPixel", ctypes.c_uint32),
                ("maxDepth", ctypes.c_uint32), ("seed", ctypes.c_uint64), ("sampleOffset", ctypes.c_uint32),
                ("totalSamples", ctypes.c_uint32), ("traversal", ctypes.c_uint32), ("flags", ctypes.c_uint32),
                ("sampleChunks", ctypes.c_uint32), ("reserved", ctypes.c_uint32 * 2), ("semantics", ctypes.c_uint32)]


class VcrtRenderStats(ctypes.Structure):
    _fields_ = [("samples", ctypes.c_uint64), ("rays", ctypes.c_uint64), ("sphereTests", ctypes.c_uint64),
                ("nodeVisits", ctypes.c_uint64), ("depthExhausted", ctypes.c_uint64), ("badMaterial", ctypes.c_uint64),
                ("candidates", ctypes.c_uint64), ("listOverflows", ctypes.c_uint64),
                ("hitsByMaterial", ctypes.c_uint64 * 4), ("pathLength", ctypes.c_uint64 * 8),
                ("renderKernelMs", ctypes.c_float), ("totalMs", ctypes.c_float), ("kernelLaunches", ctypes.c_uint32),
                ("traversalUsed", ctypes.c_uint32), ("sampleChunksUsed", ctypes.c_uint32), ("checkFailures", ctypes.c_uint32),
                ("primaryRays", ctypes.c_uint64), ("primaryKernelMs", ctypes.c_float), ("renderPasses", ctypes.c_uint32)]


class VcrtSceneRecipe(ctypes.Structure):
    _fields_ = [("gridHalf", ctypes.c_int32), ("variant", ctypes.c_uint32),
                ("lambertBelow", ctypes.c_double), ("metalBelow", ctypes.c_double)]


VCRT_SCENE_REFERENCE, VCRT_SCENE_STRESS = 0, 1
VCRT_SEMANTICS_REFERENCE, VCRT_SEMANTICS_BOOK = 0, 1


class VcrtError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"vcrt error {code}: {text}")
        self.code = code


# Every symbol include/vcrt.h and include/SceneGenerator.hpp declare; tests/test_cabi_symbols.py checks the lists agree.
PROTOTYPES = {
    "vcrtDefaultRenderParams": (None, [ctypes.POINTER(VcrtRenderParams)]),
    "vcrtDefaultCamera": (None, [ctypes.POINTER(VcrtCamera)]),
    "cudaUploadScene": (ctypes.c_int, [ctypes.POINTER(VcrtSphere), ctypes.c_uint32, ctypes.POINTER(VcrtCamera)]),
    "cudaRender": (ctypes.c_int, [ctypes.POINTER(VcrtRenderParams)]),
    "cudaDownloadImage": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t]),
    "cudaDownloadSums": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_size_t]),
    "cudaDestroyRenderer": (ctypes.c_int, []),
    "cudaSelectDevice": (ctypes.c_int, [ctypes.c_int]),
    "cudaGetRenderStats": (ctypes.c_int, [ctypes.POINTER(VcrtRenderStats)]),
    "cudaGetLastErrorString": (ctypes.c_char_p, []),
    "cudaGetDeviceImage": (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]),
    "cudaGetDeviceSums": (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_size_t)]),
    "cudaCommGetUniqueId": (ctypes.c_int, [ctypes.c_void_p]),
    "cudaCommInitRank": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]),
    "cudaCommDestroy": (ctypes.c_int, []),
    "cudaCommMaxFloat": (ctypes.c_int, [ctypes.c_float, ctypes.POINTER(ctypes.c_float)]),
    "cudaCommBarrier": (ctypes.c_int, []),
    "cudaMeasureFp32Peak": (ctypes.c_int, [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
    "cudaGetDeviceInfo": (ctypes.c_int, [ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), ctypes.c_char_p, ctypes.c_size_t]),
    "cudaEvaluateSchlick": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]),
    "cudaStreamMark": (ctypes.c_int, [ctypes.c_uint32]),
    "cudaStreamElapsed": (ctypes.c_int, [ctypes.c_uint32, ctypes.c_uint32, ctypes.POINTER(ctypes.c_float)]),
    "cudaFlushL2": (ctypes.c_int, []),
    "cudaRenderNoAlloc": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(VcrtCamera), ctypes.POINTER(VcrtRenderParams),
                                         ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    "vcrtDefaultSceneRecipe": (None, [ctypes.POINTER(VcrtSceneRecipe)]),
    "vcrtSceneCapacity": (ctypes.c_uint32, [ctypes.POINTER(VcrtSceneRecipe)]),
    "vcrtGenerateScene": (ctypes.c_int32, [ctypes.POINTER(VcrtSceneRecipe), ctypes.POINTER(VcrtSphere), ctypes.c_uint32]),
    "vcrtSceneToGlsl": (ctypes.c_size_t, [ctypes.POINTER(VcrtSphere), ctypes.c_uint32, ctypes.c_char_p, ctypes.c_size_t]),
    "vcrtSceneFromGlsl": (ctypes.c_int32, [ctypes.c_char_p, ctypes.POINTER(VcrtSphere), ctypes.c_uint32]),
}

_lib = None


def load():
    """Loads lib/libvcrt.so (built by `python -m vulkancomputeraytracing_b200.build`). Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is not built; run `python -m vulkancomputeraytracing_b200.build` "
                              "(there is no CPU fallback for the render path)")
        lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(code):
    if code != VCRT_SUCCESS:
        text = load().cudaGetLastErrorString()
        raise VcrtError(code, text.decode() if text else "")
