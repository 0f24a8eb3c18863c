/* @file vcrt.h

    C ABI of the B200 path-tracing layer: the boundary that replaces the Vulkan compute
    dispatch of the reference (Renderer.cpp:204-225 record, :678-686 submit) and its
    SPIR-V loader (Shader.cpp:34-95).  Plain pointers and sizes only; no CUDA, torch or
    C++ types cross this boundary, so cgo / JNI / ctypes / a C++ host can all bind it.
    SPDX-License-Identifier: WTFPL

    Error convention (reference: VkResult returned to main, VulkanComputeRayTracing.cpp:20-35):
    every entry point returns 0 (VCRT_SUCCESS, same value as VK_SUCCESS) or a negative code;
    nothing throws, nothing logs.  cudaGetLastErrorString() gives the text of the last failure.
    The layer is single-threaded and not re-entrant, like the reference renderer (global state,
    Renderer.cpp:13-33).  There is NO CPU fallback: without a usable CUDA device every
    compute entry point fails with VCRT_ERROR_NO_DEVICE.
*/
#ifndef VCRT_H
#define VCRT_H

#include <stddef.h>
#include <stdint.h>

#ifndef IN
#define IN  /* reference include/Common.hpp:11-12 */
#define OUT
#endif

#ifdef __cplusplus
extern "C" {
#endif

enum {
    VCRT_SUCCESS = 0,
    VCRT_ERROR_INVALID_ARGUMENT = -1,
    VCRT_ERROR_NO_DEVICE = -2,       /* no CUDA device / driver: there is no CPU path */
    VCRT_ERROR_NOT_READY = -3,       /* cudaRender before cudaUploadScene, download before render */
    VCRT_ERROR_OUT_OF_MEMORY = -4,
    VCRT_ERROR_CUDA = -5,            /* sticky CUDA error, see cudaGetLastErrorString */
    VCRT_ERROR_NCCL = -6,
    VCRT_ERROR_UNSUPPORTED = -7,
    VCRT_ERROR_SCENE = -8            /* kernel met a texture id outside 1..3 (reference: silent no-op, B13) */
};

/* One sphere, field for field the GLSL struct (shaders/include/structures.glsl:10-16).
   texture[0] = material id stored as a float: 1 lambertian, 2 metal, 3 glass
   (shaders/include/textures.glsl:10-12); texture[1] = reflect ratio / fuzz / ior;
   texture[2] unused. 40 bytes, no padding. */
typedef struct VcrtSphere {
    float center[3];
    float radius;
    float colour[3];
    float texture[3];
} VcrtSphere;

/* Camera constants of shaders/include/globals.glsl:21-24. */
typedef struct VcrtCamera {
    float lookfrom[3];
    float lookat[3];
    float vup[3];
    float vfov;  /* degrees */
} VcrtCamera;

/* How rays find their nearest sphere. */
enum {
    VCRT_TRAVERSAL_AUTO = 0,    /* the fastest for the scene: the GPU-built LBVH (measured faster than the scan at every size) */
    VCRT_TRAVERSAL_BRUTE = 1,   /* every ray tests every sphere (what the reference does, functions.glsl:77-81) */
    VCRT_TRAVERSAL_LBVH = 2,    /* GPU-built linear BVH */
    VCRT_TRAVERSAL_SIMPLE = 3   /* one thread per pixel, reference statement order; the on-device cross-check */
};

enum {
    VCRT_FLAG_ACCUMULATE = 1u << 0,   /* add this call's samples to the sums already held (progressive frames) */
    VCRT_FLAG_NO_REDUCE = 1u << 1,    /* multi-GPU: skip the NCCL reduce, keep the local partial sums */
    VCRT_FLAG_ALL_REDUCE = 1u << 2,   /* multi-GPU: every rank ends up with the full image, not just rank 0 */
    VCRT_FLAG_COUNT_STATS = 1u << 3   /* also fill the per-material / path-length counters (slower) */
};

/* What the path computes.  REFERENCE = the reference shader as it is, bugs B1/B3/B4/B4b/B5/B8 included: the
   parity mode and the default.  BOOK = the same path with those bugs fixed the way "Ray Tracing in One Weekend" writes the
   routines (centred viewport, unit-sphere scatter, face-forwarded normals, real Snell/Schlick, gamma 2 on the result). */
enum {
    VCRT_SEMANTICS_REFERENCE = 0,
    VCRT_SEMANTICS_BOOK = 1
};

/* Run-time stand-ins for the shader's compile-time constants (globals.glsl:9-17).
   Zero-initialise, then set what you need; vcrtDefaultRenderParams() gives the reference's
   values (1280x720, 1 spp, depth 50). */
typedef struct VcrtRenderParams {
    uint32_t width;            /* IMAGE_WIDTH */
    uint32_t height;           /* IMAGE_HEIGHT */
    uint32_t samplesPerPixel;  /* SAMPLES_PER_PIXEL rendered by THIS call on THIS device (0: a rank that only joins the reduce) */
    uint32_t maxDepth;         /* MAX_RECURSION_LEVEL */
    uint64_t seed;             /* Philox key */
    uint32_t sampleOffset;     /* global index of this call's first sample (spp sharding / progressive) */
    uint32_t totalSamples;     /* divisor of the final image; 0 = samples accumulated so far (after reduce: over all ranks) */
    uint32_t traversal;        /* VCRT_TRAVERSAL_* */
    uint32_t flags;            /* VCRT_FLAG_* */
    uint32_t sampleChunks;     /* 0 = auto (order-free 32.32 fixed-point sums); 1 = a pixel's samples are summed strictly in index
                                  order in binary32; n = n partial sums per pixel */
    uint32_t reserved[2];      /* must be 0 (VCRT_ERROR_INVALID_ARGUMENT otherwise) */
    uint32_t semantics;        /* VCRT_SEMANTICS_* */
} VcrtRenderParams;

/* Counters of the last cudaRender on this device. */
typedef struct VcrtRenderStats {
    uint64_t samples;          /* pixel-samples traced */
    uint64_t rays;             /* path segments traced (one per pass of functions.glsl:73) */
    uint64_t sphereTests;      /* brute force: rays * sphere count; LBVH: leaf sphere tests actually made */
    uint64_t nodeVisits;       /* LBVH only */
    uint64_t depthExhausted;   /* samples that hit maxDepth (bug B7, counted as black) */
    uint64_t badMaterial;      /* hits on a texture id outside 1..3 */
    uint64_t candidates;       /* brute force: tests that survived the conservative filter */
    uint64_t listOverflows;    /* brute force: rays that fell back to the exhaustive exact scan */
    uint64_t hitsByMaterial[4];/* VCRT_FLAG_COUNT_STATS: [0] other, [1] lambertian, [2] metal, [3] glass */
    uint64_t pathLength[8];    /* VCRT_FLAG_COUNT_STATS: paths ending in the sky after 1..7, >=8 rays */
    float renderKernelMs;      /* CUDA-event time of the path-tracing kernel(s) alone */
    float totalMs;             /* whole cudaRender: kernels + combine + reduce + normalise */
    uint32_t kernelLaunches;   /* kernels this call launched */
    uint32_t traversalUsed;    /* VCRT_TRAVERSAL_* actually run */
    uint32_t sampleChunksUsed;
    uint32_t checkFailures;    /* bounds checks that failed; always 0 unless the library was built with -DVCRT_CHECKED (tools/checked_run.py) */
    uint64_t primaryRays;      /* of `rays`: camera rays traced (and shaded) by the primary-visibility kernel; 0 where it does not run */
    float primaryKernelMs;     /* of renderKernelMs: the primary-visibility kernel(s) */
    uint32_t renderPasses;     /* launches of the path kernel this call was split into (sample sub-ranges) */
} VcrtRenderStats;

void vcrtDefaultRenderParams(OUT VcrtRenderParams* params);
void vcrtDefaultCamera(OUT VcrtCamera* camera);

/* Replaces: the baked `const sphere world[]` + camera globals (globals.glsl:21-24, :29-518) that
   the reference compiles into the shader, and vkCreateShaderModule of that shader
   (Shader.cpp:34-95, Renderer.cpp:240, :534-543).  Copies the host AoS, builds the device SoA
   (and the LBVH when asked for / needed).  The caller keeps ownership of `spheres`. */
int cudaUploadScene(IN const VcrtSphere* spheres, IN uint32_t count, IN const VcrtCamera* camera);

/* Replaces: recordComputeCommandBuffer + vkQueueSubmit + the device wait that follows a frame
   (Renderer.cpp:204-225, :678-686; platform/Linux.cpp:365).  Synchronous: returns when the
   normalised RGBA32F image (the reference's storage image, Renderer.cpp:433-453) is complete in
   device memory owned by the layer. */
int cudaRender(IN const VcrtRenderParams* params);

/* Replaces: the fragment stage's texelFetch of the result image (shaders/shader.frag:11-14).
   Copies width*height*4 floats (linear radiance, alpha = 1; no gamma, bug B8). */
int cudaDownloadImage(OUT float* rgba, IN size_t bytes);

/* Raw per-pixel SUMS (not divided) of this device, for callers that reduce themselves. */
int cudaDownloadSums(OUT float* rgba, IN size_t bytes);

/* Replaces: EndRenderingOperation's destruction of the image, memory, pipeline
   (Renderer.cpp:737-739, :752-763, :770-772).  Tolerates never-created state. */
int cudaDestroyRenderer(void);

/* CUDA analogue of picking the physical device (Environment.cpp:157-165). Default: device 0. */
int cudaSelectDevice(IN int device);

int cudaGetRenderStats(OUT VcrtRenderStats* stats);
const char* cudaGetLastErrorString(void);

/* Device pointer + size of the layer-owned result image / sum buffer, for zero-copy consumers
   in the same process (e.g. a torch tensor view). */
int cudaGetDeviceImage(OUT void** devicePointer, OUT size_t* bytes);
int cudaGetDeviceSums(OUT void** devicePointer, OUT size_t* bytes);

/* ---- multi-GPU: samples-per-pixel sharded over ranks, one NCCL reduce of the accumulation buffer ---- */
/* One process per GPU: rank 0 calls cudaCommGetUniqueId, ships the 128 bytes to every rank by
   any means (the Python host uses torch.distributed), every rank calls cudaCommInitRank.
   With a communicator, cudaRender is COLLECTIVE (unless VCRT_FLAG_NO_REDUCE): every rank must call it with the same image
   size, depth, semantics and sampleChunks, and pass totalSamples when the shards are unequal.  Errors are not collective: a
   rank whose call fails before the reduce (bad argument, out of memory) leaves the others waiting, as with any NCCL program -
   validate on every rank before the first collective call.  A rank without samples passes samplesPerPixel = 0. */
int cudaCommGetUniqueId(OUT void* id128);
int cudaCommInitRank(IN const void* id128, IN int rank, IN int worldSize);
int cudaCommDestroy(void);
/* Max over ranks of one float (timing), through the same communicator. */
int cudaCommMaxFloat(IN float value, OUT float* maxValue);
int cudaCommBarrier(void);

/* ---- measurement ---- */
/* Runs an FFMA-only kernel on the selected device and reports the FP32 FMA rate it reached:
   the denominator of the roofline (MEASURED_PEAKS.json carries HBM and bf16 only). */
int cudaMeasureFp32Peak(OUT double* tflops, OUT double* smClockMHz);
int cudaGetDeviceInfo(OUT int* smCount, OUT int* maxClockKHz, OUT char* name, IN size_t nameBytes);
/* Evaluates the device code's schlick() (functions.glsl:58-62) on host arrays: the hook of the parity test that compares it
   with the reference build on a dense set of inputs. */
int cudaEvaluateSchlick(IN const float* cosine, IN const float* ior, IN uint32_t count, OUT float* reflectProb);
/* CUDA events on the layer's own stream (the stream every kernel of cudaRender is launched on), so a caller
   can time a region of several calls on the device: record `slot` (0..15) now / milliseconds between two
   recorded slots (waits for the later one). */
int cudaStreamMark(IN uint32_t slot);
int cudaStreamElapsed(IN uint32_t fromSlot, IN uint32_t toSlot, OUT float* milliseconds);
/* Writes a scratch buffer larger than the L2 on the layer's stream (timing hygiene between timed renders). */
int cudaFlushL2(void);

/* ---- a render call that allocates nothing ---- */
/* No counterpart in the reference: added for callers that cannot call
   cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the in-order exhaustive
   kernel of VCRT_TRAVERSAL_SIMPLE (every ray tests every sphere exactly, in the caller's order: right for ANY scene), spread
   over (pixel, run of samples), with all of its scratch memory - the device copy of the scene as a structure of arrays, the
   partial-sum planes, the counters - in ONE file-scope `static __device__` array of VCRT_NOALLOC_SCRATCH_BYTES that every GPU
   which loads the library holds until the process exits.  The call makes no cudaMalloc / cudaFree and calls no allocating
   library routine; it creates no stream or event, does not synchronise and does not touch the state of the other entry points
   (no cudaUploadScene needed, cudaDownloadImage / cudaGetRenderStats do not see it).

     deviceSpheres  `count` spheres in DEVICE memory of the CURRENT device (the caller's cudaSetDevice, not cudaSelectDevice)
     camera, params HOST structs, read before the call returns.  Used: width, height, samplesPerPixel, maxDepth, seed,
                    sampleOffset, totalSamples (0 = samplesPerPixel), semantics, sampleChunks (0 or 1: a pixel's samples summed
                    strictly in index order; n: n partial sums per pixel, summed in run order).  traversal must be AUTO or SIMPLE,
                    flags and reserved must be 0.
     deviceImage    width*height*4 floats in DEVICE memory, the normalised RGBA32F image, the same bits as cudaRender with
                    the same params (traversal SIMPLE for sampleChunks <= 1, BRUTE / LBVH for n) followed by cudaDownloadImage
     stream         a cudaStream_t (NULL: the default stream).  The work is ENQUEUED there; the image is complete when the
                    stream gets past it, and a kernel fault surfaces at the caller's next synchronisation.

   Limits (VCRT_ERROR_INVALID_ARGUMENT with a text that names them): count <= VCRT_NOALLOC_MAX_SPHERES, and
   width*height * runs <= VCRT_NOALLOC_MAX_PLANE_PIXELS where runs = ceil(spp / ceil(spp / max(sampleChunks, 1))) - 1920x1080
   in one run, 1280x720 in two, 400x225 in 23.  Host pointers for deviceSpheres / deviceImage are refused.
   Every call shares the one array, which keeps the last call's contents; the call itself clears what has to start at zero.
   Two calls that OVERLAP on the GPU (different streams without an ordering between them, two host threads, two processes are
   fine - they have an array each) race on it: keep the calls of one process on one stream or order them with events.
   Hits on a texture id outside 1..3 are absorbed as in cudaRender but cannot be reported (nothing is read back).
   The first call on a device loads the module and looks the array's address up (cudaGetSymbolAddress): make it outside a
   graph capture. */
#define VCRT_NOALLOC_MAX_SPHERES 4096u
#define VCRT_NOALLOC_MAX_PLANE_PIXELS (1u << 21)
#define VCRT_NOALLOC_SCRATCH_BYTES (256u + VCRT_NOALLOC_MAX_SPHERES * 48u + VCRT_NOALLOC_MAX_PLANE_PIXELS * 16u)   /* 32.2 MiB */
int cudaRenderNoAlloc(IN const VcrtSphere* deviceSpheres, IN uint32_t count, IN const VcrtCamera* camera,
                      IN const VcrtRenderParams* params, OUT float* deviceImage, IN size_t imageBytes, IN void* stream);

#ifdef __cplusplus
}
#endif
#endif /* VCRT_H */
