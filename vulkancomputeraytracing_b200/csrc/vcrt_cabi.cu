/* @file vcrt_cabi.cu

    The C ABI of include/vcrt.h: scene upload (AoS -> device SoA blob), render dispatch,
    partial-sum combine, the NCCL reduce of the float sums, normalisation and read-back.
    Host code is C++ in the reference's C flavour: free functions, file-static state,
    integer return codes, no exceptions (reference Renderer.cpp:13-33, README.md:7).
    SPDX-License-Identifier: WTFPL
*/
#include "../../include/vcrt.h"
#include "vcrt_kernel_simple.cuh"
#include "vcrt_kernel_scan.cuh"
#include "vcrt_kernel_lbvh.cuh"
#include "vcrt_kernel_scanu.cuh"
#include "vcrt_kernel_primary.cuh"
#ifdef VCRT_EXPERIMENTS   // A/B kernels that no default call reaches: tools/experiments/README.md (build with VCRT_EXTRA_NVCC=-DVCRT_EXPERIMENTS)
#include "../../tools/experiments/vcrt_kernel_scanw.cuh"
#include "../../tools/experiments/vcrt_kernel_lbvh_gen1.cuh"
#endif

#include <cub/device/device_radix_sort.cuh>

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math.h>
#include <algorithm>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

using namespace vcrt;

static_assert(kFlagCountStats == VCRT_FLAG_COUNT_STATS, "device-side copy of the flag bit");

// ------------------------------------------------------------------ file-static state
// (reference Renderer.cpp:13-33 keeps its pipeline, image and memory handles the same way)
static int cudaComputeDevice = 0;
static bool cudaComputeDeviceReady = false;
static int cudaComputeSmCount = 0;
static cudaStream_t cudaComputeStream = nullptr;
static cudaEvent_t cudaComputeEvents[4] = {nullptr, nullptr, nullptr, nullptr};
static const uint32_t kMaxTimedPasses = 64;                 // primary-kernel time is measured for the first 64 launches of a call
static cudaEvent_t cudaComputePassEvents[2 * kMaxTimedPasses];
static uint32_t cudaComputePassEventCount = 0;

static unsigned char* cudaComputeSceneBlob = nullptr;   // device copy of the SoA blob
static SceneBlob cudaComputeSceneHeader;
static bool cudaComputeSceneReady = false;
static bool cudaComputeSceneInConstant = false;   // the filter table of the uniform-datapath scan is loaded
static bool cudaComputeSceneNonFinite = false;    // a centre or radius is NaN / inf / beyond the tame range (vcrt_device.cuh): only the in-order scan reproduces the reference
static bool cudaComputeSceneBounded = true;       // every attenuation is in [0, 1]: a sample's radiance is, too, and the 32.32 fixed-point sums cannot overflow
static unsigned char* cudaComputeHostBlob = nullptr;   // host copy of the last uploaded blob: an identical upload keeps the LBVH
static size_t cudaComputeHostBlobBytes = 0;
static size_t cudaComputeSceneBlobCapacity = 0;
struct SceneExtent { float centreMin[3], centreMax[3], radiusMin, radiusMax; };
// LBVH over the scan spheres (device indices [0, scanCount)), built on first use after an upload
static bool cudaComputeBvhReady = false;
static void* cudaComputeBvhArena = nullptr;
static size_t cudaComputeBvhArenaBytes = 0;
static BvhScene cudaComputeBvhScene;
static float4* cudaComputeHostShape = nullptr;      // host copy of shape[] (device order): the LBVH build picks its outliers from it
static VcrtCamera cudaComputeCamera;

static float4* cudaComputeSumPlanes = nullptr;          // [chunks][W*H] partial sums of the last render
static size_t cudaComputeSumPlanesBytes = 0;
static float4* cudaComputeAccumulated = nullptr;        // [W*H] sums over all samples so far (this device)
static float4* cudaComputeReduced = nullptr;            // [W*H] sums after the cross-device reduce
static long long* cudaComputeFixedSums = nullptr;       // [W*H][kFixedLanes] signed 32.32 fixed-point sums (order-independent accumulation)
static long long* cudaComputeFixedReduced = nullptr;    // the same after the cross-device reduce
static bool cudaComputeAccumulatedIsFixed = false;
static float4* cudaComputeResultImage = nullptr;        // [W*H] normalised RGBA32F (reference Renderer.cpp:433-453)
static uint32_t cudaComputeImageWidth = 0, cudaComputeImageHeight = 0;
static uint64_t cudaComputeAccumulatedSamples = 0;
static bool cudaComputeImageValid = false;
static float4* cudaComputePrimary = nullptr;            // PrimaryBuffer of one launch (vcrt_kernel_primary.cuh): plane a, then plane b
static size_t cudaComputePrimaryCapacity = 0;           // entries (32 bytes each)
static const size_t kPrimaryMaxEntries = (size_t)9 << 27;   // 36 GiB: configs[1] (1920 x 1080 x 500 = 1.04e9 samples, 33 GB) in one launch
static unsigned long long* cudaComputeCounters = nullptr;
static unsigned int* cudaComputeWorkCounter = nullptr;
static VcrtRenderStats cudaComputeLastStats;
static char cudaComputeLastError[512] = "";

static int fail(int code, const char* fmt, const char* detail) {
    snprintf(cudaComputeLastError, sizeof cudaComputeLastError, fmt, detail ? detail : "");
    return code;
}
#define CUDA_TRY(expr)                                                                    \
    do {                                                                                  \
        cudaError_t cudaTryResult = (expr);                                               \
        if (cudaTryResult != cudaSuccess)                                                 \
            return fail(cudaTryResult == cudaErrorMemoryAllocation ? VCRT_ERROR_OUT_OF_MEMORY : VCRT_ERROR_CUDA, \
                        #expr ": %s", cudaGetErrorString(cudaTryResult));                 \
    } while (0)

// Every entry point comes through here first: the layer's device is made current on EVERY call, because a host that shares
// the process with torch or another CUDA user (or calls from another thread) may have switched devices in between, and
// allocations, launches and the __constant__ table are all per device.
static int ensureDevice(void) {
    if (cudaComputeDeviceReady) { CUDA_TRY(cudaSetDevice(cudaComputeDevice)); return VCRT_SUCCESS; }
    int deviceCount = 0;
    cudaError_t e = cudaGetDeviceCount(&deviceCount);
    if (e != cudaSuccess || deviceCount == 0)
        return fail(VCRT_ERROR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU path", cudaGetErrorString(e));
    if (cudaComputeDevice >= deviceCount) return fail(VCRT_ERROR_INVALID_ARGUMENT, "device index out of range%s", "");
    CUDA_TRY(cudaSetDevice(cudaComputeDevice));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, cudaComputeDevice));
    if (prop.major < 10) return fail(VCRT_ERROR_UNSUPPORTED, "device is not sm_100 class (%s)", prop.name);
    cudaComputeSmCount = prop.multiProcessorCount;
    CUDA_TRY(cudaStreamCreateWithFlags(&cudaComputeStream, cudaStreamNonBlocking));
    for (int i = 0; i < 4; i++) CUDA_TRY(cudaEventCreate(&cudaComputeEvents[i]));
    CUDA_TRY(cudaMalloc(&cudaComputeCounters, sizeof(unsigned long long) * CTR_COUNT));
    CUDA_TRY(cudaMalloc(&cudaComputeWorkCounter, 4 * sizeof(unsigned int)));   // [0] the item dispenser, [2], [3] the cursors of the compact hand-over list
    cudaComputeDeviceReady = true;
    return VCRT_SUCCESS;
}

extern "C" const char* cudaGetLastErrorString(void) { return cudaComputeLastError; }

extern "C" void vcrtDefaultRenderParams(OUT VcrtRenderParams* params) {
    if (!params) return;
    memset(params, 0, sizeof *params);
    params->width = 1280; params->height = 720;   // globals.glsl:16-17, Common.hpp:23-24
    params->samplesPerPixel = 1;                   // globals.glsl:12
    params->maxDepth = 50;                         // globals.glsl:14
    params->traversal = VCRT_TRAVERSAL_AUTO;
}

extern "C" void vcrtDefaultCamera(OUT VcrtCamera* camera) {
    if (!camera) return;
    const VcrtCamera c = {{13.f, 2.f, 3.f}, {0.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, 20.f};  // globals.glsl:21-24
    *camera = c;
}

extern "C" int cudaSelectDevice(IN int device) {
    if (device < 0) return fail(VCRT_ERROR_INVALID_ARGUMENT, "negative device index%s", "");
    if (cudaComputeDeviceReady && device != cudaComputeDevice)
        return fail(VCRT_ERROR_NOT_READY, "cudaSelectDevice after the renderer was created; call cudaDestroyRenderer first%s", "");
    cudaComputeDevice = device;
    return ensureDevice();
}

extern "C" int cudaGetDeviceInfo(OUT int* smCount, OUT int* maxClockKHz, OUT char* name, IN size_t nameBytes) {
    int rc = ensureDevice();
    if (rc) return rc;
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, cudaComputeDevice));
    if (smCount) *smCount = prop.multiProcessorCount;
    if (maxClockKHz) CUDA_TRY(cudaDeviceGetAttribute(maxClockKHz, cudaDevAttrClockRate, cudaComputeDevice));
    if (name && nameBytes) { strncpy(name, prop.name, nameBytes - 1); name[nameBytes - 1] = 0; }
    return VCRT_SUCCESS;
}

// ------------------------------------------------------------------ scene upload
// Filter margin (see vcrt_kernel_scan.cuh).  The scan passes sphere (c, r) for ray (o, d) when
//     fl( ((c-o).u)^2 + ((c-o).v)^2 - R2 ) <= kappa*|o|^2,     R2 = r^2 + kappa*(|c|^2 + r^2)
// and the exact test (hit_exact, the reference's arithmetic) can only succeed when its own
// computed discriminant is >= 0.  Both computations differ from the real-number quantity
// a*(r^2 - D^2) (D = distance from c to the ray's line) by rounding error:
//   reference side: <= ~18 eps * (|oc|^2 + r^2)      (oc rounding, two dot products, a*c)
//   filter side   : <= ~30 eps * (|o| + |c|)^2       (basis orthonormality ~4 eps, the expanded
//                                                     form c.u - o.u, the two squares)
// with eps = 2^-24 and |oc|^2 <= 2(|o|^2+|c|^2), (|o|+|c|)^2 <= 2(|o|^2+|c|^2) the total is below
// 96 eps (|o|^2 + |c|^2 + r^2) < kappa (|o|^2 + |c|^2 + r^2) for kappa = 2^-17 = 128 eps.
// So every sphere the reference arithmetic would hit is in the candidate list; extra
// candidates cost time, never correctness.  tests/test_gpu_parity.py checks the consequence:
// VCRT_TRAVERSAL_BRUTE and VCRT_TRAVERSAL_SIMPLE images are bit-identical.
static const float kAlwaysRadius = 50.0f;   // spheres this large (the ground) skip the filter: tested exactly for every ray
static const uint32_t kMaxAlways = 64;

extern "C" int cudaUploadScene(IN const VcrtSphere* spheres, IN uint32_t count, IN const VcrtCamera* camera) {
    if ((!spheres && count) || !camera) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaUploadScene: null argument%s", "");
    int rc = ensureDevice();
    if (rc) return rc;

    uint32_t* order = new uint32_t[count ? count : 1];
    uint32_t scanCount = 0, alwaysCount = 0;
    for (uint32_t i = 0; i < count; i++) if (!(spheres[i].radius > kAlwaysRadius)) order[scanCount++] = i;
    for (uint32_t i = 0; i < count; i++) if (spheres[i].radius > kAlwaysRadius) order[scanCount + alwaysCount++] = i;
    if (alwaysCount > kMaxAlways) {   // too many huge spheres: keep the first kMaxAlways out, scan the rest like any other
        scanCount = 0; alwaysCount = 0; uint32_t kept = 0;
        for (uint32_t i = 0; i < count; i++) { bool big = spheres[i].radius > kAlwaysRadius && kept < kMaxAlways; if (big) kept++; else order[scanCount++] = i; }
        kept = 0;
        for (uint32_t i = 0; i < count; i++) if (spheres[i].radius > kAlwaysRadius && kept < kMaxAlways) { order[scanCount + alwaysCount++] = i; kept++; }
    }
    // The device arrays keep the huge spheres at the END.  That must not change the order the reference scans in (equal roots go
    // to the lower index; with NaNs about, everything depends on the order), so the split is only made when they already close
    // the caller's array, as the ground of SceneGenerator.cpp:52 does; otherwise every sphere is scanned like any other.
    for (uint32_t j = 0; j < count; j++) if (order[j] != j) { for (uint32_t i = 0; i < count; i++) order[i] = i; scanCount = count; alwaysCount = 0; break; }
    const uint32_t scanPad = (scanCount + 15u) & ~15u;   // the scan walks the list 16 (first generation: 8) spheres at a time

    SceneBlob hdr; memset(&hdr, 0, sizeof hdr);
    hdr.count = count; hdr.scanCount = scanCount; hdr.scanPad = scanPad; hdr.alwaysCount = alwaysCount;
    hdr.offHot = (uint32_t)sizeof(SceneBlob);
    hdr.offShape = hdr.offHot + scanPad * 16u;
    hdr.offSurf = hdr.offShape + count * 16u;
    hdr.offKind = hdr.offSurf + count * 16u;
    hdr.offOrig = (hdr.offKind + count * 4u + 15u) & ~15u;
    hdr.bytes = (hdr.offOrig + count * 4u + 15u) & ~15u;
    unsigned char* host = (unsigned char*)calloc(1, hdr.bytes);
    if (!host) { delete[] order; return fail(VCRT_ERROR_OUT_OF_MEMORY, "scene blob%s", ""); }
    float4* hot = (float4*)(host + hdr.offHot); float4* shape = (float4*)(host + hdr.offShape);
    float4* surf = (float4*)(host + hdr.offSurf); int* kind = (int*)(host + hdr.offKind);
    uint32_t* orig = (uint32_t*)(host + hdr.offOrig);
    cudaComputeSceneNonFinite = false;
    cudaComputeSceneBounded = true;
    for (uint32_t j = 0; j < count; j++) {
        const VcrtSphere& s = spheres[order[j]];
        orig[j] = order[j];
        // attenuation per hit: lambertian albedo * texture.y (textures.glsl:22), metal albedo (:60), glass 1; all in [0, 1] keeps
        // the radiance of a sample in [0, 1] (the sky of functions.glsl:89 is), which is what the fixed-point sums rely on
        const float ratio = (int)s.texture[0] == 1 ? s.texture[1] : 1.0f;
        if (!(s.colour[0] >= 0.f && s.colour[0] <= 1.f && s.colour[1] >= 0.f && s.colour[1] <= 1.f && s.colour[2] >= 0.f && s.colour[2] <= 1.f && ratio >= 0.f && ratio <= 1.f))
            cudaComputeSceneBounded = false;
        // (negated comparisons: a NaN lands here too)
        if (!(fabsf(s.center[0]) <= kTameSceneMax && fabsf(s.center[1]) <= kTameSceneMax && fabsf(s.center[2]) <= kTameSceneMax && fabsf(s.radius) <= kTameSceneMax))
            cudaComputeSceneNonFinite = true;
        shape[j] = make_float4(s.center[0], s.center[1], s.center[2], s.radius);
        surf[j] = make_float4(s.colour[0], s.colour[1], s.colour[2], s.texture[1]);
        kind[j] = (int)s.texture[0];   // switch(int(record.texture.x)), textures.glsl:66
        if (j < scanCount) {
            double cc = (double)s.center[0] * s.center[0] + (double)s.center[1] * s.center[1] + (double)s.center[2] * s.center[2];
            double r2 = (double)s.radius * s.radius;
            double inflated = (r2 + (double)kKappa * (cc + r2)) * (1.0 + 1.0 / 4194304.0);
            float f = (float)inflated;
            if ((double)f < inflated) f = nextafterf(f, INFINITY);
            hot[j] = make_float4(s.center[0], s.center[1], s.center[2], f);
        }
    }
    for (uint32_t j = scanCount; j < scanPad; j++) hot[j] = make_float4(0.f, 0.f, 0.f, -INFINITY);   // padding never passes
    // second-generation scan (vcrt_kernel_scanu.cuh): (cx, cy, cz, |c|^2 - R2) in the constant bank, R2 = r^2 + kappa(|c|^2 + r^2)
    cudaComputeSceneInConstant = scanPad <= (uint32_t)kMaxConstSpheres;
    if (cudaComputeSceneInConstant && scanPad) {
        float4* chot = (float4*)calloc(scanPad, sizeof(float4));
        if (!chot) { free(host); delete[] order; return fail(VCRT_ERROR_OUT_OF_MEMORY, "scene table%s", ""); }
        for (uint32_t j = 0; j < scanPad; j++) {
            if (j >= scanCount) { chot[j] = make_float4(0.f, 0.f, 0.f, INFINITY); continue; }
            const VcrtSphere& s = spheres[order[j]];
            double cc = (double)s.center[0] * s.center[0] + (double)s.center[1] * s.center[1] + (double)s.center[2] * s.center[2];
            double r2 = (double)s.radius * s.radius;
            double bound = cc - (r2 + (double)kKappaU * (cc + r2)) * (1.0 + 1.0 / 4194304.0);
            float f = (float)bound;
            if ((double)f > bound) f = nextafterf(f, -INFINITY);   // rounded DOWN: can only let more spheres pass
            chot[j] = make_float4(s.center[0], s.center[1], s.center[2], f);
        }
        cudaError_t ce = cudaMemcpyToSymbolAsync(cSceneHot, chot, (size_t)scanPad * sizeof(float4), 0, cudaMemcpyHostToDevice, cudaComputeStream);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(cudaComputeStream);
        // the primary-visibility scan (vcrt_kernel_primary.cuh): centres relative to the camera, (c', sqrt(|c'|^2 - R2)) with
        // R2 = r^2 + kappa(|c'|^2 + r^2); c' is the binary64 difference rounded once, and |c'|^2 is taken from the ROUNDED c'
        for (uint32_t j = 0; j < scanPad && ce == cudaSuccess; j++) {
            if (j >= scanCount) { chot[j] = make_float4(0.f, 0.f, 0.f, INFINITY); continue; }
            const VcrtSphere& s = spheres[order[j]];
            const float rx = (float)((double)s.center[0] - (double)camera->lookfrom[0]), ry = (float)((double)s.center[1] - (double)camera->lookfrom[1]),
                        rz = (float)((double)s.center[2] - (double)camera->lookfrom[2]);
            const double cc = (double)rx * rx + (double)ry * ry + (double)rz * rz, r2 = (double)s.radius * s.radius;
            const double bound = cc - (r2 + (double)kKappaU * (cc + r2)) * (1.0 + 1.0 / 4194304.0);
            float f = (float)bound;
            if ((double)f > bound) f = nextafterf(f, -INFINITY);   // rounded DOWN: can only let more spheres pass
            // The kernel compares |t| with the square root instead of t^2 with f (one packed multiply less per ray pair and
            // sphere).  fl(t^2) >= f > 0 implies |t| >= sqrt(f) (1 - eps), so with the root scaled by 1 - 2^-22 and rounded down
            // every ray that passed the squared form still passes; f <= 0 (the camera inside the inflated sphere) and tiny f
            // (t^2 near the subnormals) always pass.
            float root = 0.f;
            if (f > 1e-30f) {
                const double lower = sqrt((double)f) * (1.0 - 1.0 / 4194304.0);
                root = (float)lower;
                if ((double)root > lower) root = nextafterf(root, -INFINITY);
            }
            chot[j] = make_float4(rx, ry, rz, root);
        }
        if (ce == cudaSuccess) ce = cudaMemcpyToSymbolAsync(cPrimaryHot, chot, (size_t)scanPad * sizeof(float4), 0, cudaMemcpyHostToDevice, cudaComputeStream);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(cudaComputeStream);
        free(chot);
        if (ce != cudaSuccess) { free(host); delete[] order; return fail(VCRT_ERROR_CUDA, "scene table upload: %s", cudaGetErrorString(ce)); }
    }
    memcpy(host, &hdr, sizeof hdr);
    delete[] order;
    if (cudaComputeHostShape) delete[] cudaComputeHostShape;
    cudaComputeHostShape = new float4[count ? count : 1];
    memcpy(cudaComputeHostShape, shape, sizeof(float4) * count);

    // The same bytes as last time (a caller that uploads every frame, bench.py's e2e leg): the copy below still happens - the
    // caller's data is what gets rendered - but the tree built from those bytes stays (8 ms at 100 k spheres).
    const bool sameScene = cudaComputeSceneReady && cudaComputeHostBlob && cudaComputeHostBlobBytes == hdr.bytes && memcmp(cudaComputeHostBlob, host, hdr.bytes) == 0;
    cudaComputeSceneReady = false;
    cudaError_t e = cudaSuccess;
    if (hdr.bytes > cudaComputeSceneBlobCapacity) {   // the device blob keeps its address while it fits (the tree points into it)
        if (cudaComputeSceneBlob) { cudaFree(cudaComputeSceneBlob); cudaComputeSceneBlob = nullptr; }
        cudaComputeSceneBlobCapacity = 0; cudaComputeBvhReady = false;
        e = cudaMalloc(&cudaComputeSceneBlob, hdr.bytes);
        if (e == cudaSuccess) cudaComputeSceneBlobCapacity = hdr.bytes;
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(cudaComputeSceneBlob, host, hdr.bytes, cudaMemcpyHostToDevice, cudaComputeStream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaComputeStream);
    if (e != cudaSuccess) { free(host); cudaComputeBvhReady = false; return fail(e == cudaErrorMemoryAllocation ? VCRT_ERROR_OUT_OF_MEMORY : VCRT_ERROR_CUDA, "scene upload: %s", cudaGetErrorString(e)); }
    if (cudaComputeHostBlob) free(cudaComputeHostBlob);
    cudaComputeHostBlob = host; cudaComputeHostBlobBytes = hdr.bytes;   // kept for the comparison above
    cudaComputeSceneHeader = hdr;
    if (!sameScene) cudaComputeBvhReady = false;
    cudaComputeCamera = *camera;
    cudaComputeSceneReady = true;
    cudaComputeImageValid = false;
    cudaComputeAccumulatedSamples = 0;
    return VCRT_SUCCESS;
}

// shader.comp:18-39 in binary32, one rounding per operation (x86-64 host code has no FMA
// contraction without -march flags; the build also passes -ffp-contract=off).
static Frame makeFrame(const VcrtCamera& cam, uint32_t width, uint32_t height, bool book) {
    struct H3 { float x, y, z; };
    auto sub = [](H3 a, H3 b) { return H3{a.x - b.x, a.y - b.y, a.z - b.z}; };
    auto lscale = [](float s, H3 a) { return H3{s * a.x, s * a.y, s * a.z}; };
    auto divs = [](H3 a, float s) { return H3{a.x / s, a.y / s, a.z / s}; };
    auto dot = [](H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; };
    auto unit = [&](H3 a) { return divs(a, sqrtf(dot(a, a))); };
    auto cross = [](H3 a, H3 b) { return H3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; };
    auto add = [](H3 a, H3 b) { return H3{a.x + b.x, a.y + b.y, a.z + b.z}; };
    H3 lookfrom{cam.lookfrom[0], cam.lookfrom[1], cam.lookfrom[2]}, lookat{cam.lookat[0], cam.lookat[1], cam.lookat[2]};
    H3 vup{cam.vup[0], cam.vup[1], cam.vup[2]};
    float focal_length = sqrtf(dot(sub(lookfrom, lookat), sub(lookfrom, lookat)));
    float theta = cam.vfov * 0.017453292519943295f;
    float h = tanf(theta / 2);
    float viewport_height = 2 * h * focal_length;
    float viewport_width = viewport_height * (book ? (float)width / (float)height : (float)(width / height));   // bug B1: integer division
    H3 w = unit(sub(lookfrom, lookat));
    H3 u = unit(cross(vup, w));
    H3 v = cross(w, u);
    H3 viewport_u = lscale(viewport_width, u);
    H3 viewport_v = lscale(viewport_height, H3{-v.x, -v.y, -v.z});
    H3 du = divs(viewport_u, book ? (float)width : (float)height);      // bug B1: both deltas / HEIGHT
    H3 dv = divs(viewport_v, (float)height);
    H3 upper_left = sub(sub(sub(lookfrom, lscale(focal_length, w)), divs(viewport_u, 2.f)), divs(viewport_v, 2.f));
    H3 p00 = add(upper_left, lscale(0.5f, add(du, dv)));
    Frame f;
    f.center = V3{lookfrom.x, lookfrom.y, lookfrom.z}; f.pixel00 = V3{p00.x, p00.y, p00.z};
    f.du = V3{du.x, du.y, du.z}; f.dv = V3{dv.x, dv.y, dv.z};
    return f;
}

// ------------------------------------------------------------------ small kernels
// accumulated[p] = (keep ? accumulated[p] : 0) + plane[0][p] + plane[1][p] + ...   (fixed order)
__global__ void combine_planes_kernel(float4* accumulated, const float4* planes, uint32_t pixels, uint32_t chunks, int keep) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    float4 a = keep ? accumulated[p] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t c = 0; c < chunks; c++) {
        float4 s = planes[(size_t)c * pixels + p];
        a.x = __fadd_rn(a.x, s.x); a.y = __fadd_rn(a.y, s.y); a.z = __fadd_rn(a.z, s.z);
    }
    accumulated[p] = a;
}
// shader.comp:56-57: color.rgb /= SAMPLES_PER_PIXEL; alpha stays 1 (shader.comp:44).
// gamma != 0 (BOOK semantics): linear_to_gamma of the book, sqrt of the mean.
__device__ __forceinline__ float to_gamma(float v, int gamma) { return gamma ? (v > 0.f ? __fsqrt_rn(v) : 0.f) : v; }
__global__ void normalise_kernel(float4* image, const float4* sums, uint32_t pixels, float samples, int gamma) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    float4 s = sums[p];
    image[p] = make_float4(to_gamma(__fdiv_rn(s.x, samples), gamma), to_gamma(__fdiv_rn(s.y, samples), gamma), to_gamma(__fdiv_rn(s.z, samples), gamma), 1.0f);
}
// Fixed-point sums -> float sums (for cudaDownloadSums) and -> the normalised image.
__global__ void fixed_to_float_kernel(float4* sums, const long long* fixedSums, uint32_t pixels) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    const long long* f = fixedSums + kFixedLanes * (size_t)p;
    const double s = 1.0 / 4294967296.0;
    sums[p] = make_float4((float)((double)f[0] * s), (float)((double)f[1] * s), (float)((double)f[2] * s), 0.f);
}
__global__ void normalise_fixed_kernel(float4* image, const long long* fixedSums, uint32_t pixels, double samples, int gamma) {
    uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    const long long* f = fixedSums + kFixedLanes * (size_t)p;
    const double s = 1.0 / (4294967296.0 * samples);
    image[p] = make_float4(to_gamma((float)((double)f[0] * s), gamma), to_gamma((float)((double)f[1] * s), gamma), to_gamma((float)((double)f[2] * s), gamma), 1.0f);
}
// FFMA-only kernel: what the FP32 pipe of this GPU sustains when nothing else is asked of it.  512 FFMAs per trip (8
// independent chains x 64) against 3 loop instructions, so the loop overhead is below 0.6 % (round 1 had 64 per trip and read
// 4.5 % low).  bench.py reports it beside the nominal peak SMs x 128 lanes x 2 x sampled SM clock, which is the denominator.
__global__ void fp32_peak_kernel(float* out, int iters, float a, float b) {
    float x[8];
#pragma unroll
    for (int c = 0; c < 8; c++) x[c] = threadIdx.x * 1e-3f + c;
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 64; u++)
#pragma unroll
            for (int c = 0; c < 8; c++) x[c] = fmaf(x[c], a, b);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < 8; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

extern "C" int cudaMeasureFp32Peak(OUT double* tflops, OUT double* smClockMHz) {
    int rc = ensureDevice();
    if (rc) return rc;
    const int threads = 512, blocks = cudaComputeSmCount * 4, iters = 2048;
    float* out = nullptr;
    CUDA_TRY(cudaMalloc(&out, sizeof(float) * threads * blocks));
    float best = 1e30f;
    for (int rep = 0; rep < 8; rep++) {   // ~5 ms each; the first ones bring the clock up
        CUDA_TRY(cudaEventRecord(cudaComputeEvents[2], cudaComputeStream));
        fp32_peak_kernel<<<blocks, threads, 0, cudaComputeStream>>>(out, iters, 1.0001f, 1e-4f);
        CUDA_TRY(cudaEventRecord(cudaComputeEvents[3], cudaComputeStream));
        CUDA_TRY(cudaEventSynchronize(cudaComputeEvents[3]));
        float ms = 0; CUDA_TRY(cudaEventElapsedTime(&ms, cudaComputeEvents[2], cudaComputeEvents[3]));
        if (rep > 1 && ms < best) best = ms;
    }
    cudaFree(out);
    const double flop = (double)threads * blocks * iters * 512.0 * 2.0;
    if (tflops) *tflops = flop / (best * 1e-3) * 1e-12;
    // the SM clock this rate corresponds to if every FP32 lane retired one FMA per cycle (128 lanes per SM)
    if (smClockMHz) *smClockMHz = flop / (best * 1e-3) / (2.0 * 128.0 * cudaComputeSmCount) * 1e-6;
    return VCRT_SUCCESS;
}

// schlick() of the device code on a host array (tests/test_gpu_parity.py compares it with the reference build's ref_schlick).
__global__ void schlick_kernel(const float* cosine, const float* ior, uint32_t count, float* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = schlick(cosine[i], ior[i]);
}
extern "C" int cudaEvaluateSchlick(IN const float* cosine, IN const float* ior, IN uint32_t count, OUT float* reflectProb) {
    if ((!cosine || !ior || !reflectProb) && count) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaEvaluateSchlick: null argument%s", "");
    int rc = ensureDevice(); if (rc) return rc;
    if (!count) return VCRT_SUCCESS;
    float* dev = nullptr;
    CUDA_TRY(cudaMalloc(&dev, sizeof(float) * 3 * (size_t)count));
    cudaError_t e = cudaMemcpyAsync(dev, cosine, sizeof(float) * count, cudaMemcpyHostToDevice, cudaComputeStream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dev + count, ior, sizeof(float) * count, cudaMemcpyHostToDevice, cudaComputeStream);
    if (e == cudaSuccess) { schlick_kernel<<<(count + 255u) / 256u, 256, 0, cudaComputeStream>>>(dev, dev + count, count, dev + 2 * (size_t)count); e = cudaGetLastError(); }
    if (e == cudaSuccess) e = cudaMemcpyAsync(reflectProb, dev + 2 * (size_t)count, sizeof(float) * count, cudaMemcpyDeviceToHost, cudaComputeStream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(cudaComputeStream);
    cudaFree(dev);
    if (e != cudaSuccess) return fail(VCRT_ERROR_CUDA, "cudaEvaluateSchlick: %s", cudaGetErrorString(e));
    return VCRT_SUCCESS;
}

static cudaEvent_t cudaComputeMarks[16];
static bool cudaComputeMarkReady[16];
static void* cudaComputeFlushScratch = nullptr;
static const size_t kFlushBytes = 256u << 20;   // 2x the 126 MB L2

extern "C" int cudaStreamMark(IN uint32_t slot) {
    if (slot >= 16) return fail(VCRT_ERROR_INVALID_ARGUMENT, "mark slot out of range%s", "");
    int rc = ensureDevice(); if (rc) return rc;
    if (!cudaComputeMarkReady[slot]) { CUDA_TRY(cudaEventCreate(&cudaComputeMarks[slot])); cudaComputeMarkReady[slot] = true; }
    CUDA_TRY(cudaEventRecord(cudaComputeMarks[slot], cudaComputeStream));
    return VCRT_SUCCESS;
}
extern "C" int cudaStreamElapsed(IN uint32_t fromSlot, IN uint32_t toSlot, OUT float* milliseconds) {
    if (fromSlot >= 16 || toSlot >= 16 || !milliseconds) return fail(VCRT_ERROR_INVALID_ARGUMENT, "bad mark arguments%s", "");
    if (!cudaComputeMarkReady[fromSlot] || !cudaComputeMarkReady[toSlot]) return fail(VCRT_ERROR_NOT_READY, "mark never recorded%s", "");
    CUDA_TRY(cudaEventSynchronize(cudaComputeMarks[toSlot]));
    CUDA_TRY(cudaEventElapsedTime(milliseconds, cudaComputeMarks[fromSlot], cudaComputeMarks[toSlot]));
    return VCRT_SUCCESS;
}
extern "C" int cudaFlushL2(void) {
    int rc = ensureDevice(); if (rc) return rc;
    if (!cudaComputeFlushScratch) CUDA_TRY(cudaMalloc(&cudaComputeFlushScratch, kFlushBytes));
    CUDA_TRY(cudaMemsetAsync(cudaComputeFlushScratch, 0x5a, kFlushBytes, cudaComputeStream));
    return VCRT_SUCCESS;
}

// ------------------------------------------------------------------ buffers
static int ensureImageBuffers(uint32_t width, uint32_t height, uint32_t chunks) {
    const size_t pixels = (size_t)width * height;
    if (width != cudaComputeImageWidth || height != cudaComputeImageHeight) {
        if (cudaComputeAccumulated) cudaFree(cudaComputeAccumulated);
        if (cudaComputeReduced) cudaFree(cudaComputeReduced);
        if (cudaComputeResultImage) cudaFree(cudaComputeResultImage);
        if (cudaComputeFixedSums) cudaFree(cudaComputeFixedSums);
        if (cudaComputeFixedReduced) cudaFree(cudaComputeFixedReduced);
        cudaComputeFixedSums = cudaComputeFixedReduced = nullptr;
        cudaComputeAccumulated = cudaComputeReduced = cudaComputeResultImage = nullptr;
        cudaComputeImageWidth = cudaComputeImageHeight = 0;
        cudaComputeImageValid = false; cudaComputeAccumulatedSamples = 0;
        CUDA_TRY(cudaMalloc(&cudaComputeAccumulated, pixels * sizeof(float4)));
        CUDA_TRY(cudaMalloc(&cudaComputeReduced, pixels * sizeof(float4)));
        CUDA_TRY(cudaMalloc(&cudaComputeResultImage, pixels * sizeof(float4)));
        CUDA_TRY(cudaMalloc(&cudaComputeFixedSums, pixels * kFixedLanes * sizeof(long long)));
        CUDA_TRY(cudaMalloc(&cudaComputeFixedReduced, pixels * kFixedLanes * sizeof(long long)));
        cudaComputeImageWidth = width; cudaComputeImageHeight = height;
    }
    const size_t need = pixels * sizeof(float4) * chunks;
    if (need > cudaComputeSumPlanesBytes) {
        if (cudaComputeSumPlanes) cudaFree(cudaComputeSumPlanes);
        cudaComputeSumPlanes = nullptr; cudaComputeSumPlanesBytes = 0;
        CUDA_TRY(cudaMalloc(&cudaComputeSumPlanes, need));
        cudaComputeSumPlanesBytes = need;
    }
    return VCRT_SUCCESS;
}

// ------------------------------------------------------------------ NCCL (loaded lazily; single-GPU use never touches it)
#include <nccl.h>
static void* ncclLibrary = nullptr;
static ncclComm_t cudaComputeComm = nullptr;
static int cudaComputeRank = 0, cudaComputeWorldSize = 1;
static ncclResult_t (*pNcclGetUniqueId)(ncclUniqueId*) = nullptr;
static ncclResult_t (*pNcclCommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
static ncclResult_t (*pNcclCommDestroy)(ncclComm_t) = nullptr;
static ncclResult_t (*pNcclReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
static ncclResult_t (*pNcclAllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
static const char* (*pNcclGetErrorString)(ncclResult_t) = nullptr;
static float* cudaComputeCommScratch = nullptr;

static int ensureNccl(void) {
    if (ncclLibrary) return VCRT_SUCCESS;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (int i = 0; i < 2 && !ncclLibrary; i++) ncclLibrary = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
    if (!ncclLibrary) return fail(VCRT_ERROR_NCCL, "cannot load libnccl: %s", dlerror());
#define NCCL_SYM(ptr, name) *(void**)(&ptr) = dlsym(ncclLibrary, name); if (!ptr) return fail(VCRT_ERROR_NCCL, "libnccl lacks %s", name)
    NCCL_SYM(pNcclGetUniqueId, "ncclGetUniqueId"); NCCL_SYM(pNcclCommInitRank, "ncclCommInitRank");
    NCCL_SYM(pNcclCommDestroy, "ncclCommDestroy"); NCCL_SYM(pNcclReduce, "ncclReduce");
    NCCL_SYM(pNcclAllReduce, "ncclAllReduce"); NCCL_SYM(pNcclGetErrorString, "ncclGetErrorString");
#undef NCCL_SYM
    return VCRT_SUCCESS;
}
#define NCCL_TRY(expr) do { ncclResult_t ncclTryResult = (expr); if (ncclTryResult != ncclSuccess) return fail(VCRT_ERROR_NCCL, #expr ": %s", pNcclGetErrorString(ncclTryResult)); } while (0)

extern "C" int cudaCommGetUniqueId(OUT void* id128) {
    if (!id128) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null id%s", "");
    int rc = ensureNccl(); if (rc) return rc;
    ncclUniqueId id; NCCL_TRY(pNcclGetUniqueId(&id));
    memcpy(id128, &id, sizeof id);   // NCCL_UNIQUE_ID_BYTES == 128
    return VCRT_SUCCESS;
}
extern "C" int cudaCommInitRank(IN const void* id128, IN int rank, IN int worldSize) {
    if (!id128 || worldSize < 1 || rank < 0 || rank >= worldSize) return fail(VCRT_ERROR_INVALID_ARGUMENT, "bad communicator arguments%s", "");
    int rc = ensureDevice(); if (rc) return rc;
    rc = ensureNccl(); if (rc) return rc;
    if (cudaComputeComm) { pNcclCommDestroy(cudaComputeComm); cudaComputeComm = nullptr; }
    ncclUniqueId id; memcpy(&id, id128, sizeof id);
    NCCL_TRY(pNcclCommInitRank(&cudaComputeComm, worldSize, id, rank));
    cudaComputeRank = rank; cudaComputeWorldSize = worldSize;
    if (!cudaComputeCommScratch) CUDA_TRY(cudaMalloc(&cudaComputeCommScratch, 64));
    return VCRT_SUCCESS;
}
extern "C" int cudaCommDestroy(void) {
    if (cudaComputeComm && pNcclCommDestroy) pNcclCommDestroy(cudaComputeComm);
    cudaComputeComm = nullptr; cudaComputeRank = 0; cudaComputeWorldSize = 1;
    return VCRT_SUCCESS;
}
extern "C" int cudaCommMaxFloat(IN float value, OUT float* maxValue) {
    if (!maxValue) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null output%s", "");
    if (!cudaComputeComm) { *maxValue = value; return VCRT_SUCCESS; }
    CUDA_TRY(cudaMemcpyAsync(cudaComputeCommScratch, &value, sizeof value, cudaMemcpyHostToDevice, cudaComputeStream));
    NCCL_TRY(pNcclAllReduce(cudaComputeCommScratch, cudaComputeCommScratch, 1, ncclFloat, ncclMax, cudaComputeComm, cudaComputeStream));
    CUDA_TRY(cudaMemcpyAsync(maxValue, cudaComputeCommScratch, sizeof value, cudaMemcpyDeviceToHost, cudaComputeStream));
    CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));
    return VCRT_SUCCESS;
}
extern "C" int cudaCommBarrier(void) { float v = 0.f, m; return cudaCommMaxFloat(v, &m); }

// ------------------------------------------------------------------ LBVH build (vcrt_lbvh.cuh)
static size_t alignUp(size_t v) { return (v + 255) & ~(size_t)255; }
static const uint32_t kMaxBvhOutliers = 16;
static const float kBvhOutlierFactor = 3.0f;

static int ensureBvh(void) {
    if (cudaComputeBvhReady) return VCRT_SUCCESS;
    const SceneBlob& hdr = cudaComputeSceneHeader;
    BvhScene& bs = cudaComputeBvhScene; memset(&bs, 0, sizeof bs);
    bs.shape = (const float4*)(cudaComputeSceneBlob + hdr.offShape); bs.surf = (const float4*)(cudaComputeSceneBlob + hdr.offSurf);
    bs.kind = (const int*)(cudaComputeSceneBlob + hdr.offKind); bs.origIndex = (const uint32_t*)(cudaComputeSceneBlob + hdr.offOrig);
    bs.scanCount = hdr.scanCount; bs.count = hdr.count;
    // Which spheres go into the tree.  Not the "always" spheres (the ground), and not the few that are much larger than the
    // rest (RTIOW: the three unit spheres among 481 of radius 0.2): a big sphere between small neighbours in Morton order
    // inflates every ancestor box, and testing it exactly for every ray is cheaper than the visits it causes (measured:
    // 9.6 -> 8.0 node visits per ray at 485 spheres, 14.4 -> 12.7 at 10 k, +4.5 % rays/s).  Outlier = radius above
    // kBvhOutlierFactor x the median radius; at most kMaxBvhOutliers of them, the largest first.
    const uint32_t scanCount = hdr.scanCount;
    struct HostIndices { uint32_t* p; ~HostIndices() { delete[] p; } };
    HostIndices treeHolder = {new uint32_t[scanCount ? scanCount : 1]}, extraHolder = {new uint32_t[kMaxBvhOutliers + (hdr.count - scanCount) + 1]};
    uint32_t* treeIndex = treeHolder.p;
    uint32_t* extraIndex = extraHolder.p;
    uint32_t n = 0, extraCount = 0;
    {
        float* radii = new float[scanCount ? scanCount : 1];
        for (uint32_t j = 0; j < scanCount; j++) radii[j] = fabsf(cudaComputeHostShape[j].w);
        float cut = INFINITY;
        if (scanCount > 2u * kMaxBvhOutliers) {
            std::nth_element(radii, radii + scanCount / 2, radii + scanCount);
            const float limit = kBvhOutlierFactor * radii[scanCount / 2];
            uint32_t above = 0;
            for (uint32_t j = 0; j < scanCount; j++) if (fabsf(cudaComputeHostShape[j].w) > limit) above++;
            cut = limit;
            if (above > kMaxBvhOutliers) {   // keep the kMaxBvhOutliers largest out: cut = the (kMaxBvhOutliers + 1)-th largest radius
                for (uint32_t j = 0; j < scanCount; j++) radii[j] = fabsf(cudaComputeHostShape[j].w);
                std::nth_element(radii, radii + kMaxBvhOutliers, radii + scanCount, [](float a, float b) { return a > b; });
                cut = radii[kMaxBvhOutliers];
            }
        }
        delete[] radii;
        for (uint32_t j = 0; j < scanCount; j++) {
            if (fabsf(cudaComputeHostShape[j].w) > cut && extraCount < kMaxBvhOutliers) extraIndex[extraCount++] = j;
            else treeIndex[n++] = j;
        }
        for (uint32_t j = scanCount; j < hdr.count; j++) extraIndex[extraCount++] = j;
    }
    SceneExtent e; memset(&e, 0, sizeof e);
    for (uint32_t i = 0; i < n; i++) {
        const float4 sp = cudaComputeHostShape[treeIndex[i]];
        const float c[3] = {sp.x, sp.y, sp.z}, r = fabsf(sp.w);
        for (int k = 0; k < 3; k++) { e.centreMin[k] = i ? fminf(e.centreMin[k], c[k]) : c[k]; e.centreMax[k] = i ? fmaxf(e.centreMax[k], c[k]) : c[k]; }
        e.radiusMin = i ? fminf(e.radiusMin, r) : r; e.radiusMax = i ? fmaxf(e.radiusMax, r) : r;
    }
    bs.treeCount = n; bs.extraCount = extraCount;
    if (extraCount > kMaxBvhExtras) return fail(VCRT_ERROR_UNSUPPORTED, "more spheres outside the tree than the constant table holds%s", "");
    if (extraCount) {   // the records the traversal kernel reads from the constant bank (vcrt_kernel_lbvh.cuh)
        LeafRecord records[kMaxBvhExtras]; memset(records, 0, sizeof records);
        const uint32_t* hostOrig = (const uint32_t*)(cudaComputeHostBlob + hdr.offOrig);
        for (uint32_t x = 0; x < extraCount; x++) { records[x].shape = cudaComputeHostShape[extraIndex[x]]; records[x].device = extraIndex[x]; records[x].orig = hostOrig[extraIndex[x]]; }
        CUDA_TRY(cudaMemcpyToSymbolAsync(cLbvhExtras, records, sizeof(LeafRecord) * extraCount, 0, cudaMemcpyHostToDevice, cudaComputeStream));
        CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));   // `records` is on the stack
    }
    bs.centreMin = make_float3(e.centreMin[0], e.centreMin[1], e.centreMin[2]); bs.centreMax = make_float3(e.centreMax[0], e.centreMax[1], e.centreMax[2]);
    bs.radiusMin = e.radiusMin; bs.radiusMax = e.radiusMax;
    bs.halfInvRadiusMin = nextafterf(0.5f / e.radiusMin, INFINITY);
    size_t sortBytes = 0;
    if (n) CUDA_TRY(cub::DeviceRadixSort::SortKeys(nullptr, sortBytes, (const unsigned long long*)nullptr, (unsigned long long*)nullptr, (int)n, 0, 64, cudaComputeStream));
    const size_t nodes = n > 1 ? n - 1 : 1;
    size_t off = 0;
    const size_t offKeysIn = off; off += alignUp(sizeof(unsigned long long) * n);
    const size_t offKeys = off; off += alignUp(sizeof(unsigned long long) * n);
    const size_t offLeafShape = off; off += alignUp(sizeof(float4) * n);
    const size_t offLeafOrig = off; off += alignUp(sizeof(uint32_t) * n);
    const size_t offLeafDevice = off; off += alignUp(sizeof(uint32_t) * n);
    const size_t offLeafRecord = off; off += alignUp(sizeof(LeafRecord) * n);
    const size_t offNodes = off; off += alignUp(sizeof(BvhNode) * nodes);
    const size_t offOctNodes = off; off += alignUp(sizeof(BvhNode) * nodes * 8);
    const size_t offParent = off; off += alignUp(sizeof(int) * (2 * (size_t)n));
    const size_t offBounds = off; off += alignUp(sizeof(float) * 6 * (2 * (size_t)n));
    const size_t offArrivals = off; off += alignUp(sizeof(unsigned int) * nodes);
    const size_t offSort = off; off += alignUp(sortBytes);
    const size_t offTreeIndex = off; off += alignUp(sizeof(uint32_t) * (n ? n : 1));
    const size_t offExtra = off; off += alignUp(sizeof(uint32_t) * (extraCount ? extraCount : 1));
    if (off > cudaComputeBvhArenaBytes) {
        if (cudaComputeBvhArena) cudaFree(cudaComputeBvhArena);
        cudaComputeBvhArena = nullptr; cudaComputeBvhArenaBytes = 0;
        CUDA_TRY(cudaMalloc(&cudaComputeBvhArena, off));
        cudaComputeBvhArenaBytes = off;
    }
    unsigned char* arena = (unsigned char*)cudaComputeBvhArena;
    if (n) CUDA_TRY(cudaMemcpyAsync(arena + offTreeIndex, treeIndex, sizeof(uint32_t) * n, cudaMemcpyHostToDevice, cudaComputeStream));
    if (extraCount) CUDA_TRY(cudaMemcpyAsync(arena + offExtra, extraIndex, sizeof(uint32_t) * extraCount, cudaMemcpyHostToDevice, cudaComputeStream));
    CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));   // the host lists go out of scope
    bs.extra = (const uint32_t*)(arena + offExtra);
    if (n == 0) { cudaComputeBvhReady = true; return VCRT_SUCCESS; }
    BvhBuildArgs g; memset(&g, 0, sizeof g);
    g.shape = bs.shape; g.origIndex = bs.origIndex; g.count = n; g.treeIndex = (const uint32_t*)(arena + offTreeIndex);
    g.boxMin = bs.centreMin;
    auto inv = [](float lo, float hi) { return hi > lo ? 1.0f / (hi - lo) : 0.0f; };
    g.boxInvExtent = make_float3(inv(e.centreMin[0], e.centreMax[0]), inv(e.centreMin[1], e.centreMax[1]), inv(e.centreMin[2], e.centreMax[2]));
    unsigned long long* keysIn = (unsigned long long*)(arena + offKeysIn);
    g.keys = keysIn;
    g.leafShape = (float4*)(arena + offLeafShape); g.leafOrig = (uint32_t*)(arena + offLeafOrig); g.leafDevice = (uint32_t*)(arena + offLeafDevice);
    g.leafRecord = (LeafRecord*)(arena + offLeafRecord);
    g.nodes = (BvhNode*)(arena + offNodes); g.octNodes = (BvhNode*)(arena + offOctNodes); g.parent = (int*)(arena + offParent); g.bounds = (float*)(arena + offBounds);
    g.arrivals = (unsigned int*)(arena + offArrivals);
    const unsigned int blocks = (n + 255) / 256;
    lbvh_keys_kernel<<<blocks, 256, 0, cudaComputeStream>>>(g);
    CUDA_TRY(cub::DeviceRadixSort::SortKeys(arena + offSort, sortBytes, keysIn, (unsigned long long*)(arena + offKeys), (int)n, 0, 64, cudaComputeStream));
    g.keys = (unsigned long long*)(arena + offKeys);
    CUDA_TRY(cudaMemsetAsync(g.arrivals, 0, sizeof(unsigned int) * nodes, cudaComputeStream));
    lbvh_leaves_kernel<<<blocks, 256, 0, cudaComputeStream>>>(g);
    if (n > 1) {
        lbvh_hierarchy_kernel<<<blocks, 256, 0, cudaComputeStream>>>(g);
        lbvh_refit_kernel<<<blocks, 256, 0, cudaComputeStream>>>(g);
    }
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));
    bs.nodes = g.nodes; bs.octNodes = g.octNodes; bs.leafShape = g.leafShape; bs.leafOrig = g.leafOrig; bs.leafDevice = g.leafDevice; bs.leafRecord = g.leafRecord;
    cudaComputeBvhReady = true;
    return VCRT_SUCCESS;
}

template <bool FIXED>
static int launchPrimaryScan(const RenderArgs& args, uint32_t* launches) {
    const size_t smem = primary_scan_smem(cudaComputeSceneHeader.bytes);
    auto kernel = args.semantics ? primary_scan_kernel<FIXED, true> : primary_scan_kernel<FIXED, false>;
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int perSm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kPrimaryThreads, smem));
    if (perSm < 1) perSm = 1;
    kernel<<<(unsigned int)(cudaComputeSmCount * perSm), kPrimaryThreads, smem, cudaComputeStream>>>(cudaComputeSceneBlob, args, (int)cudaComputeSceneHeader.scanPad);
    CUDA_TRY(cudaGetLastError());
    *launches += 1;
    return VCRT_SUCCESS;
}

template <bool FIXED, bool STATS, bool FLAT>
static int launchLbvh(const RenderArgs& args, uint32_t* launches) {
#ifdef VCRT_EXPERIMENTS
    auto kernel = FLAT ? (args.semantics ? render_lbvh_flat_kernel<FIXED, STATS, true> : render_lbvh_flat_kernel<FIXED, STATS, false>) : render_lbvh_kernel<FIXED, STATS>;
#else
    static_assert(FLAT, "the first-generation LBVH kernel is an experiment (tools/experiments)");
    auto kernel = args.semantics ? render_lbvh_flat_kernel<FIXED, STATS, true> : render_lbvh_flat_kernel<FIXED, STATS, false>;
#endif
    int perSm = 0;
    CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, 256, 0));
    if (perSm < 1) perSm = 1;
    const unsigned int zero = 0;
    CUDA_TRY(cudaMemcpyAsync(cudaComputeWorkCounter, &zero, sizeof zero, cudaMemcpyHostToDevice, cudaComputeStream));
    kernel<<<(unsigned int)(cudaComputeSmCount * perSm), 256, 0, cudaComputeStream>>>(cudaComputeBvhScene, args);
    CUDA_TRY(cudaGetLastError());
    *launches += 1;
    return VCRT_SUCCESS;
}

// ------------------------------------------------------------------ render dispatch
template <int THREADS, int PAIRS, int CAP, bool FIXED>
static int launchScan(const RenderArgs& args, uint32_t* launches) {
    using Cfg = ScanConfig<THREADS, PAIRS, CAP>;
    const size_t smem = Cfg::smem_bytes(cudaComputeSceneHeader.bytes);
    auto kernel = render_scan_kernel<THREADS, PAIRS, CAP, FIXED>;
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned int grid = (unsigned int)cudaComputeSmCount;   // persistent: one CTA per SM
    const unsigned int firstDynamicItem = grid * Cfg::R * THREADS;
    CUDA_TRY(cudaMemcpyAsync(cudaComputeWorkCounter, &firstDynamicItem, sizeof firstDynamicItem, cudaMemcpyHostToDevice, cudaComputeStream));
    kernel<<<grid, THREADS, smem, cudaComputeStream>>>(cudaComputeSceneBlob, args);
    CUDA_TRY(cudaGetLastError());
    *launches += 1;
    return VCRT_SUCCESS;
}

template <int THREADS, int PAIRS, int LCAP, bool FIXED>
static int launchScanU(const RenderArgs& args, uint32_t* launches) {
    using Cfg = ScanUConfig<THREADS, PAIRS, LCAP>;
    const size_t smem = Cfg::smem_bytes(cudaComputeSceneHeader.bytes);
    auto kernel = args.semantics ? render_scanu_kernel<THREADS, PAIRS, LCAP, FIXED, true> : render_scanu_kernel<THREADS, PAIRS, LCAP, FIXED, false>;
    if constexpr (FIXED)   // after a primary-visibility pass the samples come from its compact list
        if (args.primary.cursors) kernel = args.semantics ? render_scanu_kernel<THREADS, PAIRS, LCAP, true, true, true> : render_scanu_kernel<THREADS, PAIRS, LCAP, true, false, true>;
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned int grid = (unsigned int)cudaComputeSmCount;   // persistent: one CTA per SM
    const unsigned int firstDynamicItem = grid * Cfg::R * THREADS;
    CUDA_TRY(cudaMemcpyAsync(cudaComputeWorkCounter, &firstDynamicItem, sizeof firstDynamicItem, cudaMemcpyHostToDevice, cudaComputeStream));
    kernel<<<grid, THREADS, smem, cudaComputeStream>>>(cudaComputeSceneBlob, args, (int)cudaComputeSceneHeader.scanPad);
    CUDA_TRY(cudaGetLastError());
    *launches += 1;
    return VCRT_SUCCESS;
}
#ifdef VCRT_EXPERIMENTS
template <int THREADS, int PAIRS, int LCAP, bool FIXED>
static int launchScanW(const RenderArgs& args, uint32_t* launches) {
    using Cfg = ScanWConfig<THREADS, PAIRS, LCAP>;
    const size_t smem = Cfg::smem_bytes(cudaComputeSceneHeader.bytes);
    auto kernel = render_scanw_kernel<THREADS, PAIRS, LCAP, FIXED>;
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned int grid = (unsigned int)cudaComputeSmCount;   // persistent: one CTA per SM
    const unsigned int firstDynamicItem = grid * Cfg::R * THREADS;
    CUDA_TRY(cudaMemcpyAsync(cudaComputeWorkCounter, &firstDynamicItem, sizeof firstDynamicItem, cudaMemcpyHostToDevice, cudaComputeStream));
    kernel<<<grid, THREADS, smem, cudaComputeStream>>>(cudaComputeSceneBlob, args, (int)cudaComputeSceneHeader.scanPad);
    CUDA_TRY(cudaGetLastError());
    *launches += 1;
    return VCRT_SUCCESS;
}
static bool scanWFits(int threads, int pairs, int lcap) {
    size_t state = (size_t)WF_COUNT * 2 * pairs * threads * 4, lists = (size_t)lcap * threads * 2, map = (size_t)2 * pairs * threads;
    return cudaComputeSceneInConstant && 16 + (size_t)cudaComputeSceneHeader.bytes + state + lists + map <= 227 * 1024;
}
#endif
static bool scanUFits(int threads, int pairs, int lcap) {
    size_t state = (size_t)SF_COUNT * 2 * pairs * threads * 4, lists = (size_t)lcap * threads * 2;
    return cudaComputeSceneInConstant && 16 + (size_t)cudaComputeSceneHeader.bytes + state + lists <= 227 * 1024;
}

static const size_t kSmemBudget = 227 * 1024;
static const uint32_t kPrimaryMinSpheres = 32;   // the primary-visibility pass pays from ~40 scan spheres on (8: 9.4 ms with, 7.4 ms without; 40: 9.8 / 9.9; 485: 38.8 / 46.0 at 1080p x 64 spp)
static bool scanFits(int threads, int pairs, int cap) {
    size_t state = (size_t)SF_COUNT * 2 * pairs * threads * 4, lists = (size_t)(cap + 2) * 2 * pairs * threads * 2;
    return 16 + (size_t)cudaComputeSceneHeader.bytes + state + lists <= kSmemBudget && cudaComputeSceneHeader.scanPad < 65536u;
}

extern "C" int cudaRender(IN const VcrtRenderParams* p) {
    if (!p) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRender: null params%s", "");
    // A rank of a multi-GPU job may have no sample of its own (total spp below the rank count: the reference's default is 1 spp,
    // globals.glsl:12).  It still has to join the reduce, or the other ranks wait for ever: 0 samples is valid there.
    const bool joinsReduce = cudaComputeComm && cudaComputeWorldSize > 1 && !(p->flags & VCRT_FLAG_NO_REDUCE);
    if (!p->width || !p->height || (!p->samplesPerPixel && !joinsReduce) || p->width > 65535u || p->height > 65535u || p->samplesPerPixel >= (1u << 24))
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRender: width/height must be 1..65535 and samplesPerPixel 1..2^24-1 (0 only for a rank that just joins the reduce)%s", "");
#ifndef VCRT_EXPERIMENTS
    if (p->reserved[0] || p->reserved[1]) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRender: reserved fields must be 0%s", "");
#endif
    int rc = ensureDevice(); if (rc) return rc;
    if (!cudaComputeSceneReady) return fail(VCRT_ERROR_NOT_READY, "cudaRender before cudaUploadScene%s", "");

    uint32_t traversal = p->traversal;
    // AUTO = the fastest, and that is the LBVH kernel at every scene size (tools/crossover.py, 1080p x 64 spp, LBVH / brute:
    // 2 spheres 4.5 / 7.4 ms, 8: 6.4 / 7.4, 40: 7.2 / 9.8, 100: 9.0 / 13.9, 485: 14.6 / 38.8): with nothing in the tree it is a
    // path tracer over the few spheres it tests directly, without the scan kernel's slot machinery
    if (traversal == VCRT_TRAVERSAL_AUTO) traversal = VCRT_TRAVERSAL_LBVH;
    // A NaN centre or radius makes hit_sphere accept the sphere with t = NaN (every comparison of functions.glsl:20-31 is false),
    // after which the reference's result depends on the ORDER of the scan; an infinite one poisons the filter / the boxes.
    // Such scenes are rendered by the in-order exhaustive kernel, whatever was asked for.
    if (cudaComputeSceneNonFinite && (traversal == VCRT_TRAVERSAL_BRUTE || traversal == VCRT_TRAVERSAL_LBVH)) traversal = VCRT_TRAVERSAL_SIMPLE;
    if (traversal == VCRT_TRAVERSAL_BRUTE && !scanFits(256, 2, 12) && !scanUFits(512, 2, 30))
        return fail(VCRT_ERROR_UNSUPPORTED, "scene does not fit shared memory for VCRT_TRAVERSAL_BRUTE%s", "");
    if (traversal > VCRT_TRAVERSAL_SIMPLE) return fail(VCRT_ERROR_INVALID_ARGUMENT, "unknown traversal%s", "");

#ifdef VCRT_EXPERIMENTS
    const bool wide = scanFits(512, 2, 12);
#endif
    const uint32_t pixels = p->width * p->height;
    // Accumulation mode.  sampleChunks == 0 (default, brute force only): 32.32 fixed-point sums, every sample added
    // with an integer RED, work items of a few samples each -> good load balance and bits that do not depend on the
    // schedule or the GPU count.  sampleChunks >= 1: float sums in strict sample order per chunk (1 = the oracle's order).
    uint32_t chunks = 1;
    bool fixedMode = false;
    if (traversal == VCRT_TRAVERSAL_BRUTE || traversal == VCRT_TRAVERSAL_LBVH) {
        chunks = p->sampleChunks;
        // the fixed-point accumulator needs radiance in [0, 1] per sample; a scene with an attenuation outside [0, 1] (or NaN)
        // gets float sums in strict sample order instead, where inf / NaN propagate as they do in the reference
        fixedMode = chunks == 0 && cudaComputeSceneBounded;
        if (chunks > p->samplesPerPixel) chunks = p->samplesPerPixel;
        if (chunks < 1u) chunks = 1u;
    }
    if (traversal == VCRT_TRAVERSAL_LBVH) { rc = ensureBvh(); if (rc) return rc; }
    // Work items.  Fixed-point mode: 1 sample per item (measured fastest: it keeps the lanes of a warp on neighbouring pixels;
    // 2 per item cost the brute-force kernel 6 % and the LBVH kernel 20 %), and where the 32-bit item numbering cannot hold
    // pixels x samples (4K x 520 spp is already 2^32) the call is split into several launches over sample sub-ranges - the
    // integer sums do not care.  Strict modes: one item per pixel and chunk, one launch.
    const uint64_t tilesAll = (uint64_t)((p->width + 7u) / 8u) * ((p->height + 3u) / 4u);
    if (tilesAll * 32u >= (1ull << 31)) return fail(VCRT_ERROR_INVALID_ARGUMENT, "image too large%s", "");
    const uint64_t kMaxItems = (1ull << 32) - (1ull << 24);
    auto planFor = [&](uint32_t spp, uint32_t perRun, ItemPlan& plan) -> bool {
        memset(&plan, 0, sizeof plan);
        plan.tilesX = (p->width + 7u) / 8u; plan.pixelsPadded = (uint32_t)(tilesAll * 32u);
        if (perRun < 1u) perRun = 1u;
        plan.sppPerRun = perRun; plan.runsPerPixel = (spp + perRun - 1u) / perRun;
        if ((uint64_t)plan.pixelsPadded * plan.runsPerPixel > kMaxItems || plan.runsPerPixel >= (1u << 22)) return false;
        plan.totalItems = plan.pixelsPadded * plan.runsPerPixel;
        const uint32_t fullBands = plan.pixelsPadded >> VCRT_BAND_SHIFT;
        plan.fullBandRuns = fullBands * plan.runsPerPixel;
        plan.lastBandFirst = fullBands << VCRT_BAND_SHIFT; plan.lastBandPixels = plan.pixelsPadded - plan.lastBandFirst;
        plan.invRuns = plan.runsPerPixel ? nextafterf(1.0f / (float)plan.runsPerPixel, 0.f) : 0.f;
        plan.invLastBand = plan.lastBandPixels ? nextafterf(1.0f / (float)plan.lastBandPixels, 0.f) : 0.f;
        plan.tiles = plan.pixelsPadded / 32u; plan.invTiles = nextafterf(1.0f / (float)plan.tiles, 0.f);
        plan.invTilesX = nextafterf(1.0f / (float)plan.tilesX, 0.f);
        plan.invWidth = nextafterf(1.0f / (float)p->width, 0.f);
        plan.tileStride = 1u;
        if (traversal == VCRT_TRAVERSAL_BRUTE && fullBands >= 1u) {
            uint32_t stride = plan.tiles / (kBandPixels / 32u);
            if (stride > 31u) stride = 31u;                       // (tile index * stride) stays below 2^31
            auto gcd = [](uint32_t a, uint32_t b) { while (b) { uint32_t t = a % b; a = b; b = t; } return a; };
            while (stride > 1u && gcd(stride, plan.tiles) != 1u) stride--;
            plan.tileStride = stride;
        }
        return true;
    };
    const bool traces = p->maxDepth != 0 && p->samplesPerPixel != 0;   // otherwise no kernel runs and the sums of this call are zero
    // Primary visibility (vcrt_kernel_primary.cuh), brute force only: the camera rays of a launch and their first scatter events
    // are done by a kernel of their own and handed to the path kernel through a buffer of 32 bytes per sample.  VCRT_NO_PRIMARY=1
    // in the environment switches it off (A/B measurements); the first-generation brute-force kernel (scenes beyond ~1200
    // spheres) does not use it, and for the LBVH it was measured to lose.
    static const bool primaryDisabled = getenv("VCRT_NO_PRIMARY") && atoi(getenv("VCRT_NO_PRIMARY")) != 0;
    bool usePrimary = traces && !primaryDisabled && traversal == VCRT_TRAVERSAL_BRUTE && scanUFits(512, 2, 30) && cudaComputeSceneHeader.scanCount >= kPrimaryMinSpheres &&
                      primary_scan_smem(cudaComputeSceneHeader.bytes) <= kSmemBudget;
#ifdef VCRT_EXPERIMENTS
    if (p->reserved[1] != 0 && p->reserved[1] != 10 && p->reserved[1] != 12) usePrimary = false;   // only the product kernels read the buffer
#endif
    uint32_t passes = 1u, perRunFixed = 1u;
    if (usePrimary) {
        const uint64_t entries = tilesAll * 32u * p->samplesPerPixel;
        if (fixedMode) passes = (uint32_t)((entries + kPrimaryMaxEntries - 1) / kPrimaryMaxEntries);
        else if (entries > kPrimaryMaxEntries) usePrimary = false;
    }
    if (fixedMode) {
#ifdef VCRT_EXPERIMENTS
        if (p->reserved[0]) perRunFixed = p->reserved[0];
#endif
        const uint64_t items = tilesAll * 32u * ((p->samplesPerPixel + perRunFixed - 1u) / perRunFixed);
        const uint32_t byItems = (uint32_t)((items + kMaxItems - 1) / kMaxItems);
        if (passes < byItems) passes = byItems;
        const uint32_t byRuns = (p->samplesPerPixel / perRunFixed) / ((1u << 22) - 2u) + 1u;   // runs per pixel and launch stay below 2^22
        if (passes < byRuns) passes = byRuns;
        if (usePrimary) {   // the compact hand-over list keeps the sample index of a launch in 21 bits
            const uint32_t bySamples = p->samplesPerPixel / (kPrimaryMaxLaunchSamples - 1u) + 1u;
            if (passes < bySamples) passes = bySamples;
        }
        if (passes < 1u) passes = 1u;
    }
    ItemPlan plan;
    if (!planFor(fixedMode ? (p->samplesPerPixel + passes - 1u) / passes : p->samplesPerPixel,
                 fixedMode ? perRunFixed : (p->samplesPerPixel + chunks - 1u) / chunks, plan))
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "too many work items: lower sampleChunks / samplesPerPixel per call%s", "");
    // Float modes: the kernels write one plane per run, and ceil(spp / ceil(spp / chunks)) can be smaller than the chunk count
    // asked for (7 samples in 5 chunks = 4 runs of 2): only the planes that are written may be summed.
    if (!fixedMode) chunks = plan.runsPerPixel;
    rc = ensureImageBuffers(p->width, p->height, chunks ? chunks : 1u); if (rc) return rc;
    if (usePrimary) {
        const uint32_t passSppMax = fixedMode ? (p->samplesPerPixel + passes - 1u) / passes : p->samplesPerPixel;
        const size_t need = (size_t)plan.pixelsPadded * passSppMax;
        if ((uint64_t)plan.tiles * passSppMax >= (1ull << 31)) usePrimary = false;
        else if (need > cudaComputePrimaryCapacity) {
            if (cudaComputePrimary) cudaFree(cudaComputePrimary);
            cudaComputePrimary = nullptr; cudaComputePrimaryCapacity = 0;
            if (cudaMalloc(&cudaComputePrimary, need * 2 * sizeof(float4)) == cudaSuccess) cudaComputePrimaryCapacity = need;
            else { cudaGetLastError(); cudaComputePrimary = nullptr; usePrimary = false; }   // no room: the path kernels trace their primary rays themselves
        }
    }
    const bool keep = (p->flags & VCRT_FLAG_ACCUMULATE) && cudaComputeAccumulatedSamples > 0;
    if (keep && fixedMode != cudaComputeAccumulatedIsFixed)
        return fail(VCRT_ERROR_UNSUPPORTED, "VCRT_FLAG_ACCUMULATE across different accumulation modes%s", "");

    RenderArgs args; memset(&args, 0, sizeof args);
    if (p->semantics > VCRT_SEMANTICS_BOOK) return fail(VCRT_ERROR_INVALID_ARGUMENT, "unknown semantics%s", "");
    const bool book = p->semantics == VCRT_SEMANTICS_BOOK;
    args.frame = makeFrame(cudaComputeCamera, p->width, p->height, book);
    args.semantics = p->semantics;
    args.width = p->width; args.height = p->height; args.spp = p->samplesPerPixel; args.sampleOffset = p->sampleOffset;
    args.maxDepth = p->maxDepth; args.seedLo = (uint32_t)p->seed; args.seedHi = (uint32_t)(p->seed >> 32);
    args.chunks = chunks; args.flags = p->flags; args.sums = cudaComputeSumPlanes;
    args.fixedSums = cudaComputeFixedSums; args.plan = plan;
    args.counters = cudaComputeCounters; args.workCounter = cudaComputeWorkCounter;
    // a thread of the primary kernel traces kPrimaryRays rays: S samples x kPrimaryRays / S tiles, S the power of two that leaves
    // the fewest sample slots empty over the launch (the larger one on a tie: the rays of one pixel share their candidates)
    args.primarySamples = 1u;
    {
        const uint32_t sppMax = fixedMode ? (p->samplesPerPixel + passes - 1u) / passes : p->samplesPerPixel;
        uint32_t bestWaste = 0xffffffffu;
        for (uint32_t sPer = (uint32_t)kPrimaryRays; sPer >= 1u; sPer >>= 1) {
            const uint32_t waste = (sppMax + sPer - 1u) / sPer * sPer - sppMax;
            if (waste < bestWaste) { bestWaste = waste; args.primarySamples = sPer; }
        }
    }
    args.primary.a = usePrimary ? cudaComputePrimary : nullptr;
    args.primary.b = usePrimary ? cudaComputePrimary + cudaComputePrimaryCapacity : nullptr;
    args.primary.entries = usePrimary ? cudaComputePrimaryCapacity : 0;
    args.primary.cursors = usePrimary && fixedMode ? cudaComputeWorkCounter + 2 : nullptr;   // fixed-point sums: the hand-over is a compact list

    uint32_t launches = 0, timedPasses = 0;
    CUDA_TRY(cudaEventRecord(cudaComputeEvents[0], cudaComputeStream));
    CUDA_TRY(cudaMemsetAsync(cudaComputeCounters, 0, sizeof(unsigned long long) * CTR_COUNT, cudaComputeStream));
    CUDA_TRY(cudaEventRecord(cudaComputeEvents[1], cudaComputeStream));
    if (fixedMode && !keep) CUDA_TRY(cudaMemsetAsync(cudaComputeFixedSums, 0, (size_t)pixels * kFixedLanes * sizeof(long long), cudaComputeStream));
    if (!traces) {   // depth 0: no pass of functions.glsl:73 runs, every sample is black (bug B7); 0 samples: nothing to add
        CUDA_TRY(cudaMemsetAsync(cudaComputeSumPlanes, 0, (size_t)pixels * sizeof(float4) * chunks, cudaComputeStream));
    } else if (traversal == VCRT_TRAVERSAL_SIMPLE) {
        DeviceScene scene; memset(&scene, 0, sizeof scene);
        scene.shape = (const float4*)(cudaComputeSceneBlob + cudaComputeSceneHeader.offShape);
        scene.surf = (const float4*)(cudaComputeSceneBlob + cudaComputeSceneHeader.offSurf);
        scene.kind = (const int*)(cudaComputeSceneBlob + cudaComputeSceneHeader.offKind);
        scene.count = (int)cudaComputeSceneHeader.count;
        dim3 grid((p->width + 15) / 16, (p->height + 7) / 8);
        render_simple_kernel<<<grid, 128, 0, cudaComputeStream>>>(scene, args);
        CUDA_TRY(cudaGetLastError());
        launches++;
    } else for (uint32_t pass = 0, done = 0; pass < passes; pass++) {
        // one launch per sample sub-range (more than one only in fixed-point mode, see the work-item plan above)
        const uint32_t passSpp = p->samplesPerPixel / passes + (pass < p->samplesPerPixel % passes ? 1u : 0u);
        if (!passSpp) continue;
        if (passes > 1u) {
            args.spp = passSpp; args.sampleOffset = p->sampleOffset + done;
            if (!planFor(passSpp, perRunFixed, args.plan)) return fail(VCRT_ERROR_INVALID_ARGUMENT, "too many work items%s", "");
        }
        done += passSpp;
        if (usePrimary) {
            const bool timed = pass < kMaxTimedPasses;
            if (timed) {
                while (cudaComputePassEventCount < 2 * (pass + 1)) { CUDA_TRY(cudaEventCreate(&cudaComputePassEvents[cudaComputePassEventCount])); cudaComputePassEventCount++; }
                CUDA_TRY(cudaEventRecord(cudaComputePassEvents[2 * pass], cudaComputeStream));
            }
            if (args.primary.cursors) CUDA_TRY(cudaMemsetAsync(args.primary.cursors, 0, 2 * sizeof(unsigned int), cudaComputeStream));
            rc = fixedMode ? launchPrimaryScan<true>(args, &launches) : launchPrimaryScan<false>(args, &launches);
            if (rc) return rc;
            if (timed) { CUDA_TRY(cudaEventRecord(cudaComputePassEvents[2 * pass + 1], cudaComputeStream)); timedPasses = pass + 1; }
        }
      if (traversal == VCRT_TRAVERSAL_LBVH) {
        const bool stats = (args.flags & kFlagCountStats) != 0;
#ifdef VCRT_EXPERIMENTS
        if (p->reserved[1] == 31)   // the first-generation nested-loop kernel
            rc = stats ? (fixedMode ? launchLbvh<true, true, false>(args, &launches) : launchLbvh<false, true, false>(args, &launches))
                       : (fixedMode ? launchLbvh<true, false, false>(args, &launches) : launchLbvh<false, false, false>(args, &launches));
        else
#endif
        rc = stats ? (fixedMode ? launchLbvh<true, true, true>(args, &launches) : launchLbvh<false, true, true>(args, &launches))
                   : (fixedMode ? launchLbvh<true, false, true>(args, &launches) : launchLbvh<false, false, true>(args, &launches));
        if (rc) return rc;
      } else {
        // 1024 threads x 2 rays (32 warps per SM hide the shading phases best; measured 94 ms vs 102 ms for 512 x 4 at
        // 1080p x 128 spp) up to ~700 spheres, 512 x 4 up to ~1200 as the scene blob grows, then the first-generation kernel
        // with 256 x 4 up to ~2400 spheres; beyond that VCRT_TRAVERSAL_BRUTE is refused (the LBVH is what large scenes are for)
        bool launched = false;
#ifdef VCRT_EXPERIMENTS
        const uint32_t shape = p->reserved[1];   // CTA shape / kernel generation for A/B runs (tools/experiments/README.md)
        launched = true;
        if (shape == 20 && scanWFits(1024, 1, 24)) rc = fixedMode ? launchScanW<1024, 1, 24, true>(args, &launches) : launchScanW<1024, 1, 24, false>(args, &launches);
        else if (shape == 21 && scanWFits(512, 2, 24)) rc = fixedMode ? launchScanW<512, 2, 24, true>(args, &launches) : launchScanW<512, 2, 24, false>(args, &launches);
        else if (shape == 10 && scanUFits(512, 2, 30)) rc = fixedMode ? launchScanU<512, 2, 30, true>(args, &launches) : launchScanU<512, 2, 30, false>(args, &launches);
        else if (shape == 12 && fixedMode && scanUFits(256, 4, 30)) rc = launchScanU<256, 4, 30, true>(args, &launches);
        else if (shape == 1) rc = fixedMode ? (wide ? launchScan<512, 2, 12, true>(args, &launches) : launchScan<256, 2, 12, true>(args, &launches))
                                            : (wide ? launchScan<512, 2, 12, false>(args, &launches) : launchScan<256, 2, 12, false>(args, &launches));
        else if (shape == 2 && fixedMode && scanFits(1024, 1, 12)) rc = launchScan<1024, 1, 12, true>(args, &launches);
        else if (shape == 3 && fixedMode && scanFits(768, 1, 12)) rc = launchScan<768, 1, 12, true>(args, &launches);
        else if (shape == 4 && fixedMode && scanFits(640, 1, 12)) rc = launchScan<640, 1, 12, true>(args, &launches);
        else launched = false;
#endif
        if (launched) { }
        else if (scanUFits(1024, 1, 30)) rc = fixedMode ? launchScanU<1024, 1, 30, true>(args, &launches) : launchScanU<1024, 1, 30, false>(args, &launches);
        else if (scanUFits(512, 2, 30)) rc = fixedMode ? launchScanU<512, 2, 30, true>(args, &launches) : launchScanU<512, 2, 30, false>(args, &launches);
        else rc = fixedMode ? launchScan<256, 2, 12, true>(args, &launches) : launchScan<256, 2, 12, false>(args, &launches);
        if (rc) return rc;
      }
    }
    CUDA_TRY(cudaEventRecord(cudaComputeEvents[2], cudaComputeStream));

    const unsigned int pixelBlocks = (pixels + 255) / 256;
    if (fixedMode) fixed_to_float_kernel<<<pixelBlocks, 256, 0, cudaComputeStream>>>(cudaComputeAccumulated, cudaComputeFixedSums, pixels);
    else combine_planes_kernel<<<pixelBlocks, 256, 0, cudaComputeStream>>>(cudaComputeAccumulated, cudaComputeSumPlanes, pixels, chunks, keep ? 1 : 0);
    launches++;
    const bool fixedSums = fixedMode;
    cudaComputeAccumulatedSamples = (keep ? cudaComputeAccumulatedSamples : 0) + p->samplesPerPixel;
    cudaComputeAccumulatedIsFixed = fixedMode;

    const float4* finalSums = cudaComputeAccumulated;
    const long long* finalFixed = cudaComputeFixedSums;
    uint64_t totalSamples = cudaComputeAccumulatedSamples;
    bool haveImage = true;
    if (cudaComputeComm && cudaComputeWorldSize > 1 && !(p->flags & VCRT_FLAG_NO_REDUCE)) {
        // The one exchange of the path: sum the per-GPU accumulation buffers over NVLink.
        const void* src = fixedSums ? (const void*)cudaComputeFixedSums : (const void*)cudaComputeAccumulated;
        void* dst = fixedSums ? (void*)cudaComputeFixedReduced : (void*)cudaComputeReduced;
        const ncclDataType_t type = fixedSums ? ncclInt64 : ncclFloat;
        const size_t elements = (size_t)pixels * (fixedSums ? (size_t)kFixedLanes : 4u);   // C3: 199 MB of int64 sums (float4 sums: 133 MB)
        if (p->flags & VCRT_FLAG_ALL_REDUCE) NCCL_TRY(pNcclAllReduce(src, dst, elements, type, ncclSum, cudaComputeComm, cudaComputeStream));
        else NCCL_TRY(pNcclReduce(src, dst, elements, type, ncclSum, 0, cudaComputeComm, cudaComputeStream));
        finalSums = cudaComputeReduced; finalFixed = cudaComputeFixedReduced;
        totalSamples *= (uint64_t)cudaComputeWorldSize;   // equal shards unless the caller says otherwise
        haveImage = (p->flags & VCRT_FLAG_ALL_REDUCE) || cudaComputeRank == 0;
    }
    if (p->totalSamples) totalSamples = p->totalSamples;
    if (haveImage) {
        if (fixedSums) normalise_fixed_kernel<<<pixelBlocks, 256, 0, cudaComputeStream>>>(cudaComputeResultImage, finalFixed, pixels, (double)totalSamples, book ? 1 : 0);
        else normalise_kernel<<<pixelBlocks, 256, 0, cudaComputeStream>>>(cudaComputeResultImage, finalSums, pixels, (float)totalSamples, book ? 1 : 0);
        launches++;
    }
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaEventRecord(cudaComputeEvents[3], cudaComputeStream));
    CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));   // the frame is complete when we return (platform/Linux.cpp:365)
    cudaComputeImageValid = haveImage;

    unsigned long long ctr[CTR_COUNT];
    CUDA_TRY(cudaMemcpy(ctr, cudaComputeCounters, sizeof ctr, cudaMemcpyDeviceToHost));
    VcrtRenderStats& st = cudaComputeLastStats; memset(&st, 0, sizeof st);
    st.samples = (uint64_t)pixels * p->samplesPerPixel;
    st.rays = ctr[CTR_RAYS]; st.depthExhausted = ctr[CTR_EXHAUSTED]; st.badMaterial = ctr[CTR_BAD_MATERIAL];
    if (p->maxDepth == 0) st.depthExhausted = st.samples;   // no kernel ran: every sample ends at the (zero) depth limit
    st.candidates = ctr[CTR_CANDIDATES]; st.listOverflows = ctr[CTR_OVERFLOWS];
    st.sphereTests = traversal == VCRT_TRAVERSAL_LBVH ? ctr[CTR_TESTS] : st.rays * cudaComputeSceneHeader.count; st.nodeVisits = ctr[CTR_NODES];
    for (int i = 0; i < 4; i++) st.hitsByMaterial[i] = ctr[CTR_HITS0 + i];
    for (int i = 0; i < 8; i++) st.pathLength[i] = ctr[CTR_PATH0 + i];
    CUDA_TRY(cudaEventElapsedTime(&st.renderKernelMs, cudaComputeEvents[1], cudaComputeEvents[2]));
    CUDA_TRY(cudaEventElapsedTime(&st.totalMs, cudaComputeEvents[0], cudaComputeEvents[3]));
    st.kernelLaunches = launches; st.traversalUsed = traversal; st.sampleChunksUsed = chunks;
    st.primaryRays = ctr[CTR_PRIMARY_RAYS]; st.renderPasses = traces && (traversal == VCRT_TRAVERSAL_BRUTE || traversal == VCRT_TRAVERSAL_LBVH) ? passes : (traces ? 1u : 0u);
    for (uint32_t i = 0; i < timedPasses; i++) { float ms = 0.f; CUDA_TRY(cudaEventElapsedTime(&ms, cudaComputePassEvents[2 * i], cudaComputePassEvents[2 * i + 1])); st.primaryKernelMs += ms; }
    st.checkFailures = (uint32_t)(ctr[CTR_CHECK_FAILED] > 0xffffffffull ? 0xffffffffull : ctr[CTR_CHECK_FAILED]);
    if (st.badMaterial) return fail(VCRT_ERROR_SCENE, "scene has texture ids outside 1..3 (hits on them were absorbed)%s", "");
    return VCRT_SUCCESS;
}

extern "C" int cudaGetRenderStats(OUT VcrtRenderStats* stats) {
    if (!stats) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null stats%s", "");
    *stats = cudaComputeLastStats;
    return VCRT_SUCCESS;
}

static int downloadBuffer(const float4* src, bool valid, float* rgba, size_t bytes) {
    if (!rgba) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null destination%s", "");
    if (!valid || !src) return fail(VCRT_ERROR_NOT_READY, "no rendered image on this device%s", "");
    const size_t need = (size_t)cudaComputeImageWidth * cudaComputeImageHeight * sizeof(float4);
    if (bytes < need) return fail(VCRT_ERROR_INVALID_ARGUMENT, "destination smaller than width*height*16 bytes%s", "");
    CUDA_TRY(cudaMemcpyAsync(rgba, src, need, cudaMemcpyDeviceToHost, cudaComputeStream));
    CUDA_TRY(cudaStreamSynchronize(cudaComputeStream));
    return VCRT_SUCCESS;
}
extern "C" int cudaDownloadImage(OUT float* rgba, IN size_t bytes) { return downloadBuffer(cudaComputeResultImage, cudaComputeImageValid, rgba, bytes); }
extern "C" int cudaDownloadSums(OUT float* rgba, IN size_t bytes) { return downloadBuffer(cudaComputeAccumulated, cudaComputeAccumulatedSamples > 0, rgba, bytes); }

extern "C" int cudaGetDeviceImage(OUT void** devicePointer, OUT size_t* bytes) {
    if (!devicePointer || !bytes) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null output%s", "");
    if (!cudaComputeImageValid) return fail(VCRT_ERROR_NOT_READY, "no rendered image on this device%s", "");
    *devicePointer = cudaComputeResultImage; *bytes = (size_t)cudaComputeImageWidth * cudaComputeImageHeight * sizeof(float4);
    return VCRT_SUCCESS;
}
extern "C" int cudaGetDeviceSums(OUT void** devicePointer, OUT size_t* bytes) {
    if (!devicePointer || !bytes) return fail(VCRT_ERROR_INVALID_ARGUMENT, "null output%s", "");
    if (!cudaComputeAccumulated || !cudaComputeAccumulatedSamples) return fail(VCRT_ERROR_NOT_READY, "nothing rendered yet%s", "");
    *devicePointer = cudaComputeAccumulated; *bytes = (size_t)cudaComputeImageWidth * cudaComputeImageHeight * sizeof(float4);
    return VCRT_SUCCESS;
}

// Tolerates never-created state, like EndRenderingOperation's null checks (Renderer.cpp:716-773).
extern "C" int cudaDestroyRenderer(void) {
    cudaCommDestroy();
    if (cudaComputeDeviceReady) {
        cudaSetDevice(cudaComputeDevice);
        if (cudaComputeStream) cudaStreamSynchronize(cudaComputeStream);
        void* buffers[] = {cudaComputePrimary, cudaComputeSceneBlob, cudaComputeSumPlanes, cudaComputeAccumulated, cudaComputeReduced, cudaComputeFixedSums, cudaComputeFixedReduced,
                           cudaComputeResultImage, cudaComputeCounters, cudaComputeWorkCounter, cudaComputeCommScratch, cudaComputeFlushScratch, cudaComputeBvhArena};
        for (void* b : buffers) if (b) cudaFree(b);
        for (int i = 0; i < 16; i++) if (cudaComputeMarkReady[i]) { cudaEventDestroy(cudaComputeMarks[i]); cudaComputeMarkReady[i] = false; }
        for (int i = 0; i < 4; i++) if (cudaComputeEvents[i]) cudaEventDestroy(cudaComputeEvents[i]);
        for (uint32_t i = 0; i < cudaComputePassEventCount; i++) cudaEventDestroy(cudaComputePassEvents[i]);
        cudaComputePassEventCount = 0;
        if (cudaComputeStream) cudaStreamDestroy(cudaComputeStream);
    }
    cudaComputeSceneBlob = nullptr; cudaComputeSumPlanes = nullptr; cudaComputeSumPlanesBytes = 0;
    cudaComputeAccumulated = cudaComputeReduced = cudaComputeResultImage = nullptr;
    cudaComputeFixedSums = cudaComputeFixedReduced = nullptr;
    cudaComputePrimary = nullptr; cudaComputePrimaryCapacity = 0;
    cudaComputeCounters = nullptr; cudaComputeWorkCounter = nullptr; cudaComputeCommScratch = nullptr; cudaComputeFlushScratch = nullptr;
    cudaComputeBvhArena = nullptr; cudaComputeBvhArenaBytes = 0; cudaComputeBvhReady = false;
    if (cudaComputeHostShape) delete[] cudaComputeHostShape;
    cudaComputeHostShape = nullptr;
    if (cudaComputeHostBlob) free(cudaComputeHostBlob);
    cudaComputeHostBlob = nullptr; cudaComputeHostBlobBytes = 0; cudaComputeSceneBlobCapacity = 0;
    cudaComputeStream = nullptr; for (int i = 0; i < 4; i++) cudaComputeEvents[i] = nullptr;
    cudaComputeImageWidth = cudaComputeImageHeight = 0; cudaComputeAccumulatedSamples = 0;
    cudaComputeImageValid = false; cudaComputeSceneReady = false; cudaComputeDeviceReady = false;
    return VCRT_SUCCESS;
}

// ------------------------------------------------------------------ cudaRenderNoAlloc: a render call that allocates nothing
// (include/vcrt.h: no counterpart in the reference; for callers that cannot call cudaMalloc.)  Nothing above this line is
// used by it except device code (hit_exact, scatter, the Philox generator), makeFrame and fail(); nothing above calls into it.
//
// ALL scratch memory of the call lives in this one array.  Every GPU that loads the library holds it until the process exits.
//     [0, 256)                              counters (Counter enum of vcrt_device.cuh, 64-bit each)
//     [256, +16 * kNoAllocMaxSpheres)       shape[]  (cx, cy, cz, r)              the scene as a structure of arrays, written
//     [.., +16 * kNoAllocMaxSpheres)        surf[]   (colour r, g, b, texture.y)  by static_scene_kernel from the caller's AoS
//     [.., +16 * kNoAllocMaxSpheres)        kind[]   int(texture.x), 4 bytes used per 16
//     [.., +16 * kNoAllocMaxPlanePixels)    planes[run][pixel]  partial sums of a pixel's runs of samples
static const uint32_t kNoAllocMaxSpheres = VCRT_NOALLOC_MAX_SPHERES, kNoAllocMaxPlanePixels = VCRT_NOALLOC_MAX_PLANE_PIXELS;
static const size_t kNoAllocOffShape = 256, kNoAllocOffSurf = kNoAllocOffShape + 16u * (size_t)kNoAllocMaxSpheres,
                    kNoAllocOffKind = kNoAllocOffSurf + 16u * (size_t)kNoAllocMaxSpheres, kNoAllocOffPlanes = kNoAllocOffKind + 16u * (size_t)kNoAllocMaxSpheres,
                    kNoAllocBytes = kNoAllocOffPlanes + 16u * (size_t)kNoAllocMaxPlanePixels;
static_assert(kNoAllocBytes == VCRT_NOALLOC_SCRATCH_BYTES, "the size include/vcrt.h states");
static_assert(sizeof(unsigned long long) * CTR_COUNT <= 256 && kNoAllocOffPlanes % 256 == 0, "scratch layout");
static __device__ __align__(256) unsigned char cudaComputeStaticScratch[kNoAllocBytes];
// In HOST code the name above is only a stand-in for the symbol, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, asked once per device with that device current (each device has its own copy of the array).
static const int kNoAllocMaxDevices = 64;
static unsigned char* cudaComputeStaticScratchAddress[kNoAllocMaxDevices];

// caller's AoS (VcrtSphere, 10 floats) -> the SoA the exact test reads; also clears the counters (atomically added to below)
__global__ void static_scene_kernel(const float* spheres, uint32_t count, unsigned char* scratch) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < (uint32_t)CTR_COUNT) reinterpret_cast<unsigned long long*>(scratch)[j] = 0ull;
    if (j >= count) return;
    const float* s = spheres + 10u * (size_t)j;
    reinterpret_cast<float4*>(scratch + kNoAllocOffShape)[j] = make_float4(s[0], s[1], s[2], s[3]);
    reinterpret_cast<float4*>(scratch + kNoAllocOffSurf)[j] = make_float4(s[4], s[5], s[6], s[8]);
    reinterpret_cast<int*>(scratch + kNoAllocOffKind)[j] = (int)s[7];   // switch(int(record.texture.x)), textures.glsl:66
}

// render_simple_kernel (vcrt_kernel_simple.cuh) with one thread per (pixel, run of samples): blockIdx.z = the run, whose
// samples [run * sppPerRun, ...) are summed in index order into planes[run][pixel] - the split cudaRender's float modes make.
__global__ void __launch_bounds__(128) render_static_kernel(DeviceScene scene, RenderArgs args) {
    const uint32_t px = blockIdx.x * 16 + (threadIdx.x & 15), py = blockIdx.y * 8 + (threadIdx.x >> 4), run = blockIdx.z;
    unsigned long long rays = 0, exhausted = 0, bad = 0;
    if (px < args.width && py < args.height) {
        const uint32_t pixel = py * args.width + px;
        const uint32_t first = run * args.plan.sppPerRun, last = min(args.spp, first + args.plan.sppPerRun);
        V3 sum = v3(0.f, 0.f, 0.f);
        for (uint32_t s = first; s < last; s++) {
            const uint32_t sample = args.sampleOffset + s;
            uint32_t rnd[4];
            philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
            Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
            V3 colour = v3(1.f, 1.f, 1.f), result = v3(0.f, 0.f, 0.f);   // bug B7: a path that runs out of depth is black
            uint32_t pass = 0;
            for (; pass < args.maxDepth; pass++) {
                rays++;
                HitRec rec; rec.t = kTMax; rec.index = -1;
                const float a = vdot(ray.d, ray.d);
                for (int i = 0; i < scene.count; i++) hit_exact(__ldg(&scene.shape[i]), i, ray, a, rec, false);
                if (rec.index < 0) { result = vmul(colour, sky(ray.d)); break; }
                philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                const float4 shape = __ldg(&scene.shape[rec.index]), surf = __ldg(&scene.surf[rec.index]);
                const int kind = __ldg(&scene.kind[rec.index]);
                int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
                if (args.semantics) {
                    BookDraws g; g.pixel = pixel; g.sample = sample; g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                    outcome = scatter_book(shape, surf, kind, rec.t, ray, colour, u, g);
                } else outcome = scatter(shape, surf, kind, rec.t, ray, colour, u) ? 1 : 0;
                if (outcome != 1) { if (outcome == 0) bad++; break; }
            }
            if (pass == args.maxDepth) exhausted++;
            sum = vadd(sum, result);
        }
        args.sums[(size_t)run * args.width * args.height + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);
    }
    for (int off = 16; off > 0; off >>= 1) {
        rays += __shfl_down_sync(0xffffffffu, rays, off);
        exhausted += __shfl_down_sync(0xffffffffu, exhausted, off);
        bad += __shfl_down_sync(0xffffffffu, bad, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&args.counters[CTR_RAYS], rays);
        if (exhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], exhausted);
        if (bad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], bad);
    }
}

// combine_planes_kernel + normalise_kernel in one: 0 + plane[0] + plane[1] + ... in run order, / samples, BOOK gamma, alpha 1
__global__ void static_finish_kernel(float4* image, const float4* planes, uint32_t pixels, uint32_t runs, float samples, int gamma) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    for (uint32_t c = 0; c < runs; c++) {
        const float4 s = planes[(size_t)c * pixels + p];
        a.x = __fadd_rn(a.x, s.x); a.y = __fadd_rn(a.y, s.y); a.z = __fadd_rn(a.z, s.z);
    }
    image[p] = make_float4(to_gamma(__fdiv_rn(a.x, samples), gamma), to_gamma(__fdiv_rn(a.y, samples), gamma), to_gamma(__fdiv_rn(a.z, samples), gamma), 1.0f);
}

static bool isDevicePointer(const void* p) {
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged;
}

extern "C" int cudaRenderNoAlloc(IN const VcrtSphere* deviceSpheres, IN uint32_t count, IN const VcrtCamera* camera,
                                 IN const VcrtRenderParams* p, OUT float* deviceImage, IN size_t imageBytes, IN void* stream) {
    if (!camera || !p || !deviceImage || (!deviceSpheres && count)) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: null argument%s", "");
    if (!p->width || !p->height || !p->samplesPerPixel || p->width > 65535u || p->height > 65535u || p->samplesPerPixel >= (1u << 24))
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: width/height must be 1..65535 and samplesPerPixel 1..2^24-1%s", "");
    if (p->flags || p->reserved[0] || p->reserved[1]) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: flags and reserved fields must be 0%s", "");
    if (p->traversal != VCRT_TRAVERSAL_AUTO && p->traversal != VCRT_TRAVERSAL_SIMPLE)
        return fail(VCRT_ERROR_UNSUPPORTED, "cudaRenderNoAlloc has the in-order exhaustive kernel only: traversal must be AUTO or SIMPLE%s", "");
    if (p->semantics > VCRT_SEMANTICS_BOOK) return fail(VCRT_ERROR_INVALID_ARGUMENT, "unknown semantics%s", "");
    char limits[160];
    if (count > kNoAllocMaxSpheres) {
        snprintf(limits, sizeof limits, "%u spheres, the static scratch array holds VCRT_NOALLOC_MAX_SPHERES = %u", count, kNoAllocMaxSpheres);
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: %s", limits);
    }
    // the split of cudaRender's float modes: runs of ceil(spp / chunks) samples, and only as many runs as that leaves
    uint32_t chunks = p->sampleChunks ? p->sampleChunks : 1u;
    if (chunks > p->samplesPerPixel) chunks = p->samplesPerPixel;
    const uint32_t perRun = (p->samplesPerPixel + chunks - 1u) / chunks, runs = (p->samplesPerPixel + perRun - 1u) / perRun;
    const uint64_t pixels = (uint64_t)p->width * p->height;
    if (pixels * runs > kNoAllocMaxPlanePixels || runs > 65535u) {
        snprintf(limits, sizeof limits, "%u x %u pixels x %u runs of samples, the static scratch array holds VCRT_NOALLOC_MAX_PLANE_PIXELS = %u (and 65535 runs)",
                 p->width, p->height, runs, kNoAllocMaxPlanePixels);
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: %s", limits);
    }
    if (imageBytes < pixels * sizeof(float4)) return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: image smaller than width*height*16 bytes%s", "");

    // set-up, once per device, with that device current: the device address of the scratch array
    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return fail(VCRT_ERROR_NO_DEVICE, "no CUDA device available (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= kNoAllocMaxDevices) return fail(VCRT_ERROR_UNSUPPORTED, "cudaRenderNoAlloc: device index beyond 63%s", "");
    if (!cudaComputeStaticScratchAddress[device]) {
        int major = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
        if (major < 10) return fail(VCRT_ERROR_UNSUPPORTED, "device is not sm_100 class%s", "");
        unsigned char* address = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void**)&address, cudaComputeStaticScratch));
        if (!address) return fail(VCRT_ERROR_CUDA, "cudaGetSymbolAddress gave no address for the scratch array%s", "");
        cudaComputeStaticScratchAddress[device] = address;
    }
    unsigned char* scratch = cudaComputeStaticScratchAddress[device];
    // a host pointer here would fault in the kernel (or, where the GPU reaches host memory, be silently wrong): refuse it
    if ((count && !isDevicePointer(deviceSpheres)) || !isDevicePointer(deviceImage))
        return fail(VCRT_ERROR_INVALID_ARGUMENT, "cudaRenderNoAlloc: deviceSpheres and deviceImage must be device memory%s", "");

    RenderArgs args; memset(&args, 0, sizeof args);
    const bool book = p->semantics == VCRT_SEMANTICS_BOOK;
    args.frame = makeFrame(*camera, p->width, p->height, book);
    args.semantics = p->semantics;
    args.width = p->width; args.height = p->height; args.spp = p->samplesPerPixel; args.sampleOffset = p->sampleOffset;
    args.maxDepth = p->maxDepth; args.seedLo = (uint32_t)p->seed; args.seedHi = (uint32_t)(p->seed >> 32);
    args.chunks = runs; args.plan.sppPerRun = perRun; args.plan.runsPerPixel = runs;
    args.sums = reinterpret_cast<float4*>(scratch + kNoAllocOffPlanes);
    args.counters = reinterpret_cast<unsigned long long*>(scratch);
    DeviceScene scene; memset(&scene, 0, sizeof scene);
    scene.shape = reinterpret_cast<const float4*>(scratch + kNoAllocOffShape);
    scene.surf = reinterpret_cast<const float4*>(scratch + kNoAllocOffSurf);
    scene.kind = reinterpret_cast<const int*>(scratch + kNoAllocOffKind);
    scene.count = (int)count;

    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t prepare = count > (uint32_t)CTR_COUNT ? count : (uint32_t)CTR_COUNT;
    static_scene_kernel<<<(prepare + 127u) / 128u, 128, 0, s>>>(reinterpret_cast<const float*>(deviceSpheres), count, scratch);
    render_static_kernel<<<dim3((p->width + 15u) / 16u, (p->height + 7u) / 8u, runs), 128, 0, s>>>(scene, args);
    static_finish_kernel<<<(unsigned int)((pixels + 255u) / 256u), 256, 0, s>>>(reinterpret_cast<float4*>(deviceImage), args.sums, (uint32_t)pixels, runs,
                                                                               (float)(p->totalSamples ? p->totalSamples : p->samplesPerPixel), book ? 1 : 0);
    CUDA_TRY(cudaGetLastError());
    return VCRT_SUCCESS;
}
