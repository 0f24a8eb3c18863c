/* @file vcrt_device.cuh

    Device-side vocabulary of the path tracer: IEEE binary32 helpers that never contract,
    Philox4x32-10, the scene / frame / work descriptors, the exact ray-sphere test and the
    three material scatter routines.
    SPDX-License-Identifier: WTFPL

    Arithmetic policy.  Everything that decides WHAT a path does (roots, hit points, normals,
    scatter directions, attenuation) is written with the *_rn intrinsics below, one rounding per
    operation in the order of the reference's GLSL statements, so nvcc can neither fuse nor
    reorder it.  Only the conservative candidate filter of the brute-force scan
    (vcrt_scan.cuh) uses fused packed FMAs, and it can only ever ADD spheres to the set that
    is then tested exactly.
*/
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define VCRT_DEV __device__ __forceinline__

namespace vcrt {

// ------------------------------------------------------------------ exact binary32
VCRT_DEV float fadd(float a, float b) { return __fadd_rn(a, b); }
VCRT_DEV float fsub(float a, float b) { return __fsub_rn(a, b); }
VCRT_DEV float fmul(float a, float b) { return __fmul_rn(a, b); }
// IEEE division and square root expand to ~10 instructions plus a slow-path call each, and the path shader uses ~50 of
// them once per ray.  Inlined they are 3-6 % faster than as shared out-of-line functions (measured on B200: brute force
// 93.8 -> 90.8 ms, LBVH 45.0 -> 42.3 ms at 1080p x 128 spp); -DVCRT_NOINLINE_DIV brings the compact form back.
#ifndef VCRT_NOINLINE_DIV
VCRT_DEV float fdiv(float a, float b) { return __fdiv_rn(a, b); }
VCRT_DEV float fsqrt(float a) { return __fsqrt_rn(a); }
#else
__device__ __noinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __noinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
#endif

typedef unsigned long long u64;

// ------------------------------------------------------------------ packed f32x2 (sm_100a FFMA2)
VCRT_DEV u64 pack2(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
VCRT_DEV void unpack2(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
VCRT_DEV u64 fmul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
VCRT_DEV u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

struct V3 { float x, y, z; };
VCRT_DEV V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
VCRT_DEV V3 vadd(V3 a, V3 b) { return v3(fadd(a.x, b.x), fadd(a.y, b.y), fadd(a.z, b.z)); }
VCRT_DEV V3 vsub(V3 a, V3 b) { return v3(fsub(a.x, b.x), fsub(a.y, b.y), fsub(a.z, b.z)); }
VCRT_DEV V3 vmul(V3 a, V3 b) { return v3(fmul(a.x, b.x), fmul(a.y, b.y), fmul(a.z, b.z)); }
VCRT_DEV V3 vscale(V3 a, float s) { return v3(fmul(a.x, s), fmul(a.y, s), fmul(a.z, s)); }
VCRT_DEV V3 vdiv(V3 a, float s) { return v3(fdiv(a.x, s), fdiv(a.y, s), fdiv(a.z, s)); }
VCRT_DEV V3 vneg(V3 a) { return v3(-a.x, -a.y, -a.z); }
// GLSL dot(): (x*x' + y*y') + z*z'
VCRT_DEV float vdot(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }
// GLSL normalize(): v / length(v)
VCRT_DEV V3 vunit(V3 a) { return vdiv(a, fsqrt(vdot(a, a))); }
// GLSL reflect(): I - 2*dot(N,I)*N
VCRT_DEV V3 vreflect(V3 i, V3 n) { return vsub(i, vscale(n, fmul(2.0f, vdot(n, i)))); }

// ------------------------------------------------------------------ Philox4x32-10
// Counter-based generator (Salmon et al., SC'11).  counter = (pixel index, global sample
// index, draw block, 0), key = the 64-bit seed.  Draw block 0 = the two pixel-jitter draws of
// shader.comp:48; block k >= 1 = the draws of the k-th scatter event (functions.glsl:42-44,
// textures.glsl:51).  The oracle's PHILOX mode uses the identical layout.
VCRT_DEV void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int round = 0; round < 10; round++) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
VCRT_DEV float u01(uint32_t bits) { return fmul((float)(bits >> 8), 5.9604644775390625e-08f); }  // [0,1), 24 bits

// ------------------------------------------------------------------ descriptors
// Scene as a structure of arrays (one float4 per sphere per array), built by cudaUploadScene
// from the VcrtSphere AoS:
//   hot[i]   = (cx, cy, cz, r^2 inflated by the filter margin)   <- the only array the scan streams
//   shape[i] = (cx, cy, cz, r)                                  <- exact test + normal
//   surf[i]  = (colour r, g, b, texture.y)                      <- attenuation + parameter
//   kind[i]  = int(texture.x)                                   <- material id
struct DeviceScene {
    const float4* hot;
    const float4* shape;
    const float4* surf;
    const int* kind;
    const int* always;   // indices of "huge" spheres kept out of the filter scan (tested exactly for every ray)
    int count;           // all spheres
    int scanCount;       // spheres in hot[] order that the filter scans (= count - alwaysCount), padded entries never pass
    int alwaysCount;
    const int* scanIndex;  // scanIndex[j] = original sphere index of hot[j]
};

// Camera frame of shader.comp:18-39, computed once on the host in the same binary32 order.
struct Frame { V3 center, pixel00, du, dv; };

// How a launch's work items (one pixel, one run of consecutive samples) are numbered; filled by the host (make_item_plan
// in vcrt_cabi.cu) so that the kernels read it from the constant bank instead of keeping it in registers.
// Items are numbered band by band: all runs of the first kBandPixels pixels, then all runs of the next band, ...  The
// paths in flight (3e5 per GPU) then touch a few MB of the accumulation buffer at a time, which stays in L2; numbering
// sample-major instead sweeps the whole buffer once per sample and made L2 write every RED through to DRAM (29.7 GB per
// C2 launch against 33 MB of image, profiles/traffic.json).  For the brute-force kernels a band is spread over the whole image
// (tileStride): a band of neighbouring tiles is all sky or all spheres, so the share of slots that need a new sample per
// round swings from band to band and the shading phases overlap the scans less well (measured 0.8 %).
#ifndef VCRT_BAND_SHIFT
#define VCRT_BAND_SHIFT 18
#endif
constexpr uint32_t kBandPixels = 1u << VCRT_BAND_SHIFT;   // padded pixels (whole 8x4 tiles); 8 MB of sums per band
struct ItemPlan {
    uint32_t tilesX;          // 8x4-pixel tiles per row; pixels are walked tile by tile so that a warp starts on one tile
    uint32_t pixelsPadded;    // tiles * 32
    uint32_t sppPerRun;       // samples per item
    uint32_t runsPerPixel;    // ceil(spp / sppPerRun)
    uint32_t totalItems;      // pixelsPadded * runsPerPixel, <= 2^32 - 2^24 (the dispenser counts a little past it)
    uint32_t fullBandRuns;    // (pixelsPadded >> VCRT_BAND_SHIFT) * runsPerPixel
    uint32_t lastBandFirst;   // first padded pixel of the last, short band
    uint32_t lastBandPixels;  // its size (may be 0)
    float invRuns, invLastBand;   // reciprocals, rounded down, for the float-estimate divisions
    uint32_t tileStride;      // consecutive tiles of the numbering are this many tiles apart in the image (coprime with the tile count)
    uint32_t tiles;           // pixelsPadded / 32
    float invTiles;
    float invTilesX;          // for tile -> (row, column) of tiles without an integer division (25 instructions)
    float invWidth;           // for pixel index -> (x, y)
};

// What the primary-visibility kernels (vcrt_kernel_primary.cuh) hand to the path kernels: every sample of a launch after its
// camera ray and first scatter event.  Two float4 planes; entry of sample s (0-based within the launch) of the pixel with padded
// tile-order id pid sits at [s * plan.pixelsPadded + pid], so the lanes of a warp read and write consecutive entries.
//   a = (origin of the scattered ray, state as bits)     state >= 0: device index of the sphere that was hit - the path goes on
//   b = (direction of the scattered ray, 0)                          with this ray, throughput = that sphere's attenuation, pass 1
//                                                         state < 0: one of the values below
//
// Fixed-point mode (the sums do not care about the order of the samples) uses the buffer as a COMPACT list instead: the primary
// kernel appends only the samples that go on (cursors[0] counts them; a sample that ended has been added to the sums and leaves
// no entry), the path kernel takes entries in list order (cursors[1]):
//   a = (origin, sample index within the launch << 11 | device index of the sphere that was hit)     b = (direction, pixel index)
// with sphere = kPrimaryCameraStart for a sample that was not pre-traced.  A finished slot of the path kernel then costs one
// atomic and two 16-byte loads - no work-item decoding, no entry for the 17 % of the samples that end at their first ray.
struct PrimaryBuffer { float4* a; float4* b; size_t entries; unsigned int* cursors; };   // entries: capacity of each plane
constexpr uint32_t kPrimarySphereBits = 11u, kPrimaryCameraStart = (1u << kPrimarySphereBits) - 1u;   // 1792 + 64 spheres at most reach these kernels
constexpr uint32_t kPrimaryMaxLaunchSamples = 1u << (32u - kPrimarySphereBits);                         // samples per pixel and launch the word can hold
constexpr int kPrimaryEnded = -1;       // the sample ended at its first ray (sky / absorbed / depth 1): b.xyz = its radiance (strict-order sums)
constexpr int kPrimaryUntraced = -2;    // not pre-traced (a ray outside the tame range, a candidate-list overflow): the path kernel starts at the camera

struct RenderArgs {
    Frame frame;
    uint32_t width, height;
    uint32_t spp;           // samples per pixel rendered by this launch
    uint32_t sampleOffset;  // global index of the first of them
    uint32_t maxDepth;
    uint32_t seedLo, seedHi;
    uint32_t chunks;        // a pixel's spp split into this many work items (partial-sum planes)
    uint32_t flags;
    float4* sums;           // float modes: [chunks][height*width] (x,y,z = sums, w unused)
    long long* fixedSums;   // fixed-point mode: [height*width][kFixedLanes] signed 32.32 sums (r, g, b)
    ItemPlan plan;
    uint32_t semantics;       // 0 = the reference shader as it is, 1 = BOOK (vcrt_shade.cuh)
    unsigned long long* counters;  // see Counter enum
    unsigned int* workCounter;     // dynamic item dispenser
    PrimaryBuffer primary;         // this launch's samples after their first scatter event; a == nullptr: every path starts at the camera
    uint32_t primarySamples;       // primary-visibility kernel: a thread's rays are this many samples (1, 2, 4, ...) of rays / this many neighbouring tiles
};

// pixel (px, py) -> its id in 8x4-tile order (the order the primary-visibility kernels write their results in)
VCRT_DEV uint32_t padded_pixel_id(uint32_t tilesX, uint32_t px, uint32_t py) {
    return (((py >> 2) * tilesX + (px >> 3)) << 5) | ((py & 3u) << 3) | (px & 7u);
}

enum Counter { CTR_RAYS = 0, CTR_EXHAUSTED, CTR_BAD_MATERIAL, CTR_CANDIDATES, CTR_OVERFLOWS, CTR_TESTS, CTR_NODES,
               CTR_HITS0, CTR_HITS1, CTR_HITS2, CTR_HITS3, CTR_PATH0, CTR_CHECK_FAILED = CTR_PATH0 + 8, CTR_PRIMARY_RAYS, CTR_COUNT };

// Bounds checks of our own (compute-sanitizer is closed on the GPU pool): with -DVCRT_CHECKED every hand-computed index into
// shared, local or global memory is compared with its bound before use and violations are counted (VcrtRenderStats.checkFailures,
// tools/checked_run.py); the access is then skipped or clamped.  The product build compiles the checks away.
#ifdef VCRT_CHECKED
#define VCRT_CHECK(counters, cond) ((cond) ? true : (atomicAdd(&(counters)[CTR_CHECK_FAILED], 1ull), false))
#else
#define VCRT_CHECK(counters, cond) (true)
#endif

// x / d given inv <= 1 / d (rounded down), for quotients below 2^22 and d below 2^24: the float estimate never exceeds the
// quotient and is at most 2 short of it
__device__ __forceinline__ uint32_t udiv_small(uint32_t x, uint32_t d, float inv, uint32_t& rem) {
    uint32_t q = __float2uint_rz(__fmul_rz(__uint2float_rz(x), inv));
    uint32_t r = x - q * d;
    if (r >= d) { q++; r -= d; }
    if (r >= d) { q++; r -= d; }
    rem = r;
    return q;
}
// item id -> (run index within the pixel, padded pixel index)
__device__ __forceinline__ void item_to_run(const ItemPlan& plan, uint32_t id, uint32_t& chunk, uint32_t& pid) {
    const uint32_t bandRun = id >> VCRT_BAND_SHIFT;            // = band * runsPerPixel + chunk while the band is a full one
    if (bandRun < plan.fullBandRuns) {
        const uint32_t band = udiv_small(bandRun, plan.runsPerPixel, plan.invRuns, chunk);
        pid = (band << VCRT_BAND_SHIFT) | (id & (kBandPixels - 1u));
    } else {                                                   // the last, short band
        chunk = udiv_small(id - plan.lastBandFirst * plan.runsPerPixel, plan.lastBandPixels, plan.invLastBand, pid);
        pid += plan.lastBandFirst;
    }
    if (plan.tileStride > 1u) {   // spread a band over the whole image: tile t of the numbering is tile (t * stride) mod tiles of the image
        uint32_t tile;
        udiv_small((pid >> 5) * plan.tileStride, plan.tiles, plan.invTiles, tile);
        pid = (tile << 5) | (pid & 31u);
    }
}

// tile index -> (column, row) of 8x4 tiles
__device__ __forceinline__ void tile_to_xy(const ItemPlan& plan, uint32_t tile, uint32_t& tx, uint32_t& ty) {
    ty = udiv_small(tile, plan.tilesX, plan.invTilesX, tx);
}

// pixel index (y * width + x) -> (x, y)
__device__ __forceinline__ void pixel_to_xy(const ItemPlan& plan, uint32_t width, uint32_t pixel, uint32_t& px, uint32_t& py) {
    py = udiv_small(pixel, width, plan.invWidth, px);
}

constexpr uint32_t kFlagCountStats = 1u << 3;   // VCRT_FLAG_COUNT_STATS of include/vcrt.h

struct Ray { V3 o, d; };
struct HitRec { float t; int index; };  // nearest accepted root and its sphere (index < 0: none)

// The "tame" range.  The candidate filter and the LBVH boxes are conservative against the ROUNDING of the reference's exact
// test; they say nothing once that test overflows (half_b^2 or a*c = inf, disc = NaN: the sphere is then "hit" at t = NaN and the
// result depends on the order of the scan).  A ray with |d|^2 in (1e-20, 1e18) and |o|^2 < 1e18 against spheres with
// |centre|, |radius| <= 1e9 keeps every intermediate below 3.7e37; every other ray takes the exhaustive in-order scan inside the
// kernels, and a scene with a sphere outside that range is rendered by the in-order kernel altogether (cudaUploadScene).
constexpr float kTameDirMin2 = 1e-20f, kTameMax2 = 1e18f, kTameSceneMax = 1e9f;
VCRT_DEV bool tame_ray(float dd, float oo) { return dd > kTameDirMin2 && dd < kTameMax2 && oo < kTameMax2; }

constexpr float kTMin = 0.001f;   // functions.glsl:75
constexpr float kTMax = 1e5f;     // globals.glsl:26 `infinity`
constexpr float kFixedOne = 4294967296.0f;   // 2^32: radiance sums are exact integers in units of 2^-32
constexpr int kFixedLanes = 3;               // r, g, b: alpha is the constant 1 of shader.comp:44 and never summed

// shader.comp:53 `color.rgb += ray_color(r)` in 32.32 fixed point: three 64-bit REDs, associative, so the image does not depend
// on the schedule or the GPU count.  The caller guarantees 0 <= result <= 1 per channel (cudaUploadScene only allows this
// accumulator for scenes whose attenuations are in [0, 1]), so a sum of 2^24 samples x 8 ranks stays far below 2^63.
VCRT_DEV void add_fixed(const RenderArgs& args, uint32_t pixel, V3 result) {
    // (checked build) the pixel is inside the image, the radiance in [0, 1] - what keeps the integer sums from overflowing
    if (!VCRT_CHECK(args.counters, pixel < args.width * args.height && result.x >= 0.f && result.x <= 1.f && result.y >= 0.f && result.y <= 1.f && result.z >= 0.f && result.z <= 1.f)) return;
    unsigned long long* acc = reinterpret_cast<unsigned long long*>(args.fixedSums) + kFixedLanes * (size_t)pixel;
    if (result.x != 0.f) atomicAdd(acc + 0, (unsigned long long)__float2ll_rn(result.x * kFixedOne));
    if (result.y != 0.f) atomicAdd(acc + 1, (unsigned long long)__float2ll_rn(result.y * kFixedOne));
    if (result.z != 0.f) atomicAdd(acc + 2, (unsigned long long)__float2ll_rn(result.z * kFixedOne));
}

}  // namespace vcrt
