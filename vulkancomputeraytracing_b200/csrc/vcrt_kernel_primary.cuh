/* @file vcrt_kernel_primary.cuh

    Primary visibility for VCRT_TRAVERSAL_BRUTE: every sample's camera ray and first scatter event, done before the path kernel runs.
    SPDX-License-Identifier: WTFPL

    43 % of all rays of the reference's loop are primary rays (2.30 rays per sample on the RTIOW scene), and they are special
    in two ways that a kernel with one path per thread cannot use:

      * they share their origin, the camera (shader.comp:50 `ray(center, ...)`).  With the sphere centres stored relative to the
        camera, c' = c - o, the conservative filter of the brute-force scan (vcrt_kernel_scanu.cuh)
            |c - o|^2 - ((c - o).dh)^2 <= R2
        needs no per-ray origin terms: t = c'.dh (3 packed FP32 operations), |t| >= sqrt(|c'|^2 - R2) (no arithmetic at all: the
        right-hand side is one scalar per sphere, the absolute value an operand modifier of the max over the thread's rays).
        3 packed operations per ray pair and sphere instead of 7 (4 while the test was t^2 >= |c'|^2 - R2), and with no path
        state to keep, a thread scans for FOUR rays at once, so the per-sphere overhead (two LDCU.64, max, compare, flag) is
        shared by 4 tests instead of 2;
      * they are coherent: the 32 lanes of a warp work on one 8x4 pixel tile, so in the shading that follows they mostly meet
        the same sphere and material: 31.8 of 32 lanes active (ncu) where the shading phase of the path kernel runs at 12 - 15.

    The kernel traces the camera ray, shades its hit exactly as the path kernel would (same routines, same Philox block,
    vcrt_shade.cuh) and writes what is left of the sample into the PrimaryBuffer (vcrt_device.cuh): the scattered ray and the
    sphere it came from, or "ended" (sky, absorbed, depth limit).  In fixed-point mode an ended sample is added to the sums
    right here and the buffer is a COMPACT list of the samples that go on (one warp-aggregated atomic per append), which the
    path kernel consumes in list order: a finished slot costs it one atomic and two 16-byte loads - no work-item decoding, and
    nothing at all for the 17 % of the samples that end at their first ray (C2: 286.8 -> 279.8 ms).  In the strict-order modes
    every sample has its entry at [sample][pixel] and the path kernel (render_scanu_kernel) starts it from there, at pass 1:
    two 16-byte loads instead of a Philox block, the camera ray and a scan.  "Every ray tests every sphere" still holds: the scan below visits all
    spheres for every primary ray, and the hit is decided by the same exact test in the same order (hit_exact over the flagged
    spheres in ascending index, then the "always" spheres), so images, ray counts and statistics do not change by a bit.

    The same split was built for the LBVH (a coherent traversal kernel for the camera rays + the per-path kernel for the rest)
    and measured: 12 - 18 % SLOWER at every scene size (485 / 10 k / 100 k spheres, stress scene).  A tree ray costs ~1000
    instructions of which only ~310 are node visits, so coherent camera rays save little, while the per-path kernel left with
    nothing but incoherent rays falls from 21 to 18.8 active lanes and pays for the hand-over.  Not kept.

    Error budget of the camera-relative filter (eps = 2^-24): the exact test can only accept a sphere whose true distance D
    from the ray's line satisfies D^2 <= r^2 + 18 eps (|oc|^2 + r^2) (vcrt_kernel_scanu.cuh).  Here c' is rounded from the
    binary64 difference (relative error eps/2 per component), |dh| = 1 +- 4 eps, the dot product and the square add <= 10 eps
    t^2, so fl(t^2) >= |oc|^2 - r^2 - 34 eps (|oc|^2 + r^2) for every acceptable sphere, and the table holds
    |c'|^2 - r^2 - kappa (|c'|^2 + r^2) rounded down with kappa = 2^-16 = 256 eps - as its square root, scaled by 1 - 2^-22 and
    rounded down again, which passes every ray the squared comparison passes (cudaUploadScene).  Extra survivors cost time,
    never correctness.

    Rays outside the tame range (vcrt_device.cuh) and threads whose candidate list overflows are handed over as "not pre-traced"
    (kPrimaryUntraced / kPrimaryCameraStart): the path kernel starts those samples at the camera itself, the old way.
*/
#pragma once
#include "vcrt_kernel_scanu.cuh"

namespace vcrt {

__constant__ float4 cPrimaryHot[kMaxConstSpheres];     // (c'x, c'y, c'z, sqrt(|c'|^2 - R2) rounded down; 0 = always passes), c' = centre - camera; padding (0,0,0,+inf)

constexpr int kPrimaryThreads = 256;
#ifndef VCRT_PRIMARY_CTAS
#define VCRT_PRIMARY_CTAS 3      // resident CTAs per SM the register budget is set for (85 registers); 4 (64 registers, 20 bytes of spills) measured 8 % slower
#endif
#ifndef VCRT_PRIMARY_RAYS
#define VCRT_PRIMARY_RAYS 4
#endif
#ifndef VCRT_PRIMARY_SHADE_UNROLL
#define VCRT_PRIMARY_SHADE_UNROLL 1
#endif
constexpr int kPrimaryShadeUnroll = VCRT_PRIMARY_SHADE_UNROLL;   // rays of a thread whose exact tests and first scatter events are interleaved: 2 / 4 measured 16 / 29 % slower (primary kernel 21.1 -> 24.5 / 27.2 ms at 1080p x 128 spp)
constexpr int kPrimaryRays = VCRT_PRIMARY_RAYS;        // samples of one pixel per thread (even)
constexpr int kPrimaryPairs = kPrimaryRays / 2;
constexpr int kPrimaryListCap = 40;

__host__ __device__ inline size_t primary_scan_smem(uint32_t blobBytes) { return 16 + (size_t)blobBytes + (size_t)kPrimaryListCap * kPrimaryThreads * 2; }

struct PrimaryCounts { uint32_t rays, exhausted, bad; };

// The first scatter event of one sample (functions.glsl:82-90 for pass 0) and its hand-over.  index < 0: the camera ray left the scene.
// appends one entry to the compact list of fixed-point mode (vcrt_device.cuh); the compiler aggregates the atomic over the warp
VCRT_DEV void primary_append(const RenderArgs& args, V3 o, V3 d, uint32_t sampleLocal, uint32_t sphere, uint32_t pixel) {
    const uint32_t at = atomicAdd(&args.primary.cursors[0], 1u);
    if (!VCRT_CHECK(args.counters, at < args.primary.entries && sampleLocal < kPrimaryMaxLaunchSamples && sphere <= kPrimaryCameraStart)) return;
    __stcs(args.primary.a + at, make_float4(o.x, o.y, o.z, __uint_as_float((sampleLocal << kPrimarySphereBits) | sphere)));
    __stcs(args.primary.b + at, make_float4(d.x, d.y, d.z, __uint_as_float(pixel)));
}

template <bool FIXED, bool BOOK>
VCRT_DEV void primary_finish(const RenderArgs& args, uint32_t pixel, uint32_t sampleLocal, size_t at, Ray ray, float t, int index,
                             float4 shape, float4 surf, int kind, PrimaryCounts& n) {
    const uint32_t sampleGlobal = args.sampleOffset + sampleLocal;
    V3 colour = v3(1.f, 1.f, 1.f), result = v3(0.f, 0.f, 0.f);   // black: bug B7 (depth exhausted) and absorbed bad materials
    bool ended = true;
    if (index < 0) {
        if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_PATH0], 1ull);
        result = vmul(colour, sky(ray.d));
    } else {
        if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_HITS0 + ((kind >= 1 && kind <= 3) ? kind : 0)], 1ull);
        uint32_t rnd[4];
        philox4x32_10(pixel, sampleGlobal, 1u, 0u, args.seedLo, args.seedHi, rnd);
        float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
        int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
        if (BOOK) {
            BookDraws g; g.pixel = pixel; g.sample = sampleGlobal; g.event = 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
            outcome = scatter_book(shape, surf, kind, t, ray, colour, u, g);
        } else outcome = scatter(shape, surf, kind, t, ray, colour, u) ? 1 : 0;
        if (outcome == 0) n.bad++;
        else if (outcome == 2) { }
        else if (1u >= args.maxDepth) n.exhausted++;
        else ended = false;
    }
    if (FIXED) {   // compact list: an ended sample goes into the sums and leaves no entry
        if (ended) add_fixed(args, pixel, result);
        else primary_append(args, ray.o, ray.d, sampleLocal, (uint32_t)index, pixel);
        return;
    }
    float4 a, b;
    if (!ended) { a = make_float4(ray.o.x, ray.o.y, ray.o.z, __int_as_float(index)); b = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f); }
    else { a = make_float4(0.f, 0.f, 0.f, __int_as_float(kPrimaryEnded)); b = make_float4(result.x, result.y, result.z, 0.f); }
    __stcs(args.primary.a + at, a);
    __stcs(args.primary.b + at, b);
}

// One warp = kPrimaryRays rays per lane: S consecutive samples (args.primarySamples, a power of two chosen by the host so that few
// sample slots stay empty: 4 at 500 spp, 1 at the reference's default of 1 spp) of the lane's pixel in kPrimaryRays / S neighbouring
// 8x4 pixel tiles.
template <bool FIXED, bool BOOK>
__global__ void __launch_bounds__(kPrimaryThreads, VCRT_PRIMARY_CTAS) primary_scan_kernel(const unsigned char* __restrict__ blobGlobal, RenderArgs args, int scanPad) {
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    unsigned char* blob = smem + 16;
    const int tid = threadIdx.x, lane = tid & 31;
    // the scene blob (exact-test and material arrays): one TMA bulk copy, everyone waits on the mbarrier
    const uint32_t blobBytes = reinterpret_cast<const SceneBlob*>(blobGlobal)->bytes;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) { mbar_expect_tx(bar, blobBytes); tma_bulk_g2s(blob, blobGlobal, blobBytes, bar); }
    mbar_wait(bar, 0);
    const SceneBlob hdr = *reinterpret_cast<const SceneBlob*>(blob);
    const float4* sShape = reinterpret_cast<const float4*>(blob + hdr.offShape);
    const float4* sSurf = reinterpret_cast<const float4*>(blob + hdr.offSurf);
    const int* sKind = reinterpret_cast<const int*>(blob + hdr.offKind);
    unsigned short* sList = reinterpret_cast<unsigned short*>(blob + blobBytes);
#define PLIST(e) sList[(e) * kPrimaryThreads + tid]
    const int scanCount = (int)hdr.scanCount, count = (int)hdr.count;
    const uint32_t tiles = args.plan.tiles;
    const uint32_t sPer = args.primarySamples, sShift = 31u - __clz(sPer), tPer = (uint32_t)kPrimaryRays >> sShift;
    const uint32_t tileGroups = (tiles + tPer - 1u) / tPer, sampleGroups = (args.spp + sPer - 1u) >> sShift;
    const uint32_t units = tileGroups * sampleGroups;   // the host keeps tiles x samples per launch below 2^31
    const uint32_t warpsPerGrid = gridDim.x * (kPrimaryThreads / 32);
    // the trip count depends on nothing but the launch, so the scan below sits in warp-uniform control flow: that is what lets
    // ptxas feed the sphere operands through uniform registers (LDCU -> UR.F32) instead of LDC into the register file
    const uint32_t trips = (units + warpsPerGrid - 1u) / warpsPerGrid;
    const V3 origin = args.frame.center;
    const float oo = fmaf(origin.x, origin.x, fmaf(origin.y, origin.y, origin.z * origin.z));
    PrimaryCounts n = {0u, 0u, 0u}; uint32_t nCand = 0;

#pragma unroll 1
    for (uint32_t trip = 0; trip < trips; trip++) {
        const uint32_t unit = trip * warpsPerGrid + blockIdx.x * (kPrimaryThreads / 32) + (tid >> 5);
        const uint32_t sampleGroup = unit / tileGroups, tileGroup = unit - sampleGroup * tileGroups;
        // ray k of the thread: sample (k mod S) of the group, tile (k / S) of the group
        auto ray_slot = [&](int k, uint32_t& tile, uint32_t& s, uint32_t& px, uint32_t& py) -> bool {
            tile = tileGroup * tPer + ((uint32_t)k >> sShift); s = (sampleGroup << sShift) + ((uint32_t)k & (sPer - 1u));
            uint32_t tx, ty;
            tile_to_xy(args.plan, tile, tx, ty);
            px = tx * 8u + (lane & 7u); py = ty * 4u + (lane >> 3);
            return unit < units && tile < tiles && s < args.spp && px < args.width && py < args.height;
        };
        // the camera rays (shader.comp:46-52) and their unit directions
        V3 d[kPrimaryRays]; float dh[kPrimaryRays][3]; uint32_t okMask = 0u;
#pragma unroll
        for (int k = 0; k < kPrimaryRays; k++) {
            uint32_t tile, s, px, py;
            dh[k][0] = dh[k][1] = dh[k][2] = 0.f; d[k] = v3(0.f, 0.f, 0.f);
            if (ray_slot(k, tile, s, px, py)) {
                const uint32_t pixel = py * args.width + px;
                uint32_t rnd[4];
                philox4x32_10(pixel, args.sampleOffset + s, 0u, 0u, args.seedLo, args.seedHi, rnd);
                const Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
                d[k] = ray.d;
                const float len2 = fmaf(ray.d.x, ray.d.x, fmaf(ray.d.y, ray.d.y, ray.d.z * ray.d.z));
                if (tame_ray(len2, oo)) {
                    const float inv = rsqrtf(len2);
                    dh[k][0] = ray.d.x * inv; dh[k][1] = ray.d.y * inv; dh[k][2] = ray.d.z * inv;
                    okMask |= 1u << k;
                }
            }
        }
        u64 dxp[kPrimaryPairs], dyp[kPrimaryPairs], dzp[kPrimaryPairs];
#pragma unroll
        for (int p = 0; p < kPrimaryPairs; p++) { dxp[p] = pack2(dh[2 * p][0], dh[2 * p + 1][0]); dyp[p] = pack2(dh[2 * p][1], dh[2 * p + 1][1]); dzp[p] = pack2(dh[2 * p][2], dh[2 * p + 1][2]); }
        // the scan: per sphere two LDCU.64, 3 packed FP32 operations per ray pair, a max of |t| over the thread's rays and one compare
        uint32_t listCount = 0u;
        auto scan_block = [&]<int U>(int at) -> uint32_t {
            uint32_t mEven = 0u, mOdd = 0u;
#pragma unroll
            for (int u = 0; u < U; u++) {
                (void)VCRT_CHECK(args.counters, at + u < scanPad && at + u < kMaxConstSpheres);
                const float4 s = cPrimaryHot[at + u];
                const u64 cx = pack2(s.x, s.x), cy = pack2(s.y, s.y), cz = pack2(s.z, s.z);
                float tmax = 0.f;   // max |t| over the thread's rays (the absolute values are operand modifiers of FMNMX)
#pragma unroll
                for (int p = 0; p < kPrimaryPairs; p++) {
                    u64 t = fmul2(cz, dzp[p]); t = ffma2(cy, dyp[p], t); t = ffma2(cx, dxp[p], t);
                    float t0, t1; unpack2(t, t0, t1);
                    tmax = p == 0 ? fmaxf(fabsf(t0), fabsf(t1)) : fmaxf(tmax, fmaxf(fabsf(t0), fabsf(t1)));
                }
                flag_if_ge(u & 1 ? mOdd : mEven, tmax, s.w, 1u << u);
            }
            return mEven | mOdd;
        };
        const int scanLim = (scanCount + 3) & ~3;
#pragma unroll 1
        for (int base = 0; base < scanLim; base += 32) {
            uint32_t mask = 0u;
            const int lim = min(32, scanLim - base);
            int kk = 0;
#pragma unroll 1
            for (; kk + 16 <= lim; kk += 16) mask |= scan_block.template operator()<16>(base + kk) << kk;
#pragma unroll 1
            for (; kk < lim; kk += 4) mask |= scan_block.template operator()<4>(base + kk) << kk;
            while (mask) {
                const int k = __ffs(mask) - 1;
                mask &= mask - 1u;
                if (listCount < (uint32_t)kPrimaryListCap && VCRT_CHECK(args.counters, base + k < scanCount)) PLIST(listCount) = (unsigned short)(base + k);
                listCount++;
            }
        }
        // exact tests (the reference's arithmetic, functions.glsl:14-33) of the survivors in ascending index, then the "always"
        // spheres; then the first scatter event
        const bool overflow = listCount > (uint32_t)kPrimaryListCap;
#pragma unroll kPrimaryShadeUnroll
        for (int k = 0; k < kPrimaryRays; k++) {
            uint32_t tile, s, px, py;
            if (!ray_slot(k, tile, s, px, py)) continue;
            const uint32_t pixel = py * args.width + px;
            const size_t at = (size_t)s * args.plan.pixelsPadded + ((size_t)tile << 5) + lane;   // strict-order modes: one entry per sample
            if (!FIXED && !VCRT_CHECK(args.counters, at < args.primary.entries)) continue;
            if (overflow || !((okMask >> k) & 1u)) {   // not pre-traced: the path kernel starts this sample at the camera
                if (FIXED) primary_append(args, origin, d[k], s, kPrimaryCameraStart, pixel);
                else __stcs(args.primary.a + at, make_float4(0.f, 0.f, 0.f, __int_as_float(kPrimaryUntraced)));
                continue;
            }
            Ray ray; ray.o = origin; ray.d = d[k];
            const float a = vdot(ray.d, ray.d);
            HitRec rec; rec.t = kTMax; rec.index = -1;
            const int nAll = (int)listCount + (count - scanCount);
            for (int e = 0; e < nAll; e++) {
                const int j = e >= (int)listCount ? scanCount + (e - (int)listCount) : (int)PLIST(e);
                if (VCRT_CHECK(args.counters, j >= 0 && j < count)) hit_exact(sShape[j], j, ray, a, rec, true);
            }
            n.rays++; nCand += listCount;
            const int j = rec.index < 0 ? 0 : rec.index;
            primary_finish<FIXED, BOOK>(args, pixel, s, at, ray, rec.t, rec.index, sShape[j], sSurf[j], sKind[j], n);
        }
    }
    // one atomic per warp per counter
    unsigned long long wRays = n.rays, wCand = nCand; uint32_t wEx = n.exhausted, wBad = n.bad;
    for (int off = 16; off > 0; off >>= 1) {
        wRays += __shfl_down_sync(0xffffffffu, wRays, off); wCand += __shfl_down_sync(0xffffffffu, wCand, off);
        wEx += __shfl_down_sync(0xffffffffu, wEx, off); wBad += __shfl_down_sync(0xffffffffu, wBad, off);
    }
    if (lane == 0) {
        atomicAdd(&args.counters[CTR_RAYS], wRays); atomicAdd(&args.counters[CTR_PRIMARY_RAYS], wRays);
        if (wCand) atomicAdd(&args.counters[CTR_CANDIDATES], wCand);
        if (wEx) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)wEx);
        if (wBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)wBad);
    }
#undef PLIST
}

}  // namespace vcrt
