/* @file vcrt_kernel_simple.cuh

    VCRT_TRAVERSAL_SIMPLE: one thread per pixel, samples and bounces in the reference's own
    loop order (shader.comp:46-54, functions.glsl:73-90), every sphere tested exactly, scene
    read through L1.  It is the on-device cross-check for the fast kernels (which must produce
    the same image bit for bit) and the any-size fallback; it is not the performance path.
    SPDX-License-Identifier: WTFPL
*/
#pragma once
#include "vcrt_shade.cuh"

namespace vcrt {

__global__ void __launch_bounds__(128) render_simple_kernel(DeviceScene scene, RenderArgs args) {
    const uint32_t px = blockIdx.x * 16 + (threadIdx.x & 15);   // 16x8 pixel tiles
    const uint32_t py = blockIdx.y * 8 + (threadIdx.x >> 4);
    unsigned long long rays = 0, exhausted = 0, bad = 0;
    if (px < args.width && py < args.height) {
        const uint32_t pixel = py * args.width + px;
        V3 sum = v3(0.f, 0.f, 0.f);
        for (uint32_t s = 0; s < args.spp; s++) {
            const uint32_t sample = args.sampleOffset + s;
            uint32_t rnd[4];
            philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
            Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
            V3 colour = v3(1.f, 1.f, 1.f);
            V3 result = v3(0.f, 0.f, 0.f);   // bug B7: depth exhaustion is black
            bool finished = false;
            for (uint32_t pass = 0; pass < args.maxDepth; pass++) {
                rays++;
                HitRec rec; rec.t = kTMax; rec.index = -1;
                const float a = vdot(ray.d, ray.d);
                for (int i = 0; i < scene.count; i++) hit_exact(__ldg(&scene.shape[i]), i, ray, a, rec, false);
                if (rec.index < 0) {
                    if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_PATH0 + min(pass, 7u)], 1ull);
                    result = vmul(colour, sky(ray.d)); finished = true; break;
                }
                if (args.flags & kFlagCountStats) { const int k = __ldg(&scene.kind[rec.index]); atomicAdd(&args.counters[CTR_HITS0 + ((k >= 1 && k <= 3) ? k : 0)], 1ull); }
                philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
                if (args.semantics) {
                    BookDraws g; g.pixel = pixel; g.sample = sample; g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                    outcome = scatter_book(__ldg(&scene.shape[rec.index]), __ldg(&scene.surf[rec.index]), __ldg(&scene.kind[rec.index]), rec.t, ray, colour, u, g);
                } else outcome = scatter(__ldg(&scene.shape[rec.index]), __ldg(&scene.surf[rec.index]), __ldg(&scene.kind[rec.index]), rec.t, ray, colour, u) ? 1 : 0;
                if (outcome != 1) { if (outcome == 0) bad++; finished = true; break; }
            }
            if (!finished) exhausted++;
            sum = vadd(sum, result);
        }
        args.sums[pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);
    }
    // one atomic per warp per counter
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

}  // namespace vcrt
