/* @file vcrt_kernel_lbvh_gen1.cuh

    EXPERIMENT (not part of the product build): the first-generation LBVH path tracer - nested loops per thread, min/max planes
    selected per node visit, one global widening per ray.  Kept for A/B runs against render_lbvh_flat_kernel
    (DESIGN.md 6.3, steps l1-l2); compiled only with -DVCRT_EXPERIMENTS (tools/experiments/README.md).
    SPDX-License-Identifier: WTFPL
*/
#pragma once
#include "../../vulkancomputeraytracing_b200/csrc/vcrt_kernel_lbvh.cuh"

namespace vcrt {

// STATS: the VCRT_FLAG_COUNT_STATS counters are a compile-time variant here (a run-time test costs this kernel 2 %)
template <bool FIXED, bool STATS>
__global__ void __launch_bounds__(256, 4) render_lbvh_kernel(BvhScene scene, RenderArgs args) {
    const uint32_t pixels = args.width * args.height;
    const uint32_t sppPerChunk = args.plan.sppPerRun;   // host-made plan (vcrt_device.cuh)
    // items walk the image in 8x4 pixel tiles so a warp starts on one tile
    const uint32_t tilesX = args.plan.tilesX, totalItems = args.plan.totalItems;
    uint32_t nRays = 0, nNodes = 0, nTests = 0, nExhausted = 0, nBad = 0;   // per thread; widened when summed
    int stack[64];

    uint32_t item = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t firstDynamic = gridDim.x * blockDim.x;
    for (;; item = firstDynamic + atomicAdd(args.workCounter, 1u)) {
        if (item >= totalItems) break;
        uint32_t chunk, pid;
        item_to_run(args.plan, item, chunk, pid);
        const uint32_t tile = pid >> 5, within = pid & 31u, ty = tile / tilesX, tx = tile - ty * tilesX;
        const uint32_t px = tx * 8u + (within & 7u), py = ty * 4u + (within >> 3);
        if (px >= args.width || py >= args.height) continue;
        const uint32_t s0 = chunk * sppPerChunk, s1 = min(args.spp, s0 + sppPerChunk);
        if (s0 >= s1) continue;
        const uint32_t pixel = py * args.width + px;
        V3 sum = v3(0.f, 0.f, 0.f);
        for (uint32_t s = s0; s < s1; s++) {
            const uint32_t sample = args.sampleOffset + s;
            uint32_t rnd[4];
            philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
            Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
            V3 colour = v3(1.f, 1.f, 1.f), result = v3(0.f, 0.f, 0.f);
            bool finished = false;
            for (uint32_t pass = 0; pass < args.maxDepth; pass++) {
                nRays++;
                const float a = vdot(ray.d, ray.d);
                OrderedHit rec; rec.t = kTMax; rec.device = -1; rec.orig = 0xffffffffu;
                const float oo = fmaf(ray.o.x, ray.o.x, fmaf(ray.o.y, ray.o.y, ray.o.z * ray.o.z));
                const bool degenerate = !tame_ray(a, oo);
                if (degenerate) {   // the reference's in-order scan, literally
                    HitRec h; h.t = kTMax; h.index = -1;
                    // original order = ascending origIndex; device order differs only by moving the always list to the end,
                    // so walk both runs merged by original index
                    uint32_t i = 0, j = scene.scanCount;
                    while (i < scene.scanCount || j < scene.count) {
                        const bool takeTree = j >= scene.count || (i < scene.scanCount && scene.origIndex[i] < scene.origIndex[j]);
                        const uint32_t k = takeTree ? i++ : j++;
                        hit_exact(scene.shape[k], (int)k, ray, a, h, false);
                        nTests++;
                    }
                    rec.t = h.t; rec.device = h.index;
                } else {
                    for (uint32_t x = 0; x < scene.extraCount; x++) { const uint32_t k = __ldg(&scene.extra[x]); hit_ordered(__ldg(&scene.shape[k]), (int)k, __ldg(&scene.origIndex[k]), ray, a, rec); nTests++; }
                    if (scene.treeCount == 1u) { hit_ordered(__ldg(&scene.leafShape[0]), (int)__ldg(&scene.leafDevice[0]), __ldg(&scene.leafOrig[0]), ray, a, rec); nTests++; }
                    else if (scene.treeCount > 1u) {
                        // per-ray widening of every box (file header)
                        const float fx = fmaxf(fabsf(ray.o.x - scene.centreMin.x), fabsf(ray.o.x - scene.centreMax.x));
                        const float fy = fmaxf(fabsf(ray.o.y - scene.centreMin.y), fabsf(ray.o.y - scene.centreMax.y));
                        const float fz = fmaxf(fabsf(ray.o.z - scene.centreMin.z), fabsf(ray.o.z - scene.centreMax.z));
                        const float F2 = fmaf(fx, fx, fmaf(fy, fy, fz * fz));
                        const float K = 3.814697265625e-06f * (F2 + scene.radiusMax * scene.radiusMax);   // 64 eps, eps = 2^-24
                        const float grow = fminf(sqrtf(K), K / (2.f * scene.radiusMin)) * 1.0001f + 9.5367431640625e-07f * (sqrtf(oo) + sqrtf(F2));
                        // slab constants per axis: t_near = near_plane * inv + bn, t_far = far_plane * inv + bf, where the near plane is
                        // the box minimum when the direction component is >= 0 (picked with one select per plane) and bn / bf carry
                        // the origin and the widening.  A component too small to invert safely leaves its axis unconstrained
                        // (inv = 0, bn = -inf, bf = +inf): conservative.
                        const bool px_ = ray.d.x >= 0.f, py_ = ray.d.y >= 0.f, pz_ = ray.d.z >= 0.f;
                        const bool ux = fabsf(ray.d.x) > 1e-30f, uy = fabsf(ray.d.y) > 1e-30f, uz = fabsf(ray.d.z) > 1e-30f;
                        const float ix = ux ? 1.0f / ray.d.x : 0.f, iy = uy ? 1.0f / ray.d.y : 0.f, iz = uz ? 1.0f / ray.d.z : 0.f;
                        const float kInf = __int_as_float(0x7f800000);
                        const float bnx = ux ? fmaf(-ray.o.x, ix, -grow * fabsf(ix)) : -kInf, bfx = ux ? fmaf(-ray.o.x, ix, grow * fabsf(ix)) : kInf;
                        const float bny = uy ? fmaf(-ray.o.y, iy, -grow * fabsf(iy)) : -kInf, bfy = uy ? fmaf(-ray.o.y, iy, grow * fabsf(iy)) : kInf;
                        const float bnz = uz ? fmaf(-ray.o.z, iz, -grow * fabsf(iz)) : -kInf, bfz = uz ? fmaf(-ray.o.z, iz, grow * fabsf(iz)) : kInf;
                        int sp = 0, node = 0;
                        for (;;) {
                            if (node >= 0) {
                                nNodes++;
                                const float4 na = __ldg(&scene.nodes[node].a), nb = __ldg(&scene.nodes[node].b), nc = __ldg(&scene.nodes[node].c);
                                const int2 ch = __ldg(reinterpret_cast<const int2*>(&scene.nodes[node].child0));
                                // child0: min (na.x, na.y, na.z) max (na.w, nb.x, nb.y); child1: min (nb.z, nb.w, nc.x) max (nc.y, nc.z, nc.w)
                                const float n0 = fmaxf(fmaxf(fmaf(px_ ? na.x : na.w, ix, bnx), fmaf(py_ ? na.y : nb.x, iy, bny)), fmaf(pz_ ? na.z : nb.y, iz, bnz));
                                const float f0 = fminf(fminf(fmaf(px_ ? na.w : na.x, ix, bfx), fmaf(py_ ? nb.x : na.y, iy, bfy)), fmaf(pz_ ? nb.y : na.z, iz, bfz));
                                const float n1 = fmaxf(fmaxf(fmaf(px_ ? nb.z : nc.y, ix, bnx), fmaf(py_ ? nb.w : nc.z, iy, bny)), fmaf(pz_ ? nc.x : nc.w, iz, bnz));
                                const float f1 = fminf(fminf(fmaf(px_ ? nc.y : nb.z, ix, bfx), fmaf(py_ ? nc.z : nb.w, iy, bfy)), fmaf(pz_ ? nc.w : nc.x, iz, bfz));
                                // accepted roots lie in (t_min, rec.t]
                                const bool h0 = !(n0 > f0) && !(f0 < 0.f) && !(n0 > rec.t);   // written so that a NaN bound passes
                                const bool h1 = !(n1 > f1) && !(f1 < 0.f) && !(n1 > rec.t);
                                if (h0 && h1) {   // nearer child first, the other one waits on the stack
                                    const bool swap = n1 < n0;
                                    node = swap ? ch.y : ch.x;
                                    stack[sp++] = swap ? ch.x : ch.y;
                                    continue;
                                }
                                if (h0 || h1) { node = h0 ? ch.x : ch.y; continue; }
                            } else {      // a leaf: the exact test
                                const int l = ~node;
                                hit_ordered(__ldg(&scene.leafShape[l]), (int)__ldg(&scene.leafDevice[l]), __ldg(&scene.leafOrig[l]), ray, a, rec);
                                nTests++;
                            }
                            if (sp == 0) break;
                            node = stack[--sp];
                        }
                    }
                }
                if (rec.device < 0) {
                    if constexpr (STATS) atomicAdd(&args.counters[CTR_PATH0 + min(pass, 7u)], 1ull);
                    result = vmul(colour, sky(ray.d)); finished = true; break;
                }
                if constexpr (STATS) { const int k = __ldg(&scene.kind[rec.device]); atomicAdd(&args.counters[CTR_HITS0 + ((k >= 1 && k <= 3) ? k : 0)], 1ull); }
                philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
                if (args.semantics) {
                    BookDraws g; g.pixel = pixel; g.sample = sample; g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                    outcome = scatter_book(__ldg(&scene.shape[rec.device]), __ldg(&scene.surf[rec.device]), __ldg(&scene.kind[rec.device]), rec.t, ray, colour, u, g);
                } else outcome = scatter(__ldg(&scene.shape[rec.device]), __ldg(&scene.surf[rec.device]), __ldg(&scene.kind[rec.device]), rec.t, ray, colour, u) ? 1 : 0;
                if (outcome != 1) { if (outcome == 0) nBad++; finished = true; break; }
            }
            if (!finished) nExhausted++;
            if (FIXED) {
                add_fixed(args, pixel, result);
            } else sum = vadd(sum, result);
        }
        if (!FIXED) args.sums[(size_t)chunk * pixels + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);
    }
    // one atomic per warp per counter
    unsigned long long wRays = nRays, wNodes = nNodes, wTests = nTests;
    for (int off = 16; off > 0; off >>= 1) {
        wRays += __shfl_down_sync(0xffffffffu, wRays, off); wNodes += __shfl_down_sync(0xffffffffu, wNodes, off);
        wTests += __shfl_down_sync(0xffffffffu, wTests, off);
        nExhausted += __shfl_down_sync(0xffffffffu, nExhausted, off); nBad += __shfl_down_sync(0xffffffffu, nBad, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&args.counters[CTR_RAYS], wRays); atomicAdd(&args.counters[CTR_NODES], wNodes); atomicAdd(&args.counters[CTR_TESTS], wTests);
        if (nExhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)nExhausted);
        if (nBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)nBad);
    }
}

}  // namespace vcrt
