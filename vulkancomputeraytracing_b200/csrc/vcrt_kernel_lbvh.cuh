/* @file vcrt_kernel_lbvh.cuh

    VCRT_TRAVERSAL_LBVH, part 2: the persistent path tracer that finds nearest hits through the
    linear BVH of vcrt_lbvh.cuh (BASELINE.json configs[3]: 10 k and 100 k spheres).
    SPDX-License-Identifier: WTFPL

    One thread = one path in flight.  A thread takes a work item (pixel, run of samples) from the
    global dispenser, traces its samples bounce by bounce and takes the next item when the run is
    done, so the grid stays full until the image is finished (no per-pixel tail).

    Same image as the brute-force kernels, bit for bit.  Why that holds:
      * the exact test (hit_exact's arithmetic, the reference's functions.glsl:14-33) runs only at
        leaves; the box test merely decides which leaves are reached, and it is conservative:
        a sphere can pass the exact test only if the distance D from its centre to the ray's line
        satisfies D^2 <= r^2 + ~18 eps (|o-c|^2 + r^2) (rounding of oc, the two dot products, a*c).
        Every box is therefore widened per ray by `grow` >= sqrt(r^2 + K) - r with
        K = 64 eps (F^2 + rmax^2), F = the largest possible |o - c| for this ray's origin, plus
        16 eps (|o| + F) for the rounding of the slab arithmetic itself;
      * the nearest hit does not depend on the visiting order: a sphere's candidate root is its
        near root if that is > t_min, else its far root (hit_exact rejects the far root whenever it
        rejects the near one against max_t), and equal roots go to the LOWER original index,
        which is what the reference's in-order scan with `max_t <= root` gives;
      * rays whose direction or origin is not finite / is zero take an in-order exhaustive scan
        (hit_exact as written), because NaN comparisons make the reference's result order-dependent.
*/
#pragma once
#include "vcrt_lbvh.cuh"

namespace vcrt {

struct BvhScene {
    const BvhNode* nodes;
    const float4* leafShape;      // sorted order
    const uint32_t* leafOrig;
    const uint32_t* leafDevice;
    const float4* shape;          // device order (exact tests of the always list, normals)
    const float4* surf;
    const int* kind;
    const uint32_t* origIndex;
    uint32_t treeCount;           // spheres in the tree = device indices [0, treeCount)
    uint32_t count;               // all spheres; [treeCount, count) are tested for every ray
    float3 centreMin, centreMax;  // bounding box of the tree spheres' centres
    float radiusMin, radiusMax;   // over the tree spheres
};

struct OrderedHit { float t; int device; uint32_t orig; };

// Candidate root of one sphere, independent of what was found before (see the file header).
VCRT_DEV void hit_ordered(float4 shape, int device, uint32_t orig, const Ray& r, float a, OrderedHit& rec) {
    V3 oc = vsub(r.o, v3(shape.x, shape.y, shape.z));
    float half_b = vdot(oc, r.d);
    float c = fsub(vdot(oc, oc), fmul(shape.w, shape.w));
    float disc = fsub(fmul(half_b, half_b), fmul(a, c));
    if (disc < 0) return;
    float sqrtd = fsqrt(disc);
    float root = fdiv(fsub(-half_b, sqrtd), a);
    if (root <= kTMin) {   // only then can the far root be the accepted one (it is never smaller than the near root)
        root = fdiv(fadd(-half_b, sqrtd), a);
        if (root <= kTMin) return;
    }
    if (!(root < rec.t || (root == rec.t && orig < rec.orig))) return;
    rec.t = root; rec.device = device; rec.orig = orig;
}

template <bool FIXED>
__global__ void __launch_bounds__(256, 2) render_lbvh_kernel(BvhScene scene, RenderArgs args) {
    const uint32_t pixels = args.width * args.height;
    const uint32_t sppPerChunk = FIXED ? args.samplesPerItem : (args.spp + args.chunks - 1u) / args.chunks;
    const uint32_t runsPerPixel = (args.spp + sppPerChunk - 1u) / sppPerChunk;
    // items walk the image in 8x4 pixel tiles so a warp starts on one tile
    const uint32_t tilesX = (args.width + 7u) / 8u, tilesY = (args.height + 3u) / 4u, pixelsPadded = tilesX * tilesY * 32u;
    const uint32_t totalItems = pixelsPadded * runsPerPixel;
    unsigned long long nRays = 0, nNodes = 0, nTests = 0;
    uint32_t nExhausted = 0, nBad = 0;
    int stack[64];

    uint32_t item = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t firstDynamic = gridDim.x * blockDim.x;
    for (;; item = firstDynamic + atomicAdd(args.workCounter, 1u)) {
        if (item >= totalItems) break;
        const uint32_t chunk = item / pixelsPadded, pid = item - chunk * pixelsPadded;
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
                const bool degenerate = !(a > 1e-30f && a < 1e30f && oo < 1e30f);
                if (degenerate) {   // the reference's in-order scan, literally
                    HitRec h; h.t = kTMax; h.index = -1;
                    // original order = ascending origIndex; device order differs only by moving the always list to the end,
                    // so walk both runs merged by original index
                    uint32_t i = 0, j = scene.treeCount;
                    while (i < scene.treeCount || j < scene.count) {
                        const bool takeTree = j >= scene.count || (i < scene.treeCount && scene.origIndex[i] < scene.origIndex[j]);
                        const uint32_t k = takeTree ? i++ : j++;
                        hit_exact(scene.shape[k], (int)k, ray, a, h);
                        nTests++;
                    }
                    rec.t = h.t; rec.device = h.index;
                } else {
                    for (uint32_t k = scene.treeCount; k < scene.count; k++) { hit_ordered(__ldg(&scene.shape[k]), (int)k, __ldg(&scene.origIndex[k]), ray, a, rec); nTests++; }
                    if (scene.treeCount == 1u) { hit_ordered(__ldg(&scene.leafShape[0]), (int)__ldg(&scene.leafDevice[0]), __ldg(&scene.leafOrig[0]), ray, a, rec); nTests++; }
                    else if (scene.treeCount > 1u) {
                        // per-ray widening of every box (file header)
                        const float fx = fmaxf(fabsf(ray.o.x - scene.centreMin.x), fabsf(ray.o.x - scene.centreMax.x));
                        const float fy = fmaxf(fabsf(ray.o.y - scene.centreMin.y), fabsf(ray.o.y - scene.centreMax.y));
                        const float fz = fmaxf(fabsf(ray.o.z - scene.centreMin.z), fabsf(ray.o.z - scene.centreMax.z));
                        const float F2 = fmaf(fx, fx, fmaf(fy, fy, fz * fz));
                        const float K = 3.814697265625e-06f * (F2 + scene.radiusMax * scene.radiusMax);   // 64 eps, eps = 2^-24
                        const float grow = fminf(sqrtf(K), K / (2.f * scene.radiusMin)) * 1.0001f + 9.5367431640625e-07f * (sqrtf(oo) + sqrtf(F2));
                        const float ix = 1.0f / ray.d.x, iy = 1.0f / ray.d.y, iz = 1.0f / ray.d.z;   // +-inf for zero components: slabs handle it
                        const float bx = -ray.o.x * ix, by = -ray.o.y * iy, bz = -ray.o.z * iz;
                        const float ex = grow * fabsf(ix), ey = grow * fabsf(iy), ez = grow * fabsf(iz);
                        int sp = 0, node = 0;
                        for (;;) {
                            nNodes++;
                            const float4 na = __ldg(&scene.nodes[node].a), nb = __ldg(&scene.nodes[node].b), nc = __ldg(&scene.nodes[node].c);
                            const int2 ch = __ldg(reinterpret_cast<const int2*>(&scene.nodes[node].child0));
                            // fminf/fmaxf drop NaN (0 * inf when the origin lies on a slab plane of a zero component): the slab then does not constrain
                            float t0, t1, n0, f0, n1, f1;
                            t0 = fmaf(na.x, ix, bx); t1 = fmaf(na.w, ix, bx); n0 = fminf(t0, t1) - ex; f0 = fmaxf(t0, t1) + ex;
                            t0 = fmaf(na.y, iy, by); t1 = fmaf(nb.x, iy, by); n0 = fmaxf(n0, fminf(t0, t1) - ey); f0 = fminf(f0, fmaxf(t0, t1) + ey);
                            t0 = fmaf(na.z, iz, bz); t1 = fmaf(nb.y, iz, bz); n0 = fmaxf(n0, fminf(t0, t1) - ez); f0 = fminf(f0, fmaxf(t0, t1) + ez);
                            t0 = fmaf(nb.z, ix, bx); t1 = fmaf(nc.y, ix, bx); n1 = fminf(t0, t1) - ex; f1 = fmaxf(t0, t1) + ex;
                            t0 = fmaf(nb.w, iy, by); t1 = fmaf(nc.z, iy, by); n1 = fmaxf(n1, fminf(t0, t1) - ey); f1 = fminf(f1, fmaxf(t0, t1) + ey);
                            t0 = fmaf(nc.x, iz, bz); t1 = fmaf(nc.w, iz, bz); n1 = fmaxf(n1, fminf(t0, t1) - ez); f1 = fminf(f1, fmaxf(t0, t1) + ez);
                            // accepted roots lie in (t_min, rec.t]; a NaN interval bound (cannot arise from finite inputs) would pass
                            bool h0 = !(n0 > f0) && !(f0 < 0.f) && !(n0 > rec.t);
                            bool h1 = !(n1 > f1) && !(f1 < 0.f) && !(n1 > rec.t);
                            int c0 = ch.x, c1 = ch.y;
                            if (h0 && h1 && n1 < n0) { int t = c0; c0 = c1; c1 = t; }   // nearer child first
                            else if (!h0) { c0 = c1; h0 = h1; h1 = false; }
                            if (h0 && c0 < 0) { const int l = ~c0; hit_ordered(__ldg(&scene.leafShape[l]), (int)__ldg(&scene.leafDevice[l]), __ldg(&scene.leafOrig[l]), ray, a, rec); nTests++; h0 = false; }
                            if (h1 && c1 < 0) { const int l = ~c1; hit_ordered(__ldg(&scene.leafShape[l]), (int)__ldg(&scene.leafDevice[l]), __ldg(&scene.leafOrig[l]), ray, a, rec); nTests++; h1 = false; }
                            if (h0) { node = c0; if (h1) stack[sp++] = c1; continue; }
                            if (h1) { node = c1; continue; }
                            if (sp == 0) break;
                            node = stack[--sp];
                        }
                    }
                }
                if (rec.device < 0) { result = vmul(colour, sky(ray.d)); finished = true; break; }
                philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                if (!scatter(__ldg(&scene.shape[rec.device]), __ldg(&scene.surf[rec.device]), __ldg(&scene.kind[rec.device]), rec.t, ray, colour, u)) {
                    nBad++; finished = true; break;
                }
            }
            if (!finished) nExhausted++;
            if (FIXED) {
                long long* acc = args.fixedSums + 4 * (size_t)pixel;
                if (result.x != 0.f) atomicAdd((unsigned long long*)acc + 0, (unsigned long long)__float2ll_rn(result.x * kFixedOne));
                if (result.y != 0.f) atomicAdd((unsigned long long*)acc + 1, (unsigned long long)__float2ll_rn(result.y * kFixedOne));
                if (result.z != 0.f) atomicAdd((unsigned long long*)acc + 2, (unsigned long long)__float2ll_rn(result.z * kFixedOne));
            } else sum = vadd(sum, result);
        }
        if (!FIXED) args.sums[(size_t)chunk * pixels + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);
    }
    // one atomic per warp per counter
    for (int off = 16; off > 0; off >>= 1) {
        nRays += __shfl_down_sync(0xffffffffu, nRays, off); nNodes += __shfl_down_sync(0xffffffffu, nNodes, off);
        nTests += __shfl_down_sync(0xffffffffu, nTests, off);
        nExhausted += __shfl_down_sync(0xffffffffu, nExhausted, off); nBad += __shfl_down_sync(0xffffffffu, nBad, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&args.counters[CTR_RAYS], nRays); atomicAdd(&args.counters[CTR_NODES], nNodes); atomicAdd(&args.counters[CTR_TESTS], nTests);
        if (nExhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)nExhausted);
        if (nBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)nBad);
    }
}

}  // namespace vcrt
