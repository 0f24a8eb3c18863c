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
        K = 32 eps (L^2 + rmax^2), L >= |o - c| for every sphere this ray can still accept, plus
        16 eps (|o| + F) for the rounding of the slab arithmetic itself (F = the largest |o - c|
        over the whole tree).
        Where 32 comes from (u = eps = 2^-24, every operation of hit_ordered rounded once, L = |o - c|):
        oc = (o - c)(1 + u);  |half_b - (o-c).d| <= 4u L|d|;  |fl(oc.oc) - L^2| <= 5u L^2;
        |cc - (L^2 - r^2)| <= 6u L^2 + 2u r^2;  |fl(half_b^2) - ((o-c).d)^2| <= 9u a L^2;
        |fl(a cc) - a(L^2 - r^2)| <= a (10u L^2 + 6u r^2);  the last subtraction keeps the sign.  So
        disc >= 0 needs D^2 <= r^2 + 19u L^2 + 6u r^2, which puts the point of closest approach inside
        the sphere's box inflated by sqrt(r^2 + K19) - r.  The traversal also needs the ACCEPTED point
        p(t_s) inside every inflated ancestor box (so that n <= t_s <= rec.t and f >= t_s > 0):
        (t_s - t_c)|d| <= 9u L + sqrt(r^2 - D^2 + K19)(1 + 5u) gives |p(t_s) - c|^2 <= r^2 + 28u L^2 + 26u r^2.
        32 covers both with room for the second-order terms.  (The first-generation kernel keeps 64
        and L = F.)
        Distance-aware widening (second-generation kernel): L = F is always valid but wasteful - at
        100 k spheres F = 230 and grow = 0.46 on spheres of radius 0.2, so every box was 3.3x too
        wide.  A sphere is only accepted with a computed root <= t0 (the best root so far: the
        ground is tested before the tree), and root >= t_c - 4 eps L/|d| - sqrt(r^2 + K)/|d| where
        t_c is the parameter of closest approach, L^2 = t_c^2 |d|^2 + D^2 and sqrt(K) <= 2^-9 (L + r);
        together L <= 1.004 (t0 |d| + 2.01 r).  Likewise the point of closest approach lies within
        D <= E = sqrt(rmax^2 + K(F)) of the centre, i.e. inside the box of the centres inflated by
        E, which the ray leaves at t_exit: L <= t_exit |d| + E.  The kernel uses
        L = min(F, 1.01 min(t0, t_exit) |d| + 2.1 rmax + 1.01 E);
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
    const BvhNode* octNodes;      // [8][treeCount-1]: the same tree with near / far planes pre-selected per direction octant
    const float4* leafShape;      // sorted order
    const uint32_t* leafOrig;
    const uint32_t* leafDevice;
    const LeafRecord* leafRecord; // (shape, device index, original index) of a sorted leaf in one 32-byte sector
    const float4* shape;          // device order (exact tests of the always list, normals)
    const float4* surf;
    const int* kind;
    const uint32_t* origIndex;
    uint32_t treeCount;           // spheres in the tree
    uint32_t count;               // all spheres
    uint32_t scanCount;           // device indices [0, scanCount) and [scanCount, count) are each in the caller's order
    const uint32_t* extra;        // device indices of the spheres outside the tree (outliers, then the always list): tested for every ray
    uint32_t extraCount;
    float3 centreMin, centreMax;  // bounding box of the tree spheres' centres
    float radiusMin, radiusMax;   // over the tree spheres
    float halfInvRadiusMin;       // >= 1 / (2 radiusMin), rounded up
};

// Resident CTAs of 256 per SM that the REGISTER BUDGET of the default variant (fixed-point sums, REFERENCE semantics) is set for.
// With 3 (<= 85 registers) ptxas still allocates 64, so four CTAs stay resident, but it schedules the kernel differently: +2.6 % at
// 485 / 10 k / 100 k spheres and on the stress scene against a budget of 4 (2: +1.6 %, 5: 92 bytes of spills, -0.6 %).  The
// strict-order and BOOK variants take 69 / 80 registers under that budget and lose 1 %, so they keep 4.
#ifndef LBVH_BLOCKS
#define LBVH_BLOCKS 3
#endif
constexpr int lbvh_min_blocks(bool fixed, bool book) { return fixed && !book ? LBVH_BLOCKS : 4; }

// The spheres outside the tree (outliers, then the always list; BvhScene::extra) as records in the constant bank: every ray tests
// all of them, so the loads are warp-uniform - LDC / LDCU instead of an index load and two dependent global loads per sphere.
constexpr uint32_t kMaxBvhExtras = 80u;   // 16 outliers + 64 always spheres (vcrt_cabi.cu)
__constant__ LeafRecord cLbvhExtras[kMaxBvhExtras];

struct OrderedHit { float t; int device; uint32_t orig; };

// Candidate root of one sphere, independent of what was found before (see the file header).
VCRT_DEV void hit_ordered(float4 shape, int device, uint32_t orig, const Ray& r, float a, OrderedHit& rec) {
    V3 oc = vsub(r.o, v3(shape.x, shape.y, shape.z));
    float half_b = vdot(oc, r.d);
    float c = fsub(vdot(oc, oc), fmul(shape.w, shape.w));
    float disc = fsub(fmul(half_b, half_b), fmul(a, c));
    if (disc < 0) return;
    if (half_b > 1e-18f && c >= 0.f) return;   // both roots <= 0 exactly, see hit_exact
    float sqrtd = fsqrt(disc);
    float root = fdiv(fsub(-half_b, sqrtd), a);
    if (root <= kTMin) {   // only then can the far root be the accepted one (it is never smaller than the near root)
        root = fdiv(fadd(-half_b, sqrtd), a);
        if (root <= kTMin) return;
    }
    if (!(root < rec.t || (root == rec.t && orig < rec.orig))) return;
    rec.t = root; rec.device = device; rec.orig = orig;
}

// ---------------------------------------------------------------------------------------------------
// Second generation: the same path tracer as ONE flat loop per thread.  Every trip of the loop traces
// exactly one ray per lane; a lane whose path has ended (sky, absorbed, depth) starts its next sample
// (or takes the next work item) in the same trip instead of idling until the longest path of its warp
// is done.  The image cannot change: a path's arithmetic is the same sequence of operations, samples
// of one run are still summed in order by the thread that owns the run (strict mode), and the
// fixed-point sums do not depend on the order at all.
#ifndef LBVH_REFILL
#define LBVH_REFILL 16   // 32 = an idle lane restarts at once; 16 measured best
#endif

#ifndef LBVH_PREFETCH
#define LBVH_PREFETCH 1   // the far child that goes onto the stack is prefetched into L1 (+2 % at 485 / 10 k / 100 k spheres)
#endif
#ifndef LBVH_K_EPS
#define LBVH_K_EPS 32.0f
#endif
constexpr float kLbvhK = LBVH_K_EPS * 5.9604644775390625e-08f;   // K = 32 eps (L^2 + rmax^2), eps = 2^-24 (file header)

VCRT_DEV float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
// an upper-bound-friendly approximate root in ONE instruction: the .ftz form needs no subnormal scaling around MUFU.SQRT (3 extra
// instructions each, 5 roots per ray), and clamping the argument to the smallest normal keeps the result >= the true root where
// .ftz would have flushed it to 0 (the roots only ever ADD to the widening: too large is conservative)
VCRT_DEV float sqrt_approx(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(fmaxf(x, 1.17549435e-38f))); return r; }

struct LbvhCounts { uint32_t rays, nodes, tests, exhausted, bad; unsigned long long* counters; };

VCRT_DEV void lbvh_nearest(const BvhScene& scene, const Ray& ray, float a, OrderedHit& rec, int* stack, LbvhCounts& n) {
    const float oo = fmaf(ray.o.x, ray.o.x, fmaf(ray.o.y, ray.o.y, ray.o.z * ray.o.z));
    const bool degenerate = !tame_ray(a, oo);
    if (degenerate) {   // the reference's in-order scan, literally
        HitRec h; h.t = kTMax; h.index = -1;
        uint32_t i = 0, j = scene.scanCount;
        while (i < scene.scanCount || j < scene.count) {
            const bool takeTree = j >= scene.count || (i < scene.scanCount && scene.origIndex[i] < scene.origIndex[j]);
            const uint32_t k = takeTree ? i++ : j++;
            hit_exact(scene.shape[k], (int)k, ray, a, h, false);
            n.tests++;
        }
        rec.t = h.t; rec.device = h.index;
        return;
    }
    for (uint32_t x = 0; x < scene.extraCount; x++) {
        if (!VCRT_CHECK(n.counters, x < kMaxBvhExtras)) break;
        hit_ordered(cLbvhExtras[x].shape, (int)cLbvhExtras[x].device, cLbvhExtras[x].orig, ray, a, rec); n.tests++;
    }
    if (scene.treeCount == 1u) { hit_ordered(__ldg(&scene.leafShape[0]), (int)__ldg(&scene.leafDevice[0]), __ldg(&scene.leafOrig[0]), ray, a, rec); n.tests++; }
    if (scene.treeCount <= 1u) return;
    // per-ray widening of every box and the slab constants: as in render_lbvh_kernel
    const float fx = fmaxf(fabsf(ray.o.x - scene.centreMin.x), fabsf(ray.o.x - scene.centreMax.x));
    const float fy = fmaxf(fabsf(ray.o.y - scene.centreMin.y), fabsf(ray.o.y - scene.centreMax.y));
    const float fz = fmaxf(fabsf(ray.o.z - scene.centreMin.z), fabsf(ray.o.z - scene.centreMax.z));
    const float F2 = fmaf(fx, fx, fmaf(fy, fy, fz * fz));
    const float rr = scene.radiusMax * scene.radiusMax;
    const bool px_ = ray.d.x >= 0.f, py_ = ray.d.y >= 0.f, pz_ = ray.d.z >= 0.f;
    const bool ux = fabsf(ray.d.x) > 1e-30f, uy = fabsf(ray.d.y) > 1e-30f, uz = fabsf(ray.d.z) > 1e-30f;
    const float ix = ux ? rcp_approx(ray.d.x) : 0.f, iy = uy ? rcp_approx(ray.d.y) : 0.f, iz = uz ? rcp_approx(ray.d.z) : 0.f;
    const float kInf = __int_as_float(0x7f800000);
    // How far away can the centre of a sphere be that this ray may still ACCEPT?  Not F (the far corner of the whole tree) but
    // L <= 1.004 (t* |d| + 2.01 rmax) with t* = min(best root so far - the ground has been tested -, parameter at which the ray
    // leaves the box of the centres inflated by E >= sqrt(rmax^2 + K(F))): file header, "distance-aware widening".  The widening
    // grows with L^2, so near the origin the boxes stay tight: at 100 k spheres F = 230 gives 0.46 on spheres of radius 0.2,
    // while a ray that meets the ground 20 units away needs 0.004.
    const float sqrtF = sqrt_approx(F2);
    const float sqrtKF = sqrt_approx(kLbvhK * (F2 + rr));
    float widen = fminf(sqrtKF, sqrtKF * sqrtKF * scene.halfInvRadiusMin);   // with L = F
    if (widen > 0.05f * scene.radiusMin) {   // a small scene needs no refinement (485 spheres: 3.7 % of the radius)
        const float E = scene.radiusMax + sqrtKF;
        const float ex = ux ? ((px_ ? scene.centreMax.x + E : scene.centreMin.x - E) - ray.o.x) * ix : kInf;
        const float ey = uy ? ((py_ ? scene.centreMax.y + E : scene.centreMin.y - E) - ray.o.y) * iy : kInf;
        const float ez = uz ? ((pz_ ? scene.centreMax.z + E : scene.centreMin.z - E) - ray.o.z) * iz : kInf;
        const float tStar = fminf(rec.t, fmaxf(fminf(fminf(ex, ey), ez), 0.f));
        const float L = fminf(sqrtF, fmaf(1.01f * tStar, sqrt_approx(a), fmaf(2.1f, scene.radiusMax, 1.01f * E)));
        const float K = kLbvhK * fmaf(L, L, rr);
        widen = fminf(sqrt_approx(K), K * scene.halfInvRadiusMin);
    }
    // the bound only has to be conservative, so it is built from the approximate unit's sqrt / rcp (relative error <= 2^-22,
    // against the safety factors above and a slab allowance of 16 eps where ~5 eps are needed, see the file header)
    const float grow = (widen + 9.5367431640625e-07f * (sqrt_approx(oo) + sqrtF)) * 1.0001f;
    const float bnx = ux ? fmaf(-ray.o.x, ix, -grow * fabsf(ix)) : -kInf, bfx = ux ? fmaf(-ray.o.x, ix, grow * fabsf(ix)) : kInf;
    const float bny = uy ? fmaf(-ray.o.y, iy, -grow * fabsf(iy)) : -kInf, bfy = uy ? fmaf(-ray.o.y, iy, grow * fabsf(iy)) : kInf;
    const float bnz = uz ? fmaf(-ray.o.z, iz, -grow * fabsf(iz)) : -kInf, bfz = uz ? fmaf(-ray.o.z, iz, grow * fabsf(iz)) : kInf;
    // planes picked once per ray: the octant's copy of the node array (vcrt_lbvh.cuh)
    const uint32_t octant = (px_ ? 0u : 1u) | (py_ ? 0u : 2u) | (pz_ ? 0u : 4u);
    uint32_t first = octant * (scene.treeCount - 1u);   // 8 (count - 1) < 2^32: the ABI caps a scene far below that
    asm volatile("" : "+r"(first));                      // keep it in a register: recomputing it per visit costs 8 instructions
    const BvhNode* nodes = scene.octNodes;
    const u64 ix2 = pack2(ix, ix), iy2 = pack2(iy, iy), iz2 = pack2(iz, iz);
    const u64 bnx2 = pack2(bnx, bnx), bny2 = pack2(bny, bny), bnz2 = pack2(bnz, bnz), bfx2 = pack2(bfx, bfx), bfy2 = pack2(bfy, bfy), bfz2 = pack2(bfz, bfz);
    int sp = 0, node = 0;
    for (;;) {
        if (node >= 0) {
            n.nodes++;
            if (!VCRT_CHECK(n.counters, (uint32_t)node < scene.treeCount - 1u)) break;
            const float4 na = __ldg(&nodes[first + (uint32_t)node].a), nb = __ldg(&nodes[first + (uint32_t)node].b), nc = __ldg(&nodes[first + (uint32_t)node].c);
            const int2 ch = __ldg(reinterpret_cast<const int2*>(&nodes[first + (uint32_t)node].child0));
#if LBVH_PREFETCH == 3
            if (ch.x >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"(&nodes[first + (uint32_t)ch.x]));
            if (ch.y >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"(&nodes[first + (uint32_t)ch.y]));
#endif
            // both children per packed FMA: na = (near0.x, near1.x, near0.y, near1.y), nb = (near0.z, near1.z, far0.x, far1.x),
            // nc = (far0.y, far1.y, far0.z, far1.z)
            float nx0, nx1, ny0, ny1, nz0, nz1, fx0, fx1, fy0, fy1, fz0, fz1;
            unpack2(ffma2(pack2(na.x, na.y), ix2, bnx2), nx0, nx1); unpack2(ffma2(pack2(na.z, na.w), iy2, bny2), ny0, ny1);
            unpack2(ffma2(pack2(nb.x, nb.y), iz2, bnz2), nz0, nz1); unpack2(ffma2(pack2(nb.z, nb.w), ix2, bfx2), fx0, fx1);
            unpack2(ffma2(pack2(nc.x, nc.y), iy2, bfy2), fy0, fy1); unpack2(ffma2(pack2(nc.z, nc.w), iz2, bfz2), fz0, fz1);
            const float n0 = fmaxf(fmaxf(nx0, ny0), nz0), f0 = fminf(fminf(fx0, fy0), fz0);
            const float n1 = fmaxf(fmaxf(nx1, ny1), nz1), f1 = fminf(fminf(fx1, fy1), fz1);
            const bool h0 = !(n0 > f0) && !(f0 < 0.f) && !(n0 > rec.t);   // written so that a NaN bound passes
            const bool h1 = !(n1 > f1) && !(f1 < 0.f) && !(n1 > rec.t);
            if (h0 && h1) {
                const bool swap = n1 < n0;
                node = swap ? ch.y : ch.x;
                const int far = swap ? ch.x : ch.y;
                if (VCRT_CHECK(n.counters, sp < 64)) stack[sp++] = far;
#if LBVH_PREFETCH >= 1
                if (far >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"(&nodes[first + (uint32_t)far]));
#if LBVH_PREFETCH == 2
                else asm volatile("prefetch.global.L1 [%0];" :: "l"(&scene.leafRecord[~far]));
#endif
#endif
                continue;
            }
            if (h0 || h1) { node = h0 ? ch.x : ch.y; continue; }
        } else {
            const int l = ~node;
            if (!VCRT_CHECK(n.counters, l >= 0 && (uint32_t)l < scene.treeCount)) break;
            const uint2 ids = __ldg(reinterpret_cast<const uint2*>(&scene.leafRecord[l].device));
            hit_ordered(__ldg(&scene.leafRecord[l].shape), (int)ids.x, ids.y, ray, a, rec);
            n.tests++;
        }
        if (sp == 0) break;
        node = stack[--sp];
    }
}

template <bool FIXED, bool STATS, bool BOOK = false>   // BOOK: VCRT_SEMANTICS_BOOK as a compile-time variant (see render_scanu_kernel)
__global__ void __launch_bounds__(256, lbvh_min_blocks(FIXED, BOOK)) render_lbvh_flat_kernel(BvhScene scene, RenderArgs args) {
    const uint32_t pixels = args.width * args.height;
    const uint32_t sppPerChunk = args.plan.sppPerRun;   // host-made plan (vcrt_device.cuh)
    const uint32_t totalItems = args.plan.totalItems;
    LbvhCounts n = {0u, 0u, 0u, 0u, 0u, args.counters};
    int stack[65];

    const uint32_t firstDynamic = gridDim.x * blockDim.x;
    uint32_t item = blockIdx.x * blockDim.x + threadIdx.x;
    bool firstItem = true, haveRun = false;
    uint32_t s = 0, s1 = 0, pixel = 0, px = 0, py = 0, chunk = 0;   // the run this thread owns
    uint32_t sample = 0, pass = 0;                                   // the path in flight
    Ray ray; ray.o = v3(0.f, 0.f, 0.f); ray.d = ray.o;
    V3 colour = v3(1.f, 1.f, 1.f), sum = v3(0.f, 0.f, 0.f);
    bool inFlight = false;

    bool drained = false;
    for (;;) {
        // idle lanes restart together once fewer than LBVH_REFILL lanes of the warp are still tracing
        const unsigned flying = __ballot_sync(0xffffffffu, inFlight);
        if (!inFlight && !drained && __popc(flying) < LBVH_REFILL) {   // start the next sample of the run, or take the next run
            for (;;) {
                if (haveRun && s < s1) break;
                if (haveRun) {
                    if (!FIXED && VCRT_CHECK(args.counters, chunk < args.chunks && pixel < pixels)) args.sums[(size_t)chunk * pixels + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);
                    haveRun = false;
                }
                if (!firstItem) item = firstDynamic + atomicAdd(args.workCounter, 1u);   // (an explicit warp-aggregated claim is 1.3 % slower)
                firstItem = false;
                if (item >= totalItems) break;
                uint32_t pid;
                item_to_run(args.plan, item, chunk, pid);
                const uint32_t tile = pid >> 5, within = pid & 31u;
                uint32_t tx, ty;
                tile_to_xy(args.plan, tile, tx, ty);
                px = tx * 8u + (within & 7u); py = ty * 4u + (within >> 3);
                if (px >= args.width || py >= args.height) continue;
                s = chunk * sppPerChunk; s1 = min(args.spp, s + sppPerChunk);
                if (s >= s1) continue;
                pixel = py * args.width + px;
                sum = v3(0.f, 0.f, 0.f);
                haveRun = true;
                if (args.maxDepth == 0u) { n.exhausted += s1 - s; s = s1; }   // no ray at all: every sample is black
            }
            if (!haveRun) drained = true;   // the dispenser is empty: this lane is done
            else {
                sample = args.sampleOffset + s; s++;
                uint32_t rnd[4];
                philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
                ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
                colour = v3(1.f, 1.f, 1.f);
                pass = 0u;
                inFlight = true;
            }
        }
        if (!__any_sync(0xffffffffu, inFlight)) break;   // nothing in flight after a refill attempt: every lane is drained
        if (!inFlight) continue;
        // one ray
        n.rays++;
        const float a = vdot(ray.d, ray.d);
        OrderedHit rec; rec.t = kTMax; rec.device = -1; rec.orig = 0xffffffffu;
        lbvh_nearest(scene, ray, a, rec, stack, n);
        V3 result = v3(0.f, 0.f, 0.f);
        if (rec.device < 0) {
            if constexpr (STATS) atomicAdd(&args.counters[CTR_PATH0 + min(pass, 7u)], 1ull);
            result = vmul(colour, sky(ray.d));
            inFlight = false;
        } else {
            if constexpr (STATS) { const int k = __ldg(&scene.kind[rec.device]); atomicAdd(&args.counters[CTR_HITS0 + ((k >= 1 && k <= 3) ? k : 0)], 1ull); }
            uint32_t rnd[4];
            philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
            float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
            int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
            if (BOOK) {
                BookDraws g; g.pixel = pixel; g.sample = sample; g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                outcome = scatter_book(__ldg(&scene.shape[rec.device]), __ldg(&scene.surf[rec.device]), __ldg(&scene.kind[rec.device]), rec.t, ray, colour, u, g);
            } else outcome = scatter(__ldg(&scene.shape[rec.device]), __ldg(&scene.surf[rec.device]), __ldg(&scene.kind[rec.device]), rec.t, ray, colour, u) ? 1 : 0;
            if (outcome != 1) { if (outcome == 0) n.bad++; inFlight = false; }
            else if (++pass >= args.maxDepth) { n.exhausted++; inFlight = false; }
        }
        if (!inFlight) {
            if (FIXED) {
                add_fixed(args, pixel, result);
            } else sum = vadd(sum, result);
        }
    }
    unsigned long long wRays = n.rays, wNodes = n.nodes, wTests = n.tests;
    uint32_t nExhausted = n.exhausted, nBad = n.bad;
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
