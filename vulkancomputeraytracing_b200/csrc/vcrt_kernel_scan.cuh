/* @file vcrt_kernel_scan.cuh

    VCRT_TRAVERSAL_BRUTE: the persistent brute-force path tracer (the hot path).
    SPDX-License-Identifier: WTFPL

    Shape of the kernel
      * one CTA per SM, resident for the whole render; the scene SoA blob is staged into shared
        memory once by a TMA bulk copy (cp.async.bulk + mbarrier) and stays there;
      * every thread owns R = 2*PAIRS ray slots.  A slot is a pixel-sample in flight: it is
        traced bounce by bounce, and the moment its path ends the slot starts the next sample
        of its work item (pixel, sample chunk), or takes a new item from a global dispenser,
        so the inner loop never runs with dead lanes (no Russian roulette, fixed max depth,
        exactly the reference's loop semantics, functions.glsl:73-90);
      * per bounce, all slots of a thread walk the whole sphere list once ("scan").  The scan
        does NOT evaluate the reference's quadratic; it runs a conservative FILTER: with (u,v)
        an orthonormal basis of the plane normal to the ray, a sphere can only be hit if
        ((c-o).u)^2 + ((c-o).v)^2 <= r^2.  That is 8 fused multiply-adds per ray per sphere,
        issued as packed FFMA2 (two rays per instruction, sphere scalars broadcast by the
        .F32 operand form).  r^2 is inflated by an error bound (see cudaUploadScene) so the
        filter can only pass too many spheres, never too few;
      * spheres that pass are appended to a per-slot candidate list in shared memory; after the
        scan each slot evaluates its few candidates with the reference's exact binary32
        arithmetic (hit_exact), so the image is bit-identical to VCRT_TRAVERSAL_SIMPLE.
    Ray state that is only touched between scans (origin, direction, throughput, ids) lives in
    shared memory as a per-thread SoA; the scan itself runs out of registers.
*/
#pragma once
#include "vcrt_shade.cuh"

namespace vcrt {

// ------------------------------------------------------------------ TMA bulk copy + mbarrier
VCRT_DEV uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
VCRT_DEV void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
VCRT_DEV void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
VCRT_DEV void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
VCRT_DEV void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra.uni WAIT_DONE;\n\t"
        "bra.uni WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// Scene blob layout in global AND shared memory (all sections 16-byte aligned), see cudaUploadScene.
struct SceneBlob {
    uint32_t count, scanCount, scanPad, alwaysCount;
    uint32_t offHot, offShape, offSurf, offKind;  // byte offsets from the blob start
    uint32_t bytes, offOrig, pad[2];   // offOrig: uint32 per device index = index in the caller's array
};

// Per-slot filter state derived from the ray (o, d):
//   ux..vz : orthonormal basis of the plane normal to d (Duff et al. 2017 construction)
//   ou, ov : -(o.u), -(o.v)  so that  (c-o).u = c.u + ou
//   m      : per-ray part of the error margin, kappa * |o|^2
struct FilterRay { float ux, uy, uz, vx, vy, vz, ou, ov, m; bool degenerate; };

constexpr float kKappa = 7.62939453125e-06f;  // 2^-17, see cudaUploadScene for the derivation

VCRT_DEV FilterRay make_filter(const Ray& r) {
    FilterRay f;
    float len2 = fmaf(r.d.x, r.d.x, fmaf(r.d.y, r.d.y, r.d.z * r.d.z));
    float oo = fmaf(r.o.x, r.o.x, fmaf(r.o.y, r.o.y, r.o.z * r.o.z));
    f.degenerate = !tame_ray(len2, oo);
    float inv = rsqrtf(len2);
    float dx = r.d.x * inv, dy = r.d.y * inv, dz = r.d.z * inv;
    float sign = copysignf(1.0f, dz);
    float a = __fdividef(-1.0f, sign + dz);   // approximate is fine: the margin covers basis error
    float b = dx * dy * a;
    f.ux = 1.0f + sign * dx * dx * a; f.uy = sign * b; f.uz = -sign * dx;
    f.vx = b; f.vy = sign + dy * dy * a; f.vz = -dy;
    f.ou = -fmaf(r.o.x, f.ux, fmaf(r.o.y, f.uy, r.o.z * f.uz));
    f.ov = -fmaf(r.o.x, f.vx, fmaf(r.o.y, f.vy, r.o.z * f.vz));
    f.m = kKappa * oo;
    return f;
}

}  // namespace vcrt

namespace vcrt {

// Shared-memory plan of one CTA (THREADS threads, R slots each):
//   [mbarrier | scene blob | slot state SoA: 16 words x R x THREADS | lists u16 CAP x R x THREADS | candidate counts]
enum SlotField { SF_OX = 0, SF_OY, SF_OZ, SF_DX, SF_DY, SF_DZ, SF_CR, SF_CG, SF_CB, SF_SUMR, SF_SUMG, SF_SUMB, SF_PIXEL, SF_SAMPLE, SF_END, SF_PASS, SF_COUNT };

template <int THREADS, int PAIRS, int CAP>
struct ScanConfig {
    static constexpr int R = 2 * PAIRS;
    static constexpr int kStateBytes = SF_COUNT * R * THREADS * 4;
    static constexpr int kListBytes = (CAP + 2) * R * THREADS * 2;  // CAP entries + one 32-bit count per slot
    static __host__ __device__ size_t smem_bytes(uint32_t blobBytes) { return 16 + (size_t)blobBytes + kStateBytes + kListBytes; }
};

constexpr uint32_t kIdle = 0xffffffffu;

// Item id -> pixel.  Items walk the image in 8x4 tiles so the 32 lanes of a warp start on one tile.
VCRT_DEV bool item_to_pixel(const ItemPlan& plan, uint32_t width, uint32_t height, uint32_t id, uint32_t& px, uint32_t& py) {
    uint32_t tile = id >> 5, within = id & 31u, tx, ty;
    tile_to_xy(plan, tile, tx, ty);
    px = tx * 8u + (within & 7u);
    py = ty * 4u + (within >> 3);
    return px < width && py < height;
}

template <int THREADS, int PAIRS, int CAP, bool FIXED>
__global__ void __launch_bounds__(THREADS, 1) render_scan_kernel(const unsigned char* __restrict__ blobGlobal, RenderArgs args) {
    using Cfg = ScanConfig<THREADS, PAIRS, CAP>;
    constexpr int R = Cfg::R;
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    unsigned char* blob = smem + 16;
    const int tid = threadIdx.x;

    // ---- stage the scene: one TMA bulk copy, everyone waits on the mbarrier
    const uint32_t blobBytes = reinterpret_cast<const SceneBlob*>(blobGlobal)->bytes;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(bar, blobBytes);
        tma_bulk_g2s(blob, blobGlobal, blobBytes, bar);
    }
    mbar_wait(bar, 0);
    const SceneBlob hdr = *reinterpret_cast<const SceneBlob*>(blob);
    const float4* sHot = reinterpret_cast<const float4*>(blob + hdr.offHot);
    const float4* sShape = reinterpret_cast<const float4*>(blob + hdr.offShape);
    const float4* sSurf = reinterpret_cast<const float4*>(blob + hdr.offSurf);
    const int* sKind = reinterpret_cast<const int*>(blob + hdr.offKind);
    uint32_t* sState = reinterpret_cast<uint32_t*>(blob + blobBytes);
    unsigned short* sList = reinterpret_cast<unsigned short*>(blob + blobBytes + Cfg::kStateBytes);
#define ST(field, r) sState[((field) * R + (r)) * THREADS + tid]
#define STF(field, r) __uint_as_float(ST(field, r))
#define LIST(e, r) sList[((e) * R + (r)) * THREADS + tid]
    uint32_t* sCount = reinterpret_cast<uint32_t*>(sList + CAP * R * THREADS);
#define CNT(r) sCount[(r) * THREADS + tid]

    // Work item = (pixel, run of consecutive samples).  Float modes: args.chunks runs per pixel, each summed in
    // sample order into its own plane.  Fixed-point mode: short runs (1 sample unless the item count needs more), every sample added to the
    // pixel's 32.32 integer sums with a RED (order-independent, so any schedule gives the same bits).
    const uint32_t sppPerChunk = args.plan.sppPerRun, totalItems = args.plan.totalItems;   // host-made plan (vcrt_device.cuh)
    const uint32_t planeStride = args.width * args.height;

    unsigned long long nRays = 0;
    uint32_t nExhausted = 0, nBad = 0, nCand = 0, nOverflow = 0;

    // ---- slot bookkeeping helpers (state lives in shared memory; each is instantiated exactly once)
    // Claims items until one maps to a real pixel with a non-empty sample range; false when the dispenser is dry.
    auto claim_item = [&](int r, uint32_t firstId, bool useFirst) -> bool {
        uint32_t id = firstId; bool have = useFirst;
        for (;;) {
            if (!have) id = atomicAdd(args.workCounter, 1u);
            have = false;
            if (id >= totalItems) { ST(SF_PIXEL, r) = kIdle; return false; }
            uint32_t chunk, pid, px, py;
            item_to_run(args.plan, id, chunk, pid);
            if (!item_to_pixel(args.plan, args.width, args.height, pid, px, py)) continue;
            uint32_t s0 = chunk * sppPerChunk, s1 = min(args.spp, s0 + sppPerChunk);
            if (s0 >= s1) continue;
            ST(SF_PIXEL, r) = py * args.width + px;
            ST(SF_SAMPLE, r) = s0;
            ST(SF_END, r) = s1;
            return true;
        }
    };
    // Starts the current sample of slot r: camera ray, throughput 1, pass 0 (shader.comp:46-52).
    auto start_sample = [&](int r) {
        uint32_t pixel = ST(SF_PIXEL, r), sample = args.sampleOffset + ST(SF_SAMPLE, r);
        uint32_t py = pixel / args.width, px = pixel - py * args.width, rnd[4];
        philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
        Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
        ST(SF_OX, r) = __float_as_uint(ray.o.x); ST(SF_OY, r) = __float_as_uint(ray.o.y); ST(SF_OZ, r) = __float_as_uint(ray.o.z);
        ST(SF_DX, r) = __float_as_uint(ray.d.x); ST(SF_DY, r) = __float_as_uint(ray.d.y); ST(SF_DZ, r) = __float_as_uint(ray.d.z);
        ST(SF_CR, r) = ST(SF_CG, r) = ST(SF_CB, r) = __float_as_uint(1.0f);
        ST(SF_PASS, r) = 0u;
    };

    const int scanPad = (int)hdr.scanPad, scanCount = (int)hdr.scanCount, count = (int)hdr.count;
    const float kNegInf = __int_as_float(0xff800000);

    // q[r] = ((c-o).u)^2 + ((c-o).v)^2 - R2 for the R rays of this thread against sphere s: 8 FFMA2 per ray pair.
#define VCRT_FILTER(s, q)                                                                                   \
    {                                                                                                       \
        const u64 cx = pack2(s.x, s.x), cy = pack2(s.y, s.y), cz = pack2(s.z, s.z), nr2 = pack2(-s.w, -s.w); \
        _Pragma("unroll") for (int p = 0; p < PAIRS; p++) {                                                 \
            u64 pu = ffma2(cz, uz[p], ou[p]); pu = ffma2(cy, uy[p], pu); pu = ffma2(cx, ux[p], pu);          \
            u64 pv = ffma2(cz, vz[p], ov[p]); pv = ffma2(cy, vy[p], pv); pv = ffma2(cx, vx[p], pv);          \
            u64 qq = ffma2(pu, pu, nr2); qq = ffma2(pv, pv, qq);                                            \
            unpack2(qq, q[2 * p], q[2 * p + 1]);                                                            \
        }                                                                                                   \
    }

    for (bool firstRound = true;; firstRound = false) {
        if (!firstRound) {
            // ---- (1) filter state of every slot, into registers
            u64 ux[PAIRS], uy[PAIRS], uz[PAIRS], vx[PAIRS], vy[PAIRS], vz[PAIRS], ou[PAIRS], ov[PAIRS];
            float m[R]; uint32_t cnt[R];
            bool anyActive = false;
#pragma unroll
            for (int p = 0; p < PAIRS; p++) {
                FilterRay f[2];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int r = 2 * p + h;
                    f[h].ux = f[h].uy = f[h].uz = f[h].vx = f[h].vy = f[h].vz = f[h].ou = f[h].ov = 0.f; f[h].m = kNegInf;
                    cnt[r] = 0u;
                    if (ST(SF_PIXEL, r) != kIdle) {
                        anyActive = true;
                        Ray ray; ray.o = v3(STF(SF_OX, r), STF(SF_OY, r), STF(SF_OZ, r)); ray.d = v3(STF(SF_DX, r), STF(SF_DY, r), STF(SF_DZ, r));
                        FilterRay g = make_filter(ray);
                        if (g.degenerate) cnt[r] = CAP + 1u;   // no usable filter: exhaustive exact scan for this ray
                        else f[h] = g;
                    }
                    m[r] = f[h].m;
                }
                ux[p] = pack2(f[0].ux, f[1].ux); uy[p] = pack2(f[0].uy, f[1].uy); uz[p] = pack2(f[0].uz, f[1].uz);
                vx[p] = pack2(f[0].vx, f[1].vx); vy[p] = pack2(f[0].vy, f[1].vy); vz[p] = pack2(f[0].vz, f[1].vz);
                ou[p] = pack2(f[0].ou, f[1].ou); ov[p] = pack2(f[0].ov, f[1].ov);
            }
            if (!anyActive) break;
            float mThread = m[0];
#pragma unroll
            for (int r = 1; r < R; r++) mThread = fmaxf(mThread, m[r]);

            // ---- (2) the scan.  Straight-line blocks of up to 32 spheres: per sphere one warp-uniform LDS.128,
            //      8 FFMA2 per ray pair and a min; survivors only set a bit.  Flagged spheres are revisited after the block.
#pragma unroll 1
            for (int base = 0; base < scanPad; base += 32) {
                uint32_t mask = 0u;
                const int lim = min(32, scanPad - base);   // scanPad is a multiple of 8
#pragma unroll 1
                for (int kk = 0; kk < lim; kk += 8) {
                    uint32_t m8 = 0u;
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const float4 s = sHot[base + kk + u];
                        float q[R];
                        VCRT_FILTER(s, q);
                        float qmin = q[0];
#pragma unroll
                        for (int r = 1; r < R; r++) qmin = fminf(qmin, q[r]);
                        if (qmin <= mThread) m8 |= 1u << u;
                    }
                    mask |= m8 << kk;
                }
                while (mask) {
                    const int k = __ffs(mask) - 1;
                    mask &= mask - 1u;
                    const int j = base + k;
                    const float4 s = sHot[j];
                    float q[R];
                    VCRT_FILTER(s, q);
#pragma unroll
                    for (int r = 0; r < R; r++)
                        if (q[r] <= m[r]) { LIST(min(cnt[r], (uint32_t)(CAP - 1)), r) = (unsigned short)j; cnt[r]++; }
                }
            }
#pragma unroll
            for (int r = 0; r < R; r++) CNT(r) = cnt[r];
        }

        // ---- (3) per slot: exact test of its candidates, shade, and (re)generate.  One copy of every routine.
#pragma unroll 1
        for (int r = 0; r < R; r++) {
            bool needClaim = firstRound, needStart = false;
            const uint32_t pixel = firstRound ? kIdle : ST(SF_PIXEL, r);
            if (pixel != kIdle) {
                Ray ray; ray.o = v3(STF(SF_OX, r), STF(SF_OY, r), STF(SF_OZ, r)); ray.d = v3(STF(SF_DX, r), STF(SF_DY, r), STF(SF_DZ, r));
                const float a = vdot(ray.d, ray.d);
                HitRec rec; rec.t = kTMax; rec.index = -1;
                const uint32_t nc = CNT(r);
                const bool exhaustive = nc > (uint32_t)CAP;
                const bool tame = tame_ray(fmaf(ray.d.x, ray.d.x, fmaf(ray.d.y, ray.d.y, ray.d.z * ray.d.z)), fmaf(ray.o.x, ray.o.x, fmaf(ray.o.y, ray.o.y, ray.o.z * ray.o.z)));   // as make_filter
                nRays++;
                if (exhaustive) nOverflow++; else nCand += nc;
                // candidates in scan order, then the "always" spheres (device indices scanCount..count-1)
                const int nScan = exhaustive ? scanCount : (int)nc, nAll = nScan + (count - scanCount);
                for (int e = 0; e < nAll; e++) {
                    const int j = e >= nScan ? scanCount + (e - nScan) : (exhaustive ? e : (int)LIST(e, r));
                    if (VCRT_CHECK(args.counters, j >= 0 && j < count && (exhaustive || e >= nScan || e < CAP))) hit_exact(sShape[j], j, ray, a, rec, tame);
                }
                V3 colour = v3(STF(SF_CR, r), STF(SF_CG, r), STF(SF_CB, r));
                V3 result = v3(0.f, 0.f, 0.f);   // black: bug B7 (depth exhausted) and absorbed bad materials
                bool ended = true;
                if (rec.index < 0) {
                    if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_PATH0 + min(ST(SF_PASS, r), 7u)], 1ull);
                    result = vmul(colour, sky(ray.d));
                } else {
                    const uint32_t pass = ST(SF_PASS, r);
                    if (args.flags & kFlagCountStats) { const int k = sKind[rec.index]; atomicAdd(&args.counters[CTR_HITS0 + ((k >= 1 && k <= 3) ? k : 0)], 1ull); }
                    uint32_t rnd[4];
                    philox4x32_10(pixel, args.sampleOffset + ST(SF_SAMPLE, r), pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                    float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                    int outcome;   // 0 = unknown material, 1 = scattered, 2 = absorbed by the material (BOOK metal)
                    if (args.semantics) {
                        BookDraws g; g.pixel = pixel; g.sample = args.sampleOffset + ST(SF_SAMPLE, r); g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                        outcome = scatter_book(sShape[rec.index], sSurf[rec.index], sKind[rec.index], rec.t, ray, colour, u, g);
                    } else outcome = scatter(sShape[rec.index], sSurf[rec.index], sKind[rec.index], rec.t, ray, colour, u) ? 1 : 0;
                    if (outcome == 0) nBad++;
                    else if (outcome == 2) { }
                    else if (pass + 1u >= args.maxDepth) nExhausted++;
                    else {
                        ended = false;
                        ST(SF_OX, r) = __float_as_uint(ray.o.x); ST(SF_OY, r) = __float_as_uint(ray.o.y); ST(SF_OZ, r) = __float_as_uint(ray.o.z);
                        ST(SF_DX, r) = __float_as_uint(ray.d.x); ST(SF_DY, r) = __float_as_uint(ray.d.y); ST(SF_DZ, r) = __float_as_uint(ray.d.z);
                        ST(SF_CR, r) = __float_as_uint(colour.x); ST(SF_CG, r) = __float_as_uint(colour.y); ST(SF_CB, r) = __float_as_uint(colour.z);
                        ST(SF_PASS, r) = pass + 1u;
                    }
                }
                if (ended) {   // shader.comp:53: color.rgb += ray_color(r), in sample order within the item
                    uint32_t s = ST(SF_SAMPLE, r);
                    const uint32_t s1 = ST(SF_END, r);
                    if (FIXED) {
                        add_fixed(args, pixel, result);
                        s++;
                    } else {
                        const uint32_t chunk = (s1 - 1u) / sppPerChunk, s0 = chunk * sppPerChunk;
                        V3 sum = (s == s0) ? v3(0.f, 0.f, 0.f) : v3(STF(SF_SUMR, r), STF(SF_SUMG, r), STF(SF_SUMB, r));
                        sum = vadd(sum, result);
                        s++;
                        if (s < s1) { ST(SF_SUMR, r) = __float_as_uint(sum.x); ST(SF_SUMG, r) = __float_as_uint(sum.y); ST(SF_SUMB, r) = __float_as_uint(sum.z); }
                        else if (VCRT_CHECK(args.counters, chunk < args.chunks && pixel < planeStride)) args.sums[(size_t)chunk * planeStride + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);   // the item's only global write
                    }
                    if (s < s1) { ST(SF_SAMPLE, r) = s; needStart = true; } else needClaim = true;
                }
            }
            if (needClaim) needStart = claim_item(r, (blockIdx.x * R + r) * THREADS + tid, firstRound);
            if (needStart) start_sample(r);
        }
    }

    atomicAdd(&args.counters[CTR_RAYS], nRays);
    if (nCand) atomicAdd(&args.counters[CTR_CANDIDATES], (unsigned long long)nCand);
    if (nExhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)nExhausted);
    if (nBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)nBad);
    if (nOverflow) atomicAdd(&args.counters[CTR_OVERFLOWS], (unsigned long long)nOverflow);
#undef VCRT_FILTER
#undef ST
#undef STF
#undef LIST
#undef CNT
}

}  // namespace vcrt
