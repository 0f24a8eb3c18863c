/* @file vcrt_kernel_scanu.cuh

    VCRT_TRAVERSAL_BRUTE, second generation: the persistent brute-force path tracer with the sphere
    stream fed through the UNIFORM datapath.
    SPDX-License-Identifier: WTFPL

    Same shape as vcrt_kernel_scan.cuh (one CTA per SM, R = 2*PAIRS pixel-samples in flight per thread,
    per bounce one scan of the whole sphere list for all of a thread's rays, exact tests of the few
    survivors, immediate regeneration of finished slots), with three changes that the ncu capture and
    the microbenchmarks of round 1 asked for:

      1. The filter's sphere operands come from the constant bank through LDCU into uniform registers
         and enter FFMA2 as `UR.F32` operands.  A warp-uniform LDS.128 costs four cycles of
         register-file write bandwidth per sphere; the uniform path costs none
         (scan loop alone: 74 % -> 84 % of the FMA pipe).
      2. The filter is 7 packed FMAs per ray pair instead of 8, and carries its per-ray margin inside:
             q'(ray, sphere) = (c.dh + e)^2 - (c.A + B)    with per ray  dh = d/|d|, e = -o.dh,
                                                           A = -2o, B = |o|^2 - kappa|o|^2
             sphere passes  <=>  q' >= |c|^2 - R2           (one scalar per sphere, rounded down on the host)
         which is |c-o|^2 - ((c-o).dh)^2 <= R2 + kappa|o|^2, the squared distance from the centre to
         the ray's line against the inflated radius (R2 = r^2 + kappa(|c|^2 + r^2), kappa = 2^-16).
      3. Survivors go to ONE list per thread (not per ray) straight from the flag mask; every ray of
         the thread then runs the exact test on that short list.  The first generation re-evaluated
         the filter for each flagged sphere to find out which ray had flagged it (10 % of all
         instructions at 7/32 active lanes).

    Error budget of the filter (eps = 2^-24; see also cudaUploadScene): the exact test can only accept
    a sphere if its computed discriminant is >= 0, i.e. a(r^2 - D^2) >= -18 eps a(|oc|^2 + r^2) with D
    the true distance from c to the line.  The filter computes |c|^2 - 2c.o + |o|^2 - (c.dh - o.dh)^2 in
    expanded form: |dh| = 1 +- 4 eps (rsqrt), each dot product is off by <= 3 eps (|c|+|o|), the square
    and the sum add <= 40 eps (|c|+|o|)^2 in total.  With (|c|+|o|)^2 <= 2(|c|^2+|o|^2) and
    |oc|^2 <= 2(|c|^2+|o|^2) the two sides together are below 116 eps (|c|^2+|o|^2+r^2) < kappa = 256 eps
    of the same quantity, which is exactly what R2 and B add.  Extra survivors cost time, never
    correctness; tests/test_gpu_parity.py checks the consequence (bit-identical images).
*/
#pragma once
#include "vcrt_kernel_scan.cuh"

namespace vcrt {

constexpr int kMaxConstSpheres = 1792;                 // 28 KB of the 64 KB constant bank (the primary-visibility scan has a table of its own)
constexpr float kKappaU = 1.52587890625e-05f;          // 2^-16
__constant__ float4 cSceneHot[kMaxConstSpheres];       // (cx, cy, cz, |c|^2 - R2 rounded down), padded with (0,0,0,+inf)

template <int THREADS, int PAIRS, int LCAP>
struct ScanUConfig {
    static constexpr int R = 2 * PAIRS;
    static constexpr int kStateBytes = SF_COUNT * R * THREADS * 4;
    static constexpr int kListBytes = LCAP * THREADS * 2;
    static __host__ __device__ size_t smem_bytes(uint32_t blobBytes) { return 16 + (size_t)blobBytes + kStateBytes + kListBytes; }
};

// m |= bit when q >= s: one compare and one predicated OR (the plain C form costs a select and an add on top of the compare)
VCRT_DEV void flag_if_ge(uint32_t& m, float q, float s, uint32_t bit) {
    asm("{\n\t.reg .pred p;\n\tsetp.ge.f32 p, %1, %2;\n\t@p or.b32 %0, %0, %3;\n\t}" : "+r"(m) : "f"(q), "f"(s), "r"(bit));
}

#ifndef VCRT_SCANU_UNROLL
#define VCRT_SCANU_UNROLL 16   // spheres per unrolled scan body (32 measured: see DESIGN.md 6.1)
#endif
// BOOK: VCRT_SEMANTICS_BOOK as a compile-time variant - with both scatter routines in one kernel the REFERENCE path carried
// 32 bytes of spills at its 64-register budget (and the LBVH kernel 24); apart, neither spills.
// COMPACT (fixed-point sums + primary-visibility pass): new samples come from the compact list of vcrt_device.cuh instead of the
// work-item dispenser - one atomic and two 16-byte loads per finished slot.
template <int THREADS, int PAIRS, int LCAP, bool FIXED, bool BOOK = false, bool COMPACT = false, int UNROLL = VCRT_SCANU_UNROLL>
__global__ void __launch_bounds__(THREADS, 1) render_scanu_kernel(const unsigned char* __restrict__ blobGlobal, RenderArgs args, int scanPad) {
    static_assert(FIXED || !COMPACT, "the compact list is the hand-over of fixed-point mode");
    using Cfg = ScanUConfig<THREADS, PAIRS, LCAP>;
    constexpr int R = Cfg::R;
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    unsigned char* blob = smem + 16;
    const int tid = threadIdx.x;

    // ---- stage the scene blob (exact-test and material arrays): one TMA bulk copy, everyone waits on the mbarrier
    const uint32_t blobBytes = reinterpret_cast<const SceneBlob*>(blobGlobal)->bytes;
    if (tid == 0) mbar_init(bar, 1);
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(bar, blobBytes);
        tma_bulk_g2s(blob, blobGlobal, blobBytes, bar);
    }
    mbar_wait(bar, 0);
    const SceneBlob hdr = *reinterpret_cast<const SceneBlob*>(blob);
    const float4* sShape = reinterpret_cast<const float4*>(blob + hdr.offShape);
    const float4* sSurf = reinterpret_cast<const float4*>(blob + hdr.offSurf);
    const int* sKind = reinterpret_cast<const int*>(blob + hdr.offKind);
    uint32_t* sState = reinterpret_cast<uint32_t*>(blob + blobBytes);
    unsigned short* sList = reinterpret_cast<unsigned short*>(blob + blobBytes + Cfg::kStateBytes);
#define ST(field, r) sState[((field) * R + (r)) * THREADS + tid]
#define STF(field, r) __uint_as_float(ST(field, r))
#define TLIST(e) sList[(e) * THREADS + tid]

    const uint32_t sppPerChunk = args.plan.sppPerRun, totalItems = args.plan.totalItems;   // host-made plan, constant bank
    const uint32_t planeStride = args.width * args.height;

    unsigned long long nRays = 0;
    uint32_t nExhausted = 0, nBad = 0, nCand = 0, nOverflow = 0;

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
    const int scanCount = (int)hdr.scanCount, count = (int)hdr.count;
    const uint32_t listed = COMPACT ? __ldcg(&args.primary.cursors[0]) : 0u;   // entries the primary kernel left (it has finished: stream order)
    const float kPosInf = __int_as_float(0x7f800000);
    uint32_t listCount = 0, exhaustMask = 0;   // candidates of this thread this round; slots that need the exhaustive exact scan

    for (bool firstRound = true;; firstRound = false) {
        if (!firstRound) {
            // ---- (1) filter state of every slot, into registers (packed: low half = even slot, high half = odd slot)
            u64 dx[PAIRS], dy[PAIRS], dz[PAIRS], ee[PAIRS], ax[PAIRS], ay[PAIRS], az[PAIRS], nb[PAIRS];
            bool anyActive = false;
            exhaustMask = 0u; listCount = 0u;
#pragma unroll
            for (int p = 0; p < PAIRS; p++) {
                float f[2][8];
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int r = 2 * p + h;
                    // a slot that must not flag anything: t = 0, -w = -inf, so q' = -inf
                    f[h][0] = f[h][1] = f[h][2] = f[h][3] = f[h][4] = f[h][5] = f[h][6] = 0.f; f[h][7] = -kPosInf;
                    if (ST(SF_PIXEL, r) != kIdle) {
                        anyActive = true;
                        const float ox = STF(SF_OX, r), oy = STF(SF_OY, r), oz = STF(SF_OZ, r);
                        const float rx = STF(SF_DX, r), ry = STF(SF_DY, r), rz = STF(SF_DZ, r);
                        const float len2 = fmaf(rx, rx, fmaf(ry, ry, rz * rz));
                        const float oo = fmaf(ox, ox, fmaf(oy, oy, oz * oz));
                        if (!tame_ray(len2, oo)) exhaustMask |= 1u << r;   // no usable filter for this ray
                        else {
                            const float inv = rsqrtf(len2);
                            f[h][0] = rx * inv; f[h][1] = ry * inv; f[h][2] = rz * inv;
                            f[h][3] = -fmaf(ox, f[h][0], fmaf(oy, f[h][1], oz * f[h][2]));   // e = -o.dh
                            f[h][4] = 2.f * ox; f[h][5] = 2.f * oy; f[h][6] = 2.f * oz;       // -A = 2o
                            f[h][7] = fmaf(kKappaU, oo, -oo);                                 // -B = kappa|o|^2 - |o|^2
                        }
                    }
                }
                dx[p] = pack2(f[0][0], f[1][0]); dy[p] = pack2(f[0][1], f[1][1]); dz[p] = pack2(f[0][2], f[1][2]); ee[p] = pack2(f[0][3], f[1][3]);
                ax[p] = pack2(f[0][4], f[1][4]); ay[p] = pack2(f[0][5], f[1][5]); az[p] = pack2(f[0][6], f[1][6]); nb[p] = pack2(f[0][7], f[1][7]);
            }
            // warp-uniform exit: keeps the scan loop in uniform control flow, which is what lets the compiler feed the sphere
            // operands through uniform registers (LDCU -> UR.F32); a finished thread idles with slots that flag nothing
            if (!__any_sync(0xffffffffu, anyActive)) break;

            // ---- (2) the scan.  Per sphere: two LDCU.64 into uniform registers, 7 FFMA2 per ray pair, a max over the
            //      thread's rays and one compare against the sphere's scalar; survivors set a bit, and after every block of
            //      32 spheres the set bits are appended to the thread's candidate list.
            //      q' = (c.dh + e)^2 - (c.A + B)  >=  |c|^2 - R2   <=>  the sphere may be hit
            // U spheres starting at `at`: bit u of the result = sphere at + u may be hit by one of the thread's rays
            auto scan_block = [&]<int U>(int at) -> uint32_t {
                uint32_t mEven = 0u, mOdd = 0u;   // two accumulators: half the dependent-OR chain
#pragma unroll
                for (int u = 0; u < U; u++) {
                    (void)VCRT_CHECK(args.counters, at + u < scanPad && at + u < kMaxConstSpheres);
                    const float4 s = cSceneHot[at + u];
                    const u64 cx = pack2(s.x, s.x), cy = pack2(s.y, s.y), cz = pack2(s.z, s.z);
                    float qmax = -kPosInf;
#pragma unroll
                    for (int p = 0; p < PAIRS; p++) {
                        u64 t = ffma2(cz, dz[p], ee[p]); t = ffma2(cy, dy[p], t); t = ffma2(cx, dx[p], t);
                        u64 w = ffma2(cz, az[p], nb[p]); w = ffma2(cy, ay[p], w); w = ffma2(cx, ax[p], w);
                        const u64 q = ffma2(t, t, w);
                        float q0, q1; unpack2(q, q0, q1);
                        qmax = fmaxf(qmax, fmaxf(q0, q1));
                    }
                    flag_if_ge(u & 1 ? mOdd : mEven, qmax, s.w, 1u << u);
                }
                return mEven | mOdd;
            };
            const int scanLim = (scanCount + 3) & ~3;   // the table is padded (never-passing entries) to a multiple of 16
#pragma unroll 1
            for (int base = 0; base < scanLim; base += 32) {
                uint32_t mask = 0u;
                const int lim = min(32, scanLim - base);
                int kk = 0;
#pragma unroll 1
                for (; kk + UNROLL <= lim; kk += UNROLL) mask |= scan_block.template operator()<UNROLL>(base + kk) << kk;
#pragma unroll 1
                for (; kk < lim; kk += 4) mask |= scan_block.template operator()<4>(base + kk) << kk;   // the tail, 4 at a time
                while (mask) {
                    const int k = __ffs(mask) - 1;
                    mask &= mask - 1u;
                    if (listCount < (uint32_t)LCAP && VCRT_CHECK(args.counters, base + k < scanCount)) TLIST(listCount) = (unsigned short)(base + k);
                    listCount++;
                }
            }
        }

        // ---- (3) per slot: exact test of the thread's candidates, shade, and (re)generate.  One copy of every routine.
        //      A new sample normally comes out of the PrimaryBuffer (vcrt_kernel_primary.cuh) as a scattered ray at pass 1.
        const bool overflow = listCount > (uint32_t)LCAP;
#pragma unroll 1
        for (int r = 0; r < R; r++) {
            uint32_t pixel = firstRound ? kIdle : ST(SF_PIXEL, r);
            if (pixel == kIdle && !firstRound) continue;   // the dispenser ran dry while this slot was looking for work
            bool itemDone = firstRound, useFirst = firstRound, haveResult = false;
            V3 result = v3(0.f, 0.f, 0.f);   // black: bug B7 (depth exhausted) and absorbed bad materials
            if (pixel != kIdle) {
                Ray ray; ray.o = v3(STF(SF_OX, r), STF(SF_OY, r), STF(SF_OZ, r)); ray.d = v3(STF(SF_DX, r), STF(SF_DY, r), STF(SF_DZ, r));
                const float a = vdot(ray.d, ray.d);
                HitRec rec; rec.t = kTMax; rec.index = -1;
                const bool tame = !((exhaustMask >> r) & 1u);
                const bool exhaustive = overflow || !tame;
                nRays++;
                if (exhaustive) nOverflow++; else nCand += listCount;
                // candidates in ascending scan order, then the "always" spheres (device indices scanCount..count-1)
                const int nScan = exhaustive ? scanCount : (int)listCount, nAll = nScan + (count - scanCount);
                for (int e = 0; e < nAll; e++) {
                    const int j = e >= nScan ? scanCount + (e - nScan) : (exhaustive ? e : (int)TLIST(e));
                    if (VCRT_CHECK(args.counters, j >= 0 && j < count && (exhaustive || e >= nScan || e < LCAP))) hit_exact(sShape[j], j, ray, a, rec, tame);
                }
                V3 colour = v3(STF(SF_CR, r), STF(SF_CG, r), STF(SF_CB, r));
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
                    if (BOOK) {
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
                if (!ended) continue;   // the slot holds a scattered ray: next scan
                haveResult = true;
            }
            // ---- the sample in the slot is finished (or there was none): add it to the sums, take the next one
            if constexpr (COMPACT) {
                if (haveResult) add_fixed(args, pixel, result);
                const uint32_t at = atomicAdd(&args.primary.cursors[1], 1u);
                if (at >= listed || !VCRT_CHECK(args.counters, at < args.primary.entries)) { ST(SF_PIXEL, r) = kIdle; continue; }   // the list is used up: the slot idles
                const float4 ea = __ldcs(args.primary.a + at), eb = __ldcs(args.primary.b + at);
                const uint32_t word = __float_as_uint(ea.w), sphere = word & kPrimaryCameraStart;
                pixel = __float_as_uint(eb.w);
                ST(SF_PIXEL, r) = pixel; ST(SF_SAMPLE, r) = word >> kPrimarySphereBits;
                if (sphere != kPrimaryCameraStart && VCRT_CHECK(args.counters, (int)sphere < count)) {   // the sample goes on after its first scatter event
                    const V3 colour = first_attenuation(sSurf[sphere], sKind[sphere], BOOK ? 1u : 0u);
                    ST(SF_OX, r) = __float_as_uint(ea.x); ST(SF_OY, r) = __float_as_uint(ea.y); ST(SF_OZ, r) = __float_as_uint(ea.z);
                    ST(SF_DX, r) = __float_as_uint(eb.x); ST(SF_DY, r) = __float_as_uint(eb.y); ST(SF_DZ, r) = __float_as_uint(eb.z);
                    ST(SF_CR, r) = __float_as_uint(colour.x); ST(SF_CG, r) = __float_as_uint(colour.y); ST(SF_CB, r) = __float_as_uint(colour.z);
                    ST(SF_PASS, r) = 1u;
                } else {   // not pre-traced (a ray outside the tame range, a candidate-list overflow): start at the camera (shader.comp:46-52)
                    uint32_t px, py, rnd[4];
                    pixel_to_xy(args.plan, args.width, pixel, px, py);
                    philox4x32_10(pixel, args.sampleOffset + (word >> kPrimarySphereBits), 0u, 0u, args.seedLo, args.seedHi, rnd);
                    const Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
                    ST(SF_OX, r) = __float_as_uint(ray.o.x); ST(SF_OY, r) = __float_as_uint(ray.o.y); ST(SF_OZ, r) = __float_as_uint(ray.o.z);
                    ST(SF_DX, r) = __float_as_uint(ray.d.x); ST(SF_DY, r) = __float_as_uint(ray.d.y); ST(SF_DZ, r) = __float_as_uint(ray.d.z);
                    ST(SF_CR, r) = ST(SF_CG, r) = ST(SF_CB, r) = __float_as_uint(1.0f);
                    ST(SF_PASS, r) = 0u;
                }
            } else for (;;) {
                if (haveResult) {   // shader.comp:53: color.rgb += ray_color(r), in sample order within the item
                    haveResult = false;
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
                    if (s < s1) ST(SF_SAMPLE, r) = s; else itemDone = true;
                }
                if (itemDone) {
                    if (!claim_item(r, (blockIdx.x * R + r) * THREADS + tid, useFirst)) break;   // the dispenser is empty: the slot idles
                    itemDone = false; useFirst = false;
                }
                pixel = ST(SF_PIXEL, r);
                const uint32_t sLocal = ST(SF_SAMPLE, r);
                uint32_t px, py;
                pixel_to_xy(args.plan, args.width, pixel, px, py);
                int state = kPrimaryUntraced;
                if (!FIXED && args.primary.a) {   // strict-order modes: one entry per sample (fixed-point mode: the COMPACT branch above)
                    const size_t at = (size_t)sLocal * args.plan.pixelsPadded + padded_pixel_id(args.plan.tilesX, px, py);
                    (void)VCRT_CHECK(args.counters, at < args.primary.entries);
                    const float4 ea = __ldcs(args.primary.a + at);
                    state = __float_as_int(ea.w);
                    if (state >= 0) {   // the sample goes on after its first scatter event
                        const float4 eb = __ldcs(args.primary.b + at);
                        const V3 colour = first_attenuation(sSurf[state], sKind[state], BOOK ? 1u : 0u);
                        ST(SF_OX, r) = __float_as_uint(ea.x); ST(SF_OY, r) = __float_as_uint(ea.y); ST(SF_OZ, r) = __float_as_uint(ea.z);
                        ST(SF_DX, r) = __float_as_uint(eb.x); ST(SF_DY, r) = __float_as_uint(eb.y); ST(SF_DZ, r) = __float_as_uint(eb.z);
                        ST(SF_CR, r) = __float_as_uint(colour.x); ST(SF_CG, r) = __float_as_uint(colour.y); ST(SF_CB, r) = __float_as_uint(colour.z);
                        ST(SF_PASS, r) = 1u;
                        break;
                    }
                    if (state == kPrimaryEnded) {
                        const float4 eb = __ldcs(args.primary.b + at);
                        result = v3(eb.x, eb.y, eb.z); haveResult = true; continue;
                    }
                }
                // no entry: the sample starts at the camera (shader.comp:46-52) and its primary ray goes through the scan
                uint32_t rnd[4];
                philox4x32_10(pixel, args.sampleOffset + sLocal, 0u, 0u, args.seedLo, args.seedHi, rnd);
                const Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
                ST(SF_OX, r) = __float_as_uint(ray.o.x); ST(SF_OY, r) = __float_as_uint(ray.o.y); ST(SF_OZ, r) = __float_as_uint(ray.o.z);
                ST(SF_DX, r) = __float_as_uint(ray.d.x); ST(SF_DY, r) = __float_as_uint(ray.d.y); ST(SF_DZ, r) = __float_as_uint(ray.d.z);
                ST(SF_CR, r) = ST(SF_CG, r) = ST(SF_CB, r) = __float_as_uint(1.0f);
                ST(SF_PASS, r) = 0u;
                break;
            }
        }
    }

    atomicAdd(&args.counters[CTR_RAYS], nRays);
    if (nCand) atomicAdd(&args.counters[CTR_CANDIDATES], (unsigned long long)nCand);
    if (nExhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)nExhausted);
    if (nBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)nBad);
    if (nOverflow) atomicAdd(&args.counters[CTR_OVERFLOWS], (unsigned long long)nOverflow);
#undef ST
#undef STF
#undef TLIST
}

}  // namespace vcrt
