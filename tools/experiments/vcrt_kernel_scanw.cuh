/* @file vcrt_kernel_scanw.cuh

    VCRT_TRAVERSAL_BRUTE, third generation: the uniform-datapath filter scan of vcrt_kernel_scanu.cuh
    with a WARP-COMPACTED shading phase.
    SPDX-License-Identifier: WTFPL

    The scan keeps every lane busy, but what follows it does not: in the second generation each thread
    walks its own slots, so a warp executes "miss: sky + accumulate + next sample", "lambertian / metal
    scatter" and "glass scatter" once per slot with whatever lanes happen to need them (ncu: 25 % of
    all instructions at 15.7 of 32 lanes = 12.8 % of the kernel's issue slots wasted).

    Here the ray state of a warp's 32 x R slots lives in shared memory, so ANY lane can shade ANY slot of
    its warp.  After the exact tests (still done by the owning thread: they need its candidate list) the
    warp ballots its slots by what they need, writes a compacted unit list and processes
        1. all hits, ordered lambertian/metal first and glass last, 32 units per pass,
        2. all path ends (misses first, then black ends), 32 units per pass: sky, accumulate, next sample,
    each with full lanes except for the last pass of a list.  Which lane shades a slot cannot change a
    bit of the result: the random draws are keyed by (pixel, sample, bounce) and all arithmetic is the
    same; tests/test_gpu_parity.py holds the kernel bit-identical to the oracle like the others.
*/
#pragma once
#include "../../vulkancomputeraytracing_b200/csrc/vcrt_kernel_scanu.cuh"

namespace vcrt {

// per-slot state, SoA in shared memory: [field][slot][thread]
enum WarpSlotField { WF_OX = 0, WF_OY, WF_OZ, WF_DX, WF_DY, WF_DZ, WF_CR, WF_CG, WF_CB, WF_SUMR, WF_SUMG, WF_SUMB,
                     WF_PIXEL, WF_SAMPLE, WF_END, WF_PASS, WF_HITT, WF_HITI, WF_COUNT };

template <int THREADS, int PAIRS, int LCAP>
struct ScanWConfig {
    static constexpr int R = 2 * PAIRS;
    static constexpr int kStateBytes = WF_COUNT * R * THREADS * 4;
    static constexpr int kListBytes = LCAP * THREADS * 2;
    static constexpr int kMapBytes = R * THREADS;            // one byte per slot: the warp's compacted unit list
    static __host__ __device__ size_t smem_bytes(uint32_t blobBytes) { return 16 + (size_t)blobBytes + kStateBytes + kListBytes + kMapBytes; }
};

constexpr uint32_t kEndedFlag = 0x80000000u;   // in WF_PASS: the path ended black while being shaded (depth exhausted, absorbed)

template <int THREADS, int PAIRS, int LCAP, bool FIXED>
__global__ void __launch_bounds__(THREADS, 1) render_scanw_kernel(const unsigned char* __restrict__ blobGlobal, RenderArgs args, int scanPad) {
    using Cfg = ScanWConfig<THREADS, PAIRS, LCAP>;
    constexpr int R = Cfg::R;
    static_assert(R <= 4, "unit list entries keep the slot in two bits");
    extern __shared__ __align__(16) unsigned char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
    unsigned char* blob = smem + 16;
    const int tid = threadIdx.x, lane = tid & 31, warpBase = tid & ~31;
    const uint32_t laneLt = (1u << lane) - 1u;

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
    unsigned char* sMap = blob + blobBytes + Cfg::kStateBytes + Cfg::kListBytes + (tid >> 5) * (R * 32);   // this warp's unit list
#define WS(field, r, t) sState[((field) * R + (r)) * THREADS + (t)]
#define WSF(field, r, t) __uint_as_float(WS(field, r, t))
#define TLIST(e) sList[(e) * THREADS + tid]

    const uint32_t sppPerChunk = args.plan.sppPerRun, totalItems = args.plan.totalItems;   // host-made plan, constant bank
    const uint32_t planeStride = args.width * args.height;

    unsigned long long nRays = 0;
    uint32_t nExhausted = 0, nBad = 0, nCand = 0, nOverflow = 0;

    // Claims items for slot (r, t) until one maps to a real pixel with a non-empty sample range; false when the dispenser is dry.
    auto claim_item = [&](int r, int t, uint32_t firstId, bool useFirst) -> bool {
        uint32_t id = firstId; bool have = useFirst;
        for (;;) {
            if (!have) id = atomicAdd(args.workCounter, 1u);
            have = false;
            if (id >= totalItems) { WS(WF_PIXEL, r, t) = kIdle; return false; }
            uint32_t chunk, pid, px, py;
            item_to_run(args.plan, id, chunk, pid);
            if (!item_to_pixel(args.plan, args.width, args.height, pid, px, py)) continue;
            uint32_t s0 = chunk * sppPerChunk, s1 = min(args.spp, s0 + sppPerChunk);
            if (s0 >= s1) continue;
            WS(WF_PIXEL, r, t) = py * args.width + px;
            WS(WF_SAMPLE, r, t) = s0;
            WS(WF_END, r, t) = s1;
            return true;
        }
    };
    auto start_sample = [&](int r, int t) {   // shader.comp:46-52
        uint32_t pixel = WS(WF_PIXEL, r, t), sample = args.sampleOffset + WS(WF_SAMPLE, r, t);
        uint32_t py = pixel / args.width, px = pixel - py * args.width, rnd[4];
        philox4x32_10(pixel, sample, 0u, 0u, args.seedLo, args.seedHi, rnd);
        Ray ray = primary_ray(args.frame, px, py, u01(rnd[0]), u01(rnd[1]));
        WS(WF_OX, r, t) = __float_as_uint(ray.o.x); WS(WF_OY, r, t) = __float_as_uint(ray.o.y); WS(WF_OZ, r, t) = __float_as_uint(ray.o.z);
        WS(WF_DX, r, t) = __float_as_uint(ray.d.x); WS(WF_DY, r, t) = __float_as_uint(ray.d.y); WS(WF_DZ, r, t) = __float_as_uint(ray.d.z);
        WS(WF_CR, r, t) = WS(WF_CG, r, t) = WS(WF_CB, r, t) = __float_as_uint(1.0f);
        WS(WF_PASS, r, t) = 0u;
    };

    const int scanCount = (int)hdr.scanCount, count = (int)hdr.count;
    const float kPosInf = __int_as_float(0x7f800000);
    uint32_t listCount = 0, exhaustMask = 0;

    // ---- first work items: a static one per slot, so the dispenser only serves the refills
#pragma unroll 1
    for (int r = 0; r < R; r++)
        if (claim_item(r, tid, (blockIdx.x * R + r) * THREADS + tid, true)) start_sample(r, tid);

    for (;;) {
        {
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
                    if (WS(WF_PIXEL, r, tid) != kIdle) {
                        anyActive = true;
                        const float ox = WSF(WF_OX, r, tid), oy = WSF(WF_OY, r, tid), oz = WSF(WF_OZ, r, tid);
                        const float rx = WSF(WF_DX, r, tid), ry = WSF(WF_DY, r, tid), rz = WSF(WF_DZ, r, tid);
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
#pragma unroll 1
            for (int base = 0; base < scanPad; base += 32) {
                uint32_t mask = 0u;
                const int lim = min(32, scanPad - base);   // scanPad is a multiple of 8
#pragma unroll 1
                for (int kk = 0; kk < lim; kk += 8) {
                    uint32_t m8 = 0u;
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const float4 s = cSceneHot[base + kk + u];
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
                        if (qmax >= s.w) m8 |= 1u << u;
                    }
                    mask |= m8 << kk;
                }
                while (mask) {
                    const int k = __ffs(mask) - 1;
                    mask &= mask - 1u;
                    if (listCount < (uint32_t)LCAP) TLIST(listCount) = (unsigned short)(base + k);
                    listCount++;
                }
            }
        }


        // ---- (3.1) the owning thread: exact tests of its candidates for each of its slots; the nearest hit goes to shared memory
        enum { CAT_IDLE = 0, CAT_MISS, CAT_HIT_LM, CAT_HIT_GLASS, CAT_BAD };
        const bool overflow = listCount > (uint32_t)LCAP;
        uint32_t catBits = 0u;   // 3 bits per slot
#pragma unroll 1
        for (int r = 0; r < R; r++) {
            int cat = CAT_IDLE;
            if (WS(WF_PIXEL, r, tid) != kIdle) {
                Ray ray; ray.o = v3(WSF(WF_OX, r, tid), WSF(WF_OY, r, tid), WSF(WF_OZ, r, tid)); ray.d = v3(WSF(WF_DX, r, tid), WSF(WF_DY, r, tid), WSF(WF_DZ, r, tid));
                const float a = vdot(ray.d, ray.d);
                HitRec rec; rec.t = kTMax; rec.index = -1;
                const bool exhaustive = overflow || ((exhaustMask >> r) & 1u);
                nRays++;
                if (exhaustive) nOverflow++; else nCand += listCount;
                // candidates in ascending scan order, then the "always" spheres (device indices scanCount..count-1)
                const int nScan = exhaustive ? scanCount : (int)listCount, nAll = nScan + (count - scanCount);
                for (int e = 0; e < nAll; e++) {
                    const int j = e >= nScan ? scanCount + (e - nScan) : (exhaustive ? e : (int)TLIST(e));
                    hit_exact(sShape[j], j, ray, a, rec, !((exhaustMask >> r) & 1u));
                }
                WS(WF_HITT, r, tid) = __float_as_uint(rec.t);
                WS(WF_HITI, r, tid) = (uint32_t)rec.index;
                if (rec.index < 0) {
                    cat = CAT_MISS;
                    if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_PATH0 + min(WS(WF_PASS, r, tid), 7u)], 1ull);
                } else {
                    const int kind = sKind[rec.index];
                    if (args.flags & kFlagCountStats) atomicAdd(&args.counters[CTR_HITS0 + ((kind >= 1 && kind <= 3) ? kind : 0)], 1ull);
                    cat = (kind == 1 || kind == 2) ? CAT_HIT_LM : (kind == 3 ? CAT_HIT_GLASS : CAT_BAD);
                    if (cat == CAT_BAD) nBad++;
                }
            }
            catBits |= (uint32_t)cat << (3 * r);
        }

        // ---- (3.2) all hits of the warp, lambertian/metal first, glass last: scatter (textures.glsl:65-71)
        uint32_t total = 0;
#pragma unroll
        for (int g = CAT_HIT_LM; g <= CAT_HIT_GLASS; g++)
#pragma unroll
            for (int r = 0; r < R; r++) {
                const bool mine = ((catBits >> (3 * r)) & 7u) == (uint32_t)g;
                const uint32_t m = __ballot_sync(0xffffffffu, mine);
                if (mine) sMap[total + __popc(m & laneLt)] = (unsigned char)((r << 5) | lane);
                total += __popc(m);
            }
        __syncwarp();
#pragma unroll 1
        for (uint32_t base = 0; base < total; base += 32u) {
            const uint32_t k = base + (uint32_t)lane;
            if (k < total) {
                const uint32_t entry = sMap[k];
                const int r = (int)(entry >> 5), t = warpBase + (int)(entry & 31u);
                Ray ray; ray.o = v3(WSF(WF_OX, r, t), WSF(WF_OY, r, t), WSF(WF_OZ, r, t)); ray.d = v3(WSF(WF_DX, r, t), WSF(WF_DY, r, t), WSF(WF_DZ, r, t));
                V3 colour = v3(WSF(WF_CR, r, t), WSF(WF_CG, r, t), WSF(WF_CB, r, t));
                const uint32_t pixel = WS(WF_PIXEL, r, t), sample = args.sampleOffset + WS(WF_SAMPLE, r, t), pass = WS(WF_PASS, r, t);
                const float hitT = WSF(WF_HITT, r, t);
                const int hitI = (int)WS(WF_HITI, r, t);
                uint32_t rnd[4];
                philox4x32_10(pixel, sample, pass + 1u, 0u, args.seedLo, args.seedHi, rnd);
                float u[4] = {u01(rnd[0]), u01(rnd[1]), u01(rnd[2]), u01(rnd[3])};
                int outcome;   // 1 = scattered, 2 = absorbed by the material (BOOK metal); unknown materials never get here
                if (args.semantics) {
                    BookDraws g; g.pixel = pixel; g.sample = sample; g.event = pass + 1u; g.seedLo = args.seedLo; g.seedHi = args.seedHi;
                    outcome = scatter_book(sShape[hitI], sSurf[hitI], sKind[hitI], hitT, ray, colour, u, g);
                } else outcome = scatter(sShape[hitI], sSurf[hitI], sKind[hitI], hitT, ray, colour, u) ? 1 : 0;
                if (outcome == 1 && pass + 1u >= args.maxDepth) { nExhausted++; outcome = 2; }   // bug B7: black
                if (outcome == 1) {
                    WS(WF_OX, r, t) = __float_as_uint(ray.o.x); WS(WF_OY, r, t) = __float_as_uint(ray.o.y); WS(WF_OZ, r, t) = __float_as_uint(ray.o.z);
                    WS(WF_DX, r, t) = __float_as_uint(ray.d.x); WS(WF_DY, r, t) = __float_as_uint(ray.d.y); WS(WF_DZ, r, t) = __float_as_uint(ray.d.z);
                    WS(WF_CR, r, t) = __float_as_uint(colour.x); WS(WF_CG, r, t) = __float_as_uint(colour.y); WS(WF_CB, r, t) = __float_as_uint(colour.z);
                    WS(WF_PASS, r, t) = pass + 1u;
                } else WS(WF_PASS, r, t) = pass | kEndedFlag;
            }
        }
        __syncwarp();

        // ---- (3.3) all path ends of the warp, misses first: sky or black, accumulate (shader.comp:53), next sample / next item
        total = 0;
#pragma unroll
        for (int g = 0; g < 2; g++)
#pragma unroll
            for (int r = 0; r < R; r++) {
                const uint32_t cat = (catBits >> (3 * r)) & 7u;
                const bool black = cat == CAT_BAD || ((cat == CAT_HIT_LM || cat == CAT_HIT_GLASS) && (WS(WF_PASS, r, tid) & kEndedFlag));
                const bool mine = g == 0 ? cat == CAT_MISS : black;
                const uint32_t m = __ballot_sync(0xffffffffu, mine);
                if (mine) sMap[total + __popc(m & laneLt)] = (unsigned char)((g << 7) | (r << 5) | lane);
                total += __popc(m);
            }
        __syncwarp();
#pragma unroll 1
        for (uint32_t base = 0; base < total; base += 32u) {
            const uint32_t k = base + (uint32_t)lane;
            if (k < total) {
                const uint32_t entry = sMap[k];
                const int r = (int)((entry >> 5) & 3u), t = warpBase + (int)(entry & 31u);
                const uint32_t pixel = WS(WF_PIXEL, r, t);
                V3 result = v3(0.f, 0.f, 0.f);   // black: bug B7 (depth exhausted), absorbed, unknown material
                if (!(entry & 0x80u)) {
                    const V3 d = v3(WSF(WF_DX, r, t), WSF(WF_DY, r, t), WSF(WF_DZ, r, t));
                    result = vmul(v3(WSF(WF_CR, r, t), WSF(WF_CG, r, t), WSF(WF_CB, r, t)), sky(d));
                }
                uint32_t s = WS(WF_SAMPLE, r, t);
                const uint32_t s1 = WS(WF_END, r, t);
                if (FIXED) {
                    add_fixed(args, pixel, result);
                    s++;
                } else {
                    const uint32_t chunk = (s1 - 1u) / sppPerChunk, s0 = chunk * sppPerChunk;
                    V3 sum = (s == s0) ? v3(0.f, 0.f, 0.f) : v3(WSF(WF_SUMR, r, t), WSF(WF_SUMG, r, t), WSF(WF_SUMB, r, t));
                    sum = vadd(sum, result);
                    s++;
                    if (s < s1) { WS(WF_SUMR, r, t) = __float_as_uint(sum.x); WS(WF_SUMG, r, t) = __float_as_uint(sum.y); WS(WF_SUMB, r, t) = __float_as_uint(sum.z); }
                    else args.sums[(size_t)chunk * planeStride + pixel] = make_float4(sum.x, sum.y, sum.z, 0.f);   // the item's only global write
                }
                bool start = true;
                if (s < s1) WS(WF_SAMPLE, r, t) = s;
                else start = claim_item(r, t, 0u, false);
                if (start) start_sample(r, t);
            }
        }
        __syncwarp();
    }

    atomicAdd(&args.counters[CTR_RAYS], nRays);
    if (nCand) atomicAdd(&args.counters[CTR_CANDIDATES], (unsigned long long)nCand);
    if (nExhausted) atomicAdd(&args.counters[CTR_EXHAUSTED], (unsigned long long)nExhausted);
    if (nBad) atomicAdd(&args.counters[CTR_BAD_MATERIAL], (unsigned long long)nBad);
    if (nOverflow) atomicAdd(&args.counters[CTR_OVERFLOWS], (unsigned long long)nOverflow);
#undef WS
#undef WSF
#undef TLIST
}

}  // namespace vcrt
