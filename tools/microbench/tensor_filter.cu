// SPDX-License-Identifier: WTFPL
// Proof of concept for round 2: the brute-force filter's two dot products per (sphere, ray) pair,
//     t = c.dh + e      w = c.(2o) + nb      (then q' = t*t + w >= thr on the CUDA cores)
// as a tcgen05 TF32 MMA with FP32 accuracy recovered by a hi/lo split packed into the K dimension:
//     A row (sphere)  = [c_hi(4) | c_lo(4)]   B column = [b_hi(4) | b_hi(4)]     -> c_hi.b_hi + c_lo.b_hi
//     A row           = [c_hi(4) |   0    ]   B column = [b_lo(4) |   0    ]     -> c_hi.b_lo        (accumulated)
// Tile: 128 spheres x 128 rays (N = 256 columns: t and w per ray), 2 MMAs m128 n256 k8 per tile, accumulators in TMEM
// (double buffered), epilogue by 4 warps with tcgen05.ld.  Reports pairs/s and the error of t, w against double.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tensor_filter tensor_filter.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count)); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {   // bounded: a wrong descriptor must not hang the GPU
    for (uint32_t spin = 0; spin < (1u << 24); spin++) {
        uint32_t done;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (done) return;
    }
    printf("mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
    __trap();
}
// K-major, no swizzle: core matrices of 8 rows x 16 bytes; LBO = distance between the two K halves, SBO = between 8-row groups
__device__ __forceinline__ uint64_t make_desc(const void* p, uint32_t lboBytes, uint32_t sboBytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(p) >> 4) & 0x3fff);
    d |= (uint64_t)((lboBytes >> 4) & 0x3fff) << 16;
    d |= (uint64_t)((sboBytes >> 4) & 0x3fff) << 32;
    d |= (uint64_t)1 << 46;   // descriptor version for sm_100
    return d;                 // base offset 0, layout type 0 = no swizzle
}
__device__ __forceinline__ void mma_tf32(uint32_t tmemD, uint64_t descA, uint64_t descB, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(tmemD), "l"(descA), "l"(descB), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                   "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
                   "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]),
                   "=r"(r[30]), "=r"(r[31]) : "r"(taddr) : "memory");
}

constexpr int kRows = 128, kCols = 256, kOperandBytes = (kRows + kRows + kCols + kCols) * 32;   // A1 A2 B1 B2, 32 bytes per row
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kCols >> 3) << 17) | ((uint32_t)(kRows >> 4) << 24);

// operands: already in core-matrix order (host), thr[128]; out: D of the first tile [128][256]; passes: per-CTA count
template <int G, int MODE>   // G epilogue warps per TMEM lane quarter; MODE 0 = full, 1 = MMA only (epilogue just releases the buffer)
__global__ void __launch_bounds__(128 * G + 32, 1) k_filter(const unsigned char* __restrict__ operands, const float* __restrict__ thr,
                                                   float* __restrict__ out, unsigned long long* __restrict__ passes, int iters) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA1 = smem; unsigned char* sA2 = sA1 + kRows * 32; unsigned char* sB1 = sA2 + kRows * 32; unsigned char* sB2 = sB1 + kCols * 32;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kOperandBytes);   // [2]
    uint64_t* empty = full + 2;                                           // [2]
    uint32_t* tmemBase = reinterpret_cast<uint32_t*>(empty + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int i = tid; i < kOperandBytes / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = reinterpret_cast<const uint4*>(operands)[i];
    if (tid == 0) { mbar_init(&full[0], 1); mbar_init(&full[1], 1); mbar_init(&empty[0], 4 * G); mbar_init(&empty[1], 4 * G); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy smem writes -> visible to the tensor core
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(smem_u32(tmemBase)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmemBase;

    if (MODE == 4) {   // TMEM read rate alone: every epilogue warp reads its 2 x 32 columns `iters` times, no MMA, no barriers
        if (warp < 4 * G) {
            const int quarter = warp & 3, group = warp >> 2;
            uint32_t x = 0;
            for (int it = 0; it < iters; it++) {
                uint32_t ra[32], rb[32];
                const uint32_t at = tmem + ((uint32_t)(quarter * 32) << 16) + (it & 1) * kCols + group * (kCols / G);
                tmem_ld32(at, ra); tmem_ld32(at + 32, rb);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                for (int j = 0; j < 32; j++) x |= ra[j] & rb[j];
            }
            if (x == 0x12345678u) atomicAdd(passes + blockIdx.x, 1ull);
        }
    } else
    if (warp == 4 * G) {   // ---- the MMA issuer: one lane
        if (lane == 0) {
            const uint64_t dA1 = make_desc(sA1, 128, 256), dA2 = make_desc(sA2, 128, 256), dB1 = make_desc(sB1, 128, 256), dB2 = make_desc(sB2, 128, 256);
            for (int it = 0; it < iters; it++) {
                const int stage = it & 1;
                if (it >= 2) mbar_wait(&empty[stage], ((it >> 1) - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                mma_tf32(tmem + stage * kCols, dA1, dB1, kIdesc, 0u);
                mma_tf32(tmem + stage * kCols, dA2, dB2, kIdesc, 1u);
                mma_commit(&full[stage]);
            }
        }
    } else {           // ---- epilogue: warp w owns TMEM lanes 32 (w % 4) .. + 31 = spheres, and every G-th chunk of 32 columns
        const int quarter = warp & 3, group = warp >> 2;
        const float myThr = thr[quarter * 32 + lane];
        unsigned long long count = 0;
        for (int it = 0; it < iters; it++) {
            const int stage = it & 1;
            if (lane == 0) mbar_wait(&full[stage], (it >> 1) & 1);
            __syncwarp();
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (MODE == 3) {   // loads only
                uint32_t ra[32], rb[32];
                const uint32_t at = tmem + ((uint32_t)(quarter * 32) << 16) + stage * kCols + group * (kCols / G);
                tmem_ld32(at, ra); if (G <= 4) tmem_ld32(at + 32, rb); else { for (int j = 0; j < 32; j++) rb[j] = 0; }
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                uint32_t x = 0;
#pragma unroll
                for (int j = 0; j < 32; j++) x |= ra[j] & rb[j];
                if (x == 0x12345678u) count++;
            }
            if (MODE == 0) {
#pragma unroll 1
                for (int c = group * 32; c < kCols; c += 32 * G) {
                    uint32_t r[32];
                    tmem_ld32(tmem + ((uint32_t)(quarter * 32) << 16) + stage * kCols + c, r);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (it == 0) {
#pragma unroll
                        for (int j = 0; j < 32; j++) out[(size_t)blockIdx.x * kRows * kCols + (quarter * 32 + lane) * kCols + c + j] = __uint_as_float(r[j]);
                    }
                    // q' = t*t + w per ray; a coarse max over 4 rays first, the per-ray test only when that passes
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        const float q0 = fmaf(__uint_as_float(r[j]), __uint_as_float(r[j]), __uint_as_float(r[j + 1]));
                        const float q1 = fmaf(__uint_as_float(r[j + 2]), __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                        const float q2 = fmaf(__uint_as_float(r[j + 4]), __uint_as_float(r[j + 4]), __uint_as_float(r[j + 5]));
                        const float q3 = fmaf(__uint_as_float(r[j + 6]), __uint_as_float(r[j + 6]), __uint_as_float(r[j + 7]));
                        if (fmaxf(fmaxf(q0, q1), fmaxf(q2, q3)) >= myThr) count += (q0 >= myThr) + (q1 >= myThr) + (q2 >= myThr) + (q3 >= myThr);
                    }
                }
            }
            if (MODE == 2) {   // both loads of the warp's 64 columns in flight together, FFMA2 on (t0,t1)(w0,w1), one compare per 16 rays
                static_assert(MODE != 2 || G == 4, "MODE 2: 4 column groups of 64");
                uint32_t ra[32], rb[32];
                const uint32_t at = tmem + ((uint32_t)(quarter * 32) << 16) + stage * kCols + group * 64;
                tmem_ld32(at, ra); tmem_ld32(at + 32, rb);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                auto q2 = [](uint32_t t0, uint32_t t1, uint32_t w0, uint32_t w1) -> unsigned long long {
                    unsigned long long t, w, q;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(t) : "r"(t0), "r"(t1)); asm("mov.b64 %0, {%1, %2};" : "=l"(w) : "r"(w0), "r"(w1));
                    asm("fma.rn.f32x2 %0, %1, %1, %2;" : "=l"(q) : "l"(t), "l"(w));
                    return q;
                };
                auto hmax = [](unsigned long long q) -> float { float a, b; asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(q)); return fmaxf(a, b); };
#pragma unroll
                for (int half = 0; half < 2; half++) {
                    uint32_t* r = half ? rb : ra;
                    float m = -3e38f;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) m = fmaxf(m, hmax(q2(r[j], r[j + 1], r[j + 2], r[j + 3])));
                    if (m >= myThr) {   // rare: look again, ray by ray
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            count += (fmaf(__uint_as_float(r[j]), __uint_as_float(r[j]), __uint_as_float(r[j + 2])) >= myThr);
                            count += (fmaf(__uint_as_float(r[j + 1]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 3])) >= myThr);
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[stage]);
        }
        for (int off = 16; off > 0; off >>= 1) count += __shfl_down_sync(0xffffffffu, count, off);
        if (lane == 0) atomicAdd(passes + blockIdx.x, count);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(tmem) : "memory");
}

static float tf32_hi(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xffffe000u; float r; memcpy(&r, &u, 4); return r; }
// row-major [rows][8 floats] -> core-matrix order (8 rows x 16 B blocks; K half stride 128 B, 8-row group stride 256 B)
static void to_core(const std::vector<float>& rowMajor, int rows, unsigned char* dst) {
    for (int r = 0; r < rows; r++) for (int k = 0; k < 8; k++) {
        size_t off = (size_t)(r / 8) * 256 + (size_t)(k / 4) * 128 + (size_t)(r % 8) * 16 + (size_t)(k % 4) * 4;
        memcpy(dst + off, &rowMajor[(size_t)r * 8 + k], 4);
    }
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); const int sms = p.multiProcessorCount;
    srand(7);
    auto rnd = [] { return (float)rand() / RAND_MAX; };
    std::vector<float> sph(kRows * 4), ray(kCols * 4), thr(kRows);
    for (int i = 0; i < kRows; i++) { sph[i * 4] = 22.f * rnd() - 11.f; sph[i * 4 + 1] = 0.2f; sph[i * 4 + 2] = 22.f * rnd() - 11.f; sph[i * 4 + 3] = 1.f;
        thr[i] = sph[i * 4] * sph[i * 4] + 0.04f + sph[i * 4 + 2] * sph[i * 4 + 2] - 0.04f; }
    for (int j = 0; j < kCols / 2; j++) {   // ray j: origin o, unit direction dh
        float o[3] = {13.f * rnd(), 2.f * rnd(), 3.f * rnd()}, d[3] = {-rnd(), -0.2f * rnd(), -rnd()};
        float n = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]); for (float& x : d) x /= n;
        float e = -(o[0] * d[0] + o[1] * d[1] + o[2] * d[2]), oo = o[0] * o[0] + o[1] * o[1] + o[2] * o[2];
        float* ct = &ray[(4 * (j / 2) + (j & 1)) * 4]; float* cw = &ray[(4 * (j / 2) + 2 + (j & 1)) * 4];   // columns: t0 t1 w0 w1 | t2 t3 w2 w3 ...
        ct[0] = d[0]; ct[1] = d[1]; ct[2] = d[2]; ct[3] = e;
        cw[0] = 2 * o[0]; cw[1] = 2 * o[1]; cw[2] = 2 * o[2]; cw[3] = -oo;
    }
    std::vector<float> A1(kRows * 8), A2(kRows * 8, 0.f), B1(kCols * 8), B2(kCols * 8, 0.f);
    for (int i = 0; i < kRows; i++) for (int k = 0; k < 4; k++) { float v = sph[i * 4 + k], hi = tf32_hi(v), lo = tf32_hi(v - hi);
        A1[i * 8 + k] = hi; A1[i * 8 + 4 + k] = lo; A2[i * 8 + k] = hi; }
    for (int j = 0; j < kCols; j++) for (int k = 0; k < 4; k++) { float v = ray[j * 4 + k], hi = tf32_hi(v), lo = tf32_hi(v - hi);
        B1[j * 8 + k] = hi; B1[j * 8 + 4 + k] = hi; B2[j * 8 + k] = lo; }
    std::vector<unsigned char> ops(kOperandBytes);
    to_core(A1, kRows, ops.data()); to_core(A2, kRows, ops.data() + kRows * 32);
    to_core(B1, kCols, ops.data() + 2 * kRows * 32); to_core(B2, kCols, ops.data() + 2 * kRows * 32 + kCols * 32);

    unsigned char* dOps; float *dThr, *dOut; unsigned long long* dPass;
    CK(cudaMalloc(&dOps, kOperandBytes)); CK(cudaMemcpy(dOps, ops.data(), kOperandBytes, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dThr, kRows * 4)); CK(cudaMemcpy(dThr, thr.data(), kRows * 4, cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dOut, (size_t)sms * kRows * kCols * 4)); CK(cudaMemset(dOut, 0, (size_t)sms * kRows * kCols * 4));
    CK(cudaMalloc(&dPass, sms * 8)); CK(cudaMemset(dPass, 0, sms * 8));
    const size_t smemBytes = kOperandBytes + 64;
    k_filter<1, 0><<<1, 160, smemBytes>>>(dOps, dThr, dOut, dPass, 4);
    CK(cudaDeviceSynchronize());
    std::vector<float> D(kRows * kCols); CK(cudaMemcpy(D.data(), dOut, D.size() * 4, cudaMemcpyDeviceToHost));
    double maxErr = 0, maxRef = 0; int bad = 0;
    for (int i = 0; i < kRows; i++) for (int j = 0; j < kCols; j++) {
        double ref = 0; for (int k = 0; k < 4; k++) ref += (double)sph[i * 4 + k] * (double)ray[j * 4 + k];
        double err = fabs(ref - (double)D[i * kCols + j]);
        if (err > maxErr) maxErr = err; if (fabs(ref) > maxRef) maxRef = fabs(ref);
        if (err > 1e-2 && bad < 5) { printf("  mismatch row %d col %d: got %g want %g\n", i, j, D[i * kCols + j], ref); bad++; }
    }
    printf("{\"check\": \"D vs double\", \"max_abs_err\": %.3e, \"max_abs_ref\": %.3e, \"rel\": %.3e}\n", maxErr, maxRef, maxErr / maxRef);

    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    const int iters = 20000;
#define RUN(G, MODE, NAME) { float best = 1e30f; \
    for (int rep = 0; rep < 3; rep++) { CK(cudaMemset(dPass, 0, sms * 8)); \
        CK(cudaEventRecord(a)); k_filter<G, MODE><<<sms, 128 * G + 32, smemBytes>>>(dOps, dThr, dOut, dPass, iters); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); \
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms; } \
    CK(cudaGetLastError()); unsigned long long pass0 = 0; CK(cudaMemcpy(&pass0, dPass, 8, cudaMemcpyDeviceToHost)); \
    const double pairs = (double)sms * iters * kRows * (kCols / 2); \
    printf("{\"bench\": \"%s\", \"ms\": %.3f, \"tpairs_s\": %.3f, \"mma_tflops\": %.1f, \"cycles_per_tile_at_1.87GHz\": %.0f, \"passes_per_tile\": %.1f}\n", \
           NAME, best, pairs / best * 1e-9, (double)sms * iters * 2.0 * kRows * kCols * 8 * 2 / best * 1e-9, best * 1e-3 / iters * 1.87e9, (double)pass0 / iters); }
    RUN(1, 1, "mma only"); RUN(4, 1, "mma only, 16 idle epilogue warps"); RUN(4, 3, "16 warps, loads only (2 x LDTM.x32 each)"); RUN(4, 4, "TMEM reads alone, 16 warps"); RUN(1, 4, "TMEM reads alone, 4 warps"); RUN(4, 2, "16 warps, packed epilogue"); RUN(4, 2, "16 epilogue warps, packed epilogue");
    return 0;
}
