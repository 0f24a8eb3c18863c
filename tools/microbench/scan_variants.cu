// SPDX-License-Identifier: WTFPL
// Variants of the brute-force scan inner loop (2 ray pairs per thread, 8 FFMA2 per pair per sphere)
// to find what keeps the FMA pipe below ~70 % in the straightforward form.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__constant__ float4 cscene[512];

#define PAIRS 2
#define FILTER(s, q) { const u64 cx = pk(s.x, s.x), cy = pk(s.y, s.y), cz = pk(s.z, s.z), nr2 = pk(-s.w, -s.w); \
    _Pragma("unroll") for (int p = 0; p < PAIRS; p++) { \
        u64 pu = ffma2(cz, uz[p], ou[p]); pu = ffma2(cy, uy[p], pu); pu = ffma2(cx, ux[p], pu); \
        u64 pv = ffma2(cz, vz[p], ov[p]); pv = ffma2(cy, vy[p], pv); pv = ffma2(cx, vx[p], pv); \
        u64 qq = ffma2(pu, pu, nr2); qq = ffma2(pv, pv, qq); q[p] = qq; } }

// MODE 0: mask per sphere (FMNMX3+FMNMX+FSETP+LOP3), as the kernel does
// MODE 1: no ALU work: q folded into an accumulator by one more FFMA2 per pair (counted as extra FMA work)
// MODE 2: like 0 but the scene comes from constant memory (no LDS)
// MODE 3: integer sign trick: OR of (q - m) sign bits accumulated with LOP3, one test per 8 spheres
// MODE 4: like 0, min over 2 spheres per compare (FMNMX3 x2 ... ) i.e. 8 values -> 1 flag per 2 spheres
template <int MODE, int UNROLL>
__global__ void __launch_bounds__(512, 1) k_scan(float* out, const float4* __restrict__ scene, int n, int reps) {
    extern __shared__ float4 s[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = scene[i];
    __syncthreads();
    const float* rs = out + 7 * (blockIdx.x * blockDim.x + threadIdx.x) % 4096;
    u64 ux[PAIRS], uy[PAIRS], uz[PAIRS], vx[PAIRS], vy[PAIRS], vz[PAIRS], ou[PAIRS], ov[PAIRS];
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        float t = threadIdx.x * .01f + p;
        ux[p] = pk(.6f + rs[p], .59f + rs[p+1]); uy[p] = pk(0.f + rs[p+2], .01f + rs[p+3]); uz[p] = pk(-.8f + rs[p+4], -.81f + rs[p+5]);
        vx[p] = pk(-.1f + rs[p+6], -.11f + rs[p+7]); vy[p] = pk(.98f + rs[p+8], .97f + rs[p+9]); vz[p] = pk(-.05f + rs[p+10], -.04f + rs[p+11]);
        ou[p] = pk(t + rs[p+12], t + .5f + rs[p+13]); ov[p] = pk(1.f + p + rs[p+14], 2.f + p + rs[p+15]);
    }
    const float m = 1e-4f + rs[20];
    unsigned total = 0; u64 acc[PAIRS] = {0, 0};
#pragma unroll 1
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int base = 0; base < n; base += 32) {
            unsigned mask = 0; const int lim = min(32, n - base);
#pragma unroll 1
            for (int kk = 0; kk < lim; kk += UNROLL) {
                unsigned m8 = 0; unsigned sg = 0;
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const float4 sp = (MODE == 2) ? cscene[base + kk + u] : s[base + kk + u];
                    u64 q[PAIRS]; FILTER(sp, q);
                    if (MODE == 1) { acc[0] = ffma2(q[0], pk(1.f, 1.f), acc[0]); acc[1] = ffma2(q[1], pk(1.f, 1.f), acc[1]); }
                    else if (MODE == 3) {
                        float a, b, c, d; upk(q[0], a, b); upk(q[1], c, d);
                        sg |= (__float_as_uint(a) | __float_as_uint(b)) | (__float_as_uint(c) | __float_as_uint(d));   // m folded into R2 by the caller
                    } else if (MODE == 4) {
                        float a, b, c, d; upk(q[0], a, b); upk(q[1], c, d);
                        float mn = fminf(fminf(a, b), fminf(c, d));
                        if (u & 1) { float prev = __uint_as_float(sg); if (fminf(prev, mn) <= m) m8 |= 3u << (u - 1); } else sg = __float_as_uint(mn);
                    } else {
                        float a, b, c, d; upk(q[0], a, b); upk(q[1], c, d);
                        float mn = fminf(fminf(a, b), fminf(c, d));
                        if (mn <= m) m8 |= 1u << u;
                    }
                }
                if (MODE == 3) { if ((int)sg < 0) m8 = 0xffu; }
                mask |= m8 << kk;
            }
            total += __popc(mask);
        }
#pragma unroll
        for (int p = 0; p < PAIRS; p++) ou[p] = ffma2(ou[p], pk(1.f, 1.f), pk(1e-3f, 1e-3f));
    }
    float lo, hi; upk(acc[0], lo, hi); float lo2, hi2; upk(acc[1], lo2, hi2);
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)total + lo + hi + lo2 + hi2;
}
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
template <typename F> static float timeit(F f) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f(); CK(cudaDeviceSynchronize()); float best = 1e30f;
    for (int i = 0; i < 5; i++) { CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms; }
    CK(cudaGetLastError()); return best;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); int sms = p.multiProcessorCount;
    float* out; CK(cudaMalloc(&out, 4 * sms * 2048)); CK(cudaMemset(out, 0, 4 * sms * 2048));
    const int n = 488; float4 h[512];
    for (int i = 0; i < 512; i++) h[i] = make_float4((i % 22) - 11.f, 0.2f, (i / 22) - 11.f, i < 484 ? 0.04f : -1e30f);
    float4* scene; CK(cudaMalloc(&scene, sizeof h)); CK(cudaMemcpy(scene, h, sizeof h, cudaMemcpyHostToDevice));
    CK(cudaMemcpyToSymbol(cscene, h, sizeof h));
    const int reps = 64, threads = 512, blocks = sms; size_t sh = sizeof(float4) * 512;
#define RUN(M, U, NAME) { float ms = timeit([&] { k_scan<M, U><<<blocks, threads, sh>>>(out, scene, n, reps); }); \
    double tests = (double)blocks * threads * 4 * n * reps; double fma = tests * (M == 1 ? 9 : 8); \
    printf("{\"bench\": \"%s\", \"ms\": %.4f, \"gtests_s\": %.1f, \"fma_pipe_frac\": %.3f}\n", NAME, ms, tests / ms * 1e-6, fma / ms * 1e-9 / 36.96); }
    RUN(0, 8, "mask_u8"); RUN(0, 4, "mask_u4"); RUN(0, 16, "mask_u16"); RUN(1, 8, "noalu_u8"); RUN(2, 8, "const_u8"); RUN(3, 8, "signor_u8"); RUN(4, 8, "pairmin_u8");
    return 0;
}
