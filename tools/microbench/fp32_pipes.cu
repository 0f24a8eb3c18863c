// SPDX-License-Identifier: WTFPL
// Micro-benchmarks that size the FP32 roofline of a B200 SM for the path tracer:
//   FFMA / FFMA2 (packed f32x2, sm_100a) issue rates, the .F32 broadcast operand form,
//   and the cost of a warp-uniform LDS.128 next to the FMA pipe.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o fp32_pipes fp32_pipes.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>

typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ u64 fadd2(u64 a, u64 b) { u64 r; asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 fmul2(u64 a, u64 b) { u64 r; asm volatile("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

#define CHAINS 8

// 1: scalar FFMA, CHAINS independent chains
__global__ void k_ffma(float* out, int iters, float a, float b) {
    float x[CHAINS];
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = threadIdx.x * 1e-3f + c;
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int c = 0; c < CHAINS; c++) x[c] = fmaf(x[c], a, b);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) s += x[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 2: packed FFMA2, CHAINS independent chains of 64-bit registers
__global__ void k_ffma2(float* out, int iters, float a, float b) {
    u64 x[CHAINS]; u64 pa = pk(a, a * 1.0001f), pb = pk(b, b * 1.0001f);
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = pk(threadIdx.x * 1e-3f + c, c);
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int c = 0; c < CHAINS; c++) x[c] = ffma2(x[c], pa, pb);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) { float lo, hi; upk(x[c], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 3: FFMA2 with a scalar-broadcast (.F32) multiplicand
__global__ void k_ffma2_bcast(float* out, int iters, float a, float b) {
    u64 x[CHAINS]; u64 pb = pk(b, b * 1.0001f);
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = pk(threadIdx.x * 1e-3f + c, c);
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int c = 0; c < CHAINS; c++) x[c] = ffma2(pk(a, a), x[c], pb);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) { float lo, hi; upk(x[c], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 4: FADD2 / FMUL2
__global__ void k_fadd2(float* out, int iters, float a) {
    u64 x[CHAINS]; u64 pa = pk(a, a * 1.0001f);
#pragma unroll
    for (int c = 0; c < CHAINS; c++) x[c] = pk(threadIdx.x * 1e-3f + c, c);
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int c = 0; c < CHAINS; c++) x[c] = fadd2(x[c], pa);
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS; c++) { float lo, hi; upk(x[c], lo, hi); s += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// 5: mixed FFMA (scalar) + FFMA2 to see whether they share one pipe
__global__ void k_mix(float* out, int iters, float a, float b) {
    u64 x[4]; float y[4]; u64 pa = pk(a, a * 1.0001f), pb = pk(b, b * 1.0001f);
#pragma unroll
    for (int c = 0; c < 4; c++) { x[c] = pk(threadIdx.x * 1e-3f + c, c); y[c] = c + threadIdx.x; }
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 8; u++)
#pragma unroll
            for (int c = 0; c < 4; c++) { x[c] = ffma2(x[c], pa, pb); y[c] = fmaf(y[c], a, b); }
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < 4; c++) { float lo, hi; upk(x[c], lo, hi); s += lo + hi + y[c]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// 6: scan skeletons. Scene float4 (cx,cy,cz,r2) in shared memory, warp-uniform LDS.128.
// 6a: scalar direct form, RAYS rays per thread, 10 FMA-pipe instr / test (a == 1)
template <int RAYS>
__global__ void k_scan_scalar(float* out, const float4* __restrict__ scene, int n, int reps) {
    const float* rs = out + 7 * (blockIdx.x * blockDim.x + threadIdx.x) % 4096;
    extern __shared__ float4 s[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = scene[i];
    __syncthreads();
    float ox[RAYS], oy[RAYS], oz[RAYS], dx[RAYS], dy[RAYS], dz[RAYS]; int cnt[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) { ox[r] = threadIdx.x * .01f + r + rs[0]; oy[r] = 2.f + rs[1+r]; oz[r] = 3.f + r + rs[2+r]; dx[r] = -.9f + rs[3+r]; dy[r] = -.1f * r + rs[4+r]; dz[r] = -.2f + rs[5+r]; cnt[r] = 0; }
#pragma unroll 1
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 4
        for (int i = 0; i < n; i++) {
            float4 sp = s[i];
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float x = ox[r] - sp.x, y = oy[r] - sp.y, z = oz[r] - sp.z;
                float hb = fmaf(z, dz[r], fmaf(y, dy[r], x * dx[r]));
                float c = fmaf(z, z, fmaf(y, y, fmaf(x, x, -sp.w)));
                float disc = fmaf(hb, hb, -c);
                if (disc >= 0.f) cnt[r] += i;
            }
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) ox[r] += 1e-3f;
    }
    int t = 0;
#pragma unroll
    for (int r = 0; r < RAYS; r++) t += cnt[r];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)t;
}
// 6b: packed 2-D basis filter: per ray pair 8 FFMA2 / test, PAIRS pairs per thread
template <int PAIRS>
__global__ void k_scan_packed(float* out, const float4* __restrict__ scene, int n, int reps) {
    const float* rs = out + 7 * (blockIdx.x * blockDim.x + threadIdx.x) % 4096;
    extern __shared__ float4 s[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = scene[i];
    __syncthreads();
    u64 ux[PAIRS], uy[PAIRS], uz[PAIRS], vx[PAIRS], vy[PAIRS], vz[PAIRS], ou[PAIRS], ov[PAIRS]; float m[PAIRS]; int cnt[PAIRS];
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        float t = threadIdx.x * .01f + p;
        ux[p] = pk(.6f + rs[p], .59f + rs[p+1]); uy[p] = pk(0.f + rs[p+2], .01f + rs[p+3]); uz[p] = pk(-.8f + rs[p+4], -.81f + rs[p+5]);
        vx[p] = pk(-.1f + rs[p+6], -.11f + rs[p+7]); vy[p] = pk(.98f + rs[p+8], .97f + rs[p+9]); vz[p] = pk(-.05f + rs[p+10], -.04f + rs[p+11]);
        ou[p] = pk(t + rs[p+12], t + .5f + rs[p+13]); ov[p] = pk(1.f + p + rs[p+14], 2.f + p + rs[p+15]); m[p] = 1e-4f + rs[p+16]; cnt[p] = 0;
    }
#pragma unroll 1
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 4
        for (int i = 0; i < n; i++) {
            float4 sp = s[i];
            u64 cx = pk(sp.x, sp.x), cy = pk(sp.y, sp.y), cz = pk(sp.z, sp.z), nr2 = pk(-sp.w, -sp.w);
#pragma unroll
            for (int p = 0; p < PAIRS; p++) {
                u64 pu = ffma2(cz, uz[p], ou[p]); pu = ffma2(cy, uy[p], pu); pu = ffma2(cx, ux[p], pu);
                u64 pv = ffma2(cz, vz[p], ov[p]); pv = ffma2(cy, vy[p], pv); pv = ffma2(cx, vx[p], pv);
                u64 q = ffma2(pu, pu, nr2); q = ffma2(pv, pv, q);
                float lo, hi; upk(q, lo, hi);
                if (fminf(lo, hi) <= m[p]) cnt[p] += i;
            }
        }
#pragma unroll
        for (int p = 0; p < PAIRS; p++) ou[p] = fadd2(ou[p], pk(1e-3f, 1e-3f));
    }
    int t = 0;
#pragma unroll
    for (int p = 0; p < PAIRS; p++) t += cnt[p];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)t;
}
// 6c: scalar 2-D basis filter (8 FFMA / test), RAYS rays per thread
template <int RAYS>
__global__ void k_scan_basis_scalar(float* out, const float4* __restrict__ scene, int n, int reps) {
    const float* rs = out + 7 * (blockIdx.x * blockDim.x + threadIdx.x) % 4096;
    extern __shared__ float4 s[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) s[i] = scene[i];
    __syncthreads();
    float ux[RAYS], uy[RAYS], uz[RAYS], vx[RAYS], vy[RAYS], vz[RAYS], ou[RAYS], ov[RAYS]; int cnt[RAYS];
#pragma unroll
    for (int r = 0; r < RAYS; r++) {
        float t = threadIdx.x * .01f + r;
        ux[r] = .6f + rs[r]; uy[r] = .01f * r + rs[r+1]; uz[r] = -.8f + rs[r+2]; vx[r] = -.1f + rs[r+3]; vy[r] = .98f + rs[r+4]; vz[r] = -.05f + rs[r+5]; ou[r] = t + rs[r+6]; ov[r] = 1.f + r + rs[r+7]; cnt[r] = 0;
    }
#pragma unroll 1
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 4
        for (int i = 0; i < n; i++) {
            float4 sp = s[i];
#pragma unroll
            for (int r = 0; r < RAYS; r++) {
                float pu = fmaf(sp.x, ux[r], fmaf(sp.y, uy[r], fmaf(sp.z, uz[r], ou[r])));
                float pv = fmaf(sp.x, vx[r], fmaf(sp.y, vy[r], fmaf(sp.z, vz[r], ov[r])));
                float q = fmaf(pv, pv, fmaf(pu, pu, -sp.w));
                if (q <= 1e-4f) cnt[r] += i;
            }
        }
#pragma unroll
        for (int r = 0; r < RAYS; r++) ou[r] += 1e-3f;
    }
    int t = 0;
#pragma unroll
    for (int r = 0; r < RAYS; r++) t += cnt[r];
    out[blockIdx.x * blockDim.x + threadIdx.x] = (float)t;
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <typename F>
static float timeit(F f, int reps = 5) {
    cudaEvent_t a, b; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
    f(); CK(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int i = 0; i < reps; i++) {
        CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
        float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main(int argc, char** argv) {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int sms = p.multiProcessorCount; int clk_khz = 0; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz_attr\": %d}\n", p.name, sms, clk_khz);
    float* out; CK(cudaMalloc(&out, sizeof(float) * sms * 8 * 1024)); CK(cudaMemset(out, 0, sizeof(float) * sms * 8 * 1024));
    const int iters = 4096;
    for (int threads : {256, 512, 1024}) {
        int blocks = sms * (2048 / threads);
        double lanes = (double)blocks * threads;
        float ms;
        ms = timeit([&] { k_ffma<<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); });
        printf("{\"bench\": \"ffma\", \"threads\": %d, \"ms\": %.4f, \"tflops\": %.2f}\n", threads, ms, lanes * iters * 8 * CHAINS * 2 / ms * 1e-9);
        ms = timeit([&] { k_ffma2<<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); });
        printf("{\"bench\": \"ffma2\", \"threads\": %d, \"ms\": %.4f, \"tflops\": %.2f}\n", threads, ms, lanes * iters * 8 * CHAINS * 4 / ms * 1e-9);
        ms = timeit([&] { k_ffma2_bcast<<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); });
        printf("{\"bench\": \"ffma2_bcast\", \"threads\": %d, \"ms\": %.4f, \"tflops\": %.2f}\n", threads, ms, lanes * iters * 8 * CHAINS * 4 / ms * 1e-9);
        ms = timeit([&] { k_fadd2<<<blocks, threads>>>(out, iters, 1e-4f); });
        printf("{\"bench\": \"fadd2\", \"threads\": %d, \"ms\": %.4f, \"tinst_lane\": %.2f}\n", threads, ms, lanes * iters * 8 * CHAINS * 2 / ms * 1e-9);
        ms = timeit([&] { k_mix<<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f); });
        printf("{\"bench\": \"mix_ffma_ffma2\", \"threads\": %d, \"ms\": %.4f, \"tflops\": %.2f}\n", threads, ms, lanes * iters * 8 * 4 * 6 / ms * 1e-9);
    }
    // scan skeletons
    const int n = 485; float4* scene; CK(cudaMalloc(&scene, sizeof(float4) * n));
    float4* h = (float4*)malloc(sizeof(float4) * n);
    for (int i = 0; i < n; i++) h[i] = make_float4((i % 22) - 11.f, 0.2f, (i / 22) - 11.f, 0.04f);
    CK(cudaMemcpy(scene, h, sizeof(float4) * n, cudaMemcpyHostToDevice));
    const int reps = 64; size_t sh = sizeof(float4) * n;
#define SCAN(NAME, KERN, RPT, THREADS) do { int threads = THREADS; int blocks = sms * (1024 / threads < 1 ? 1 : 1024 / threads); \
        float ms = timeit([&] { KERN<<<blocks, threads, sh>>>(out, scene, n, reps); }); \
        double tests = (double)blocks * threads * (RPT) * n * reps; \
        printf("{\"bench\": \"%s\", \"threads\": %d, \"blocks\": %d, \"rays_per_thread\": %d, \"ms\": %.4f, \"gtests_s\": %.1f, \"alg_tflops17\": %.2f}\n", NAME, threads, blocks, RPT, ms, tests / ms * 1e-6, tests * 17 / ms * 1e-9); } while (0)
    SCAN("scan_scalar_r1", k_scan_scalar<1>, 1, 512);
    SCAN("scan_scalar_r2", k_scan_scalar<2>, 2, 512);
    SCAN("scan_scalar_r4", k_scan_scalar<4>, 4, 512);
    SCAN("scan_scalar_r4_256", k_scan_scalar<4>, 4, 256);
    SCAN("scan_basis_scalar_r2", k_scan_basis_scalar<2>, 2, 512);
    SCAN("scan_basis_scalar_r4", k_scan_basis_scalar<4>, 4, 512);
    SCAN("scan_packed_p1", k_scan_packed<1>, 2, 512);
    SCAN("scan_packed_p2", k_scan_packed<2>, 4, 512);
    SCAN("scan_packed_p2_256", k_scan_packed<2>, 4, 256);
    SCAN("scan_packed_p3", k_scan_packed<3>, 6, 512);
    SCAN("scan_packed_p4", k_scan_packed<4>, 8, 256);
    SCAN("scan_packed_p4_512", k_scan_packed<4>, 8, 512);
    return 0;
}
