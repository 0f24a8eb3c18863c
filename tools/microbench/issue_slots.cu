// SPDX-License-Identifier: WTFPL
// Does an FFMA2 (2 FMA-pipe passes) leave its second issue cycle free for other pipes?
// Streams of FFMA2 interleaved with ALU (LOP3/IADD3), FMNMX or warp-uniform LDS.128 work.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0,%1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }

// MODE 0: 8 FFMA2 chains only. 1: + one LOP3 per FFMA2. 2: + one FMNMX per FFMA2.
// 3: + one warp-uniform LDS.128 per 8 FFMA2. 4: + LDS.128 per 16 FFMA2. 5: + 2 LOP3 per FFMA2.
// 6: scalar FFMA x8 + LOP3 per FFMA (reference for a 1-pass op)
template <int MODE>
__global__ void k(float* out, int iters, float a, float b, const float4* __restrict__ g) {
    __shared__ float4 s[512];
    for (int i = threadIdx.x; i < 512; i += blockDim.x) s[i] = g[i];
    __syncthreads();
    u64 x[8]; float y[8]; u64 pb = pk(b, b * 1.0001f);
    unsigned v[8]; float f[8];
#pragma unroll
    for (int c = 0; c < 8; c++) { x[c] = pk(threadIdx.x * 1e-3f + c, c); v[c] = threadIdx.x * 77 + c; f[c] = c * 0.5f + threadIdx.x; y[c] = c + threadIdx.x * 0.25f; }
    float4 acc = make_float4(0, 0, 0, 0);
#pragma unroll 1
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (MODE == 3 || (MODE == 4 && (u & 1) == 0)) { float4 t = s[(i * 4 + u) & 511]; acc.x += t.x; acc.y = fmaxf(acc.y, t.y); acc.z = fmaxf(acc.z, t.z); acc.w = fmaxf(acc.w, t.w); }
#pragma unroll
            for (int c = 0; c < 8; c++) {
                if (MODE == 6) y[c] = fmaf(y[c], a, b); else x[c] = ffma2(pk(a, a), x[c], pb);
                if (MODE == 1 || MODE == 5 || MODE == 6) v[c] = (v[c] ^ (v[c] >> 3)) & 0x7fffffffu;   // SHF + LOP3
                if (MODE == 5) v[c] = v[c] + (v[c] << 1);
                if (MODE == 2) f[c] = fminf(f[c], f[(c + 1) & 7] + 0.f) ;
            }
        }
    }
    float sum = acc.x + acc.y + acc.z + acc.w;
#pragma unroll
    for (int c = 0; c < 8; c++) { float lo, hi; upk(x[c], lo, hi); sum += lo + hi + (float)v[c] + f[c] + y[c]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
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
    float* out; CK(cudaMalloc(&out, 4 * sms * 2048)); float4* g; CK(cudaMalloc(&g, 16 * 512)); CK(cudaMemset(g, 0, 16 * 512));
    const int iters = 4096, threads = 512, blocks = sms * 4;
    double n2 = (double)blocks * threads * iters * 4 * 8;  // FFMA2 (or FFMA) instr count x lanes
#define RUN(M, NAME) { float ms = timeit([&] { k<M><<<blocks, threads>>>(out, iters, 1.0001f, 1e-4f, g); }); \
    printf("{\"bench\": \"%s\", \"ms\": %.4f, \"fma_lane_inst_T\": %.2f}\n", NAME, ms, n2 * (M == 6 ? 1 : 2) / ms * 1e-9); }
    RUN(0, "ffma2_only"); RUN(1, "ffma2+shf+lop3"); RUN(5, "ffma2+shf+lop3+lea"); RUN(2, "ffma2+fadd+fmnmx");
    RUN(3, "ffma2x8+lds128"); RUN(4, "ffma2x16+lds128"); RUN(6, "ffma+shf+lop3");
    return 0;
}
