// SPDX-License-Identifier: WTFPL
// The scan loop of render_scanu_kernel in isolation (constant-bank sphere stream, 7 FFMA2 per ray pair per sphere,
// one flag per sphere per thread): what does the loop alone reach, and what would sharing the sphere operands between
// more ray pairs per thread buy?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scanu_loop scanu_loop.cu
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ void flag_if_ge(uint32_t& m, float q, float s, uint32_t bit) {
    asm("{\n\t.reg .pred p;\n\tsetp.ge.f32 p, %1, %2;\n\t@p or.b32 %0, %0, %3;\n\t}" : "+r"(m) : "f"(q), "f"(s), "r"(bit));
}
__constant__ float4 cscene[512];

// MODE 0: as the kernel (max over the thread's rays, compare, predicated OR)
// MODE 1: no flag work at all (q summed into an accumulator with one extra FFMA2 per pair; counted as FMA work)
// MODE 2: flags by sign: threshold folded in with one more FFMA2 per pair (q - thr), then max + funnel shift of the sign bit
template <int THREADS, int PAIRS, int MODE, int UNROLL>
__global__ void __launch_bounds__(THREADS, 1) k_scan(float* out, int n, int reps) {
    const int tid = threadIdx.x;
    u64 dx[PAIRS], dy[PAIRS], dz[PAIRS], ee[PAIRS], ax[PAIRS], ay[PAIRS], az[PAIRS], nb[PAIRS], acc[PAIRS];
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        float f = 0.001f * (tid + 7 * p);
        dx[p] = pk(0.6f + f, 0.59f - f); dy[p] = pk(-0.1f, -0.12f + f); dz[p] = pk(0.79f, 0.8f - f); ee[p] = pk(-9.f + f, -9.1f);
        ax[p] = pk(26.f, 26.f + f); ay[p] = pk(4.f, 4.f - f); az[p] = pk(6.f + f, 6.f); nb[p] = pk(-182.f, -182.f + f);
        acc[p] = pk(0.f, 0.f);
    }
    uint32_t total = 0;
    for (int rep = 0; rep < reps; rep++) {
#pragma unroll 1
        for (int base = 0; base < n; base += 32) {
            uint32_t mask = 0u;
#pragma unroll 1
            for (int kk = 0; kk < 32; kk += UNROLL) {
                uint32_t mEven = 0u, mOdd = 0u;
#pragma unroll
                for (int u = 0; u < UNROLL; u++) {
                    const float4 s = cscene[base + kk + u];
                    const u64 cx = pk(s.x, s.x), cy = pk(s.y, s.y), cz = pk(s.z, s.z);
                    float qmax = -__int_as_float(0x7f800000);
#pragma unroll
                    for (int p = 0; p < PAIRS; p++) {
                        u64 t = ffma2(cz, dz[p], ee[p]); t = ffma2(cy, dy[p], t); t = ffma2(cx, dx[p], t);
                        u64 w = ffma2(cz, az[p], nb[p]); w = ffma2(cy, ay[p], w); w = ffma2(cx, ax[p], w);
                        if (MODE == 2) w = ffma2(pk(s.w, s.w), pk(-1.f, -1.f), w);
                        const u64 q = ffma2(t, t, w);
                        if (MODE == 1) acc[p] = ffma2(q, pk(1e-6f, 1e-6f), acc[p]);
                        else { float q0, q1; upk(q, q0, q1); qmax = fmaxf(qmax, fmaxf(q0, q1)); }
                    }
                    if (MODE == 0) flag_if_ge(u & 1 ? mOdd : mEven, qmax, s.w, 1u << u);
                    if (MODE == 2) mEven = __funnelshift_l(__float_as_uint(qmax), mEven, 1);
                }
                if (MODE == 2) mask = (mask << UNROLL) | (~mEven & ((1u << UNROLL) - 1u)); else mask |= (mEven | mOdd) << kk;
            }
            total += __popc(mask);
        }
#pragma unroll
        for (int p = 0; p < PAIRS; p++) ee[p] = ffma2(ee[p], pk(1.f, 1.f), pk(1e-3f, 1e-3f));
    }
    float sum = (float)total;
#pragma unroll
    for (int p = 0; p < PAIRS; p++) { float lo, hi; upk(acc[p], lo, hi); sum += lo + hi; }
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
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0)); const int sms = p.multiProcessorCount;
    float* out; CK(cudaMalloc(&out, 4 * sms * 2048)); CK(cudaMemset(out, 0, 4 * sms * 2048));
    const int n = 480; float4 h[512];
    for (int i = 0; i < 512; i++) h[i] = make_float4((i % 22) - 11.f, 0.2f, (i / 22) - 11.f, i < 484 ? (float)(i % 22) * (i % 22) + 1.f : 3e38f);
    CK(cudaMemcpyToSymbol(cscene, h, sizeof h));
    int khz = 0; CK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
    const int reps = 64;
#define RUN(T, P, M, U, NAME) { float ms = timeit([&] { k_scan<T, P, M, U><<<sms, T>>>(out, n, reps); }); \
    double tests = (double)sms * T * 2 * P * n * reps; double fma2 = tests / 2 * (M == 0 ? 7 : 8); \
    double pipe = fma2 * 2 / 32 / (sms * 4.0) / (ms * 1e-3);   /* FMA-pipe cycles per second per scheduler */ \
    printf("{\"bench\": \"%s\", \"ms\": %.4f, \"gtests_s\": %.1f, \"fma_pipe_ghz\": %.3f, \"of_attr_clock\": %.3f}\n", NAME, ms, tests / ms * 1e-6, pipe * 1e-9, pipe / (khz * 1e3)); }
    RUN(1024, 1, 0, 16, "1024x1 flags u16"); RUN(1024, 1, 0, 8, "1024x1 flags u8"); RUN(1024, 1, 1, 16, "1024x1 noflags u16"); RUN(1024, 1, 2, 16, "1024x1 signshift u16");
    RUN(512, 2, 0, 16, "512x2 flags u16"); RUN(512, 2, 0, 8, "512x2 flags u8"); RUN(512, 2, 1, 8, "512x2 noflags u8");
    RUN(1024, 2, 0, 8, "1024x2 flags u8 (64 regs)"); RUN(768, 2, 0, 8, "768x2 flags u8"); RUN(768, 2, 0, 16, "768x2 flags u16"); RUN(512, 3, 0, 8, "512x3 flags u8"); RUN(512, 4, 0, 8, "512x4 flags u8"); RUN(256, 4, 0, 8, "256x4 flags u8");
    return 0;
}
