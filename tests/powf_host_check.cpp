// SPDX-License-Identifier: WTFPL
// TEST INFRASTRUCTURE: compiles the product's csrc/vcrt_powf.cuh for the HOST (the header is written for nvcc and for a plain
// C++ compiler alike) so that tests/test_powf.py can compare the restated powf(x, 5) with the C library of the machine at hand
// and with the golden vectors of the reference build, without a GPU.  Built by the test with g++ -O2 -ffp-contract=off.
#include "../vulkancomputeraytracing_b200/csrc/vcrt_powf.cuh"
#include <cmath>
#include <cstdint>

extern "C" {
// x^5 by the restatement and by libm for n inputs; returns the number of differing results (NaN == NaN), first one in *bad.
long powf5_compare(const float* x, long n, float* bad) {
    long mismatches = 0;
    for (long i = 0; i < n; i++) {
        const float mine = vcrt::powf5(x[i]), libm = powf(x[i], 5.0f);
        const bool same = vcrt::pw_bits(mine) == vcrt::pw_bits(libm) || (mine != mine && libm != libm);
        if (!same && mismatches++ == 0 && bad) *bad = x[i];
    }
    return mismatches;
}
// functions.glsl:58-62 with the device code's operation order (vcrt_shade.cuh schlick) around powf5
void schlick_many(const float* cosine, const float* ior, long n, float* out) {
    for (long i = 0; i < n; i++) {
        volatile float r0 = (1.f - ior[i]) / (1.f + ior[i]);
        r0 = r0 * r0;
        volatile float p = vcrt::powf5(1.f - cosine[i]);
        volatile float t = (1.f - r0) * p;
        out[i] = r0 + t;
    }
}
}
