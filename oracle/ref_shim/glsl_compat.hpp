// SPDX-License-Identifier: WTFPL
// TEST INFRASTRUCTURE (oracle/): a minimal GLSL-450 vocabulary for C++20 so that the
// reference's compute-shader sources (shaders/shader.comp + shaders/include/*.glsl, read
// from /root/reference at build time by oracle/build_ref.py, never copied into this
// repo) compile and run on the CPU as the reference's own arithmetic.
// Everything is IEEE binary32; build with -ffp-contract=off so no FMA is formed.
#pragma once
#include <cmath>
#include <cstdint>

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    vec2(float a, float b) : x(a), y(b) {}
    vec2 xy() const { return *this; }
};
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec2 xy() const { return vec2(x, y); }
    vec2 xz() const { return vec2(x, z); }
    vec2 yz() const { return vec2(y, z); }
};
struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec3 rgb() const { return vec3(x, y, z); }
    void set_rgb(const vec3& v) { x = v.x; y = v.y; z = v.z; }
};
struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
};
struct uvec2 { unsigned x, y; };
struct uvec3 {
    unsigned x, y, z;
    uvec2 xy() const { return uvec2{x, y}; }
};
inline ivec2 make_ivec2(uvec2 v) { return ivec2((int)v.x, (int)v.y); }

inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3& operator*=(vec3& a, vec3 b) { a = a * b; return a; }
inline vec3& operator+=(vec3& a, vec3 b) { a = a + b; return a; }

inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
inline vec3 normalize(vec3 a) { return a / length(a); }
inline vec3 cross(vec3 a, vec3 b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline vec3 reflect(vec3 i, vec3 n) { return i - 2.0f * dot(n, i) * n; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline float radians(float d) { return d * 0.017453292519943295f; }
inline float fract(float v) { return v - std::floor(v); }
inline float tan(float v) { return std::tan(v); }
inline float sin(float v) { return std::sin(v); }
inline float sqrt(float v) { return std::sqrt(v); }
inline float pow(float b, float e) { return std::pow(b, e); }  // B6: C powf semantics

// Built-in inputs/outputs of the compute stage.
struct image2D { float* texels; int width, height; };
extern thread_local uvec3 gl_GlobalInvocationID;
extern image2D OutputImage;
inline void imageStore(image2D& img, ivec2 p, vec4 c) {
    if (p.x < 0 || p.y < 0 || p.x >= img.width || p.y >= img.height) return;
    float* t = img.texels + 4 * ((size_t)p.y * img.width + p.x);
    t[0] = c.x; t[1] = c.y; t[2] = c.z; t[3] = c.w;
}
// Run-time stand-ins for the shader's compile-time #defines (globals.glsl:9-17).
extern int ref_samples_per_pixel, ref_max_recursion_level, ref_image_width, ref_image_height;
// Instrumentation (does not alter arithmetic).
extern thread_local uint64_t ref_rays_traced, ref_depth_exhausted;
