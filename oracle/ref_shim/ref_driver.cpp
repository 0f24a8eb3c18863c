// SPDX-License-Identifier: WTFPL
// TEST INFRASTRUCTURE (oracle/): host loop that executes the reference compute shader
// (translated to C++ by oracle/build_ref.py into oracle/_ref/gen/) once per pixel, the way
// vkCmdDispatch does (reference Renderer.cpp:221), but over EVERY pixel (bug B9 defined).
#include "glsl_compat.hpp"
#include <omp.h>
#include <cstring>

thread_local uvec3 gl_GlobalInvocationID;
image2D OutputImage;
int ref_samples_per_pixel = 1, ref_max_recursion_level = 50, ref_image_width = 1280, ref_image_height = 720;
thread_local uint64_t ref_rays_traced = 0, ref_depth_exhausted = 0;

#include "gen/shader.comp.hpp"  // defines shader_main() and everything it includes

extern "C" {
// Renders W x H x spp with the reference's own shader arithmetic (hash RNG). rgba: W*H*4 floats.
// stats[0] = rays traced, stats[1] = samples that ran out of depth.
int ref_render(int width, int height, int spp, int depth, int threads, float* rgba, uint64_t* stats) {
    ref_image_width = width; ref_image_height = height;
    ref_samples_per_pixel = spp; ref_max_recursion_level = depth;
    OutputImage = image2D{rgba, width, height};
    uint64_t rays = 0, exhausted = 0;
    if (threads > 0) omp_set_num_threads(threads);
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : rays, exhausted)
    for (int y = 0; y < height; y++) {
        ref_rays_traced = 0; ref_depth_exhausted = 0;
        for (int x = 0; x < width; x++) {
            gl_GlobalInvocationID = uvec3{(unsigned)x, (unsigned)y, 0u};
            shader_main();
        }
        rays += ref_rays_traced; exhausted += ref_depth_exhausted;
    }
    if (stats) { stats[0] = rays; stats[1] = exhausted; }
    return 0;
}
int ref_world_count(void) { return (int)(sizeof(world) / sizeof(world[0])); }
// Copies the baked scene as 10 floats per sphere: centre xyz, radius, colour rgb, texture xyz.
void ref_world(float* out) {
    int n = ref_world_count();
    for (int i = 0; i < n; i++) {
        const sphere& s = world[i]; float* o = out + 10 * i;
        o[0] = s.center.x; o[1] = s.center.y; o[2] = s.center.z; o[3] = s.radius;
        o[4] = s.colour.x; o[5] = s.colour.y; o[6] = s.colour.z;
        o[7] = s.texture.x; o[8] = s.texture.y; o[9] = s.texture.z;
    }
}
float ref_rand(float x, float y) { return rand(vec2(x, y)); }
float ref_schlick(float cosine, float ior) { return schlick(cosine, ior); }
void ref_schlick_many(const float* cosine, const float* ior, int n, float* out) { for (int i = 0; i < n; i++) out[i] = schlick(cosine[i], ior[i]); }
int ref_refract(const float* v, const float* n, float eta, float* out) {
    vec3 r; bool ok = modified_refract(vec3(v[0], v[1], v[2]), vec3(n[0], n[1], n[2]), eta, r);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; return ok ? 1 : 0;
}
void ref_unit_vector(const float* seed, float* out) {
    vec3 r = random_in_unit_sphere(vec3(seed[0], seed[1], seed[2]));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
// First hit of an arbitrary ray against the baked scene: returns sphere index or -1, *t = root.
int ref_first_hit(const float* o, const float* d, float* t) {
    ray r = ray(vec3(o[0], o[1], o[2]), vec3(d[0], d[1], d[2]));
    hit_record rec; rec.min_t = 0.001f; rec.max_t = infinity; int best = -1;
    for (int i = 0; i < ref_world_count(); i++) if (hit_sphere(world[i], r, rec)) best = i;
    *t = rec.max_t; return best;
}
}
