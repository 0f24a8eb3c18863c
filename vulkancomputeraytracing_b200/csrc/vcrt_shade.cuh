/* @file vcrt_shade.cuh

    Exact ray-sphere test, primary-ray generation, the three materials and the sky:
    the parts of the path that are evaluated once per ray (not once per ray per sphere).
    SPDX-License-Identifier: WTFPL
*/
#pragma once
#include "vcrt_device.cuh"
#include "vcrt_powf.cuh"

namespace vcrt {

// functions.glsl:14-33 for one sphere, `a` = dot(d,d) hoisted by the caller (same value every
// time the reference recomputes it).  Updates rec when this sphere gives a nearer accepted root.
// `shortcut`: the ray is tame (tame_ray, vcrt_device.cuh), which is what the early reject below is proven for; the exhaustive
// scans of non-tame rays and the in-order cross-check kernel pass false and run the reference's statements literally.
VCRT_DEV void hit_exact(float4 shape, int index, const Ray& r, float a, HitRec& rec, bool shortcut) {
    V3 oc = vsub(r.o, v3(shape.x, shape.y, shape.z));
    float half_b = vdot(oc, r.d);
    float c = fsub(vdot(oc, oc), fmul(shape.w, shape.w));
    float disc = fsub(fmul(half_b, half_b), fmul(a, c));
    if (disc < 0) return;
    // Moving away from a sphere we are not inside (half_b > 0, c >= 0): a*c >= 0 gives disc <= fl(half_b^2), and
    // sqrt(fl(x*x)) == |x| exactly in round-to-nearest binary arithmetic, so sqrtd <= half_b and both numerators are <= 0:
    // the reference rejects both roots against t_min.  Same outcome, without the square root and the two divisions
    // (this is every ray leaving the surface it just scattered from, and every sphere behind the origin).  Not for rays whose
    // half_b^2 or a*c may overflow: inf - inf = NaN passes `disc < 0`, the reference then accepts the sphere at t = NaN.
    if (shortcut && half_b > 1e-18f && c >= 0.f) return;   // 1e-18: half_b^2 stays a normal number, which the sqrt identity needs
    float sqrtd = fsqrt(disc);
    float root = fdiv(fsub(-half_b, sqrtd), a);
    if (root <= kTMin || rec.t <= root) {
        root = fdiv(fadd(-half_b, sqrtd), a);
        if (root <= kTMin || rec.t <= root) return;
    }
    rec.t = root;
    rec.index = index;
}

// shader.comp:42-52: the camera ray of pixel (px,py), sample jitter (ju,jv) in [0,1).
VCRT_DEV Ray primary_ray(const Frame& f, uint32_t px, uint32_t py, float ju, float jv) {
    V3 pixel_center = vadd(vadd(f.pixel00, vscale(f.du, (float)px)), vscale(f.dv, (float)py));
    V3 random_square = vadd(vscale(f.du, fadd(-0.5f, ju)), vscale(f.dv, fadd(-0.5f, jv)));
    V3 pixel_sample = vadd(pixel_center, random_square);
    Ray r; r.o = f.center; r.d = vsub(pixel_sample, f.center);  // never normalised (bug B3)
    return r;
}

// functions.glsl:85-89: sky colour seen along d.  mix(vec3(1), vec3(.5,.7,1), a)
VCRT_DEV V3 sky(V3 d) {
    // normalize(d).y is the only component used: d.y / length(d), the same two operations the full normalize applies to y
    float unit_y = fdiv(d.y, fsqrt(vdot(d, d)));
    float a = fmul(0.5f, fadd(unit_y, 1.0f));
    float ia = fsub(1.0f, a);
    return v3(fadd(fmul(1.f, ia), fmul(.5f, a)), fadd(fmul(1.f, ia), fmul(.7f, a)), fadd(fmul(1.f, ia), fmul(1.f, a)));
}

// functions.glsl:42-44 with the three hash draws replaced by Philox draws (bug B4 kept:
// a unit vector in the positive octant).
VCRT_DEV V3 random_in_unit_sphere(const float u[4]) { return vunit(v3(u[0], u[1], u[2])); }

// functions.glsl:46-56
VCRT_DEV bool modified_refract(V3 v, V3 n, float ni_over_nt, V3& refracted) {
    float dt = vdot(v, n);
    float disc = fsub(1.f, fmul(fmul(ni_over_nt, ni_over_nt), fsub(1.f, fmul(dt, dt))));
    if (disc > 0.f) {
        V3 lhs = vscale(vsub(v, vscale(n, dt)), ni_over_nt);
        refracted = vsub(lhs, vscale(n, fsqrt(disc)));
        return true;
    }
    return false;
}

// functions.glsl:58-62.  pow(x, 5.) carries the bits of C powf as the reference build evaluates it (vcrt_powf.cuh): three
// binary32 multiplies differ from that by an ulp on half of the inputs, which flips `rand < reflect_prob` now and then.
// A negative base (bug B6: 90 % of the calls, the cosine is an unnormalised dot product) gives a negative power, so the
// caller's test is false and the ray refracts, as in the oracle.  One binary64 chain per refracting glass hit.
// Out of line: the binary64 chain would otherwise raise the register count of every kernel for a routine that one ray in
// twenty reaches (inlined: 44 bytes of spills in the brute-force kernel at its 64-register budget).
__device__ __noinline__ float powf5_call(float x) { return powf5(x); }
VCRT_DEV float schlick(float cosine, float ior) {
    float r0 = fdiv(fsub(1.f, ior), fadd(1.f, ior));
    r0 = fmul(r0, r0);
    return fadd(r0, fmul(fsub(1.f, r0), powf5_call(fsub(1.f, cosine))));
}

// textures.glsl:19-71.  Scatters `ray` at a hit on sphere `index` at parameter t; multiplies
// `colour` (the path throughput) by the attenuation.  Returns false for a texture id outside
// 1..3 (reference: silent no-op that re-hits until the depth runs out; here: absorbed + counted).
VCRT_DEV bool scatter(const float4 shape, const float4 surf, int kind, float t, Ray& ray, V3& colour, const float u[4]) {
    V3 point = vadd(vscale(ray.d, t), ray.o);                                   // functions.glsl:35
    V3 normal = vdiv(vsub(point, v3(shape.x, shape.y, shape.z)), shape.w);      // functions.glsl:36, outward
    V3 albedo = v3(surf.x, surf.y, surf.z);
    float param = surf.w;
    if (kind == 1 || kind == 2) {
        // lambertian (textures.glsl:19-25; bug B4b: extra * texture.y) and metal (textures.glsl:58-63) share their one expensive
        // step, the unit vector of functions.glsl:42-44, so a warp with both materials walks it once instead of twice
        const V3 rv = random_in_unit_sphere(u);
        const bool lambert = kind == 1;
        const V3 base = lambert ? normal : vreflect(ray.d, normal);
        const V3 offset = lambert ? rv : vscale(rv, param);
        const V3 tint = vmul(colour, albedo);
        colour = lambert ? vscale(tint, param) : tint;
        ray.o = point; ray.d = vadd(base, offset);
        return true;
    }
    if (kind == 3) {  // glass, textures.glsl:27-56
        V3 outward_normal, refracted = v3(0.f, 0.f, 0.f);
        V3 reflected = vreflect(ray.d, normal);
        float ni_over_nt, reflect_prob, cosine;
        float dn = vdot(ray.d, normal);
        if (dn > 0.f) {
            outward_normal = vneg(normal);
            ni_over_nt = param;
            cosine = fsqrt(fsub(1.f, fmul(fmul(param, param), fsub(1.f, fmul(dn, dn)))));
        } else {
            outward_normal = normal;
            ni_over_nt = fdiv(1.f, param);
            cosine = -dn;
        }
        if (modified_refract(ray.d, outward_normal, ni_over_nt, refracted)) reflect_prob = schlick(cosine, param);
        else reflect_prob = 1.f;
        ray.o = point;
        ray.d = (u[0] < reflect_prob) ? reflected : refracted;
        return true;
    }
    return false;
}

// Throughput of a path after its FIRST scatter event on a sphere with these material words: what scatter() / scatter_book() make of
// colour = (1, 1, 1).  The path kernels use it when they take over a sample from the primary-visibility kernels, which hand over
// the scattered ray and the sphere index but not the colour (same operations, so the same bits: 1 * x is x).
VCRT_DEV V3 first_attenuation(float4 surf, int kind, uint32_t semantics) {
    V3 colour = v3(1.f, 1.f, 1.f);
    if (kind == 1 || kind == 2) {
        const V3 tint = vmul(colour, v3(surf.x, surf.y, surf.z));
        colour = (kind == 1 && !semantics) ? vscale(tint, surf.w) : tint;   // bug B4b: the reference's lambertian multiplies by texture.y too
    }
    return colour;
}

// ------------------------------------------------------------------ BOOK semantics (VCRT_SEMANTICS_BOOK)
// The same path with the reference's matched bugs fixed the way "Ray Tracing in One Weekend" (v4.0.1) writes the routines:
// face-forwarded normals, uniform-on-the-sphere scatter by rejection sampling, lambertian attenuation = albedo, metal
// reflects the direction and absorbs below-surface scatter, dielectric with unit directions, Snell and Schlick.
// Draws: Philox block (pixel, sample, event, attempt); see the oracle's BOOK section for the layout.
struct BookDraws { uint32_t pixel, sample, event, seedLo, seedHi; };

VCRT_DEV V3 book_unit_vector(const BookDraws& g, const float first[4]) {
    float u0 = first[0], u1 = first[1], u2 = first[2];
    for (uint32_t attempt = 0;; attempt++) {
        V3 p = v3(fsub(fmul(2.f, u0), 1.f), fsub(fmul(2.f, u1), 1.f), fsub(fmul(2.f, u2), 1.f));
        float lensq = vdot(p, p);
        if (1e-30f < lensq && lensq <= 1.f) return vdiv(p, fsqrt(lensq));
        if (attempt == 63u) return v3(1.f, 0.f, 0.f);
        uint32_t rnd[4];
        philox4x32_10(g.pixel, g.sample, g.event, attempt + 1u, g.seedLo, g.seedHi, rnd);
        u0 = u01(rnd[0]); u1 = u01(rnd[1]); u2 = u01(rnd[2]);
    }
}

// Returns 0 = unknown material (absorbed + counted), 1 = scattered, 2 = absorbed by the material (metal below the surface).
VCRT_DEV int scatter_book(const float4 shape, const float4 surf, int kind, float t, Ray& ray, V3& colour, const float u[4], const BookDraws& g) {
    V3 point = vadd(vscale(ray.d, t), ray.o);
    V3 outward = vdiv(vsub(point, v3(shape.x, shape.y, shape.z)), shape.w);
    const bool front_face = vdot(ray.d, outward) < 0.f;                  // hittable.h set_face_normal
    V3 normal = front_face ? outward : vneg(outward);
    V3 albedo = v3(surf.x, surf.y, surf.z);
    if (kind == 1 || kind == 2) {
        const V3 rv = book_unit_vector(g, u);
        V3 direction;
        if (kind == 1) {                                                 // lambertian::scatter
            direction = vadd(normal, rv);
            const float s = 1e-8f;
            if (fabsf(direction.x) < s && fabsf(direction.y) < s && fabsf(direction.z) < s) direction = normal;
        } else {                                                         // metal::scatter
            const float fuzz = surf.w < 1.f ? surf.w : 1.f;
            V3 reflected = vsub(ray.d, vscale(normal, fmul(2.f, vdot(ray.d, normal))));
            direction = vadd(vunit(reflected), vscale(rv, fuzz));
        }
        colour = vmul(colour, albedo);
        ray.o = point; ray.d = direction;
        return (kind == 1 || vdot(direction, normal) > 0.f) ? 1 : 2;
    }
    if (kind == 3) {                                                     // dielectric::scatter
        const float ri = front_face ? fdiv(1.0f, surf.w) : surf.w;
        V3 ud = vunit(ray.d);
        const float cos_theta = fminf(vdot(vneg(ud), normal), 1.0f);
        const float sin_theta = fsqrt(fsub(1.0f, fmul(cos_theta, cos_theta)));
        float r0 = fdiv(fsub(1.f, ri), fadd(1.f, ri));
        r0 = fmul(r0, r0);
        const float x = fsub(1.f, cos_theta), x2 = fmul(x, x);
        const float reflectance = fadd(r0, fmul(fsub(1.f, r0), fmul(fmul(x2, x2), x)));
        V3 direction;
        if (fmul(ri, sin_theta) > 1.0f || reflectance > u[3]) direction = vsub(ud, vscale(normal, fmul(2.f, vdot(ud, normal))));
        else {
            V3 perp = vscale(vadd(ud, vscale(normal, cos_theta)), ri);
            V3 parallel = vscale(normal, -fsqrt(fabsf(fsub(1.0f, vdot(perp, perp)))));
            direction = vadd(perp, parallel);
        }
        ray.o = point; ray.d = direction;
        return 1;
    }
    return 0;
}

}  // namespace vcrt
