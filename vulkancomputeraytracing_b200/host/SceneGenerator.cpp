/* @file SceneGenerator.cpp

    Scene producer of the render path as a library call (reference: the stand-alone tool
    SceneGenerator.cpp:23-56, whose stdout is pasted into globals.glsl by hand).
    SPDX-License-Identifier: WTFPL

    What has to be kept for the output to equal the baked table:
      * the generator is a default-constructed std::mt19937 feeding one
        std::uniform_real_distribution<double>(0,1) (SceneGenerator.cpp:17-21);
      * the reference writes several random_double() calls as ARGUMENTS of one call
        (point3(...) and printf(...)); their order is unspecified in C++ and g++/x86-64 runs them
        last argument first.  That order is spelled out below, so any compiler gives the table;
      * every value goes through "%.2f" text and back through a binary32 GLSL literal.
*/
#include "../../include/SceneGenerator.hpp"

#include <math.h>
#include <random>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static std::mt19937 sceneRandomEngine;
static std::uniform_real_distribution<double> sceneRandomDistribution(0.0, 1.0);

static double NextRandomDouble(void) { return sceneRandomDistribution(sceneRandomEngine); }

// "%.2f" and back: what printing + pasting into GLSL does to a value.
static float ThroughTwoDecimals(IN double value)
{
    char text[64];
    snprintf(text, sizeof text, "%.2f", value);
    return strtof(text, nullptr);
}

static void StoreSphere(OUT VcrtSphere* sphere, float x, float y, float z, float radius, float r, float g, float b,
                        float kind, float parameter)
{
    *sphere = VcrtSphere{ .center = { x, y, z }, .radius = radius, .colour = { r, g, b }, .texture = { kind, parameter, 0.0f } };
}

extern "C" void vcrtDefaultSceneRecipe(OUT VcrtSceneRecipe* recipe)
{
    if (!recipe) return;
    *recipe = VcrtSceneRecipe{ .gridHalf = 11, .variant = VCRT_SCENE_REFERENCE, .lambertBelow = 0.8, .metalBelow = 0.95 };
}

extern "C" uint32_t vcrtSceneCapacity(IN const VcrtSceneRecipe* recipe)
{
    if (!recipe || recipe->gridHalf < 0 || recipe->gridHalf > 4096) return 0;
    const uint32_t side = 2u * (uint32_t)recipe->gridHalf;
    return side * side + 4u;
}

extern "C" int32_t vcrtGenerateScene(IN const VcrtSceneRecipe* recipe, OUT VcrtSphere* spheres, IN uint32_t capacity)
{
    if (!recipe || !spheres || recipe->gridHalf < 0 || recipe->gridHalf > 4096 || recipe->variant > VCRT_SCENE_STRESS)
        return VCRT_ERROR_INVALID_ARGUMENT;
    if (capacity < vcrtSceneCapacity(recipe)) return VCRT_ERROR_INVALID_ARGUMENT;
    sceneRandomEngine.seed(std::mt19937::default_seed);   // every call restarts the sequence: same recipe, same scene
    sceneRandomDistribution.reset();
    const double lambertBelow = recipe->lambertBelow, metalBelow = recipe->metalBelow;
    uint32_t count = 0;
    for (int a = -recipe->gridHalf; a < recipe->gridHalf; a++) {
        for (int b = -recipe->gridHalf; b < recipe->gridHalf; b++) {
            const double chooseMaterial = NextRandomDouble();
            const double jitterZ = NextRandomDouble();   // third constructor argument is evaluated first
            const double jitterX = NextRandomDouble();
            const float x = (float)(a + 0.9 * jitterX), y = (float)0.2, z = (float)(b + 0.9 * jitterZ);
            const float dx = x - 4.0f, dy = y - 0.2f, dz = z - 0.0f;
            if (dx * dx + dy * dy + dz * dz < 0.9) continue;   // the reference's length() is the SQUARED length
            VcrtSphere* out = &spheres[count];
            if (chooseMaterial < lambertBelow || chooseMaterial < metalBelow) {
                const double fourth = NextRandomDouble(), third = NextRandomDouble(), second = NextRandomDouble(),
                             first = NextRandomDouble();   // printf arguments, last one first
                StoreSphere(out, ThroughTwoDecimals(x), ThroughTwoDecimals(y), ThroughTwoDecimals(z), 0.2f,
                            ThroughTwoDecimals(first), ThroughTwoDecimals(second), ThroughTwoDecimals(third),
                            chooseMaterial < lambertBelow ? 1.0f : 2.0f, ThroughTwoDecimals(fourth));
            } else {
                StoreSphere(out, ThroughTwoDecimals(x), ThroughTwoDecimals(y), ThroughTwoDecimals(z), 0.2f, 1.0f, 1.0f, 1.0f, 3.0f, 1.5f);
            }
            count++;
        }
    }
    const bool stress = recipe->variant == VCRT_SCENE_STRESS;
    StoreSphere(&spheres[count++], 0, 1, 0, 1.0f, 1.0f, 1.0f, 1.0f, 3.0f, 1.5f);
    if (stress) StoreSphere(&spheres[count++], -4, 1, 0, 1.0f, 0.4f, 0.2f, 0.1f, 2.0f, 0.0f);
    else StoreSphere(&spheres[count++], -4, 1, 0, 1.0f, 0.4f, 0.2f, 0.1f, 1.0f, 1.0f);
    StoreSphere(&spheres[count++], 4, 1, 0, 1.0f, 0.7f, 0.6f, 0.5f, 2.0f, 1.0f);
    // globals.glsl:517: the ground is hand-added after the generated lines, so it is LAST
    if (stress) StoreSphere(&spheres[count++], 0, -1000, 0, 1000.0f, 0.5f, 0.5f, 0.5f, 2.0f, 0.1f);
    else StoreSphere(&spheres[count++], 0, -1000, 0, 1000.0f, 0.5f, 0.5f, 0.5f, 1.0f, 1.0f);
    return (int32_t)count;
}

static const char* materialNames[4] = { "0", "TEXTURE_LAMBERTIAN", "TEXTURE_METAL", "TEXTURE_GLASS" };

extern "C" size_t vcrtSceneToGlsl(IN const VcrtSphere* spheres, IN uint32_t count, OUT char* text, IN size_t capacity)
{
    if (!spheres && count) return 0;
    size_t used = 0;
    char line[256];
    for (uint32_t i = 0; i < count; i++) {
        const VcrtSphere& s = spheres[i];
        const int kind = (int)s.texture[0];
        int n;
        if (s.radius == 0.2f && kind == 3)   // the three shapes of line the reference tool prints
            n = snprintf(line, sizeof line, "sphere(vec3(%.2f,%.2f,%.2f), 0.2, vec3(1.0,1.0,1.0), vec3(TEXTURE_GLASS,1.5,0.0)),\n",
                         s.center[0], s.center[1], s.center[2]);
        else if (s.radius == 0.2f && (kind == 1 || kind == 2))
            n = snprintf(line, sizeof line, "sphere(vec3(%.2f,%.2f,%.2f), 0.2, vec3(%.2f,%.2f,%.2f), vec3(%s,%.2f,0.0)),\n",
                         s.center[0], s.center[1], s.center[2], s.colour[0], s.colour[1], s.colour[2], materialNames[kind], s.texture[1]);
        else
            n = snprintf(line, sizeof line, "sphere(vec3(%.9g,%.9g,%.9g), %.9g, vec3(%.9g,%.9g,%.9g), vec3(%s,%.9g,0.0)),\n",
                         s.center[0], s.center[1], s.center[2], s.radius, s.colour[0], s.colour[1], s.colour[2],
                         (kind >= 1 && kind <= 3) ? materialNames[kind] : materialNames[0], s.texture[1]);
        if (text && used + (size_t)n < capacity) memcpy(text + used, line, (size_t)n);
        used += (size_t)n;
    }
    if (text && used < capacity) text[used] = 0;
    return used;
}

// Accepts `sphere(vec3(x,y,z), r, vec3(r,g,b), vec3(KIND,p,q))` with optional trailing comma / spaces; skips blank lines
// and // comments.  KIND is one of the three TEXTURE_* names (textures.glsl:10-12) or a number.
extern "C" int32_t vcrtSceneFromGlsl(IN const char* text, OUT VcrtSphere* spheres, IN uint32_t capacity)
{
    if (!text || (!spheres && capacity)) return VCRT_ERROR_INVALID_ARGUMENT;
    uint32_t count = 0;
    const char* cursor = text;
    while ((cursor = strstr(cursor, "sphere(")) != nullptr) {
        const char* lineStart = cursor;
        while (lineStart > text && lineStart[-1] != '\n') lineStart--;
        const char* comment = strstr(lineStart, "//");
        if (comment && comment < cursor) { cursor += 7; continue; }
        float v[9]; char kindText[32]; float q;
        int fields = sscanf(cursor, "sphere( vec3( %f , %f , %f ) , %f , vec3( %f , %f , %f ) , vec3( %31[A-Za-z0-9_.] , %f , %f ) )",
                            &v[0], &v[1], &v[2], &v[3], &v[4], &v[5], &v[6], kindText, &v[7], &q);
        if (fields != 10) return VCRT_ERROR_SCENE;
        float kind;
        if (!strcmp(kindText, "TEXTURE_LAMBERTIAN")) kind = 1.0f;
        else if (!strcmp(kindText, "TEXTURE_METAL")) kind = 2.0f;
        else if (!strcmp(kindText, "TEXTURE_GLASS")) kind = 3.0f;
        else kind = strtof(kindText, nullptr);
        if (count >= capacity) return VCRT_ERROR_INVALID_ARGUMENT;
        StoreSphere(&spheres[count], v[0], v[1], v[2], v[3], v[4], v[5], v[6], kind, v[7]);
        spheres[count].texture[2] = q;
        count++;
        cursor += 7;
    }
    return (int32_t)count;
}
