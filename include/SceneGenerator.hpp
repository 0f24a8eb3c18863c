/* @file SceneGenerator.hpp

    Run-time form of the reference's offline scene tool.  The reference prints GLSL text that is
    pasted into the shader by hand (SceneGenerator.cpp:23-56 -> shaders/include/globals.glsl:30);
    here the same draw sequence fills VcrtSphere records that go straight to cudaUploadScene.
    SPDX-License-Identifier: WTFPL
*/
#ifndef SCENE_GENERATOR_HPP
#define SCENE_GENERATOR_HPP

#include "vcrt.h"

#ifdef __cplusplus
extern "C" {
#endif

enum {
    VCRT_SCENE_REFERENCE = 0,   /* SceneGenerator.cpp as written + the hand-added ground sphere (globals.glsl:517) */
    VCRT_SCENE_STRESS = 1       /* divergence stress: no lambertian at all (big lambertian sphere -> metal, ground -> metal fuzz 0.1) */
};

/* gridHalf 11, lambertBelow 0.8, metalBelow 0.95, VCRT_SCENE_REFERENCE reproduce the table baked
   into globals.glsl:31-517 (484 generated spheres, then the ground LAST). */
typedef struct VcrtSceneRecipe {
    int32_t gridHalf;        /* cells span [-gridHalf, gridHalf) in x and z (SceneGenerator.cpp:24-25) */
    uint32_t variant;        /* VCRT_SCENE_* */
    double lambertBelow;     /* choose_mat < lambertBelow -> lambertian (SceneGenerator.cpp:33) */
    double metalBelow;       /* choose_mat < metalBelow   -> metal      (SceneGenerator.cpp:38) */
} VcrtSceneRecipe;

void vcrtDefaultSceneRecipe(OUT VcrtSceneRecipe* recipe);

/* Upper bound of the sphere count a recipe can produce (cells + 3 big spheres + ground). */
uint32_t vcrtSceneCapacity(IN const VcrtSceneRecipe* recipe);

/* Fills `spheres` (capacity entries) and returns the number written, or a negative VCRT_ERROR_*. */
int32_t vcrtGenerateScene(IN const VcrtSceneRecipe* recipe, OUT VcrtSphere* spheres, IN uint32_t capacity);

/* Writes the scene as the GLSL initialiser text the reference tool prints (one sphere(...) per line,
   the generated part byte-identical to SceneGenerator.cpp's stdout).  Returns bytes written
   (excluding the terminating 0), or the size needed when `text` is NULL / too small. */
size_t vcrtSceneToGlsl(IN const VcrtSphere* spheres, IN uint32_t count, OUT char* text, IN size_t capacity);

/* Parses that text form (globals.glsl:31-517 style lines) back into records; returns the count or a negative code. */
int32_t vcrtSceneFromGlsl(IN const char* text, OUT VcrtSphere* spheres, IN uint32_t capacity);

#ifdef __cplusplus
}
#endif
#endif /* SCENE_GENERATOR_HPP */
