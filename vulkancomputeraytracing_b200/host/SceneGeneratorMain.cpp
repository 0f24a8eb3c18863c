/* @file SceneGeneratorMain.cpp

    Command-line form of the scene producer: prints the GLSL initialiser lines the reference's tool
    prints (SceneGenerator.cpp:23-56), so its stdout can be diffed against globals.glsl:31-515.
    SPDX-License-Identifier: WTFPL
*/
#include "../../include/SceneGenerator.hpp"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

int main(int argc, char** argv)
{
    VcrtSceneRecipe recipe;
    vcrtDefaultSceneRecipe(&recipe);
    bool withGround = false;
    for (int i = 1; i < argc; i++) {
        if (!strcmp(argv[i], "--grid") && i + 1 < argc) recipe.gridHalf = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--stress")) { recipe.variant = VCRT_SCENE_STRESS; recipe.lambertBelow = 0.0; recipe.metalBelow = 0.5; }
        else if (!strcmp(argv[i], "--ground")) withGround = true;
        else { fprintf(stderr, "usage: SceneGenerator [--grid G] [--stress] [--ground]\n"); return -1; }
    }
    const uint32_t capacity = vcrtSceneCapacity(&recipe);
    VcrtSphere* spheres = new VcrtSphere[capacity ? capacity : 1];
    const int32_t count = vcrtGenerateScene(&recipe, spheres, capacity);
    if (count < 4) return -1;
    // the reference tool prints the grid spheres, a blank line, then the three radius-1 spheres; the ground is not its output
    const uint32_t grid = (uint32_t)count - 4u;
    const char* big[3] = {
        "sphere(vec3(0, 1, 0),1.0, vec3(1.0,1.0,1.0), vec3(TEXTURE_GLASS,1.5,0.0)),\n",
        "sphere(vec3(-4, 1, 0),1.0, vec3(0.4, 0.2, 0.1), vec3(TEXTURE_LAMBERTIAN,1.0,0.0)),\n",
        "sphere(vec3(4, 1, 0),1.0, vec3(0.7, 0.6, 0.5), vec3(TEXTURE_METAL,1.0,0.0)),\n",
    };
    const size_t bytes = vcrtSceneToGlsl(spheres, grid, nullptr, 0);
    char* text = new char[bytes + 1];
    vcrtSceneToGlsl(spheres, grid, text, bytes + 1);
    fputs(text, stdout);
    printf("\n");
    if (recipe.variant == VCRT_SCENE_REFERENCE) for (int i = 0; i < 3; i++) fputs(big[i], stdout);
    else { char line[1024]; vcrtSceneToGlsl(spheres + grid, 3, line, sizeof line); fputs(line, stdout); }
    if (withGround) { char line[512]; vcrtSceneToGlsl(spheres + grid + 3, 1, line, sizeof line); fputs(line, stdout); }
    delete[] text;
    delete[] spheres;
    return 0;
}
