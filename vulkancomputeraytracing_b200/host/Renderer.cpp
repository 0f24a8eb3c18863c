/* @file Renderer.cpp

    Renderer of the headless program.  The reference's Renderer.cpp builds two Vulkan pipelines and
    submits a compute dispatch per frame; all of that is replaced by three calls into the C ABI
    (include/vcrt.h): cudaUploadScene in Begin, cudaRender in DrawNextFrame, cudaDestroyRenderer in
    End.  Entry-point names, order and the VkResult convention are the reference's
    (include/Renderer.hpp:14-20).
    SPDX-License-Identifier: WTFPL
*/
#include "../../include/Renderer.hpp"
#include "../../include/SceneGenerator.hpp"

#include <stdio.h>
#include <string.h>

uint32_t rendererFrameIndex = 0;
uint32_t rendererAccumulatedSamples = 0;
VcrtRenderStats rendererLastStats;

static VcrtSphere* rendererSceneSpheres = nullptr;   // host AoS, owned here like the reference owns its handles (Renderer.cpp:13-33)
static uint32_t rendererSceneCount = 0;
static bool rendererSceneUploaded = false;

static VkResult translateResult(IN int code)
{
    switch (code) {
    case VCRT_SUCCESS: return VK_SUCCESS;
    case VCRT_ERROR_NO_DEVICE: return VK_ERROR_INITIALIZATION_FAILED;
    case VCRT_ERROR_OUT_OF_MEMORY: return VK_ERROR_OUT_OF_DEVICE_MEMORY;
    case VCRT_ERROR_UNSUPPORTED: return VK_ERROR_FEATURE_NOT_PRESENT;
    case VCRT_ERROR_CUDA: case VCRT_ERROR_NCCL: return VK_ERROR_DEVICE_LOST;
    default: return VK_ERROR_UNKNOWN;
    }
}

static VkResult loadSceneFromFile(IN const char* path)
{
    FILE* file = fopen(path, "rb");
    if (!file) return VK_ERROR_INITIALIZATION_FAILED;
    fseek(file, 0, SEEK_END);
    const long size = ftell(file);
    fseek(file, 0, SEEK_SET);
    char* text = new char[(size_t)size + 1];
    const size_t got = fread(text, 1, (size_t)size, file);
    fclose(file);
    text[got] = 0;
    uint32_t capacity = 1;
    for (const char* c = text; *c; c++) capacity += *c == '\n';
    rendererSceneSpheres = new VcrtSphere[capacity];
    const int32_t count = vcrtSceneFromGlsl(text, rendererSceneSpheres, capacity);
    delete[] text;
    if (count <= 0) return VK_ERROR_INITIALIZATION_FAILED;
    rendererSceneCount = (uint32_t)count;
    return VK_SUCCESS;
}

VkResult BeginRenderingOperation(void)
{
    const RendererConfiguration& config = rendererConfiguration;
    VkResult result = translateResult(cudaSelectDevice(config.device));
    if (result != VK_SUCCESS) return result;

    // The scene: where the reference compiles `const sphere world[]` into the shader (globals.glsl:29-518).
    if (config.sceneFile) {
        result = loadSceneFromFile(config.sceneFile);
        if (result != VK_SUCCESS) return result;
    } else {
        VcrtSceneRecipe recipe;
        vcrtDefaultSceneRecipe(&recipe);
        recipe.gridHalf = (int32_t)config.sceneGridHalf;
        recipe.variant = config.sceneVariant;
        if (config.sceneVariant == VCRT_SCENE_STRESS) { recipe.lambertBelow = 0.0; recipe.metalBelow = 0.5; }
        const uint32_t capacity = vcrtSceneCapacity(&recipe);
        rendererSceneSpheres = new VcrtSphere[capacity ? capacity : 1];
        const int32_t count = vcrtGenerateScene(&recipe, rendererSceneSpheres, capacity);
        if (count <= 0) return VK_ERROR_INITIALIZATION_FAILED;
        rendererSceneCount = (uint32_t)count;
    }
    VcrtCamera camera;
    vcrtDefaultCamera(&camera);   // globals.glsl:21-24
    result = translateResult(cudaUploadScene(rendererSceneSpheres, rendererSceneCount, &camera));
    if (result != VK_SUCCESS) return result;
    rendererSceneUploaded = true;
    rendererFrameIndex = 0;
    rendererAccumulatedSamples = 0;
    memset(&rendererLastStats, 0, sizeof rendererLastStats);
    return VK_SUCCESS;
}

VkResult DrawNextFrame(void)
{
    if (!rendererSceneUploaded) return VK_NOT_READY;
    const RendererConfiguration& config = rendererConfiguration;
    VcrtRenderParams params;
    vcrtDefaultRenderParams(&params);
    params.width = config.width;
    params.height = config.height;
    params.samplesPerPixel = config.samplesPerFrame;
    params.maxDepth = config.maxDepth;
    params.seed = config.seed;
    params.traversal = config.traversal;
    params.semantics = config.semantics;
    if (config.progressive) {
        // frame k adds sample indices [k*spp, (k+1)*spp) to the running sums; the image is their mean so far
        params.sampleOffset = rendererFrameIndex * config.samplesPerFrame;
        params.flags |= VCRT_FLAG_ACCUMULATE;
    }
    const VkResult result = translateResult(cudaRender(&params));   // synchronous, like the device wait after each frame (platform/Linux.cpp:365)
    if (result != VK_SUCCESS) return result;
    cudaGetRenderStats(&rendererLastStats);
    rendererAccumulatedSamples = config.progressive ? rendererAccumulatedSamples + config.samplesPerFrame : config.samplesPerFrame;
    rendererFrameIndex++;
    return VK_SUCCESS;
}

VkResult EndRenderingOperation(void)
{
    const int code = cudaDestroyRenderer();   // tolerates never-created state
    if (rendererSceneSpheres) delete[] rendererSceneSpheres;
    rendererSceneSpheres = nullptr;
    rendererSceneCount = 0;
    rendererSceneUploaded = false;
    return translateResult(code);
}
