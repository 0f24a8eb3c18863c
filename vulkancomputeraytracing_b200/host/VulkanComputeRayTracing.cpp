/* @file VulkanComputeRayTracing.cpp

    Entry point of the headless program: front-end -> begin -> loop -> end, the same order the
    reference's main walks (VulkanComputeRayTracing.cpp:17-42), with the Vulkan environment steps gone.
    SPDX-License-Identifier: WTFPL

    usage: VulkanComputeRayTracing [--width W] [--height H] [--spp N] [--depth D] [--frames F] [--progressive]
                                   [--seed S] [--device I] [--traversal auto|brute|lbvh|simple]
                                   [--grid G] [--stress] [--book] [--scene file.glsl] [--out prefix]
*/
#include "../../include/Frontend.hpp"
#include "../../include/Platform.hpp"
#include "../../include/Renderer.hpp"
#include "../../include/SceneGenerator.hpp"

#include <iostream>
#include <stdlib.h>
#include <string.h>
#include <string>

using std::cerr;
using std::cout;
using std::endl;

const char* applicationNameNarrow = "Vulkan Compute Raytracing";

RendererConfiguration rendererConfiguration = {
    .width = WINDOW_WIDTH, .height = WINDOW_HEIGHT, .samplesPerFrame = 1, .maxDepth = 50, .frames = 1, .progressive = 0,
    .traversal = VCRT_TRAVERSAL_AUTO, .semantics = VCRT_SEMANTICS_REFERENCE, .sceneGridHalf = 11, .sceneVariant = VCRT_SCENE_REFERENCE, .seed = 0, .device = 0,
    .sceneFile = nullptr, .outputPrefix = nullptr,
};

static int parseArguments(int argc, char** argv)
{
    RendererConfiguration& config = rendererConfiguration;
    for (int i = 1; i < argc; i++) {
        const char* flag = argv[i];
        const char* value = i + 1 < argc ? argv[i + 1] : nullptr;
        auto number = [&](uint32_t& target) { if (!value) return false; target = (uint32_t)strtoul(value, nullptr, 10); i++; return true; };
        bool ok = true;
        if (!strcmp(flag, "--width")) ok = number(config.width);
        else if (!strcmp(flag, "--height")) ok = number(config.height);
        else if (!strcmp(flag, "--spp")) ok = number(config.samplesPerFrame);
        else if (!strcmp(flag, "--depth")) ok = number(config.maxDepth);
        else if (!strcmp(flag, "--frames")) ok = number(config.frames);
        else if (!strcmp(flag, "--grid")) ok = number(config.sceneGridHalf);
        else if (!strcmp(flag, "--progressive")) config.progressive = 1;
        else if (!strcmp(flag, "--stress")) config.sceneVariant = VCRT_SCENE_STRESS;
        else if (!strcmp(flag, "--book")) config.semantics = VCRT_SEMANTICS_BOOK;
        else if (!strcmp(flag, "--seed") && value) { config.seed = strtoull(value, nullptr, 10); i++; }
        else if (!strcmp(flag, "--device") && value) { config.device = atoi(value); i++; }
        else if (!strcmp(flag, "--scene") && value) { config.sceneFile = value; i++; }
        else if (!strcmp(flag, "--out") && value) { config.outputPrefix = value; i++; }
        else if (!strcmp(flag, "--traversal") && value) {
            if (!strcmp(value, "auto")) config.traversal = VCRT_TRAVERSAL_AUTO;
            else if (!strcmp(value, "brute")) config.traversal = VCRT_TRAVERSAL_BRUTE;
            else if (!strcmp(value, "lbvh")) config.traversal = VCRT_TRAVERSAL_LBVH;
            else if (!strcmp(value, "simple")) config.traversal = VCRT_TRAVERSAL_SIMPLE;
            else ok = false;
            i++;
        } else ok = false;
        if (!ok) { cerr << "Bad argument: " << flag << endl; return -1; }
    }
    return 0;
}

int main(int argc, char** argv)
{
    if (parseArguments(argc, argv)) return -1;
    cout << "Hello CUDA." << endl;
    if (CreateVulkanWindowFrontend() != VK_SUCCESS) {
        cerr << "Cannot create headless frontend." << endl;
        return -1;
    }
    if (BeginRenderingOperation() != VK_SUCCESS) {
        cerr << "Cannot begin rendering operation: " << cudaGetLastErrorString() << endl;
        return -1;
    }
    if (PlatformEnterEventLoop() != VK_SUCCESS) {
        cerr << "Rendering failed: " << cudaGetLastErrorString() << endl;
        EndRenderingOperation();
        return -1;
    }
    if (rendererConfiguration.outputPrefix) {
        const std::string prefix = rendererConfiguration.outputPrefix;
        if (FrontendAcquireImage() != VK_SUCCESS || FrontendWritePfm((prefix + ".pfm").c_str()) != VK_SUCCESS ||
            FrontendWritePpm((prefix + ".ppm").c_str()) != VK_SUCCESS || FrontendWritePng((prefix + ".png").c_str()) != VK_SUCCESS) {
            cerr << "Cannot write the image." << endl;
            EndRenderingOperation();
            return -1;
        }
    }
    EndRenderingOperation();
    DestroyVulkanWindowFrontend();
    cout << "Bye CUDA." << endl;
    return 0;
}
