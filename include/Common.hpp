/* @file Common.hpp

    Shared definitions of the headless host program.  Same role as the reference's
    include/Common.hpp:11-25 (IN/OUT markers, window size, RENDER_ITERATION), minus <vulkan/vulkan.h>:
    there is no Vulkan on this path, so VkResult is a local integer type whose values keep the
    reference's spelling and sign convention (VK_SUCCESS = 0, errors negative).
    SPDX-License-Identifier: WTFPL
*/
#ifndef COMMON_HPP
#define COMMON_HPP

#include "vcrt.h"   /* defines IN / OUT */

#include <cstddef>
#include <cstdint>

typedef int32_t VkResult;
enum : int32_t {
    VK_SUCCESS = 0,
    VK_NOT_READY = 1,
    VK_ERROR_OUT_OF_HOST_MEMORY = -1,
    VK_ERROR_OUT_OF_DEVICE_MEMORY = -2,
    VK_ERROR_INITIALIZATION_FAILED = -3,      /* reference use: platform/Win32.cpp:55 */
    VK_ERROR_DEVICE_LOST = -4,
    VK_ERROR_FEATURE_NOT_PRESENT = -8,        /* reference use: platform/Linux.cpp:197 */
    VK_ERROR_UNKNOWN = -13,                   /* reference use: Renderer.cpp:430 */
    VK_ERROR_OUT_OF_DATE_KHR = -1000001004    /* reference use: platform/Win32.cpp:100 */
};

/* Defaults = the reference's compile-time constants (Common.hpp:23-25, globals.glsl:12-17). */
constexpr uint32_t WINDOW_WIDTH = 1280;
constexpr uint32_t WINDOW_HEIGHT = 720;
constexpr uint32_t RENDER_ITERATION = 100;   /* dead in the reference; here: frames of progressive accumulation */

/* Run-time stand-in for the constants the reference bakes into host AND shader by hand. */
typedef struct RendererConfiguration {
    uint32_t width, height;        /* WINDOW_WIDTH x WINDOW_HEIGHT == IMAGE_WIDTH x IMAGE_HEIGHT */
    uint32_t samplesPerFrame;      /* SAMPLES_PER_PIXEL of one DrawNextFrame */
    uint32_t maxDepth;             /* MAX_RECURSION_LEVEL */
    uint32_t frames;               /* event-loop iterations of the headless platform */
    uint32_t progressive;          /* 0: every frame re-renders from scratch (what the reference does); 1: frames accumulate */
    uint32_t traversal;            /* VCRT_TRAVERSAL_* */
    uint32_t semantics;            /* VCRT_SEMANTICS_*: the reference shader as it is, or the BOOK-corrected path */
    uint32_t sceneGridHalf;        /* 11 = the reference scene */
    uint32_t sceneVariant;         /* VCRT_SCENE_* */
    uint64_t seed;
    int32_t device;
    const char* sceneFile;         /* optional GLSL-text scene (globals.glsl:31-517 form) instead of the generator */
    const char* outputPrefix;      /* "<prefix>.pfm" (linear), "<prefix>.ppm" and "<prefix>.png" (sRGB-encoded) are written at the end; NULL = none */
} RendererConfiguration;

extern RendererConfiguration rendererConfiguration;
extern const char* applicationNameNarrow;

#endif
