/* @file Platform.cpp

    Headless platform layer: the reference's loop is `while(!quit){ DrawNextFrame(); handleEvent();
    vkDeviceWaitIdle(); }` (platform/Linux.cpp:356-367); with no window the quit condition is a frame count.
    SPDX-License-Identifier: WTFPL
*/
#include "../../include/Platform.hpp"
#include "../../include/Renderer.hpp"

#include <stdio.h>

VkResult PlatformCreateWindow(void) { return VK_SUCCESS; }

VkResult PlatformEnterEventLoop(void)
{
    for (uint32_t frame = 0; frame < rendererConfiguration.frames; frame++) {
        const VkResult result = DrawNextFrame();
        if (result != VK_SUCCESS) return result;   // the reference ignores this on Linux (platform/Linux.cpp:363); we stop
        const VcrtRenderStats& stats = rendererLastStats;
        printf("frame %u: %u spp in image, %.2f ms, %.1f Mrays/s, %.3f T sphere tests/s\n", frame, rendererAccumulatedSamples,
               stats.totalMs, stats.totalMs > 0 ? (double)stats.rays / stats.totalMs * 1e-3 : 0.0,
               stats.totalMs > 0 ? (double)stats.sphereTests / stats.totalMs * 1e-9 : 0.0);
    }
    return VK_SUCCESS;
}
