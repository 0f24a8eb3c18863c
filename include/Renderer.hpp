/* @file Renderer.hpp

    The render entry of the program: the three calls main() and the platform loop make
    (reference include/Renderer.hpp:14-20, same names, arity and return convention).
    SPDX-License-Identifier: WTFPL
*/
#ifndef RENDERER_HPP
#define RENDERER_HPP

#include "Common.hpp"

/* Upload the scene and camera to the device (replaces pipeline + result-image creation, Renderer.cpp:228-652). */
VkResult BeginRenderingOperation(void);

/* Path-trace one frame into the device RGBA32F image (replaces Renderer.cpp:654-711); called by the platform loop. */
VkResult DrawNextFrame(void);

/* Release everything Begin created (replaces Renderer.cpp:714-774); safe when Begin never ran. */
VkResult EndRenderingOperation(void);

/* Frames drawn since Begin, samples per pixel in the current image and the counters of the last frame. */
extern uint32_t rendererFrameIndex;
extern uint32_t rendererAccumulatedSamples;
extern VcrtRenderStats rendererLastStats;

#endif
