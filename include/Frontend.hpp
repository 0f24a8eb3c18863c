/* @file Frontend.hpp

    Headless front-end: where the reference owns a swapchain that the fragment stage fills from
    the result image (Frontend.hpp:13-22, shaders/shader.frag:11-14), this one owns a host copy of
    the RGBA32F image and writes it to files.
    SPDX-License-Identifier: WTFPL
*/
#ifndef FRONTEND_HPP
#define FRONTEND_HPP

#include "Common.hpp"

extern float* frontendHostImage;        /* width*height*4 floats, linear radiance, alpha 1 */
extern uint32_t frontendImageWidth, frontendImageHeight;

VkResult CreateVulkanWindowFrontend(void);
VkResult DestroyVulkanWindowFrontend(void);

/* "Present": read the device image back (the texelFetch of shader.frag:14). */
VkResult FrontendAcquireImage(void);
/* Linear float image, bottom-up rows as the PFM format wants. */
VkResult FrontendWritePfm(IN const char* path);
/* 8-bit image through the linear->sRGB transfer the reference's B8G8R8A8_SRGB swapchain applies (Frontend.cpp:43). */
VkResult FrontendWritePpm(IN const char* path);
/* The same 8-bit image as a PNG (stored deflate blocks, no compression library; carries an sRGB chunk). */
VkResult FrontendWritePng(IN const char* path);

#endif
