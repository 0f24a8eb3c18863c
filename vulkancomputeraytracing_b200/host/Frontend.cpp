/* @file Frontend.cpp

    Headless front-end (reference Frontend.cpp:81-177 creates a swapchain; shader.frag:11-14 copies
    the result image into it).  Here "presenting" is a device->host copy and, on request, two files.
    SPDX-License-Identifier: WTFPL
*/
#include "../../include/Frontend.hpp"

#include <math.h>
#include <stdio.h>

float* frontendHostImage = nullptr;
uint32_t frontendImageWidth = 0, frontendImageHeight = 0;

VkResult CreateVulkanWindowFrontend(void)
{
    frontendImageWidth = rendererConfiguration.width;
    frontendImageHeight = rendererConfiguration.height;
    if (!frontendImageWidth || !frontendImageHeight) return VK_ERROR_INITIALIZATION_FAILED;
    frontendHostImage = new float[(size_t)frontendImageWidth * frontendImageHeight * 4];
    return VK_SUCCESS;
}

VkResult DestroyVulkanWindowFrontend(void)
{
    if (frontendHostImage) delete[] frontendHostImage;
    frontendHostImage = nullptr;
    frontendImageWidth = frontendImageHeight = 0;
    return VK_SUCCESS;
}

VkResult FrontendAcquireImage(void)
{
    if (!frontendHostImage) return VK_NOT_READY;
    const size_t bytes = (size_t)frontendImageWidth * frontendImageHeight * 4 * sizeof(float);
    return cudaDownloadImage(frontendHostImage, bytes) == VCRT_SUCCESS ? VK_SUCCESS : VK_ERROR_UNKNOWN;
}

VkResult FrontendWritePfm(IN const char* path)
{
    if (!frontendHostImage || !path) return VK_NOT_READY;
    FILE* file = fopen(path, "wb");
    if (!file) return VK_ERROR_UNKNOWN;
    fprintf(file, "PF\n%u %u\n-1.0\n", frontendImageWidth, frontendImageHeight);   // negative scale = little endian
    float* row = new float[(size_t)frontendImageWidth * 3];
    for (uint32_t y = frontendImageHeight; y-- > 0;) {
        const float* source = frontendHostImage + (size_t)y * frontendImageWidth * 4;
        for (uint32_t x = 0; x < frontendImageWidth; x++)
            for (int c = 0; c < 3; c++) row[3 * x + c] = source[4 * x + c];
        fwrite(row, sizeof(float), (size_t)frontendImageWidth * 3, file);
    }
    delete[] row;
    fclose(file);
    return VK_SUCCESS;
}

static unsigned char encodeSrgb(IN float linear)
{
    if (!(linear > 0.0f)) return 0;   // also catches NaN
    if (linear >= 1.0f) return 255;
    const float encoded = linear <= 0.0031308f ? 12.92f * linear : 1.055f * powf(linear, 1.0f / 2.4f) - 0.055f;
    return (unsigned char)lrintf(encoded * 255.0f);
}

VkResult FrontendWritePpm(IN const char* path)
{
    if (!frontendHostImage || !path) return VK_NOT_READY;
    FILE* file = fopen(path, "wb");
    if (!file) return VK_ERROR_UNKNOWN;
    fprintf(file, "P6\n%u %u\n255\n", frontendImageWidth, frontendImageHeight);
    unsigned char* row = new unsigned char[(size_t)frontendImageWidth * 3];
    for (uint32_t y = 0; y < frontendImageHeight; y++) {
        const float* source = frontendHostImage + (size_t)y * frontendImageWidth * 4;
        for (uint32_t x = 0; x < frontendImageWidth; x++)
            for (int c = 0; c < 3; c++) {
                const float value = source[4 * x + c];
                // BOOK images are already gamma-2 encoded by the layer (the book's linear_to_gamma): quantise only
                row[3 * x + c] = rendererConfiguration.semantics == VCRT_SEMANTICS_BOOK
                                     ? (unsigned char)(255.999f * (value > 0.0f ? (value < 0.999f ? value : 0.999f) : 0.0f))
                                     : encodeSrgb(value);
            }
        fwrite(row, 1, (size_t)frontendImageWidth * 3, file);
    }
    delete[] row;
    fclose(file);
    return VK_SUCCESS;
}
