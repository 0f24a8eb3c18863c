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

/* One 8-bit RGB row of the image as the swapchain would hold it. */
static void encodeRow(IN uint32_t y, OUT unsigned char* row)
{
    const float* source = frontendHostImage + (size_t)y * frontendImageWidth * 4;
    for (uint32_t x = 0; x < frontendImageWidth; x++)
        for (int c = 0; c < 3; c++) {
            const float value = source[4 * x + c];
            // BOOK images are already gamma-2 encoded by the layer (the book's linear_to_gamma): quantise only
            row[3 * x + c] = rendererConfiguration.semantics == VCRT_SEMANTICS_BOOK
                                 ? (unsigned char)(255.999f * (value > 0.0f ? (value < 0.999f ? value : 0.999f) : 0.0f))
                                 : encodeSrgb(value);
        }
}

/* PNG without a compression library: the zlib stream is made of stored (uncompressed) deflate blocks. */
static uint32_t pngCrcTable[256];
static bool pngCrcTableReady = false;
static uint32_t pngCrc(IN uint32_t crc, IN const unsigned char* data, IN size_t bytes)
{
    if (!pngCrcTableReady) {
        for (uint32_t n = 0; n < 256; n++) {
            uint32_t c = n;
            for (int k = 0; k < 8; k++) c = (c & 1u) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            pngCrcTable[n] = c;
        }
        pngCrcTableReady = true;
    }
    for (size_t i = 0; i < bytes; i++) crc = pngCrcTable[(crc ^ data[i]) & 0xffu] ^ (crc >> 8);
    return crc;
}
static void pngBigEndian(OUT unsigned char* out, IN uint32_t value)
{
    out[0] = (unsigned char)(value >> 24); out[1] = (unsigned char)(value >> 16); out[2] = (unsigned char)(value >> 8); out[3] = (unsigned char)value;
}
static bool pngChunk(IN FILE* file, IN const char* type, IN const unsigned char* data, IN uint32_t bytes)
{
    unsigned char word[4];
    pngBigEndian(word, bytes);
    if (fwrite(word, 1, 4, file) != 4 || fwrite(type, 1, 4, file) != 4) return false;
    if (bytes && fwrite(data, 1, bytes, file) != bytes) return false;
    uint32_t crc = pngCrc(0xffffffffu, (const unsigned char*)type, 4);
    crc = pngCrc(crc, data, bytes) ^ 0xffffffffu;
    pngBigEndian(word, crc);
    return fwrite(word, 1, 4, file) == 4;
}

VkResult FrontendWritePng(IN const char* path)
{
    if (!frontendHostImage || !path) return VK_NOT_READY;
    const size_t rowBytes = 1 + (size_t)frontendImageWidth * 3;            // filter byte 0 + RGB
    if (rowBytes > 65535) return VK_ERROR_FEATURE_NOT_PRESENT;             // one stored block per row
    FILE* file = fopen(path, "wb");
    if (!file) return VK_ERROR_UNKNOWN;
    static const unsigned char signature[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    unsigned char header[13];
    pngBigEndian(header, frontendImageWidth); pngBigEndian(header + 4, frontendImageHeight);
    header[8] = 8; header[9] = 2; header[10] = 0; header[11] = 0; header[12] = 0;   // 8 bits, RGB, deflate, no filter, no interlace
    const unsigned char srgbIntent = 0;                                     // the samples are sRGB-encoded (B8G8R8A8_SRGB swapchain)
    const size_t streamBytes = 2 + (size_t)frontendImageHeight * (5 + rowBytes) + 4;
    unsigned char* stream = new unsigned char[streamBytes];
    size_t at = 0;
    stream[at++] = 0x78; stream[at++] = 0x01;                               // zlib header: deflate, 32 K window, no preset dictionary
    uint32_t adlerA = 1, adlerB = 0;
    for (uint32_t y = 0; y < frontendImageHeight; y++) {
        stream[at++] = y + 1 == frontendImageHeight ? 1 : 0;                // BFINAL, BTYPE = 00 (stored)
        stream[at++] = (unsigned char)(rowBytes & 0xff); stream[at++] = (unsigned char)(rowBytes >> 8);
        stream[at++] = (unsigned char)(~rowBytes & 0xff); stream[at++] = (unsigned char)((~rowBytes >> 8) & 0xff);
        unsigned char* row = stream + at;
        row[0] = 0;
        encodeRow(y, row + 1);
        for (size_t i = 0; i < rowBytes; i++) { adlerA = (adlerA + row[i]) % 65521u; adlerB = (adlerB + adlerA) % 65521u; }
        at += rowBytes;
    }
    pngBigEndian(stream + at, (adlerB << 16) | adlerA); at += 4;
    bool ok = fwrite(signature, 1, 8, file) == 8 && pngChunk(file, "IHDR", header, 13);
    if (rendererConfiguration.semantics != VCRT_SEMANTICS_BOOK) ok = ok && pngChunk(file, "sRGB", &srgbIntent, 1);
    ok = ok && streamBytes <= 0x7fffffffu && pngChunk(file, "IDAT", stream, (uint32_t)at) && pngChunk(file, "IEND", nullptr, 0);
    delete[] stream;
    fclose(file);
    return ok ? VK_SUCCESS : VK_ERROR_UNKNOWN;
}

VkResult FrontendWritePpm(IN const char* path)
{
    if (!frontendHostImage || !path) return VK_NOT_READY;
    FILE* file = fopen(path, "wb");
    if (!file) return VK_ERROR_UNKNOWN;
    fprintf(file, "P6\n%u %u\n255\n", frontendImageWidth, frontendImageHeight);
    unsigned char* row = new unsigned char[(size_t)frontendImageWidth * 3];
    for (uint32_t y = 0; y < frontendImageHeight; y++) {
        encodeRow(y, row);
        fwrite(row, 1, (size_t)frontendImageWidth * 3, file);
    }
    delete[] row;
    fclose(file);
    return VK_SUCCESS;
}
