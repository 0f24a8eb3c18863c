# SPDX-License-Identifier: WTFPL
"""The headless front-end's file writers (host/Frontend.cpp), on a synthetic image: no GPU involved.
A small C++ harness is compiled against Frontend.cpp the way the host program uses it."""
import os
import struct
import subprocess
import zlib

import numpy as np

from vulkancomputeraytracing_b200 import build, cabi

HARNESS = r"""
#include "Frontend.hpp"
#include <stdio.h>
#include <stdlib.h>
RendererConfiguration rendererConfiguration = {};
int main(int argc, char** argv)
{
    rendererConfiguration.width = (uint32_t)atoi(argv[1]); rendererConfiguration.height = (uint32_t)atoi(argv[2]);
    rendererConfiguration.semantics = (uint32_t)atoi(argv[3]);
    if (CreateVulkanWindowFrontend() != VK_SUCCESS) return 2;
    FILE* f = fopen(argv[4], "rb");
    size_t n = (size_t)frontendImageWidth * frontendImageHeight * 4;
    if (!f || fread(frontendHostImage, sizeof(float), n, f) != n) return 3;
    fclose(f);
    char path[1024];
    snprintf(path, sizeof path, "%s.pfm", argv[5]); if (FrontendWritePfm(path) != VK_SUCCESS) return 4;
    snprintf(path, sizeof path, "%s.ppm", argv[5]); if (FrontendWritePpm(path) != VK_SUCCESS) return 5;
    snprintf(path, sizeof path, "%s.png", argv[5]); if (FrontendWritePng(path) != VK_SUCCESS) return 6;
    return DestroyVulkanWindowFrontend() == VK_SUCCESS ? 0 : 7;
}
"""


def decode_png(data):
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        size, kind = struct.unpack(">I4s", data[at:at + 8])
        body = data[at + 8:at + 8 + size]
        assert struct.unpack(">I", data[at + 8 + size:at + 12 + size])[0] == zlib.crc32(kind + body)   # every chunk's CRC
        chunks.append((kind, body)); at += 12 + size
    kinds = [k for k, _ in chunks]
    assert kinds[0] == b"IHDR" and kinds[-1] == b"IEND"
    w, h, depth, colour, comp, filt, lace = struct.unpack(">IIBBBBB", chunks[0][1])
    assert (depth, colour, comp, filt, lace) == (8, 2, 0, 0, 0)
    raw = zlib.decompress(b"".join(b for k, b in chunks if k == b"IDAT"))                              # checks the Adler-32 too
    rows = np.frombuffer(raw, np.uint8).reshape(h, 1 + 3 * w)
    assert (rows[:, 0] == 0).all()                                                                      # filter type "none"
    return rows[:, 1:].reshape(h, w, 3), kinds


def srgb8(linear):
    x = np.clip(np.nan_to_num(linear.astype(np.float64), nan=0.0), 0, 1)
    enc = np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)
    return enc * 255.0


def run_harness(tmp_path, image, semantics):
    build.build()
    lib = os.path.dirname(cabi.LIB_PATH)
    src, exe = tmp_path / "harness.cpp", tmp_path / "harness"
    src.write_text(HARNESS)
    frontend = os.path.join(build.HOST, "Frontend.cpp")
    subprocess.check_call([build.GXX] + build.GXX_FLAGS + ["-o", str(exe), str(src), frontend, "-L", lib, "-lvcrt", f"-Wl,-rpath,{lib}"])
    raw, prefix = tmp_path / "image.f32", str(tmp_path / f"out{semantics}")
    image.astype("<f4").tofile(raw)
    h, w = image.shape[:2]
    subprocess.check_call([str(exe), str(w), str(h), str(semantics), str(raw), prefix])
    return prefix


def test_png_ppm_pfm_of_a_synthetic_image(tmp_path):
    rng = np.random.default_rng(7)
    h, w = 37, 53
    image = np.ones((h, w, 4), np.float32)
    image[..., :3] = rng.uniform(0, 1, (h, w, 3)) ** 3
    image[0, 0, :3] = [0.0, 1.0, 2.5]; image[0, 1, :3] = [-1.0, np.nan, 0.0031308]; image[1, 0, :3] = [1e-9, 0.5, 0.999999]
    prefix = run_harness(tmp_path, image, 0)
    with open(prefix + ".pfm", "rb") as f:
        assert f.readline() == b"PF\n" and f.readline() == b"%d %d\n" % (w, h) and f.readline() == b"-1.0\n"
        assert np.array_equal(np.frombuffer(f.read(), "<f4").reshape(h, w, 3)[::-1], image[..., :3], equal_nan=True)
    with open(prefix + ".ppm", "rb") as f:
        assert f.readline() == b"P6\n" and f.readline() == b"%d %d\n" % (w, h) and f.readline() == b"255\n"
        ppm = np.frombuffer(f.read(), np.uint8).reshape(h, w, 3)
    png, kinds = decode_png(open(prefix + ".png", "rb").read())
    assert np.array_equal(png, ppm) and b"sRGB" in kinds                 # the same 8-bit image in both containers
    assert np.abs(ppm.astype(np.float64) - srgb8(image[..., :3])).max() <= 0.5 + 1e-3   # the B8G8R8A8_SRGB transfer, rounded to nearest
    assert tuple(ppm[0, 0]) == (0, 255, 255) and tuple(ppm[0, 1][:2]) == (0, 0)


def test_book_mode_files_are_quantised_only(tmp_path):
    # BOOK images are gamma-encoded by the layer already: the writers must not apply the sRGB curve again
    h, w = 5, 9
    image = np.ones((h, w, 4), np.float32)
    image[..., :3] = np.linspace(0, 1, h * w * 3, dtype=np.float32).reshape(h, w, 3)
    prefix = run_harness(tmp_path, image, 1)
    png, kinds = decode_png(open(prefix + ".png", "rb").read())
    expect = (255.999 * np.clip(image[..., :3].astype(np.float64), 0, 0.999)).astype(np.uint8)
    assert np.abs(png.astype(int) - expect.astype(int)).max() <= 1 and b"sRGB" not in kinds
