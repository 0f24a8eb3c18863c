#!/usr/bin/env python3
# SPDX-License-Identifier: WTFPL
"""Generates tests/golden/* from the REFERENCE's own code (oracle/_ref: the reference's shader
sources compiled as C++ by oracle/build_ref.py, and its SceneGenerator.cpp compiled unmodified).
Run in the container where /root/reference is mounted:

    make -C oracle && python tests/golden/make_golden.py

Files written:
  ref_world.npy              (485,10) f32  the scene baked into globals.glsl:31-517
  ref_scene_text.sha256      SHA-256 of SceneGenerator.cpp's stdout
  ref_hash_64x36_s4.npy      (36,64,4) f32 image rendered by the reference shader code, 4 spp, depth 50
  ref_hash_96x54_s2_d3.npy   (54,96,4) f32 the same at 2 spp, depth 3 (exercises bug B7: depth exhaustion)
  ref_kat.json               known answers: rand / schlick / refract / unit vector / first hits / image statistics
  ref_hash_400x225_s256_u16.npz   (225,400,3) u16: the reference shader code at 400x225, 256 spp, depth 50, RGB quantised to
                             2^-16 (error <= 7.7e-6, against an RMSE tolerance of 1e-2): the high-spp acceptance image
  ref_schlick_dense.json     CRC-32 per block of 2^14 of ref_schlick over oracle_harness.schlick_inputs() (3 x 2^20 inputs)
"""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_harness as oh   # noqa: E402

VP = oh.VP


def main():
    ref = oh.Reference.open()
    assert ref is not None, "build oracle/_ref first (make -C oracle)"
    np.save(os.path.join(HERE, "ref_world.npy"), ref.world())
    text = subprocess.check_output([os.path.join(oh.ORACLE_DIR, "_ref", "SceneGenerator")])
    with open(os.path.join(HERE, "ref_scene_text.sha256"), "w") as f:
        f.write(hashlib.sha256(text).hexdigest() + "\n")

    img, st = ref.render(64, 36, 4, 50)
    np.save(os.path.join(HERE, "ref_hash_64x36_s4.npy"), img)
    img2, st2 = ref.render(96, 54, 2, 3)
    np.save(os.path.join(HERE, "ref_hash_96x54_s2_d3.npy"), img2)

    kat = {"render_64x36_s4": st, "render_96x54_s2_d3": st2}
    kat["rand"] = [[x, y, float(ref.lib.ref_rand(x, y))] for x, y in
                   [(0, 0), (1, 1), (2, 2), (0.5, 0.25), (15, 15), (16, 16), (-0.3, 0.7), (0.123, -4.5), (100.0, 100.0), (499, 499), (500, 500)]]
    kat["schlick"] = [[c, i, float(ref.lib.ref_schlick(c, i))] for c, i in
                      [(0.5, 1.5), (1.0, 1.5), (0.0, 1.5), (3.7, 1.5), (-2.0, 1.5), (13.2, 1.5), (0.25, 2.4)]]
    refr = []
    for v, n, eta in [((0.70710678, -0.70710678, 0), (0, 1, 0), 1 / 1.5), ((0.70710678, -0.70710678, 0), (0, 1, 0), 1.5),
                      ((3.0, -9.0, 2.0), (0, 1, 0), 1 / 1.5), ((0.1, -0.05, 0.99), (0.6, 0.8, 0), 1.5)]:
        va = np.array(v, np.float32); na = np.array(n, np.float32); out = np.zeros(3, np.float32)
        ok = ref.lib.ref_refract(va.ctypes.data_as(VP), na.ctypes.data_as(VP), eta, out.ctypes.data_as(VP))
        refr.append({"v": va.tolist(), "n": na.tolist(), "eta": float(np.float32(eta)), "ok": int(ok), "out": out.tolist()})
    kat["refract"] = refr
    unit = []
    for seed in [(0.3, -0.2, 0.9), (-12.5, -1.9, -2.7), (1, 2, 3)]:
        s = np.array(seed, np.float32); out = np.zeros(3, np.float32)
        ref.lib.ref_unit_vector(s.ctypes.data_as(VP), out.ctypes.data_as(VP))
        unit.append({"seed": s.tolist(), "out": out.tolist()})
    kat["unit_vector"] = unit
    # first hits of pixel-centre rays (no jitter), through the ORACLE's camera frame (itself checked
    # against SURVEY.md's camera known answers in tests) and the REFERENCE's hit_sphere loop
    orc = oh.Oracle()
    hits = []
    for (w, h, pts) in [(400, 225, [(0, 0), (112, 112), (200, 112), (150, 112), (399, 224), (100, 200), (300, 150), (37, 181), (250, 90)]),
                        (1280, 720, [(360, 360), (640, 360), (1279, 719), (5, 700), (900, 400)])]:
        f = orc.camera(w, h)
        for (px, py) in pts:
            pc = (f["pixel00"] + np.float32(px) * f["du"]).astype(np.float32)
            pc = (pc + np.float32(py) * f["dv"]).astype(np.float32)
            d = (pc - f["center"]).astype(np.float32)
            idx, t = ref.first_hit(f["center"], d)
            hits.append({"w": w, "h": h, "px": px, "py": py, "o": f["center"].tolist(), "d": d.tolist(), "index": idx, "t": t})
    kat["first_hit"] = hits
    # image statistics of the reference shader code at BASELINE.json configs[0] (400x225, 16 spp, depth 50)
    big, bst = ref.render(400, 225, 16, 50)
    kat["render_400x225_s16"] = {"rays": bst["rays"], "exhausted": bst["exhausted"],
                                  "mean_rgb": [float(x) for x in big[..., :3].reshape(-1, 3).astype(np.float64).mean(0)],
                                  "min": float(big[..., :3].min()), "max": float(big[..., :3].max())}
    # 8x-downsampled 400x225x16 image: a low-noise fixture for statistical parity of other RNGs (50x28 blocks of 8x8 px,
    # the ragged last row of 1 pixel dropped)
    blocks = big[:224, :400, :3].astype(np.float64).reshape(28, 8, 50, 8, 3).mean((1, 3)).astype(np.float32)
    np.save(os.path.join(HERE, "ref_hash_400x225_s16_blocks8.npy"), blocks)
    with open(os.path.join(HERE, "ref_kat.json"), "w") as f:
        json.dump(kat, f, indent=1)
    # the high-spp acceptance image (RMSE < 1e-2 wherever both sides have >= 128 spp)
    hi, hst = ref.render(400, 225, 256, 50)
    assert hi[..., :3].min() >= 0 and hi[..., :3].max() <= 1 and (hi[..., 3] == 1).all()
    np.savez_compressed(os.path.join(HERE, "ref_hash_400x225_s256_u16.npz"),
                        rgb=np.round(hi[..., :3].astype(np.float64) * 65535.0).astype(np.uint16),
                        rays=np.uint64(hst["rays"]), exhausted=np.uint64(hst["exhausted"]),
                        mean_rgb=hi[..., :3].reshape(-1, 3).astype(np.float64).mean(0))
    # schlick on a dense input set: what functions.glsl:58-62 returns in the reference build (C powf of this image's libm)
    cos, ior = oh.schlick_inputs()
    out = ref.schlick_many(cos, ior)
    with open(os.path.join(HERE, "ref_schlick_dense.json"), "w") as f:
        json.dump({"inputs": "oracle_harness.schlick_inputs(1 << 20)", "count": int(out.size), "block": 1 << 14, "crc32": oh.block_crcs(out),
                   "negative_results": int((out < 0).sum()), "nan_results": int(np.isnan(out).sum())}, f)
    print("golden written:", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
