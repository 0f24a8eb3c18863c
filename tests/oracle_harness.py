# SPDX-License-Identifier: WTFPL
"""TEST INFRASTRUCTURE: ctypes access to the CPU oracle (oracle/_build) and, where it was built,
the reference's own shader code compiled as C++ (oracle/_ref).  Only tests/, smoke() and bench.py's
CPU-baseline legs import this."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
VP = ctypes.c_void_p
RNG_HASH, RNG_PHILOX = 0, 1
CAMERA = np.array([13, 2, 3, 0, 0, 0, 0, 1, 0, 20], np.float32)   # globals.glsl:21-24


def schlick_inputs(per_ior=1 << 20):
    """The dense input set of the schlick parity test: per_ior cosines for each ior in (1.5, 2.4, 0.5).  Half of them are spread
    over (-15, 2), a quarter lies in [0, 1] (1 - cosine is the positive base a refracting ray sees) and a quarter in (1, 15): the
    cosine is an unnormalised dot product, so most calls of a render have a NEGATIVE base (bug B6).  Deterministic (PCG64 seed
    20251105); used by make_golden.py and the tests alike."""
    rng = np.random.default_rng(20251105)
    cos, ior = [], []
    for value in (1.5, 2.4, 0.5):
        wide = rng.uniform(-15.0, 2.0, per_ior // 2).astype(np.float32)
        unit = rng.uniform(0.0, 1.0, per_ior // 4).astype(np.float32)
        above = rng.uniform(1.0, 15.0, per_ior - per_ior // 2 - per_ior // 4).astype(np.float32)
        cos.append(np.concatenate([wide, unit, above])); ior.append(np.full(per_ior, value, np.float32))
    return np.concatenate(cos), np.concatenate(ior)


def block_crcs(values, block=1 << 14):
    import zlib
    raw = np.ascontiguousarray(values, np.float32)
    return [zlib.crc32(raw[i:i + block].tobytes()) for i in range(0, raw.size, block)]


def build_oracle():
    so = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
    src = os.path.join(ORACLE_DIR, "vcrt_oracle.cpp")
    fast = os.path.join(ORACLE_DIR, "_build", "liboracle_fast.so")
    if not (os.path.exists(so) and os.path.exists(fast)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "_build/liboracle.so", "_build/liboracle_fast.so"])
    return so, fast


class Oracle:
    def __init__(self, fast=False):
        so, fast_so = build_oracle()
        self.lib = L = ctypes.CDLL(fast_so if fast else so)
        L.oracle_generate_scene.argtypes = [ctypes.c_int, ctypes.c_double, ctypes.c_double, VP, ctypes.c_int]
        L.oracle_generate_scene.restype = ctypes.c_int
        L.oracle_render.argtypes = [VP, ctypes.c_int, VP] + [ctypes.c_int] * 6 + [ctypes.c_uint64, ctypes.c_int, ctypes.c_int, VP, VP]
        L.oracle_hash_rand.argtypes = [ctypes.c_float, ctypes.c_float]; L.oracle_hash_rand.restype = ctypes.c_float
        L.oracle_schlick.argtypes = [ctypes.c_float, ctypes.c_float]; L.oracle_schlick.restype = ctypes.c_float
        L.oracle_refract.argtypes = [VP, VP, ctypes.c_float, VP]
        L.oracle_philox.argtypes = [VP, VP, VP]
        L.oracle_camera.argtypes = [VP, ctypes.c_int, ctypes.c_int, VP]
        L.oracle_first_hit.argtypes = [VP, ctypes.c_int, VP, VP, VP]
        L.oracle_render_book.argtypes = [VP, ctypes.c_int, VP] + [ctypes.c_int] * 5 + [ctypes.c_uint64, ctypes.c_int, ctypes.c_int, VP, VP]
        L.oracle_camera_book.argtypes = [VP, ctypes.c_int, ctypes.c_int, VP]

    def scene(self, grid_half=11, lambert_below=0.8, metal_below=0.95):
        cap = 4 * grid_half * grid_half + 4
        table = np.zeros((cap, 10), np.float32)
        n = self.lib.oracle_generate_scene(grid_half, lambert_below, metal_below, table.ctypes.data_as(VP), cap)
        assert n > 0
        return np.ascontiguousarray(table[:n])

    def render(self, spheres, width, height, spp, depth=50, rng=RNG_PHILOX, seed=0, sample_offset=0, normalise=True,
               threads=0, camera=CAMERA):
        spheres = np.ascontiguousarray(spheres, np.float32)
        img = np.zeros((height, width, 4), np.float32)
        st = np.zeros(14, np.uint64)
        cam = np.ascontiguousarray(camera, np.float32)
        rc = self.lib.oracle_render(spheres.ctypes.data_as(VP), spheres.shape[0], cam.ctypes.data_as(VP), width, height, spp,
                                    sample_offset, depth, rng, seed, 1 if normalise else 0, threads, img.ctypes.data_as(VP),
                                    st.ctypes.data_as(VP))
        assert rc == 0
        stats = {"rays": int(st[0]), "exhausted": int(st[1]), "hits": [int(x) for x in st[2:6]],
                 "path_hist": [int(x) for x in st[6:14]]}
        return img, stats

    def render_book(self, spheres, width, height, spp, depth=50, seed=0, sample_offset=0, normalise=True, threads=0, camera=CAMERA):
        """BOOK semantics: the book's routines in binary32; parity unpinned (see the oracle source)."""
        spheres = np.ascontiguousarray(spheres, np.float32)
        img = np.zeros((height, width, 4), np.float32)
        st = np.zeros(14, np.uint64)
        cam = np.ascontiguousarray(camera, np.float32)
        rc = self.lib.oracle_render_book(spheres.ctypes.data_as(VP), spheres.shape[0], cam.ctypes.data_as(VP), width, height, spp,
                                         sample_offset, depth, seed, 1 if normalise else 0, threads, img.ctypes.data_as(VP),
                                         st.ctypes.data_as(VP))
        assert rc == 0
        return img, {"rays": int(st[0]), "exhausted": int(st[1]), "hits": [int(x) for x in st[2:6]], "path_hist": [int(x) for x in st[6:14]]}

    def camera(self, width, height, camera=CAMERA, book=False):
        out = np.zeros(12, np.float32)
        cam = np.ascontiguousarray(camera, np.float32)
        (self.lib.oracle_camera_book if book else self.lib.oracle_camera)(cam.ctypes.data_as(VP), width, height, out.ctypes.data_as(VP))
        return {"center": out[0:3], "pixel00": out[3:6], "du": out[6:9], "dv": out[9:12]}

    def first_hit(self, spheres, origin, direction):
        spheres = np.ascontiguousarray(spheres, np.float32)
        o = np.ascontiguousarray(origin, np.float32); d = np.ascontiguousarray(direction, np.float32)
        t = np.zeros(1, np.float32)
        idx = self.lib.oracle_first_hit(spheres.ctypes.data_as(VP), spheres.shape[0], o.ctypes.data_as(VP), d.ctypes.data_as(VP),
                                        t.ctypes.data_as(VP))
        return idx, float(t[0])

    def philox(self, ctr, key):
        c = np.array(ctr, np.uint32); k = np.array(key, np.uint32); out = np.zeros(4, np.uint32)
        self.lib.oracle_philox(c.ctypes.data_as(VP), k.ctypes.data_as(VP), out.ctypes.data_as(VP))
        return out

    def refract(self, v, n, eta):
        v = np.array(v, np.float32); n = np.array(n, np.float32); out = np.zeros(3, np.float32)
        ok = self.lib.oracle_refract(v.ctypes.data_as(VP), n.ctypes.data_as(VP), eta, out.ctypes.data_as(VP))
        return bool(ok), out


class Reference:
    """oracle/_ref/libvcrt_ref.so: the reference's GLSL compiled as C++ (see oracle/build_ref.py)."""

    @staticmethod
    def open(fast=False):
        so = os.path.join(ORACLE_DIR, "_ref", "libvcrt_ref_fast.so" if fast else "libvcrt_ref.so")
        if not os.path.exists(so):
            return None
        return Reference(so)

    def __init__(self, so):
        self.lib = L = ctypes.CDLL(so)
        L.ref_render.argtypes = [ctypes.c_int] * 5 + [VP, VP]
        L.ref_rand.argtypes = [ctypes.c_float, ctypes.c_float]; L.ref_rand.restype = ctypes.c_float
        L.ref_schlick.argtypes = [ctypes.c_float, ctypes.c_float]; L.ref_schlick.restype = ctypes.c_float
        L.ref_refract.argtypes = [VP, VP, ctypes.c_float, VP]
        if hasattr(L, "ref_schlick_many"):
            L.ref_schlick_many.argtypes = [VP, VP, ctypes.c_int, VP]
        L.ref_unit_vector.argtypes = [VP, VP]
        L.ref_first_hit.argtypes = [VP, VP, VP]
        L.ref_world.argtypes = [VP]

    def world(self):
        n = self.lib.ref_world_count()
        out = np.zeros((n, 10), np.float32)
        self.lib.ref_world(out.ctypes.data_as(VP))
        return out

    def render(self, width, height, spp, depth=50, threads=0):
        img = np.zeros((height, width, 4), np.float32)
        st = np.zeros(2, np.uint64)
        assert self.lib.ref_render(width, height, spp, depth, threads, img.ctypes.data_as(VP), st.ctypes.data_as(VP)) == 0
        return img, {"rays": int(st[0]), "exhausted": int(st[1])}

    def schlick_many(self, cosine, ior):
        cosine = np.ascontiguousarray(cosine, np.float32); ior = np.ascontiguousarray(ior, np.float32); out = np.empty_like(cosine)
        self.lib.ref_schlick_many(cosine.ctypes.data_as(VP), ior.ctypes.data_as(VP), cosine.size, out.ctypes.data_as(VP))
        return out

    def first_hit(self, origin, direction):
        o = np.ascontiguousarray(origin, np.float32); d = np.ascontiguousarray(direction, np.float32)
        t = np.zeros(1, np.float32)
        idx = self.lib.ref_first_hit(o.ctypes.data_as(VP), d.ctypes.data_as(VP), t.ctypes.data_as(VP))
        return idx, float(t[0])
