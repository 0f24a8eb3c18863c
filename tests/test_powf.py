# SPDX-License-Identifier: WTFPL
"""pow(x, 5.) of functions.glsl:61 (CPU part): the product's restatement of glibc's powf (csrc/vcrt_powf.cuh), compiled for the
host, against (a) the C library of this machine and (b) the golden vectors of the reference build (tests/golden/
ref_schlick_dense.json, generated from oracle/_ref).  The GPU part is tests/test_gpu_parity.py::test_schlick_dense."""
import ctypes
import json
import os
import subprocess

import numpy as np
import pytest

import oracle_harness as oh

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def powf_host(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("powf") / "libpowf_host.so")
    subprocess.check_call(["/usr/bin/g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", so, os.path.join(HERE, "powf_host_check.cpp")])
    lib = ctypes.CDLL(so)
    lib.powf5_compare.argtypes = [oh.VP, ctypes.c_long, oh.VP]; lib.powf5_compare.restype = ctypes.c_long
    lib.schlick_many.argtypes = [oh.VP, oh.VP, ctypes.c_long, oh.VP]
    return lib


def test_powf5_equals_this_machines_libm(powf_host):
    # 6e6 bases: what schlick() feeds it (1 - cosine, cosine in (-15, 2)), log-uniform magnitudes over the whole main path of
    # powf (2^-25 .. 2^25, both signs), and the special values.  glibc 2.39 x86-64 with FMA (the image's libm): 0 mismatches
    # (8e7 inputs were run once by hand).  Another libm may round differently - then this test says so, and the goldens decide.
    rng = np.random.default_rng(5)
    x = np.concatenate([
        (1.0 - rng.uniform(-15.0, 2.0, 3_000_000)).astype(np.float32),
        (np.exp2(rng.uniform(-25.0, 25.0, 3_000_000)) * rng.choice([-1.0, 1.0], 3_000_000)).astype(np.float32),
        np.array([0.0, -0.0, 1.0, -1.0, 2.0, np.inf, -np.inf, np.nan, 1e-40, -1e-40, 3e8, -3e8, 1e-9, 16.0, -14.0], np.float32)])
    bad = ctypes.c_float()
    n = powf_host.powf5_compare(x.ctypes.data_as(oh.VP), x.size, ctypes.byref(bad))
    assert n == 0, f"{n} of {x.size} differ from powf of this machine's libm, first at x = {bad.value!r}"


def test_schlick_restatement_equals_reference_build_golden(powf_host):
    with open(os.path.join(HERE, "golden", "ref_schlick_dense.json")) as f:
        want = json.load(f)
    cos, ior = oh.schlick_inputs()
    assert cos.size == want["count"]
    out = np.empty_like(cos)
    powf_host.schlick_many(cos.ctypes.data_as(oh.VP), ior.ctypes.data_as(oh.VP), cos.size, out.ctypes.data_as(oh.VP))
    got = oh.block_crcs(out, want["block"])
    differing = sum(a != b for a, b in zip(got, want["crc32"]))
    assert differing == 0, f"{differing} of {len(got)} blocks differ from ref_schlick of the reference build"
    assert int((out < 0).sum()) == want["negative_results"] > cos.size // 5      # bug B6: negative bases give negative "probabilities"


def test_schlick_golden_is_what_the_live_reference_build_returns(reference):
    with open(os.path.join(HERE, "golden", "ref_schlick_dense.json")) as f:
        want = json.load(f)
    cos, ior = oh.schlick_inputs()
    assert oh.block_crcs(reference.schlick_many(cos, ior), want["block"]) == want["crc32"]


def test_three_multiplies_would_not_do(powf_host):
    # why the device does not compute x^5 as (x^2)^2 x in binary32: half of the inputs come out one ulp off powf
    x = (1.0 - np.random.default_rng(6).uniform(0.0, 1.0, 100_000)).astype(np.float32)
    x2 = x * x
    x5 = (x2 * x2) * x
    libm = np.array([np.float32(v) for v in np.power(x.astype(np.float64), 5.0)], np.float32)   # correctly rounded x^5
    assert 0.3 < (x5 != libm).mean() < 0.7
