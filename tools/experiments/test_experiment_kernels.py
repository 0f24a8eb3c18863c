# SPDX-License-Identifier: WTFPL
"""A/B parity of the EXPERIMENT kernels (tools/experiments/*.cuh) - not part of the product or of the driver's test run.

    VCRT_EXTRA_NVCC=-DVCRT_EXPERIMENTS python -m vulkancomputeraytracing_b200.build --force
    python -m pytest tools/experiments/test_experiment_kernels.py -q -p no:cacheprovider --rootdir tests -c /dev/null

The product build refuses non-zero `reserved` fields; with -DVCRT_EXPERIMENTS reserved[1] selects a kernel generation / CTA
shape and reserved[0] the samples per work item."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_harness as oh
from vulkancomputeraytracing_b200 import cabi, host
from test_gpu_parity import assert_same, BRUTE, LBVH


@pytest.fixture(scope="module")
def oracle():
    return oh.Oracle()


@pytest.fixture(scope="module")
def vcrt():
    return cabi.load()


@pytest.mark.parametrize("shape", [1, 10, 11, 20, 21])
def test_every_brute_force_kernel_generation_agrees(oracle, vcrt, shape):
    # development switch `shape`: 1 = first generation (LDS-fed filter), 10 / 11 = uniform-datapath scan (512x4 / 1024x2 rays),
    # 20 / 21 = warp-compacted shading (1024x2 / 512x4 rays).  Same bits from all of them, in both semantics and accumulators.
    for scene, cam in ((host.generate_scene(), None), (host.stress_scene(), None)):
        host.upload_scene(scene, cam)
        try:
            cpu, cst = oracle.render(scene, 203, 113, 3, depth=9)
            gpu, st = host.render(203, 113, 3, depth=9, traversal=BRUTE, sample_chunks=1, shape=shape)
            assert_same(gpu, st, cpu, cst)
            fixed, sf = host.render(203, 113, 3, depth=9, traversal=BRUTE, shape=shape)
            assert sf.rays == cst["rays"] and np.abs(fixed - cpu).max() <= 2.0 ** -22
            if shape != 1:
                book, cb = oracle.render_book(scene, 96, 54, 2)
                gb, sb = host.render(96, 54, 2, traversal=BRUTE, sample_chunks=1, shape=shape, semantics=cabi.VCRT_SEMANTICS_BOOK)
                assert sb.rays == cb["rays"] and sb.depthExhausted == cb["exhausted"] and np.array_equal(gb, book)
        finally:
            host.destroy()


def test_both_lbvh_kernel_generations_agree(oracle, vcrt):
    # development switch shape 31: the first-generation LBVH kernel (nested sample / bounce loops); the default is the flat
    # loop whose idle lanes restart together.  Same bits as the oracle in strict order, same fixed-point image, both semantics.
    for scene in (host.generate_scene(), host.stress_scene(), host.generate_scene(30)):
        host.upload_scene(scene)
        try:
            for shape in (0, 31):
                for depth in (9, 1, 0):
                    cpu, cst = oracle.render(scene, 203, 113, 3, depth=depth)
                    gpu, st = host.render(203, 113, 3, depth=depth, traversal=LBVH, sample_chunks=1, shape=shape)
                    assert_same(gpu, st, cpu, cst)
                book, cb = oracle.render_book(scene, 96, 54, 2)
                gb, sb = host.render(96, 54, 2, traversal=LBVH, sample_chunks=1, shape=shape, semantics=cabi.VCRT_SEMANTICS_BOOK)
                assert sb.rays == cb["rays"] and sb.depthExhausted == cb["exhausted"] and np.array_equal(gb, book)
            nested, sn = host.render(203, 113, 5, traversal=LBVH, shape=31)
            flat, sf = host.render(203, 113, 5, traversal=LBVH)
            assert sf.rays == sn.rays and 0 < sf.nodeVisits <= sn.nodeVisits and np.array_equal(flat, nested)   # the flat kernel widens its boxes less
        finally:
            host.destroy()


