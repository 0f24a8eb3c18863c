# SPDX-License-Identifier: WTFPL
"""The product's scene producer (host/SceneGenerator.cpp, in libvcrt.so) vs the reference's baked
table and vs the oracle's independent restatement.  CPU only (no device call)."""
import ctypes
import hashlib
import os
import subprocess

import numpy as np

from vulkancomputeraytracing_b200 import cabi, host

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_reference_recipe_reproduces_baked_world(vcrt):
    np.testing.assert_array_equal(host.generate_scene(), np.load(os.path.join(GOLDEN, "ref_world.npy")))


def test_tool_stdout_is_byte_identical_to_reference_tool():
    exe = os.path.join(os.path.dirname(cabi.LIB_PATH), "SceneGenerator")
    text = subprocess.check_output([exe])
    with open(os.path.join(GOLDEN, "ref_scene_text.sha256")) as f:
        assert hashlib.sha256(text).hexdigest() == f.read().strip()


def test_scaled_and_stress_recipes_agree_with_oracle(vcrt, oracle):
    for grid in (3, 50):
        np.testing.assert_array_equal(host.generate_scene(grid), oracle.scene(grid))
    assert host.generate_scene(50).shape[0] == 10003          # BASELINE.json configs[3]: "10k"
    stress = host.stress_scene()
    base = oracle.scene(11, 0.0, 0.5)
    np.testing.assert_array_equal(stress[:-4], base[:-4])     # grid part: same draws, thresholds 0 / 0.5
    assert not (stress[:, 7] == 1).any()                      # no lambertian anywhere
    assert stress[-1, 3] == 1000 and stress[-1, 7] == 2 and stress[-1, 8] == np.float32(0.1)


def test_glsl_text_round_trip(vcrt):
    scene = host.generate_scene()
    ptr = scene.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere))
    need = vcrt.vcrtSceneToGlsl(ptr, scene.shape[0], None, 0)
    buf = ctypes.create_string_buffer(need + 1)
    assert vcrt.vcrtSceneToGlsl(ptr, scene.shape[0], buf, need + 1) == need
    back = np.zeros((600, 10), np.float32)
    n = vcrt.vcrtSceneFromGlsl(buf.value, back.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), 600)
    assert n == 485
    np.testing.assert_array_equal(back[:n], scene)
    # comments and blank lines are skipped; a malformed line is an error, not a silent drop
    text = b"// sphere(vec3(0,0,0), 1.0, vec3(1,1,1), vec3(TEXTURE_GLASS,1.5,0.0)),\n\n" + buf.value[:200].rsplit(b"\n", 1)[0] + b"\n"
    assert vcrt.vcrtSceneFromGlsl(text, back.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), 600) == text.count(b"\nsphere(")
    assert vcrt.vcrtSceneFromGlsl(b"sphere(vec3(1,2), 1.0)", back.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere)), 600) == cabi.VCRT_ERROR_SCENE


def test_bad_recipes_are_rejected(vcrt):
    recipe = cabi.VcrtSceneRecipe()
    vcrt.vcrtDefaultSceneRecipe(ctypes.byref(recipe))
    assert (recipe.gridHalf, recipe.lambertBelow, recipe.metalBelow, recipe.variant) == (11, 0.8, 0.95, 0)
    out = np.zeros((10, 10), np.float32)
    ptr = out.ctypes.data_as(ctypes.POINTER(cabi.VcrtSphere))
    assert vcrt.vcrtGenerateScene(ctypes.byref(recipe), ptr, 10) == cabi.VCRT_ERROR_INVALID_ARGUMENT   # capacity too small
    recipe.gridHalf = -1
    assert vcrt.vcrtGenerateScene(ctypes.byref(recipe), ptr, 10) == cabi.VCRT_ERROR_INVALID_ARGUMENT
    recipe.gridHalf = 0
    assert vcrt.vcrtGenerateScene(ctypes.byref(recipe), ptr, 10) == 4    # empty grid: the three big spheres + ground
