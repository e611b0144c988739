"""Host tree builders (c-raytracer_b200/csrc/bvh_build.cpp): the binned-SAH
binary tree (one primitive per leaf) and its 8-wide collapse into the 128-byte device nodes, checked by a small C++
program (tests/bvh_check.cpp) on seeded boxes -- every primitive in exactly one
leaf slot (the shadow query tints once per transparent primitive, reference
accel.c:369-373), the stored half-precision boxes contain what is below them, children and primitives of a node are
contiguous, the depth fits the device stacks, and the multi-threaded build
gives the single-threaded tree bit for bit."""
import os
import subprocess

import pytest

from _util import ROOT


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("bvh") / "bvh_check")
    csrc = os.path.join(ROOT, "c-raytracer_b200", "csrc")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", csrc, os.path.join(ROOT, "tests", "bvh_check.cpp"),
                    os.path.join(csrc, "bvh_build.cpp"), "-o", exe, "-lpthread"], check=True)
    return exe


@pytest.mark.parametrize("n,seed,kind", [
    (0, 1, 0), (1, 1, 0), (2, 2, 0), (5, 3, 0), (9, 4, 2), (1000, 5, 0), (1000, 6, 1), (4000, 7, 2),
    (60000, 8, 0), (60000, 9, 1), (150000, 10, 2),
    (150, 11, 3), (600, 12, 3), (120, 13, 4), (500, 14, 4),  # deep chains; surfaces that overflow FP32 (once a stack overflow in the collapse)
])
def test_tree_invariants(checker, n, seed, kind):
    r = subprocess.run([checker, str(n), str(seed), str(kind)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok")


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_node_test_is_conservative(tmp_path_factory, seed):
    """The device's half-precision node test, emulated bit for bit on the CPU (tests/node_test_emu.cpp), never
    misses a slot whose content the ray segment truly meets: random, axis-parallel, face / edge / corner grazing
    from near and far, origins inside the node (with kWideSlack set to 0 the same program fails at once)."""
    exe = str(tmp_path_factory.mktemp("emu") / "node_test_emu")
    csrc = os.path.join(ROOT, "c-raytracer_b200", "csrc")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", csrc, os.path.join(ROOT, "tests", "node_test_emu.cpp"),
                    os.path.join(csrc, "bvh_build.cpp"), "-o", exe, "-lpthread"], check=True)
    r = subprocess.run([exe, "20000", "30000", str(seed)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.startswith("ok")


@pytest.mark.parametrize("seed", [1, 2])
def test_rest_box_cull_is_conservative(tmp_path_factory, seed):
    """The test that lets a shadow ray skip the tree (crt_device.cuh ray_misses_box against DEmitter.rest_lo / rest_hi as
    crt_api.cu pads them), restated in the same FP32 operations (tests/rest_box_emu.cpp): no ray that meets its box
    within the light distance is culled -- random, axis-parallel, origins up to 10^6 box sizes away, flat boxes, points
    on faces / edges / corners, light barely beyond the occluder; a NaN ray is never culled.  (Without the margins the
    same program reports a false cull within 12 000 cases.)"""
    exe = str(tmp_path_factory.mktemp("emu") / "rest_box_emu")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", os.path.join(ROOT, "tests", "rest_box_emu.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe, "3000000", str(seed)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 false culls" in r.stdout
    # and it is worth having: nearly every box that lies beside the ray is recognised
    culled = float(r.stdout.split("(")[-1].split("%")[0])
    assert culled > 90.0
