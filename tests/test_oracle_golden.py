"""Pins the oracle (oracle/crt_oracle.c, the CPU restatement of the trace loop)
against golden vectors produced by the UNMODIFIED reference compiled in place
(tests/golden/make_golden.py).  The restatement keeps the reference's
floating-point operation order, so images must agree bit for bit except for the
documented tie / node-epsilon cases; a tiny allowance covers
those."""
import os

import numpy as np
import pytest

from _util import GOLDEN, golden_images, load_golden_image, open_case, rel_rmse, scene_path
from c_raytracer_b200 import _abi, host
from oracle import oracle


@pytest.mark.parametrize("name", golden_images())
def test_oracle_image_equals_reference(name):
    case = load_golden_image(name)
    sc, opts = open_case(case)
    st = _abi.Stats()
    raster, z = oracle.OracleScene(sc).render(opts, st)
    # ray counts: cast_ray() invocations and is_light_blocked() calls of the reference
    assert (st.rays_closest, st.rays_shadow) == (case["cast_ray"], case["shadow"])
    differing = int(np.any(raster != case["raster"], axis=2).sum())
    assert differing <= max(1, raster.shape[0] * raster.shape[1] // 1000), differing
    assert rel_rmse(raster, case["raster"]) <= 1e-6
    np.testing.assert_array_equal(z, case["z"])


def test_oracle_brute_force_equals_bvh():
    """The oracle's own BVH does not change results (topology independence)."""
    for name in ("scenetest", "scene6s_ambient"):
        case = load_golden_image(name + "_ambient" if name == "scenetest" else name)
        sc, opts = open_case(case)
        a, za = oracle.OracleScene(sc).render(opts)
        b, zb = oracle.OracleScene(sc, brute_force=True).render(opts)
        np.testing.assert_array_equal(za, zb)
        assert rel_rmse(a, b) <= 1e-6


@pytest.mark.parametrize("scene", ["scene1", "scene3", "scene4", "scenetest", "scenetest2", "scene5_small", "scene6_small"])
def test_oracle_ray_batches(scene):
    """get_closest_intersection() / is_light_blocked() of the reference on 1500 rays."""
    g = np.load(os.path.join(GOLDEN, f"rays_{scene}.npz"))
    orc = oracle.OracleScene(host.Scene(scene_path(scene)))
    obj, t, normal = orc.trace_closest(g["rays"])
    same = obj == g["object"]
    assert same.mean() >= 0.999  # exact-distance ties may pick the other primitive
    np.testing.assert_array_equal(t[same], g["t"][same])
    np.testing.assert_array_equal(normal[same], g["normal"][same])
    blocked, light = orc.trace_shadow(g["rays"], g["dist"], g["emitter"], g["light_in"])
    np.testing.assert_array_equal(blocked, g["blocked"].astype(np.uint8))
    free = blocked == 0
    np.testing.assert_allclose(light[free], g["light_out"][free], rtol=1e-6)  # kt products commute up to rounding


def test_oracle_textures():
    g = np.load(os.path.join(GOLDEN, "textures.npz"))
    keys = sorted({k.rsplit(":", 1)[0] for k in g.files})
    assert len(keys) >= 8
    kinds = set()
    for key in keys:
        scene, mi = key.split(":")
        sc = host.Scene(scene_path(scene))
        kinds.add(int(sc.materials["tex_type"][int(mi)]))
        got = oracle.OracleScene(sc).texture(int(mi), g[key + ":points"])
        np.testing.assert_array_equal(got, g[key + ":colors"])
    assert kinds == {0, 1, 2, 3}  # uniform, checkerboard, brick, noisy periodic all covered


def test_oracle_simplex_noise():
    g = np.load(os.path.join(GOLDEN, "noise.npz"))
    np.testing.assert_array_equal(oracle.simplex_noise(g["points"]), g["values"])
    assert np.abs(g["values"]).max() <= 1.0


def test_oracle_primitive_kats():
    """moller_trumbore() and line_intersects_sphere() known answers, incl. edge,
    tangent, inside-sphere and behind-origin cases."""
    import ctypes as C
    g = np.load(os.path.join(GOLDEN, "primitives.npz"))
    rec = g["records"]
    L = oracle.lib()
    hits = np.zeros(len(rec), dtype=np.int32)
    ts = np.zeros(len(rec), dtype=np.float32)
    for i, r in enumerate(rec):
        o = _abi.Object()
        kind = int(r[0])
        o.type = _abi.CRT_OBJ_TRIANGLE if kind == 0 else _abi.CRT_OBJ_SPHERE
        o.epsilon = float(r[1])
        o.a[:] = [float(x) for x in r[2:5]]
        o.b[:] = [float(x) for x in r[5:8]]
        o.c[:] = [float(x) for x in r[8:11]]
        org = (C.c_float * 3)(*[float(x) for x in r[17:20]])
        d = (C.c_float * 3)(*[float(x) for x in r[14:17]])
        t = C.c_float()
        hits[i] = L.oracle_primitive_hit(kind, C.byref(o), org, d, C.byref(t))
        ts[i] = t.value if hits[i] else 0.0
    np.testing.assert_array_equal(hits, g["hit"])
    np.testing.assert_array_equal(ts, g["t"])
    assert 0.3 < g["hit"].mean() < 0.8
