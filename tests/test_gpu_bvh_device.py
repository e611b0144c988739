"""The tree built on the device (crt_scene_desc.build = CRT_BUILD_DEVICE, crt_bvh_device.cu;
SURVEY 8(f) N3; reference accel_init(), accel.c:266-315) against the host builder's: the same
invariants on the records the kernels read, the same answers to the same rays, the same frames."""
import os

import numpy as np
import pytest

from _util import ROOT, scene_path
import _pkg

_pkg.load()
from c_raytracer_b200 import _abi, device, host, parity  # noqa: E402

pytestmark = pytest.mark.gpu

SLACK = 3.5  # bvh_build.h kWideSlack


@pytest.fixture(scope="module")
def gpu():
    if device.device_count() < 1:
        pytest.skip("no CUDA device")
    return 0


def half_to_float(bits):
    return bits.astype(np.uint16).view(np.float16).astype(np.float64)


def true_boxes(prims):
    """(lo, hi) of every primitive record, no padding"""
    info = prims[:, 7].view(np.uint32)
    sphere = (info & 0x80000000) != 0
    v0 = prims[:, 0:3].astype(np.float64)
    p1 = v0 + prims[:, 4:7].astype(np.float64)
    p2 = v0 + prims[:, 8:11].astype(np.float64)
    lo = np.minimum(v0, np.minimum(p1, p2))
    hi = np.maximum(v0, np.maximum(p1, p2))
    r = prims[:, 4].astype(np.float64)[:, None]
    lo[sphere] = (v0 - r)[sphere]
    hi[sphere] = (v0 + r)[sphere]
    return lo, hi


def check_tree(nodes, prims):
    """The invariants of bvh_build.h on the records as the kernels read them; returns the depth."""
    n_nodes, n_prims = len(nodes), len(prims)
    assert n_nodes >= 1
    origin = nodes[:, 0:3].view(np.float32).astype(np.float64)
    meta = nodes[:, 3]
    child_base = nodes[:, 4].view(np.int32).astype(np.int64)
    prim_base = nodes[:, 5].view(np.int32).astype(np.int64)
    parent = nodes[:, 6].view(np.int32)
    scale = np.ldexp(1.0, (meta & 0xff).astype(np.int64) - 127)
    lmask, imask = (meta >> 8) & 0xff, (meta >> 16) & 0xff
    nl = np.array([bin(int(x)).count("1") for x in lmask])
    ni = np.array([bin(int(x)).count("1") for x in imask])
    # leaves first, then inner children, then empty slots
    assert np.array_equal(lmask, (1 << nl) - 1)
    assert np.array_equal(imask, ((1 << ni) - 1) << nl)
    assert (nl + ni >= 1).all() and (nl + ni <= 8).all()
    planes = nodes[:, 8:32].copy().view(np.uint16).reshape(n_nodes, 3, 2, 8)
    centre, half = half_to_float(planes[:, :, 0, :]), half_to_float(planes[:, :, 1, :])
    lo_w = origin[:, :, None] + (centre - (half - SLACK)) * scale[:, None, None]   # (node, axis, slot)
    hi_w = origin[:, :, None] + (centre + (half - SLACK)) * scale[:, None, None]
    plo, phi = true_boxes(prims)
    seen_prim = np.zeros(n_prims, dtype=np.int32)
    seen_node = np.zeros(n_nodes, dtype=np.int32)
    sub_lo = np.full((n_nodes, 3), np.inf)
    sub_hi = np.full((n_nodes, 3), -np.inf)
    # empty slots can never be hit: half extent -infinity
    for j in range(8):
        empty = ((lmask | imask) >> j) & 1 == 0
        assert np.isneginf(half[empty][:, :, j]).all()
    # leaf slots: the slot's box holds its primitive; every primitive in exactly one slot
    for j in range(8):
        has = np.nonzero((lmask >> j) & 1)[0]
        p = prim_base[has] + j
        assert (p >= 0).all() and (p < n_prims).all()
        np.add.at(seen_prim, p, 1)
        assert (lo_w[has, :, j] <= plo[p]).all() and (hi_w[has, :, j] >= phi[p]).all(), f"leaf slot {j} does not hold its primitive"
        np.minimum.at(sub_lo, has, plo[p])
        np.maximum.at(sub_hi, has, phi[p])
    assert (seen_prim == 1).all(), "a primitive is in no leaf slot or in several"
    # inner slots: child links, parents; then bottom-up: a slot's box holds everything below its child
    kids = {}
    for j in range(8):
        has = np.nonzero((imask >> j) & 1)[0]
        c = child_base[has] + j
        assert (c > has).all() and (c < n_nodes).all(), "children come after their parent"
        np.add.at(seen_node, c, 1)
        assert np.array_equal(parent[c], has.astype(np.int32))
        kids[j] = (has, c)
    assert seen_node[0] == 0 and (seen_node[1:] == 1).all() and parent[0] == -1
    depth = np.ones(n_nodes, dtype=np.int64)
    order = np.argsort(-np.arange(n_nodes))  # children have larger indices: reverse order is bottom-up
    slot_of_child = np.zeros(n_nodes, dtype=np.int64)
    for j, (has, c) in kids.items():
        slot_of_child[c] = j
    for k in order[:-1] if n_nodes > 1 else []:
        p, j = parent[k], slot_of_child[k]
        assert (lo_w[p, :, j] <= sub_lo[k]).all() and (hi_w[p, :, j] >= sub_hi[k]).all(), f"node {p} slot {j} does not hold its subtree"
        sub_lo[p] = np.minimum(sub_lo[p], sub_lo[k])
        sub_hi[p] = np.maximum(sub_hi[p], sub_hi[k])
        depth[p] = max(depth[p], depth[k] + 1)
    return int(depth[0])


def random_rays(prims, n, seed):
    rng = np.random.default_rng(seed)
    plo, phi = true_boxes(prims)
    lo, hi = plo.min(axis=0), phi.max(axis=0)
    span = np.maximum(hi - lo, 1e-3)
    org = (lo - 0.5 * span + rng.random((n, 3)) * 2.0 * span).astype(np.float32)
    tgt = (lo + rng.random((n, 3)) * span).astype(np.float32)
    d = tgt - org
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([d.astype(np.float32), org], axis=1)


@pytest.mark.parametrize("name", ["scene1", "scene4", "scene5_small", "scene6_small", "scenetest"])
def test_device_tree_invariants_and_rays(gpu, name):
    sc = host.Scene(scene_path(name))
    on_host = device.DeviceScene(sc, 0, build="host")
    on_dev = device.DeviceScene(sc, 0, build="device")
    assert on_dev.info.built_on_device == 1 and on_host.info.built_on_device == 0
    assert on_dev.info.num_bounded == on_host.info.num_bounded
    for d in (on_host, on_dev):
        nodes, prims = d.read_tree()
        assert check_tree(nodes, prims) == d.info.max_depth
    rays = random_rays(prims, 200_000, 11)
    obj_h, t_h, n_h = on_host.trace_closest(rays)
    obj_d, t_d, n_d = on_dev.trace_closest(rays)
    # the closest hit does not depend on the tree (ties between coincident primitives aside)
    same = obj_h == obj_d
    assert same.mean() > 0.9999
    np.testing.assert_array_equal(t_h[same], t_d[same])
    assert np.abs(t_h[~same] - t_d[~same]).max(initial=0.0) <= 1e-6 * max(1.0, float(np.abs(t_h).max()))
    dist = np.full(len(rays), 1e30, dtype=np.float32)
    emitter = np.full(len(rays), -1, dtype=np.int32)
    light = np.ones((len(rays), 3), dtype=np.float32)
    b_h, l_h = on_host.trace_shadow(rays, dist, emitter, light.copy())
    b_d, l_d = on_dev.trace_shadow(rays, dist, emitter, light.copy())
    np.testing.assert_array_equal(b_h, b_d)
    # kt factors are multiplied in the order the walk meets them: the product agrees to rounding
    np.testing.assert_allclose(l_h[b_h == 0], l_d[b_d == 0], rtol=2e-6, atol=0)


def test_frames_agree_with_the_host_built_tree(gpu):
    for name, w, h, flags in [("scene4", 320, 180, ()), ("scene5_small", 240, 135, ("-g", "path", "-n", "4"))]:
        sc = host.Scene(scene_path(name))
        a, za = device.DeviceScene(sc, 0, build="host").render(sc.options(w, h, *flags))
        b, zb = device.DeviceScene(sc, 0, build="device").render(sc.options(w, h, *flags))
        np.testing.assert_array_equal(za, zb)
        fp = parity.frame_parity(b, a)
        assert fp["ok"] and fp["rel_rmse_all"] < 1e-5, fp


def test_engine_flag(gpu, tmp_path):
    """engine --bvh device: the same 8-bit image as the default build (scene4 has no transparent surfaces,
    so not even the order of kt factors differs)."""
    import subprocess
    a, b = str(tmp_path / "h.tif"), str(tmp_path / "d.tif")
    common = [os.path.join(ROOT, "engine"), "scenes/scene4.json"]
    tail = ["160", "90"]
    out = subprocess.run(common + [a] + tail, check=True, cwd=ROOT, capture_output=True, text=True).stdout
    assert "built on the host" in out
    out = subprocess.run(common + [b] + tail + ["--bvh", "device"], check=True, cwd=ROOT, capture_output=True, text=True).stdout
    assert "built on the device" in out
    assert open(a, "rb").read() == open(b, "rb").read()


def test_full_dragon_builds_in_milliseconds(gpu):
    """The benchmark mesh: 871 402 triangles; the device build must beat the host build by a wide margin."""
    path = os.path.join(ROOT, "scenes", "scene5.json")
    if not os.path.exists(os.path.join(ROOT, "meshes", "dragon.stl")):
        pytest.skip("meshes/dragon.stl not generated")
    sc = host.Scene(path)
    device.DeviceScene(sc, 0, build="device").close()  # first call pays CUDA module loading
    on_dev = device.DeviceScene(sc, 0, build="device")
    on_host = device.DeviceScene(sc, 0, build="host")
    print(f"dragon: device build {on_dev.info.build_ms:.1f} ms ({on_dev.info.num_nodes} nodes, depth {on_dev.info.max_depth}); "
          f"host build {on_host.info.build_ms:.1f} ms ({on_host.info.num_nodes} nodes, depth {on_host.info.max_depth})")
    assert on_dev.info.build_ms < 0.25 * on_host.info.build_ms
    nodes, prims = on_dev.read_tree()
    assert check_tree(nodes, prims) == on_dev.info.max_depth
