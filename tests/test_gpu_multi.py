"""Several devices, one frame: crt_multi_* (include/crt.h) -- the scene replicated, ONE work queue that
every resident warp of every device drains with system-scope atomics, pixels written straight into the
common frame (the reference runs the rows on OpenMP threads inside render(), render.c:349-352).

The frame must not depend on who renders which unit: every configuration here is compared bit for
bit with the single-device frame.  A one-GPU box exercises the same code with two replicas on device
0 (both persistent kernels pull from the one counter); with two or more GPUs the replicas sit on
different devices and reach frame and counter by peer access over NVLink."""
import os
import subprocess

import numpy as np
import pytest

from _util import ROOT, scene_path
from c_raytracer_b200 import _abi, device, host

pytestmark = pytest.mark.gpu


def single_frame(sc, opts):
    dev = device.DeviceScene(sc, 0)
    st = _abi.Stats()
    raster, z = dev.render(opts, st)
    dev.close()
    return raster, z, st


CASES = [("scene3", 333, 187, ()), ("scene5_small", 160, 90, ("-g", "path", "-n", "3")), ("scene4", 96, 54, ("-g", "path", "-n", "2"))]


@pytest.mark.parametrize("scene,w,h,flags", CASES)
@pytest.mark.parametrize("static", [False, True])
def test_two_replicas_on_one_device(scene, w, h, flags, static, monkeypatch):
    if device.device_count() < 1:
        pytest.fail("no CUDA device: the trace loop has no CPU fallback")
    sc = host.Scene(scene_path(scene))
    opts = sc.options(w, h, *flags)
    opts.flags |= _abi.CRT_RENDER_COUNT_STATS
    want, want_z, want_st = single_frame(sc, opts)
    if static:
        monkeypatch.setenv("CRT_MULTI_STATIC", "1")
    multi = device.MultiScene(sc, [0, 0, 0] if static else [0, 0])
    assert multi.dynamic == (not static)
    st = _abi.Stats()
    for _ in range(2):  # twice: the work counter and the scratch are reused
        raster, z, ms = multi.render(opts, st)
        np.testing.assert_array_equal(raster, want)
        np.testing.assert_array_equal(z, want_z)
    assert (st.rays_closest, st.rays_shadow) == (want_st.rays_closest, want_st.rays_shadow)
    assert len(ms) == len(multi.devices) and all(m > 0 for m in ms)
    multi.close()


def test_all_devices_share_one_frame():
    n = device.device_count()
    if n < 2:
        pytest.skip("needs two CUDA devices")
    sc = host.Scene(scene_path("scene5_small"))
    opts = sc.options(640, 360, "-g", "path", "-n", "4")
    want, want_z, _ = single_frame(sc, opts)
    multi = device.MultiScene(sc, list(range(n)))
    assert multi.dynamic, "no peer access between the devices of this box"
    raster, z, ms = multi.render(opts)
    np.testing.assert_array_equal(raster, want)
    np.testing.assert_array_equal(z, want_z)
    # every device took part
    assert all(m > 0 for m in ms)
    multi.close()


def test_multi_rejects_bad_arguments():
    sc = host.Scene(scene_path("scene1"))
    with pytest.raises(device.CrtError):
        device.MultiScene(sc, [])
    with pytest.raises(device.CrtError):
        device.MultiScene(sc, [0, 9999])
    multi = device.MultiScene(sc, [0, 0])
    bad = sc.options(32, 32)
    bad.shard_index, bad.shard_count = 0, 2
    with pytest.raises(device.CrtError, match="shard"):
        multi.render(bad)
    multi.close()


def test_engine_gpus_flag_writes_the_same_raw_frame(tmp_path):
    """engine ... --devices 0,0 -f == engine ... -f, byte for byte (the C host drives crt_multi_render)."""
    a, b = str(tmp_path / "one.tif"), str(tmp_path / "two.tif")
    common = [os.path.join(ROOT, "engine"), "scenes/scene3.json"]
    tail = ["200", "112", "-f", "-g", "path", "-n", "2"]
    subprocess.run(common + [a] + tail, check=True, cwd=ROOT, capture_output=True)
    out = subprocess.run(common + [b] + tail + ["--devices", "0,0"], check=True, cwd=ROOT, capture_output=True, text=True).stdout
    assert "Using 2 CUDA devices." in out and "shared work queue" in out
    assert open(a, "rb").read() == open(b, "rb").read()
    n = device.device_count()
    if n >= 2:
        c = str(tmp_path / "all.tif")
        out = subprocess.run(common + [c] + tail + ["--gpus", "max"], check=True, cwd=ROOT, capture_output=True, text=True).stdout
        assert f"Using {n} CUDA devices." in out
        assert open(a, "rb").read() == open(c, "rb").read()
