"""The `engine` executable end to end on the GPU box: same command line as the
reference engine (reference src/raytracer/main.c:23-89), C host -> C ABI -> CUDA,
raw `-f` TIFF (FP32 raster + z-buffer in tag 65000, reference image.c:64-85) and
8-bit TIFF (truncating quantiser, image.c:96-98) equal to what the library call
returns for the same options."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from _util import ROOT
from c_raytracer_b200 import _abi, device, host

pytestmark = pytest.mark.gpu


class _Img(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("bits", C.c_uint32), ("rgb8", C.c_void_p),
                ("raster", C.POINTER(C.c_float)), ("z", C.POINTER(C.c_float)), ("z_count", C.c_uint32)]


def _run_engine(args):
    exe = os.path.join(ROOT, "engine")
    assert os.path.exists(exe), "engine not built: run __graft_entry__.build()"
    out = subprocess.run([exe] + args, capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 0, out.stdout + out.stderr
    return out.stdout


def test_engine_raw_output_equals_library_frame(tmp_path):
    if device.device_count() < 1:
        pytest.fail("no CUDA device")
    w, h = 80, 45
    flags = ["-g", "path", "-n", "3", "-b", "4", "--seed", "5"]
    p = str(tmp_path / "out.tif")
    log = _run_engine(["scenes/scene3.json", p, str(w), str(h), "-f", "--stats"] + flags)
    assert "Commencing raytracing." in log and "Saving image." in log  # the reference's phase markers
    L = host.lib()
    img = _Img()
    err = C.create_string_buffer(512)
    assert L.rt_tiff_read(p.encode(), C.byref(img), err) == 0, err.value
    assert (img.width, img.height, img.bits, img.z_count) == (w, h, 32, w * h)
    raster = np.ctypeslib.as_array(img.raster, shape=(h, w, 3)).copy()
    z = np.ctypeslib.as_array(img.z, shape=(h, w)).copy()
    L.rt_tiff_free(C.byref(img))
    sc = host.Scene(os.path.join(ROOT, "scenes", "scene3.json"))
    want, want_z = device.DeviceScene(sc, 0).render(sc.options(w, h, *flags))
    np.testing.assert_array_equal(raster, want)
    np.testing.assert_array_equal(z, want_z)


def test_engine_8bit_output(tmp_path):
    Image = pytest.importorskip("PIL.Image")
    w, h = 64, 36
    p = str(tmp_path / "out.tif")
    _run_engine(["scenes/scene1.json", p, str(w), str(h), "-m", "max"])  # -m is accepted and ignored
    sc = host.Scene(os.path.join(ROOT, "scenes", "scene1.json"))
    want, _ = device.DeviceScene(sc, 0).render(sc.options(w, h))
    np.testing.assert_array_equal(np.array(Image.open(p)), host.quantize(want))


def test_engine_reports_missing_device_flag(tmp_path):
    exe = os.path.join(ROOT, "engine")
    out = subprocess.run([exe, "scenes/scene1.json", str(tmp_path / "o.tif"), "8", "8", "--device", "99"],
                         capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 1 and "device" in out.stdout.lower()


def test_engine_stats_json(tmp_path):
    """--stats-json: the counters of the frame as JSON; they equal the library's own for the same options."""
    import json
    if device.device_count() < 1:
        pytest.fail("no CUDA device")
    w, h = 96, 54
    flags = ["-g", "path", "-n", "2", "--seed", "3"]
    p, j = str(tmp_path / "out.tif"), str(tmp_path / "stats.json")
    log = _run_engine(["scenes/scene4.json", p, str(w), str(h), "--stats-json", j] + flags)
    assert "Statistics written" in log and "Rays:" in log  # --stats-json implies the counted kernel instance
    d = json.load(open(j))
    sc = host.Scene(os.path.join(ROOT, "scenes", "scene4.json"))
    o = sc.options(w, h, *flags)
    o.flags |= _abi.CRT_RENDER_COUNT_STATS
    st = _abi.Stats()
    dev = device.DeviceScene(sc, 0)
    dev.render(o, st)
    assert (d["width"], d["height"], d["samples_per_pixel"], d["seed"]) == (w, h, 2, 3)
    assert (d["rays_closest"], d["rays_shadow"], d["box_tests"], d["plane_tests"]) == (st.rays_closest, st.rays_shadow, st.box_tests, st.plane_tests)
    assert d["tri_tests"] == list(st.tri_tests) and d["sphere_tests"] == list(st.sphere_tests)
    assert d["flops"] == pytest.approx(_abi.algorithmic_flops(st), rel=1e-12)
    rays = st.rays_closest + st.rays_shadow
    assert d["mrays_per_s"] == pytest.approx(rays / (d["kernel_ms"] * 1e3), rel=1e-5)
    assert d["box_tests_per_ray"] == pytest.approx(st.box_tests / rays, rel=1e-5)
    r = d["roofline"]
    assert r["bound"] == "fp32" and 30.0 < r["peak"] < 90.0 and 0.0 < r["frac"] < 1.0
    assert r["frac"] == pytest.approx(r["achieved"] / r["peak"], abs=2e-6)
    assert d["bvh"]["nodes"] == dev.info.num_nodes and d["bvh"]["primitives"] == dev.info.num_bounded and d["bvh"]["built_on"] == "host"
    assert d["devices"] == [{"device": 0, "kernel_ms": d["kernel_ms"]}] and d["work_queue"] == "single device"
