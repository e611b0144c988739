"""Parity of the CUDA post-processing path (crt_postprocess[_device] of
include/crt.h, c-raytracer_b200/csrc/crt_post.cu) with (a) the golden rasters
of the UNMODIFIED reference postprocess() (tests/golden/post_*.npz) and (b) the
CPU oracle on seeded frames, plus properties at BASELINE.json's frame sizes.

The bar is BIT-EXACT: brighten and mist are single-rounded maps, and the
depth-of-field gather adds every pixel's contributions in the order of the
reference's serial scatter, so no tolerance is needed."""
import os
import subprocess

import numpy as np
import pytest

from _util import ROOT
from c_raytracer_b200 import _abi, device, host
from oracle import oracle
from test_oracle_post import load_post_case, post_cases, seeded_frame

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    if device.device_count() < 1:
        pytest.fail("no CUDA device: the post-processing path has no CPU fallback")
    return 0


@pytest.mark.parametrize("name", post_cases())
def test_matches_reference_golden(gpu, name):
    raster, z, flags, want = load_post_case(name)
    st = _abi.PostStats()
    got = device.postprocess(host.post_options(*flags), raster, z, gpu, st)
    np.testing.assert_array_equal(got, want)
    assert st.launches >= 1 and st.kernel_ms > 0


@pytest.mark.parametrize("w,h,seed,flags", [
    (333, 187, 11, ("--dof", "0.9", "-5")),
    (333, 187, 12, ("-b", "1.3", "--dof", "2.2", "-13", "--mist", "3", "5", "quad", "0.2", "0.3", "0.4")),
    (640, 47, 13, ("--dof", "-0.7", "3.1", "--mist", "0", "11", "lin", "0.5", "0.5", "0.5")),
    (31, 257, 14, ("--dof", "1.5", "-8")),
    (1, 1, 15, ("--dof", "3", "1", "-b", "2")),
    (1027, 3, 16, ("--mist", "2", "7", "inv-quad", "0.9", "0.8", "0.7", "-b", "0.5")),
    (2, 2, 17, ("-b", "3")),
])
def test_matches_oracle_on_seeded_frames(gpu, w, h, seed, flags):
    raster, z = seeded_frame(w, h, seed)
    o = host.post_options(*flags)
    st, ost = _abi.PostStats(), _abi.PostStats()
    got = device.postprocess(o, raster, z, gpu, st)
    want = oracle.postprocess(o, raster, z, ost)
    np.testing.assert_array_equal(got, want)
    assert st.dof_max_radius == ost.dof_max_radius
    assert (st.dof_scale, st.dof_bias) == (ost.dof_scale, ost.dof_bias)


def test_dof_camera(gpu):
    raster, z = seeded_frame(200, 120, 21, sky=0)
    o = host.post_options("--dof-camera", "12", "0.5", "4.5")
    st, ost = _abi.PostStats(), _abi.PostStats()
    got = device.postprocess(o, raster, z, gpu, st)
    np.testing.assert_array_equal(got, oracle.postprocess(o, raster, z, ost))
    assert (st.dof_scale, st.dof_bias, st.dof_max_radius) == (ost.dof_scale, ost.dof_bias, ost.dof_max_radius)


def test_errors(gpu):
    raster, z = seeded_frame(40, 30, 1, sky=0.3)
    with pytest.raises(device.CrtError, match="dof-camera"):  # z_min = 0
        device.postprocess(host.post_options("--dof-camera", 9, 0.6, 5.5), raster, z)
    with pytest.raises(device.CrtError, match="circle of confusion"):
        device.postprocess(host.post_options("--dof", 1e6, 0), raster, np.maximum(z, 1))
    bad = host.post_options("-b", 2)
    bad.ops = 64
    with pytest.raises(device.CrtError, match="unknown"):
        device.postprocess(bad, raster, z)
    bad = host.post_options("--mist", 1, 2, "lin", 0, 0, 0)
    bad.mist_falloff = 7
    with pytest.raises(device.CrtError, match="falloff"):
        device.postprocess(bad, raster, z)
    with pytest.raises(device.CrtError):
        device.postprocess(host.post_options("-b", 2), raster, z, device=99)
    # no operation selected: the frame comes back untouched
    np.testing.assert_array_equal(device.postprocess(host.post_options(), raster, z), raster)


@pytest.mark.parametrize("w,h", [(1920, 1080), (3840, 2160)])
def test_full_size_properties(gpu, w, h):
    """BASELINE.json's frame sizes, through properties that need no CPU reference run."""
    raster, z = seeded_frame(w, h, 31)
    # identities (see tests/test_oracle_post.py::test_identities)
    np.testing.assert_array_equal(device.postprocess(host.post_options("--dof", 0, 0), raster, z), raster)
    np.testing.assert_array_equal(device.postprocess(host.post_options("--mist", 1e6, 1, "lin", 1, 1, 1), raster, z), raster)
    # brighten by a power of two is exact and invertible
    up = device.postprocess(host.post_options("-b", 4), raster, z)
    np.testing.assert_array_equal(device.postprocess(host.post_options("-b", 0.25), up, z), raster)
    # depth of field: deterministic, a convex combination of the frame, and flat frames stay flat
    o = host.post_options("--dof", 0.6, -3)
    st = _abi.PostStats()
    a = device.postprocess(o, raster, z, gpu, st)
    np.testing.assert_array_equal(device.postprocess(o, raster, z), a)
    assert 2 <= st.dof_max_radius <= 8
    assert a.min() >= raster.min() - 1e-6 and a.max() <= raster.max() + 1e-6
    flat = device.postprocess(o, np.full_like(raster, 0.375), z)
    np.testing.assert_allclose(flat, 0.375, rtol=2e-6)
    # the pointwise maps against the oracle at full size (the oracle does these in a second)
    o = host.post_options("-b", 1.7, "--mist", 3, 6, "quad", 0.3, 0.4, 0.5)
    np.testing.assert_array_equal(device.postprocess(o, raster, z), oracle.postprocess(o, raster, z))


def test_full_hd_depth_of_field_equals_oracle(gpu):
    raster, z = seeded_frame(1920, 1080, 41)
    o = host.post_options("--dof", 1.6, -8, "--mist", 5, 8, "lin", 0.6, 0.6, 0.7)
    st = _abi.PostStats()
    got = device.postprocess(o, raster, z, gpu, st)
    assert st.dof_max_radius >= 4
    np.testing.assert_array_equal(got, oracle.postprocess(o, raster, z))


def test_device_pointer_entry_point_after_render(gpu):
    """crt_render_device -> crt_postprocess_device on the same device buffers and stream:
    the frame never leaves the GPU between the trace loop and the post-processor."""
    torch = pytest.importorskip("torch")
    from _util import scene_path
    sc = host.Scene(scene_path("scene3"))
    dev = device.DeviceScene(sc, gpu)
    w, h = 160, 90
    opts = sc.options(w, h)
    raster = torch.empty((h, w, 3), dtype=torch.float32, device="cuda:0")
    z = torch.empty((h, w), dtype=torch.float32, device="cuda:0")
    stream = torch.cuda.Stream()
    dev.render_device(opts, raster.data_ptr(), z.data_ptr(), stream.cuda_stream)
    dev.sync()
    frame, depth = raster.cpu().numpy(), z.cpu().numpy()
    o = host.post_options("-b", 1.5, "--dof-camera", 9, 0.6, 5.5, "--mist", 2, 12, "quad", 0.1, 0.1, 0.2)
    st = _abi.PostStats()
    device.postprocess_device(o, w, h, raster.data_ptr(), z.data_ptr(), gpu, stream.cuda_stream, st)
    stream.synchronize()
    np.testing.assert_array_equal(raster.cpu().numpy(), oracle.postprocess(o, frame, depth))
    np.testing.assert_array_equal(z.cpu().numpy(), depth)  # the z-buffer is read only
    # an unaligned view takes the scalar kernel
    base = torch.zeros(w * h * 3 + 1, dtype=torch.float32, device="cuda:0")
    view = base[1:]
    view.copy_(torch.from_numpy(frame).reshape(-1))
    device.postprocess_device(host.post_options("-b", 3), w, h, view.data_ptr(), z.data_ptr(), gpu)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(view.cpu().numpy().reshape(h, w, 3), oracle.postprocess(host.post_options("-b", 3), frame, depth))


def test_postprocess_executable(gpu, tmp_path):
    """engine -f | postprocess: the reference's two-step command line (README), on the GPU."""
    Image = pytest.importorskip("PIL.Image")
    for exe in ("engine", "postprocess"):
        assert os.path.exists(os.path.join(ROOT, exe)), f"{exe} not built: run __graft_entry__.build()"
    raw, out = str(tmp_path / "raw.tif"), str(tmp_path / "out.tif")
    subprocess.run([os.path.join(ROOT, "engine"), "scenes/scene3.json", raw, "96", "54", "-f"], check=True, cwd=ROOT, capture_output=True)
    flags = ["--dof", "0.9", "-6", "--mist", "2", "12", "lin", "0.3", "0.3", "0.4", "-b", "2"]
    r = subprocess.run([os.path.join(ROOT, "postprocess"), raw, out] + flags, cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    for marker in ("Loading image.", "Commencing Postprocessing", "Applying depth of field", "Saving image.", "Terminating."):
        assert marker in r.stdout
    frame, depth = host.load_raw(raw)
    want = oracle.postprocess(host.post_options(*flags), frame, depth)
    np.testing.assert_array_equal(np.array(Image.open(out)), host.quantize(want))
    # -f keeps the processed frame raw, ready for another pass
    r = subprocess.run([os.path.join(ROOT, "postprocess"), raw, out, "-f"] + flags, cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0
    got, got_z = host.load_raw(out)
    np.testing.assert_array_equal(got, want)
    np.testing.assert_array_equal(got_z, depth)
    # errors end like the reference's error(): a log line and exit status 1
    r = subprocess.run([os.path.join(ROOT, "postprocess"), raw, out, "--mist", "1", "2", "cubic", "0", "0", "0"], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 1 and "falloff" in r.stdout
    r = subprocess.run([os.path.join(ROOT, "postprocess"), str(tmp_path / "missing.tif"), out], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 1
