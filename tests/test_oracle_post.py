"""CPU suite of the post-processing path: the oracle restatement
(oracle/crt_oracle_post.c) against golden rasters produced by the UNMODIFIED
reference postprocess() (tests/golden/post_*.npz, made by make_golden_post.py),
bit for bit; the flag parser; the raw-frame reader; and -- where
oracle/_ref/ref_post exists -- live runs of the reference on seeded frames."""
import glob
import os

import numpy as np
import pytest

from _util import GOLDEN
from c_raytracer_b200 import _abi, host
from oracle import oracle


def post_cases():
    return sorted(os.path.basename(p)[len("post_"):-4] for p in glob.glob(os.path.join(GOLDEN, "post_*.npz")))


def load_post_case(name):
    g = np.load(os.path.join(GOLDEN, f"post_{name}.npz"))
    img = np.load(os.path.join(GOLDEN, f"image_{str(g['image'])}.npz"))
    return img["raster"], img["z"], [str(f) for f in g["flags"]], g["raster"]


def seeded_frame(w, h, seed, sky=0.2):
    """A synthetic frame with the structure of a rendered one: smooth depth ramps, a few
    nearer blobs, background pixels with z = 0 (nothing hit, reference image.c:45-46)."""
    rng = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    z = (4 + 6 * ys / h + 2 * np.sin(xs / 9)).astype(np.float32)
    for _ in range(5):
        cx, cy, r = rng.uniform(0, w), rng.uniform(0, h), rng.uniform(3, max(4, w / 6))
        z = np.where((xs - cx) ** 2 + (ys - cy) ** 2 < r * r, np.float32(rng.uniform(1.5, 5)), z)
    if sky:
        z = np.where(rng.random((h, w)) < sky, np.float32(0), z)
    raster = rng.uniform(0, 1.2, size=(h, w, 3)).astype(np.float32)
    return raster, z.astype(np.float32)


@pytest.mark.parametrize("name", post_cases())
def test_oracle_matches_reference_golden(name):
    raster, z, flags, want = load_post_case(name)
    got = oracle.postprocess(host.post_options(*flags), raster, z)
    np.testing.assert_array_equal(got, want)


def test_golden_cases_exist():
    assert len(post_cases()) >= 12


def test_flag_parser():
    o = host.post_options("--mist", 3, 9, "inv-quad", .5, .25, 1, "-b", "1.5", "--dof", 2, -4, "--dof-camera", 1, 2, 3)
    assert o.ops == _abi.CRT_POST_BRIGHTEN | _abi.CRT_POST_DOF | _abi.CRT_POST_MIST  # --dof wins (postproc.c:47-52)
    assert (o.brighten, o.dof_scale, o.dof_bias) == (1.5, 2.0, -4.0)
    assert (o.mist_start, o.mist_depth, o.mist_falloff, list(o.mist_color)) == (3.0, 9.0, _abi.CRT_FALLOFF_INV_QUADRATIC, [.5, .25, 1.0])
    assert host.post_options().ops == 0
    assert host.post_options("--dof", 1).ops == 0  # a flag whose values run past argv is absent (argv.c:38-55)
    assert host.post_options("--dof-camera", 2, 0.5, 6).ops == _abi.CRT_POST_DOF_CAMERA
    with pytest.raises(host.SceneError, match="falloff"):
        host.post_options("--mist", 3, 9, "cubic", 0, 0, 0)


def test_oracle_rejects_what_the_reference_cannot_compute():
    raster, z = seeded_frame(40, 30, 1, sky=0.3)
    with pytest.raises(ValueError):  # z_min = 0 -> division by zero in postproc.c:63-64
        oracle.postprocess(host.post_options("--dof-camera", 9, 0.6, 5.5), raster, z)
    with pytest.raises(ValueError):  # radius^2 overflows the reference's int
        oracle.postprocess(host.post_options("--dof", 1e6, 0), raster, np.maximum(z, 1))


def test_identities():
    raster, z = seeded_frame(61, 35, 2)
    # radius 0 everywhere: alpha = min(1/0, 1) = 1, every pixel gathers only itself
    np.testing.assert_array_equal(oracle.postprocess(host.post_options("--dof", 0, 0), raster, z), raster)
    # mist that starts behind everything has opacity 0
    np.testing.assert_array_equal(oracle.postprocess(host.post_options("--mist", 1e6, 1, "lin", 1, 1, 1), raster, z), raster)
    np.testing.assert_array_equal(oracle.postprocess(host.post_options("-b", 1), raster, z), raster)
    st = _abi.PostStats()
    out = oracle.postprocess(host.post_options("--dof", 0.5, -2), np.full_like(raster, 0.25), z, st)
    assert st.dof_max_radius >= 1
    np.testing.assert_allclose(out, 0.25, rtol=1e-5)  # a flat frame stays flat (weights are normalised)


def test_raw_frame_round_trip(tmp_path):
    raster, z = seeded_frame(33, 17, 3)
    p = str(tmp_path / "f.tif")
    host.save_tiff(p, raster, z, raw=True)
    r2, z2 = host.load_raw(p)
    np.testing.assert_array_equal(r2, raster)
    np.testing.assert_array_equal(z2, z)
    host.save_tiff(p, raster, z, raw=False)
    with pytest.raises(host.SceneError):
        host.load_raw(p)


LIVE = pytest.mark.skipif(not os.path.exists(oracle.REF_POST), reason="oracle/_ref/ref_post not built (needs /root/reference)")


@LIVE
@pytest.mark.parametrize("w,h,seed,flags", [
    (97, 53, 5, ("--dof", "0.8", "-4")),
    (97, 53, 6, ("-b", "1.3", "--dof", "1.4", "-9", "--mist", "3", "5", "quad", "0.2", "0.3", "0.4")),
    (128, 31, 7, ("--dof", "-0.7", "3.1")),
    (64, 64, 8, ("--mist", "2", "7", "inv-quad", "0.9", "0.8", "0.7")),
])
def test_live_reference(tmp_path, w, h, seed, flags):
    raster, z = seeded_frame(w, h, seed)
    p = str(tmp_path / "in.tif")
    host.save_tiff(p, raster, z, raw=True)
    want = oracle.ref_postprocess(p, flags)
    np.testing.assert_array_equal(oracle.postprocess(host.post_options(*flags), raster, z), want)


@LIVE
def test_live_reference_dof_camera(tmp_path):
    raster, z = seeded_frame(80, 45, 9, sky=0)
    p = str(tmp_path / "in.tif")
    host.save_tiff(p, raster, z, raw=True)
    flags = ("--dof-camera", "12", "0.5", "4.5")
    st = _abi.PostStats()
    got = oracle.postprocess(host.post_options(*flags), raster, z, st)
    assert st.dof_max_radius > 0
    np.testing.assert_array_equal(got, oracle.ref_postprocess(p, flags))
