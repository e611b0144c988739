"""In-tracer thin lens and fog (crt_render_opts.lens_* / fog_*).  EXTENSIONS: the
reference has no such code (its depth of field and mist are post-processes, postproc.c:47-188), so
parity is UNPINNED here -- these tests hold the properties the definitions in include/crt.h imply:
off means the reference's frame bit for bit, fog converges to its colour, the plane in focus is the
sharp one."""
import numpy as np
import pytest

from _util import scene_path
import _pkg

_pkg.load()
from c_raytracer_b200 import device, host  # noqa: E402

pytestmark = pytest.mark.gpu

W, H = 320, 180


@pytest.fixture(scope="module")
def scene():
    if device.device_count() < 1:
        pytest.skip("no CUDA device")
    sc = host.Scene(scene_path("scene1"))
    return sc, device.DeviceScene(sc, 0)


def test_off_is_the_reference_frame(scene):
    sc, dev = scene
    base, z = dev.render(sc.options(W, H))
    o = sc.options(W, H)
    o.lens_samples, o.focus_distance = 8, 3.0          # no radius: still a pinhole
    o.fog_color[0], o.fog_color[1], o.fog_color[2] = 0.5, 0.6, 0.7  # no density: clear air
    got, gz = dev.render(o)
    np.testing.assert_array_equal(got, base)
    np.testing.assert_array_equal(gz, z)


def test_fog_converges_to_its_colour(scene):
    sc, dev = scene
    base, z = dev.render(sc.options(W, H))
    fog = np.array([0.5, 0.6, 0.7], dtype=np.float32)
    dist = []
    for density in (0.02, 0.1, 0.5, 50.0):
        o = sc.options(W, H)
        o.fog_density = density
        for k in range(3):
            o.fog_color[k] = float(fog[k])
        got, gz = dev.render(o)
        np.testing.assert_array_equal(gz, z)  # the z-buffer is geometry
        # a primary ray that leaves the scene carries exactly the fog colour
        miss = z == 0
        if miss.any():
            np.testing.assert_array_equal(got[miss], np.broadcast_to(fog, got[miss].shape))
        dist.append(float(np.abs(got - fog).mean()))
    assert dist[0] > dist[1] > dist[2] > dist[3]
    assert dist[3] < 1e-3


def gradient(frame):
    lum = frame.astype(np.float64).sum(axis=2)
    gy, gx = np.gradient(lum)
    return np.hypot(gx, gy)


def test_the_plane_in_focus_is_the_sharp_one(scene):
    """Edges of the pinhole frame in a near and in a far depth band: focusing on a band keeps more of
    its edge contrast than focusing on the other one does."""
    sc, dev = scene
    # -b 2: the reference does not renormalise a reflected direction (render.c:313-314), and the error of a
    # chain of mirror bounces between the spheres grows by |t| / radius per bounce; one lens path in millions
    # reaches a specular term of 1e16 after six bounces.  That is the reference's arithmetic, not the lens.
    flags = ("-b", "2")
    base, z = dev.render(sc.options(W, H, *flags))
    hit = z > 0
    zs = np.sort(z[hit])
    z_near, z_far = float(zs[len(zs) // 10]), float(zs[-len(zs) // 10])
    assert z_far > 1.5 * z_near
    g0 = gradient(base)
    bands = {}
    for name, zc in (("near", z_near), ("far", z_far)):
        band = hit & (np.abs(z - zc) < 0.2 * zc)
        bands[name] = band & (g0 >= np.quantile(g0[band], 0.9))  # the band's strongest edges
        assert bands[name].sum() > 50
    kept = {}
    for name, focus in (("near", z_near), ("far", z_far)):
        o = sc.options(W, H, *flags)
        o.lens_radius, o.focus_distance, o.lens_samples = 0.3, focus, 64
        frame, gz = dev.render(o)
        np.testing.assert_array_equal(gz, z)  # lens sample 0 is the pinhole ray
        again, _ = dev.render(o)
        np.testing.assert_array_equal(again, frame)  # deterministic
        assert abs(float(frame.mean()) / float(base.mean()) - 1.0) < 0.1
        g = gradient(frame)
        kept[name] = {b: float(g[m].mean() / g0[m].mean()) for b, m in bands.items()}
    print("edge contrast kept (focus -> band):", kept)
    assert kept["near"]["near"] > 1.15 * kept["far"]["near"]
    assert kept["far"]["far"] > 1.15 * kept["near"]["far"]
    # out of focus is blurred: well below the pinhole frame's contrast
    assert kept["far"]["near"] < 0.8 and kept["near"]["far"] < 0.8
