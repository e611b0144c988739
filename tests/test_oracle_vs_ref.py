"""Live check of the oracle against the compiled reference (oracle/_ref/ref_engine).
Runs only where that binary exists and the current machine can execute it."""
import numpy as np
import pytest

from _util import rel_rmse, scene_path
from c_raytracer_b200 import _abi, host
from oracle import oracle

pytestmark = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/ref_engine not built (needs /root/reference)")


@pytest.mark.parametrize("scene,w,h,flags", [
    ("scene1", 80, 45, ()),
    ("scene3", 80, 45, ("-g", "path", "-n", "2", "-s", "blinn")),
    ("scenetest2", 64, 36, ("-l", "lin", "--seed", "3")),
    ("scene5_small", 64, 36, ("-g", "path", "-n", "2")),
])
def test_live_image(scene, w, h, flags):
    seed = int(flags[flags.index("--seed") + 1]) if "--seed" in flags else 1
    ref, ref_z, cnt = oracle.ref_render(scene_path(scene), w, h, flags, seed=seed)
    sc = host.Scene(scene_path(scene))
    st = _abi.Stats()
    raster, z = oracle.OracleScene(sc).render(sc.options(w, h, *flags), st)
    assert (st.rays_closest, st.rays_shadow) == (cnt["cast_ray"], cnt["shadow"])
    np.testing.assert_array_equal(z, ref_z)
    assert rel_rmse(raster, ref) <= 1e-6


def test_seed_changes_the_reference_image():
    a, _, _ = oracle.ref_render(scene_path("scene1"), 48, 27, (), seed=1)
    b, _, _ = oracle.ref_render(scene_path("scene1"), 48, 27, (), seed=2)
    assert 1e-5 < rel_rmse(a, b) < 0.05  # area-light noise floor


def test_reference_own_render_loop_matches_driver_loop(tmp_path):
    """The driver's pixel loop (oracle/ref_engine.c) restates render.c:352-366; with an
    emitter that casts no shadow samples the image is deterministic and the
    reference's own render() must produce the same bytes."""
    import json
    import os
    import subprocess
    s = json.load(open(scene_path("scene1")))
    for o in s["Objects"]:
        o["parameters"].pop("lights", None)
    p = tmp_path / "nolights.json"
    p.write_text(json.dumps(s))
    outs = []
    for extra in ([], ["--stock-loop"]):
        out = str(tmp_path / ("a.bin" if not extra else "b.bin"))
        subprocess.run([oracle.REF_ENGINE, "render", str(p), out, "64", "36", "-m", "1"] + extra, check=True,
                       cwd=os.path.dirname(oracle.HERE), stdout=subprocess.DEVNULL)
        outs.append(oracle.read_dump(out))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])
    assert outs[0][0].max() > 0
