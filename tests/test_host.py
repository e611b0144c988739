"""Host C code (src/core, src/raytracer, lib/json): JSON reader, scene loader
against the reference's loader (fixtures in tests/golden/scenes.json, produced by
the compiled reference), error behaviour, image plane, quantiser, TIFF files."""
import ctypes as C
import hashlib
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from _util import GOLDEN, ROOT, scene_path
from c_raytracer_b200 import host


# ---- lib/json ---------------------------------------------------------------

def _json():
    L = host.lib()
    L.json_parse.restype = C.c_void_p
    L.json_parse.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    L.json_free.argtypes = [C.c_void_p]
    L.json_get.restype = C.c_void_p
    L.json_get.argtypes = [C.c_void_p, C.c_char_p]
    L.json_at.restype = C.c_void_p
    L.json_at.argtypes = [C.c_void_p, C.c_size_t]
    L.json_size.restype = C.c_size_t
    L.json_size.argtypes = [C.c_void_p]
    L.json_number.restype = C.c_double
    L.json_number.argtypes = [C.c_void_p]
    L.json_int.argtypes = [C.c_void_p]
    L.json_string.restype = C.c_char_p
    L.json_string.argtypes = [C.c_void_p]
    L.json_typeof.argtypes = [C.c_void_p]
    return L


def _parse(text):
    L = _json()
    err = C.create_string_buffer(128)
    v = L.json_parse(text.encode(), err, 128)
    return v, err.value.decode()


def test_json_values():
    L = _json()
    v, err = _parse('{"a": [1, -2.5e1, 3e400], "s": "x\\n\\u00e9\\ud83d\\ude00", "o": {"k": true, "n": null}, "a2": [], "dup": 1, "dup": 2}')
    assert v, err
    a = L.json_get(v, b"a")
    assert L.json_size(a) == 3
    assert L.json_number(L.json_at(a, 0)) == 1.0
    assert L.json_number(L.json_at(a, 1)) == -25.0
    assert L.json_number(L.json_at(a, 2)) == float("inf")
    assert L.json_int(L.json_at(a, 2)) == 2 ** 31 - 1  # saturating int view, like cJSON's valueint
    assert L.json_string(L.json_get(v, b"s")).decode() == "x\né\U0001F600"
    assert L.json_size(v) == 6  # members counted like cJSON_GetArraySize on an object
    assert L.json_number(L.json_get(v, b"dup")) == 1.0  # first match wins
    assert L.json_get(v, b"A") is None  # case-sensitive
    assert L.json_size(L.json_get(v, b"a2")) == 0
    L.json_free(v)


@pytest.mark.parametrize("text", ['', '{', '{"a":}', '[1,]', '{"a" 1}', '"unterminated', '[1] x', '{"a":01}', 'nul', '"\\q"', '[1 2]', '"\\ud800"'])
def test_json_rejects_malformed(text):
    v, err = _parse(text)
    assert not v
    assert "at byte" in err


# ---- scene loader -----------------------------------------------------------------

FIXTURES = json.load(open(os.path.join(GOLDEN, "scenes.json")))


@pytest.mark.parametrize("key", sorted(FIXTURES))
def test_scene_loader_matches_reference(key):
    """Objects (scene order, meshes expanded, resolved epsilons, -r scaling) and
    materials are byte-identical to what the reference's scene_load() builds."""
    name, extra = key.split(":")
    rescale = extra.split()[1] if extra else None
    fx = FIXTURES[key]
    sc = host.Scene(scene_path(name), rescale=rescale)
    assert len(sc.objects) == fx["num_objects"]
    assert len(sc.materials) == fx["num_materials"]
    assert hashlib.sha256(sc.materials.tobytes()).hexdigest() == fx["materials_sha256"]
    assert hashlib.sha256(sc.objects.tobytes()).hexdigest() == fx["objects_sha256"]
    assert list(sc.desc.ambient) == fx["ambient"]


def _write(tmp_path, obj):
    p = tmp_path / "scene.json"
    p.write_text(json.dumps(obj) if not isinstance(obj, str) else obj)
    return str(p)


def _base_scene():
    return json.load(open(scene_path("scene1")))


def test_scene_errors(tmp_path):
    with pytest.raises(host.SceneError, match="Unable to open scene file"):
        host.Scene(str(tmp_path / "missing.json"))
    with pytest.raises(host.SceneError, match="Failed to parse scene"):
        host.Scene(_write(tmp_path, "{ not json"))
    with pytest.raises(host.SceneError, match="Expected parent token of type Object"):
        host.Scene(_write(tmp_path, "[1, 2]"))
    s = _base_scene()
    del s["Materials"]
    with pytest.raises(host.SceneError, match=r"Expected token \[Materials\] of type \[Array\]"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Camera"]["extra"] = 1
    with pytest.raises(host.SceneError, match=r"Expected token \[Camera\] to contain 5 elements"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Camera"]["fov"] = 180
    with pytest.raises(host.SceneError, match="Expected camera fov"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Camera"]["position"] = [0, 1]
    with pytest.raises(host.SceneError, match=r"Expected token \[position\] of length \[3\]"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Objects"][0]["parameters"]["material"] = 12345
    with pytest.raises(host.SceneError, match=r"Failed to get material id \[12345\]"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Materials"][0]["texture"]["type"] = "marble"
    with pytest.raises(host.SceneError, match=r"Unrecognized token \[marble\] in texture"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    for m in s["Materials"]:
        m["ke"] = [0, 0, 0]
    with pytest.raises(host.SceneError, match="Expected non-zero number of emittant objects"):
        host.Scene(_write(tmp_path, s))
    s = _base_scene()
    s["Objects"].append({"type": "Mesh", "parameters": {"material": s["Materials"][0]["id"], "filename": "meshes/nope.stl",
                                                       "position": [0, 0, 0], "rotation": [0, 0, 0], "scale": 1}})
    with pytest.raises(host.SceneError, match="Failed to open mesh file meshes/nope.stl"):
        host.Scene(_write(tmp_path, s))


def test_ascii_stl_rejected(tmp_path):
    stl = tmp_path / "a.stl"
    stl.write_bytes(b"solid ascii" + b" " * 100)
    s = _base_scene()
    s["Objects"].append({"type": "Mesh", "parameters": {"material": s["Materials"][0]["id"], "filename": str(stl),
                                                       "position": [0, 0, 0], "rotation": [0, 0, 0], "scale": 1}})
    with pytest.raises(host.SceneError, match="does not use binary encoding"):
        host.Scene(_write(tmp_path, s))


def test_stl_triangle_count_is_held_against_the_file(tmp_path):
    """A mesh header may claim any triangle count: truncated files and counts beyond the file are refused before
    memory is set aside for them (a 134-byte file claiming 2^32 - 1 triangles once asked for 256 GB)."""
    import struct
    one = struct.pack("<12fH", 0, 0, 1, 0, 0, 0, 1, 0, 0, 0, 1, 0, 0)
    s = _base_scene()
    stl = tmp_path / "m.stl"
    s["Objects"].append({"type": "Mesh", "parameters": {"material": s["Materials"][0]["id"], "filename": str(stl),
                                                       "position": [0, 0, 0], "rotation": [0, 0, 0], "scale": 1}})
    for count, body, ok in ((1, one, True), (2, one + one, True), (2, one, False), (0xffffffff, one, False), (0x80000000, one + one, False),
                            (1, one[:-1], False), (0, b"", True)):
        stl.write_bytes(b"\0" * 80 + struct.pack("<I", count) + body)
        if ok:
            sc = host.Scene(_write(tmp_path, s))
            assert (sc.objects["type"] == 1).sum() == count
        else:
            with pytest.raises(host.SceneError, match="Failed to read triangle in mesh file"):
                host.Scene(_write(tmp_path, s))
    stl.write_bytes(b"\0" * 83)
    with pytest.raises(host.SceneError, match="Failed to read header of mesh file"):
        host.Scene(_write(tmp_path, s))


def test_emitter_without_lights_and_defaults(tmp_path):
    """`lights` absent -> 0 samples; `epsilon` absent -> automatic rules."""
    s = _base_scene()
    for o in s["Objects"]:
        o["parameters"].pop("lights", None)
        o["parameters"].pop("epsilon", None)
    sc = host.Scene(_write(tmp_path, s))
    objs = sc.objects
    assert (objs["num_lights"] == 0).all()
    sph = objs[objs["type"] == 0]
    np.testing.assert_array_equal(sph["epsilon"], (sph["b"][:, 0] * np.float32(0.0003)).astype(np.float32))
    assert (objs[objs["type"] == 2]["epsilon"] == np.float32(1e-6)).all()


# ---- image plane, quantiser, TIFF --------------------------------------------------

def test_image_plane_matches_reference_formula():
    sc = host.Scene(scene_path("scene5_small"))
    o = sc.options(1920, 1080)
    cam = FIXTURES["scene5_small:"]["camera"]
    pos, ax, ay, az = (np.array(cam[i:i + 3], dtype=np.float32) for i in (0, 3, 6, 9))
    fov, focal = np.float32(cam[12]), np.float32(cam[13])
    f32 = np.float32
    size_x = f32(f32(f32(2) * focal) * f32(np.tan(f32(f32(fov * f32(3.1415927)) / f32(360.0)), dtype=np.float32)))
    size_y = f32(f32(size_x * f32(1080)) / f32(1920))
    vx = ax * f32(size_x / f32(1920))
    vy = ay * f32(size_y / f32(1080))
    centre = az * focal + pos
    corner = (centre + vx * f32(0.5 - 1920 / 2.0)) + vy * f32(0.5 - 1080 / 2.0)
    np.testing.assert_allclose(np.array(o.vx[:]), vx, rtol=1e-6)
    np.testing.assert_allclose(np.array(o.vy[:]), vy, rtol=1e-6)
    np.testing.assert_allclose(np.array(o.corner[:]), corner, rtol=1e-6, atol=1e-6)
    np.testing.assert_array_equal(np.array(o.cam_pos[:]), pos)
    assert (o.max_bounces, o.samples_per_pixel, o.seed, o.shard_count) == (10, 1, 1, 1)
    assert o.min_intensity_sqr == pytest.approx(1e-4)


def test_options_parse_flags():
    sc = host.Scene(scene_path("scene1"))
    o = sc.options(8, 8, "-b", "3", "-a", "0.5", "-s", "blinn", "-g", "path", "-n", "16", "-l", "lin", "-o", "2.5", "--seed", "9")
    assert (o.max_bounces, o.reflection_model, o.gi_model, o.samples_per_pixel, o.attenuation, o.seed) == (3, 1, 1, 16, 1, 9)
    assert o.min_intensity_sqr == pytest.approx(0.25)
    assert o.attenuation_offset == pytest.approx(2.5)
    # unknown values are ignored, a flag without its value is ignored (argv.c:50-55)
    o = sc.options(8, 8, "-s", "gouraud", "-g")
    assert (o.reflection_model, o.gi_model) == (0, 0)


def test_quantiser_truncates():
    v = np.array([[[-1.0, 0.0, 0.0039], [0.5, 0.999, 1.0], [2.0, 0.00392157, 0.99609375]]], dtype=np.float32)
    q = host.quantize(v)
    expect = np.floor(np.clip(v * np.float32(255.0), 0, 255)).astype(np.uint8)
    np.testing.assert_array_equal(q, expect)
    assert q[0, 1, 0] == 127  # truncation, not rounding (image.c:96-98)


def _parse_tiff(path):
    b = open(path, "rb").read()
    assert b[:4] == b"II*\x00"
    ifd = struct.unpack_from("<I", b, 4)[0]
    n = struct.unpack_from("<H", b, ifd)[0]
    tags = {}
    for i in range(n):
        tag, typ, cnt, val = struct.unpack_from("<HHII", b, ifd + 2 + 12 * i)
        tags[tag] = (typ, cnt, val)
    return b, tags


def test_tiff_raw_layout_matches_reference_contract(tmp_path):
    """What the reference post-processor's image_load() checks (postprocess/image.c:40-70):
    3 samples, 32 bits, contiguous, z-buffer tag 65000 with W*H floats; one row per strip."""
    rng = np.random.default_rng(0)
    h, w = 7, 5
    raster = rng.random((h, w, 3), dtype=np.float32)
    z = rng.random((h, w), dtype=np.float32)
    p = str(tmp_path / "raw.tif")
    host.save_tiff(p, raster, z, raw=True)
    b, tags = _parse_tiff(p)
    assert list(tags) == sorted(tags)
    assert tags[256][2] == w and tags[257][2] == h
    assert tags[277][2] & 0xffff == 3 and tags[284][2] & 0xffff == 1 and tags[262][2] & 0xffff == 2
    assert tags[278][2] == 1 and tags[274][2] & 0xffff == 1
    assert 339 not in tags  # no SampleFormat, as the reference
    typ, cnt, off = tags[258]
    assert (typ, cnt) == (3, 3) and struct.unpack_from("<3H", b, off) == (32, 32, 32)
    typ, cnt, off = tags[65000]
    assert (typ, cnt) == (11, w * h)
    np.testing.assert_array_equal(np.frombuffer(b, dtype="<f4", count=w * h, offset=off).reshape(h, w), z)
    _, nstrips, offs = tags[273]
    assert nstrips == h
    rows = [np.frombuffer(b, dtype="<f4", count=w * 3, offset=struct.unpack_from("<I", b, offs + 4 * r)[0]) for r in range(h)]
    np.testing.assert_array_equal(np.stack(rows).reshape(h, w, 3), raster)


def test_tiff_8bit_readable_by_pillow(tmp_path):
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(1)
    raster = rng.random((9, 13, 3), dtype=np.float32) * 1.2
    p = str(tmp_path / "img.tif")
    host.save_tiff(p, raster, raw=False)
    img = np.array(Image.open(p))
    np.testing.assert_array_equal(img, host.quantize(raster))


def test_tiff_reader_roundtrip(tmp_path):
    L = host.lib()

    class Img(C.Structure):
        _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("bits", C.c_uint32), ("rgb8", C.c_void_p),
                    ("raster", C.POINTER(C.c_float)), ("z", C.POINTER(C.c_float)), ("z_count", C.c_uint32)]
    rng = np.random.default_rng(2)
    raster = rng.random((4, 6, 3), dtype=np.float32)
    z = rng.random((4, 6), dtype=np.float32)
    p = str(tmp_path / "raw.tif")
    host.save_tiff(p, raster, z, raw=True)
    img = Img()
    err = C.create_string_buffer(512)
    assert L.rt_tiff_read(p.encode(), C.byref(img), err) == 0, err.value
    assert (img.width, img.height, img.bits, img.z_count) == (6, 4, 32, 24)
    np.testing.assert_array_equal(np.ctypeslib.as_array(img.raster, shape=(4, 6, 3)), raster)
    np.testing.assert_array_equal(np.ctypeslib.as_array(img.z, shape=(4, 6)), z)
    L.rt_tiff_free(C.byref(img))


def test_engine_cli_help_and_usage():
    exe = os.path.join(ROOT, "engine")
    if not os.path.exists(exe):
        pytest.skip("engine not built")
    out = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    assert out.stdout.startswith("Render a scene using raytracing.\n")
    assert "[-f]                             : DEFAULT = OFF     : save raw output for post-processing." in out.stdout
    out = subprocess.run([exe, "a", "b"], capture_output=True, text=True)
    assert out.returncode == 1 and out.stdout.startswith("Too few arguments.")
    out = subprocess.run([exe, "missing.json", "o.tif", "8", "8"], capture_output=True, text=True, cwd=ROOT)
    assert out.returncode == 1 and "Unable to open scene file [missing.json]." in out.stdout


def test_parity_tiers():
    """c-raytracer_b200/parity.py: the deterministic tier bounds the whole frame, the stochastic
    tier bounds the whole frame by twice the reference's seed-to-seed rel-RMSE on top of the
    outlier-excluded check."""
    from c_raytracer_b200 import parity
    rng = np.random.default_rng(0)
    ref = rng.random((100, 200, 3)).astype(np.float32)
    ok = parity.frame_parity(ref * np.float32(1 + 1e-5), ref)
    assert ok["ok"] and ok["tier"] == "deterministic" and ok["allowed"] == 2 and ok["outliers"] == 0
    one_bad = ref.copy()
    one_bad[3, 4] += 0.02  # 1 of 20000 pixels: within the 1e-4 allowance, and the whole frame stays below 1e-3
    assert parity.frame_parity(one_bad, ref)["ok"]
    one_bad[3, 4] += 0.5    # ... but not when that one pixel lifts the whole-frame error above 1e-3
    assert not parity.frame_parity(one_bad, ref)["ok"]
    many_bad = ref.copy()
    many_bad[:3, :3] += 0.5  # 9 outliers > 2 allowed
    assert not parity.frame_parity(many_bad, ref)["ok"]
    # stochastic tier: the same 9 flipped pixels pass (<= 1e-3 of 20000 = 20) while the whole frame stays within 2 x seed-to-seed
    st = parity.frame_parity(many_bad, ref, seed_to_seed=0.1)
    assert st["ok"] and st["tier"] == "stochastic" and st["bound_all"] == 0.2 and st["rel_rmse_all"] > st["rel_rmse"]
    # ... and fail when the whole-frame error exceeds that bound although the outlier count is fine
    assert not parity.frame_parity(many_bad, ref, seed_to_seed=1e-3)["ok"]
    small = ref[:10, :10]
    assert parity.frame_parity(small, small)["allowed"] == 0  # deterministic tier, 100 pixels: no outlier allowed


def test_bench_goldens_are_complete():
    """tests/golden/bench_*.npz (the benchmark flags on the benchmark geometry, rendered by the
    unmodified reference) carry the frame, the ray counts, the mesh hash and the Tier-2 number."""
    import glob
    files = sorted(glob.glob(os.path.join(GOLDEN, "bench_*.npz")))
    assert len(files) >= 2
    for f in files:
        g = np.load(f)
        h, w = int(g["height"]), int(g["width"])
        assert g["raster"].shape == (h, w, 3) and g["z"].shape == (h, w)
        assert 0.01 < float(g["seed_to_seed_rel_rmse"]) < 1.0
        assert len(str(g["mesh_sha256"])) == 64 and int(g["shadow"]) > 0


def test_loaders_survive_mutated_inputs(tmp_path):
    """tools/fuzz: the scene, STL and TIFF loaders built with ASan + UBSan on seeded mutations of the stock inputs
    (truncations, flipped bytes, extreme numbers, counts beyond the file, deep nesting): a case may be refused, never
    crash.  A short fixed-seed run here; tools/fuzz/run.sh <seed> <cases> for longer ones."""
    import shutil
    if not shutil.which("gcc"):
        pytest.skip("gcc missing")
    env = dict(os.environ, CRT_FUZZ_DIR=str(tmp_path))
    r = subprocess.run(["bash", os.path.join(ROOT, "tools", "fuzz", "run.sh"), "3", "120"], capture_output=True, text=True, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("bad 0") == 3
