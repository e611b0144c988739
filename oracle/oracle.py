"""TEST INFRASTRUCTURE (oracle).  ctypes binding of oracle/liboracle.so (the CPU
restatement, crt_oracle.c) and a runner for oracle/_ref/ref_engine (the
unmodified reference compiled in place).  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs may import this module."""
import ctypes as C
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
ORACLE_SO = os.path.join(HERE, "liboracle.so")
REF_ENGINE = os.path.join(HERE, "_ref", "ref_engine")
ENGINE_REF = os.path.join(HERE, "_ref", "engine_ref")
ENGINE_REF_STOCK = os.path.join(HERE, "_ref", "engine_ref_stock")
REF_POST = os.path.join(HERE, "_ref", "ref_post")                 # reference post-processor, FP32 raster dumped
POSTPROCESS_REF = os.path.join(HERE, "_ref", "postprocess_ref")   # the reference's own postprocess executable

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_SO):
            subprocess.run(["make", "-C", HERE, "liboracle.so"], check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(ORACLE_SO)
        L.oracle_scene_create.restype = C.c_void_p
        L.oracle_scene_create.argtypes = [C.c_void_p, C.c_int]
        L.oracle_scene_destroy.argtypes = [C.c_void_p]
        L.oracle_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_trace_shadow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.oracle_texture_eval.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_size_t, C.c_void_p]
        L.oracle_simplex_noise.restype = C.c_float
        L.oracle_simplex_noise.argtypes = [C.c_float, C.c_float, C.c_float]
        L.oracle_primitive_hit.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_postprocess.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


class OracleScene:
    """CPU restatement over a host.Scene (or anything with a .desc SceneDesc)."""

    def __init__(self, scene, brute_force=False):
        self.scene = scene  # keeps the host arrays alive
        self._h = lib().oracle_scene_create(C.addressof(scene.desc), int(brute_force))

    def close(self):
        if getattr(self, "_h", None):
            lib().oracle_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def render(self, opts, stats=None):
        w, h = opts.width, opts.height
        raster = np.zeros((h, w, 3), dtype=np.float32)
        z = np.zeros((h, w), dtype=np.float32)
        lib().oracle_render(self._h, C.addressof(opts), raster.ctypes.data, z.ctypes.data,
                            C.addressof(stats) if stats is not None else None)
        return raster, z

    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = len(rays)
        obj = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        normal = np.zeros((n, 3), dtype=np.float32)
        lib().oracle_trace_closest(self._h, rays.ctypes.data, n, obj.ctypes.data, t.ctypes.data, normal.ctypes.data)
        return obj, t, normal

    def trace_shadow(self, rays, dist, emitter, light):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = len(rays)
        dist = np.ascontiguousarray(dist, dtype=np.float32)
        emitter = np.ascontiguousarray(emitter, dtype=np.int32)
        light = np.array(light, dtype=np.float32).reshape(n, 3).copy()
        blocked = np.zeros(n, dtype=np.uint8)
        lib().oracle_trace_shadow(self._h, rays.ctypes.data, dist.ctypes.data, emitter.ctypes.data, n,
                                  light.ctypes.data, blocked.ctypes.data)
        return blocked, light

    def texture(self, material, points):
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        out = np.zeros_like(pts)
        if lib().oracle_texture_eval(self._h, material, pts.ctypes.data, len(pts), out.ctypes.data):
            raise ValueError("bad material index")
        return out


def simplex_noise(points):
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    return np.array([lib().oracle_simplex_noise(*map(float, p)) for p in pts], dtype=np.float32)


# ---- the compiled reference (oracle/_ref) ---------------------------------

def postprocess(opts, raster, z_buffer, stats=None):
    """CPU restatement of the reference's postprocess() (crt_oracle_post.c); returns a new raster.
    Raises ValueError with the crt status on rejected options."""
    z = np.ascontiguousarray(z_buffer, dtype=np.float32)
    out = np.array(raster, dtype=np.float32, order="C", copy=True)
    h, w = z.shape
    rc = lib().oracle_postprocess(C.addressof(opts), w, h, out.ctypes.data, z.ctypes.data,
                                  C.addressof(stats) if stats is not None else None)
    if rc:
        raise ValueError(rc)
    return out


def ref_postprocess(raw_tiff, flags):
    """FP32 raster left by the UNMODIFIED reference postprocess() (oracle/_ref/ref_post) run on
    a raw TIFF with the reference's own flags."""
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "out.f32")
        _run([REF_POST, raw_tiff, out] + [str(f) for f in flags])
        b = open(out, "rb").read()
    w, h = (int(x) for x in np.frombuffer(b[:8], dtype=np.uint32))
    return np.frombuffer(b[8:], dtype=np.float32).reshape(h, w, 3).copy()


def have_ref():
    return os.path.exists(REF_ENGINE)


def _run(args):
    subprocess.run(args, check=True, cwd=ROOT, stdout=subprocess.DEVNULL)


def read_dump(path):
    b = open(path, "rb").read()
    hdr = np.frombuffer(b[:16], dtype=np.uint32)
    assert hdr[0] == 0x44545243
    w, h = int(hdr[1]), int(hdr[2])
    o = 16
    raster = np.frombuffer(b[o:o + w * h * 12], dtype=np.float32).reshape(h, w, 3).copy()
    o += w * h * 12
    z = np.frombuffer(b[o:o + w * h * 4], dtype=np.float32).reshape(h, w).copy()
    o += w * h * 4
    cnt = np.frombuffer(b[o:o + 32], dtype=np.uint64)
    secs = float(np.frombuffer(b[o + 32:o + 40], dtype=np.float64)[0])
    return raster, z, {"cast_ray": int(cnt[0]), "closest_full": int(cnt[1]), "shadow": int(cnt[2]), "box_tests": int(cnt[3]), "seconds": secs}


def ref_render(scene_path, width, height, flags=(), seed=1, threads="max"):
    """Image of the UNMODIFIED reference cast_ray() with the structural RNG stream."""
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "out.bin")
        _run([REF_ENGINE, "render", scene_path, out, str(width), str(height), "-m", str(threads),
              "--seed", str(seed)] + [str(f) for f in flags])
        return read_dump(out)


def ref_rays(scene_path, rays, dist, emitter, light):
    rays = np.asarray(rays, dtype=np.float32).reshape(-1, 6)
    n = len(rays)
    rec = np.zeros((n, 11), dtype=np.float32)
    rec[:, :6] = rays
    rec[:, 6] = dist
    rec[:, 7] = emitter
    rec[:, 8:11] = np.asarray(light, dtype=np.float32).reshape(n, 3)
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.uint32(n).tobytes())
            f.write(rec.tobytes())
        _run([REF_ENGINE, "rays", scene_path, fin, fout])
        dt = np.dtype([("object", "<i4"), ("t", "<f4"), ("normal", "<f4", 3), ("blocked", "<i4"), ("light", "<f4", 3)])
        return np.fromfile(fout, dtype=dt)


def ref_texture(scene_path, material, points):
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        pts.tofile(fin)
        _run([REF_ENGINE, "tex", scene_path, str(material), fin, fout])
        return np.fromfile(fout, dtype=np.float32).reshape(-1, 3)


def ref_noise(points):
    pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        pts.tofile(fin)
        _run([REF_ENGINE, "noise", fin, fout])
        return np.fromfile(fout, dtype=np.float32)


def ref_kat(records):
    """records: (n, 20) float32, see oracle/ref_engine.c mode_kat."""
    rec = np.ascontiguousarray(records, dtype=np.float32).reshape(-1, 20)
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.uint32(len(rec)).tobytes())
            f.write(rec.tobytes())
        _run([REF_ENGINE, "kat", fin, fout])
        return np.fromfile(fout, dtype=np.dtype([("hit", "<i4"), ("t", "<f4"), ("t2", "<f4")]))
