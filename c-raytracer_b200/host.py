"""Host-side mirror of the reference's scene / image interface, bound to the C
host library build/libcrt_host.so (src/core, src/raytracer).  Nothing here
touches the GPU."""
import ctypes as C
import os

import numpy as np

from . import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_HOST_SO = os.path.join(ROOT, "build", "libcrt_host.so")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_HOST_SO):
            raise RuntimeError(f"{_HOST_SO} is missing: run `make host` (or __graft_entry__.build())")
        L = C.CDLL(_HOST_SO)
        L.rt_host_scene_open.restype = C.c_void_p
        L.rt_host_scene_open.argtypes = [C.c_char_p, C.c_char_p]
        L.rt_host_scene_close.argtypes = [C.c_void_p]
        L.rt_host_last_error.restype = C.c_char_p
        L.rt_host_scene_desc.argtypes = [C.c_void_p, C.POINTER(_abi.SceneDesc)]
        L.rt_host_opts.argtypes = [C.c_void_p, C.c_uint, C.c_uint, C.POINTER(_abi.RenderOpts)]
        L.rt_host_opts_parse.argtypes = [C.POINTER(_abi.RenderOpts), C.c_int, C.POINTER(C.c_char_p)]
        L.rt_host_save.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_int]
        L.rt_host_save.restype = C.c_int
        L.rt_host_quantize.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.rt_host_post_opts_parse.argtypes = [C.POINTER(_abi.PostOpts), C.c_int, C.POINTER(C.c_char_p)]
        L.rt_host_load_raw.argtypes = [C.c_char_p, C.POINTER(C.c_uint), C.POINTER(C.c_uint),
                                       C.POINTER(C.POINTER(C.c_float)), C.POINTER(C.POINTER(C.c_float))]
        L.rt_host_free.argtypes = [C.c_void_p]
        L.rt_host_free.restype = None
        _lib = L
    return _lib


class SceneError(RuntimeError):
    pass


class Scene:
    """A loaded scene file (the state the reference keeps in globals after
    scene_load(), reference src/raytracer/scene.c:70-122).  Mesh paths are
    resolved relative to `cwd` (default: the repository root)."""

    def __init__(self, path, rescale=None, cwd=ROOT):
        path = os.path.abspath(path)
        old = os.getcwd()
        os.chdir(cwd)
        try:
            self._h = lib().rt_host_scene_open(path.encode(), rescale.encode() if rescale else None)
        finally:
            os.chdir(old)
        if not self._h:
            raise SceneError(lib().rt_host_last_error().decode())
        self.desc = _abi.SceneDesc()
        lib().rt_host_scene_desc(self._h, C.byref(self.desc))
        self.path = path

    def close(self):
        if getattr(self, "_h", None):
            lib().rt_host_scene_close(self._h)
            self._h = None

    def __del__(self):
        self.close()

    @property
    def objects(self):
        n = self.desc.num_objects
        buf = (C.c_char * (n * 64)).from_address(C.addressof(self.desc.objects.contents))
        return np.frombuffer(buf, dtype=_abi.OBJECT_DTYPE, count=n)

    @property
    def materials(self):
        n = self.desc.num_materials
        buf = (C.c_char * (n * 128)).from_address(C.addressof(self.desc.materials.contents))
        return np.frombuffer(buf, dtype=_abi.MATERIAL_DTYPE, count=n)

    def options(self, width, height, *flags):
        """Render options for a width x height frame; `flags` are engine
        command-line flags, e.g. "-g", "path", "-n", "16"."""
        o = _abi.RenderOpts()
        lib().rt_host_opts(self._h, width, height, C.byref(o))
        if flags:
            argv = [b"engine"] + [str(f).encode() for f in flags]
            arr = (C.c_char_p * len(argv))(*argv)
            lib().rt_host_opts_parse(C.byref(o), len(argv), arr)
        return o


def post_options(*flags):
    """The post-processor's command-line flags as a PostOpts, e.g.
    post_options("-b", 1.5, "--dof", 2, -4, "--mist", 3, 10, "lin", .5, .5, .6)
    (reference src/postprocess/postproc.c:36-92)."""
    o = _abi.PostOpts()
    argv = [b"postprocess"] + [str(f).encode() for f in flags]
    arr = (C.c_char_p * len(argv))(*argv)
    if lib().rt_host_post_opts_parse(C.byref(o), len(argv), arr):
        raise SceneError(lib().rt_host_last_error().decode())
    return o


def load_raw(path):
    """(raster (H, W, 3), z_buffer (H, W)) of a raw frame (-f): what the post-processor reads
    (reference src/postprocess/image.c:30-73)."""
    w, h = C.c_uint(), C.c_uint()
    pr, pz = C.POINTER(C.c_float)(), C.POINTER(C.c_float)()
    if lib().rt_host_load_raw(os.fsencode(path), C.byref(w), C.byref(h), C.byref(pr), C.byref(pz)):
        raise SceneError(lib().rt_host_last_error().decode())
    try:
        raster = np.ctypeslib.as_array(pr, shape=(h.value, w.value, 3)).copy()
        z = np.ctypeslib.as_array(pz, shape=(h.value, w.value)).copy()
    finally:
        lib().rt_host_free(pr)
        lib().rt_host_free(pz)
    return raster, z


def quantize(raster):
    """The reference's truncating 8-bit quantiser (image.c:96-98)."""
    r = np.ascontiguousarray(raster, dtype=np.float32)
    out = np.empty(r.shape, dtype=np.uint8)
    lib().rt_host_quantize(r.ctypes.data, r.size, out.ctypes.data)
    return out


def save_tiff(path, raster, z_buffer=None, raw=False):
    r = np.ascontiguousarray(raster, dtype=np.float32)
    h, w = r.shape[:2]
    z = np.ascontiguousarray(z_buffer if z_buffer is not None else np.zeros((h, w)), dtype=np.float32)
    if lib().rt_host_save(os.fsencode(path), w, h, r.ctypes.data, z.ctypes.data, int(raw)):
        raise RuntimeError(lib().rt_host_last_error().decode())
