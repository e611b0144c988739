"""Binding of the device library csrc/libcrt.so (the C ABI of include/crt.h).

This is the product path: hand-written sm_100a CUDA behind a C ABI.  There is
no fallback -- if the library is missing or no GPU is usable, calls raise."""
import ctypes as C
import os

import numpy as np

from . import _abi

HERE = os.path.dirname(os.path.abspath(__file__))
# CRT_LIBCRT: build-variant override used by tools/ experiments only
LIB_PATH = os.environ.get("CRT_LIBCRT") or os.path.join(HERE, "csrc", "libcrt.so")

# every symbol include/crt.h declares
SYMBOLS = [
    "crt_last_error", "crt_abi_version", "crt_device_count",
    "crt_scene_create", "crt_scene_destroy", "crt_scene_get_info",
    "crt_render", "crt_render_device", "crt_render_device_sync",
    "crt_scene_resident_warps", "crt_render_device_shared",
    "crt_shared_alloc", "crt_shared_open", "crt_shared_release", "crt_host_register", "crt_host_unregister", "crt_scene_read_tree",
    "crt_multi_create", "crt_multi_destroy", "crt_multi_num_devices", "crt_multi_is_dynamic", "crt_multi_get_info", "crt_multi_render",
    "crt_trace_closest", "crt_trace_shadow", "crt_texture_eval", "crt_measure_fp32_peak",
    "crt_postprocess", "crt_postprocess_device", "crt_postprocess_device_noalloc",
]

_lib = None


class CrtError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CrtError(f"{LIB_PATH} is missing: build it with `make device` or __graft_entry__.build(); "
                           "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.crt_last_error.restype = C.c_char_p
        L.crt_scene_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.crt_scene_destroy.argtypes = [C.c_void_p]
        L.crt_scene_destroy.restype = None
        L.crt_scene_get_info.argtypes = [C.c_void_p, C.c_void_p]
        L.crt_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.crt_render_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.crt_render_device_sync.argtypes = [C.c_void_p, C.c_void_p]
        L.crt_scene_resident_warps.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32)]
        L.crt_render_device_shared.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.crt_shared_alloc.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]
        L.crt_shared_open.argtypes = [C.c_int, C.c_void_p, C.POINTER(C.c_void_p)]
        L.crt_shared_release.argtypes = [C.c_int, C.c_void_p, C.c_int]
        L.crt_scene_read_tree.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
        L.crt_host_register.argtypes = [C.c_void_p, C.c_size_t]
        L.crt_host_unregister.argtypes = [C.c_void_p]
        L.crt_multi_create.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        L.crt_multi_destroy.argtypes = [C.c_void_p]
        L.crt_multi_destroy.restype = None
        L.crt_multi_num_devices.argtypes = [C.c_void_p]
        L.crt_multi_is_dynamic.argtypes = [C.c_void_p]
        L.crt_multi_get_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.crt_multi_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]
        L.crt_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
        L.crt_trace_shadow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
        L.crt_texture_eval.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_size_t, C.c_void_p]
        L.crt_measure_fp32_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
        L.crt_postprocess.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.crt_postprocess_device.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
        L.crt_postprocess_device_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise CrtError(f"crt error {rc}: {lib().crt_last_error().decode()}")


def device_count():
    return lib().crt_device_count()


def measure_fp32_peak(device=0):
    v = C.c_double()
    _check(lib().crt_measure_fp32_peak(device, C.byref(v)))
    return v.value


def postprocess(opts, raster, z_buffer, device=0, stats=None):
    """crt_postprocess: brighten / depth of field / mist on a host frame (H, W, 3) + (H, W);
    returns the processed raster (a new array)."""
    z = np.ascontiguousarray(z_buffer, dtype=np.float32)
    out = np.array(raster, dtype=np.float32, order="C", copy=True)
    h, w = z.shape
    assert out.shape == (h, w, 3)
    _check(lib().crt_postprocess(device, C.addressof(opts), w, h, out.ctypes.data, z.ctypes.data,
                                 C.addressof(stats) if stats is not None else None))
    return out


def postprocess_device(opts, width, height, d_raster_ptr, d_z_ptr, device=0, stream=None, stats=None):
    """crt_postprocess_device: in place on device pointers (e.g. the buffers crt_render_device filled)."""
    _check(lib().crt_postprocess_device(device, C.addressof(opts), width, height, C.c_void_p(d_raster_ptr), C.c_void_p(d_z_ptr),
                                        C.c_void_p(stream) if stream else None,
                                        C.addressof(stats) if stats is not None else None))


def postprocess_device_noalloc(opts, width, height, d_raster_ptr, d_z_ptr, device=0, stream=None):
    """crt_postprocess_device_noalloc: the depth-of-field path on device pointers with the library's fixed
    scratch array (no allocation inside the call; frames up to NOALLOC_MAX_PIXELS)."""
    _check(lib().crt_postprocess_device_noalloc(device, C.addressof(opts), width, height, C.c_void_p(d_raster_ptr),
                                                C.c_void_p(d_z_ptr), C.c_void_p(stream) if stream else None))


NOALLOC_MAX_PIXELS = 1 << 21  # CRT_POST_NOALLOC_MAX_PIXELS


class DeviceScene:
    """A scene resident on one GPU (crt_scene*).  `scene` is a host.Scene or
    anything exposing a `desc` (_abi.SceneDesc)."""

    def __init__(self, scene, device=0, build=None):
        """build: None (what scene.desc says: the host), "host" or "device" (crt_scene_desc.build)."""
        self.scene = scene
        h = C.c_void_p()
        desc = scene.desc
        if build is not None:
            desc = _abi.SceneDesc.from_buffer_copy(scene.desc)
            desc.build = {"host": _abi.CRT_BUILD_HOST, "device": _abi.CRT_BUILD_DEVICE}[build]
        _check(lib().crt_scene_create(C.addressof(desc), device, C.byref(h)))
        self._h = h
        self.info = _abi.SceneInfo()
        _check(lib().crt_scene_get_info(self._h, C.addressof(self.info)))

    def read_tree(self):
        """(nodes, prims): the device tree as numpy arrays -- nodes (num_nodes, 32) uint32 words of the
        128-byte records, prims (num_bounded, 16) float32."""
        nodes = np.empty((self.info.num_nodes, 32), dtype=np.uint32)
        prims = np.empty((self.info.num_bounded, 16), dtype=np.float32)
        _check(lib().crt_scene_read_tree(self._h, nodes.ctypes.data, nodes.nbytes, prims.ctypes.data, prims.nbytes))
        return nodes, prims

    def close(self):
        if getattr(self, "_h", None):
            lib().crt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def render(self, opts, stats=None):
        """crt_render: host buffers out (the reference-facing call)."""
        w, h = opts.width, opts.height
        raster = np.empty((h, w, 3), dtype=np.float32)
        z = np.empty((h, w), dtype=np.float32)
        _check(lib().crt_render(self._h, C.addressof(opts), raster.ctypes.data, z.ctypes.data,
                                C.addressof(stats) if stats is not None else None))
        return raster, z

    def render_into(self, opts, raster, z, stats=None):
        """crt_render into caller-owned host arrays (e.g. pinned memory)."""
        _check(lib().crt_render(self._h, C.addressof(opts), raster.ctypes.data, z.ctypes.data,
                                C.addressof(stats) if stats is not None else None))

    def render_device(self, opts, d_raster_ptr, d_z_ptr, stream=None):
        """crt_render_device: device pointers, enqueued on `stream` (int handle or None)."""
        _check(lib().crt_render_device(self._h, C.addressof(opts), C.c_void_p(d_raster_ptr), C.c_void_p(d_z_ptr),
                                       C.c_void_p(stream) if stream else None))

    def sync(self, stats=None):
        _check(lib().crt_render_device_sync(self._h, C.addressof(stats) if stats is not None else None))

    def resident_warps(self, opts=None):
        w = C.c_uint32()
        _check(lib().crt_scene_resident_warps(self._h, C.addressof(opts) if opts is not None else None, C.byref(w)))
        return w.value

    def render_device_shared(self, opts, d_raster_ptr, d_z_ptr, stream, share):
        """crt_render_device_shared: this scene as one participant of a frame several devices / processes render."""
        _check(lib().crt_render_device_shared(self._h, C.addressof(opts), C.c_void_p(d_raster_ptr), C.c_void_p(d_z_ptr),
                                              C.c_void_p(stream) if stream else None,
                                              C.addressof(share) if share is not None else None))

    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = len(rays)
        obj = np.empty(n, dtype=np.int32)
        t = np.empty(n, dtype=np.float32)
        normal = np.zeros((n, 3), dtype=np.float32)
        _check(lib().crt_trace_closest(self._h, rays.ctypes.data, n, obj.ctypes.data, t.ctypes.data, normal.ctypes.data))
        return obj, t, normal

    def trace_shadow(self, rays, dist, emitter, light):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        n = len(rays)
        dist = np.ascontiguousarray(dist, dtype=np.float32)
        emitter = np.ascontiguousarray(emitter, dtype=np.int32)
        light = np.array(light, dtype=np.float32).reshape(n, 3).copy()
        blocked = np.zeros(n, dtype=np.uint8)
        _check(lib().crt_trace_shadow(self._h, rays.ctypes.data, dist.ctypes.data, emitter.ctypes.data, n,
                                      light.ctypes.data, blocked.ctypes.data))
        return blocked, light

    def texture(self, material, points):
        pts = np.ascontiguousarray(points, dtype=np.float32).reshape(-1, 3)
        out = np.zeros_like(pts)
        _check(lib().crt_texture_eval(self._h, material, pts.ctypes.data, len(pts), out.ctypes.data))
        return out


def shared_alloc(dev, nbytes):
    """-> (device pointer, 64-byte IPC handle) of fresh device memory other processes can map"""
    p = C.c_void_p()
    h = (C.c_ubyte * 64)()
    _check(lib().crt_shared_alloc(dev, nbytes, C.byref(p), h))
    return p.value, bytes(h)


def shared_open(dev, handle):
    p = C.c_void_p()
    h = (C.c_ubyte * 64).from_buffer_copy(handle)
    _check(lib().crt_shared_open(dev, h, C.byref(p)))
    return p.value


def source_fingerprint():
    """SHA-256 over the sources libcrt.so is built from (csrc/, include/, the nvcc flags in the Makefile): the
    library itself is not reproducible bit for bit, so profiles that belong to one build are tied to this."""
    import hashlib
    root = os.path.dirname(HERE)
    files = sorted(os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc")) if f.endswith((".cu", ".cuh", ".h", ".cpp")))
    files += [os.path.join(root, "include", f) for f in sorted(os.listdir(os.path.join(root, "include")))]
    h = hashlib.sha256()
    for f in files:
        h.update(os.path.basename(f).encode() + b"\0" + open(f, "rb").read())
    for line in open(os.path.join(root, "Makefile")):
        if line.startswith("NVFLAGS") or line.startswith("\t-Xcompiler"):
            h.update(line.encode())
    return h.hexdigest()


def host_register(array):
    """crt_host_register on a numpy array the caller keeps alive; undo with host_unregister()."""
    _check(lib().crt_host_register(C.c_void_p(array.ctypes.data), array.nbytes))


def host_unregister(array):
    _check(lib().crt_host_unregister(C.c_void_p(array.ctypes.data)))


def shared_release(dev, ptr, owner):
    _check(lib().crt_shared_release(dev, C.c_void_p(ptr), 1 if owner else 0))


class MultiScene:
    """A scene replicated on several GPUs of this process (crt_multi*): one frame, one dynamic work queue."""

    def __init__(self, scene, devices):
        self.scene = scene
        self.devices = list(devices)
        arr = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        _check(lib().crt_multi_create(C.addressof(scene.desc), arr, len(self.devices), C.byref(h)))
        self._h = h
        self.dynamic = bool(lib().crt_multi_is_dynamic(self._h))
        self.info = _abi.SceneInfo()
        _check(lib().crt_multi_get_info(self._h, 0, C.addressof(self.info)))

    def close(self):
        if getattr(self, "_h", None):
            lib().crt_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        self.close()

    def render(self, opts, stats=None, into=None):
        """crt_multi_render -> (raster, z, per-device kernel ms)"""
        w, h = opts.width, opts.height
        raster, z = into if into is not None else (np.empty((h, w, 3), dtype=np.float32), np.empty((h, w), dtype=np.float32))
        ms = (C.c_double * len(self.devices))()
        _check(lib().crt_multi_render(self._h, C.addressof(opts), raster.ctypes.data, z.ctypes.data,
                                      C.addressof(stats) if stats is not None else None, ms))
        return raster, z, list(ms)
