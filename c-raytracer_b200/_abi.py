"""ctypes mirrors of the structures in include/crt.h (the C ABI).

Field order and types must match include/crt.h exactly; tests/test_abi.py checks
the sizes against the C compiler.
"""
import ctypes as C

import numpy as np

CRT_ABI_VERSION = 2
CRT_BUILD_HOST, CRT_BUILD_DEVICE = 0, 1

CRT_OBJ_SPHERE, CRT_OBJ_TRIANGLE, CRT_OBJ_PLANE = 0, 1, 2
CRT_TEX_UNIFORM, CRT_TEX_CHECKERBOARD, CRT_TEX_BRICK, CRT_TEX_NOISY_PERIODIC = 0, 1, 2, 3
CRT_MAT_REFLECTIVE, CRT_MAT_TRANSPARENT, CRT_MAT_EMITTANT = 1, 2, 4
CRT_REFLECTION_PHONG, CRT_REFLECTION_BLINN = 0, 1
CRT_GI_AMBIENT, CRT_GI_PATH = 0, 1
CRT_ATTEN_NONE, CRT_ATTEN_LINEAR, CRT_ATTEN_SQUARE = 0, 1, 2
CRT_RENDER_COUNT_STATS = 1
CRT_FUNC_SIN, CRT_FUNC_SAW, CRT_FUNC_TRIANGLE, CRT_FUNC_SQUARE = 0, 1, 2, 3
CRT_OK, CRT_ERR_INVALID, CRT_ERR_NO_DEVICE, CRT_ERR_CUDA, CRT_ERR_NOMEM, CRT_ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5

CRT_POST_BRIGHTEN, CRT_POST_DOF, CRT_POST_DOF_CAMERA, CRT_POST_MIST = 1, 2, 4, 8
CRT_FALLOFF_QUADRATIC, CRT_FALLOFF_LINEAR, CRT_FALLOFF_INV_QUADRATIC = 0, 1, 2

f3 = C.c_float * 3


class Material(C.Structure):
    _fields_ = [
        ("ks", f3), ("ka", f3), ("kr", f3), ("kt", f3), ("ke", f3),
        ("shininess", C.c_float), ("refractive_index", C.c_float),
        ("flags", C.c_uint32), ("tex_type", C.c_int32), ("tex_func", C.c_int32),
        ("color", f3 * 2), ("scale", C.c_float), ("param1", C.c_float), ("param2", C.c_float),
        ("id", C.c_int32), ("reserved", C.c_float * 2),
    ]


class Object(C.Structure):
    _fields_ = [
        ("type", C.c_int32), ("material", C.c_int32), ("num_lights", C.c_uint32), ("epsilon", C.c_float),
        ("a", f3), ("b", f3), ("c", f3), ("n", f3),
    ]


class SceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32), ("num_objects", C.c_uint32), ("objects", C.POINTER(Object)),
        ("num_materials", C.c_uint32), ("materials", C.POINTER(Material)),
        ("ambient", f3), ("build", C.c_uint32),
    ]


class RenderOpts(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32),
        ("cam_pos", f3), ("corner", f3), ("vx", f3), ("vy", f3),
        ("max_bounces", C.c_uint32), ("min_intensity_sqr", C.c_float),
        ("reflection_model", C.c_int32), ("gi_model", C.c_int32), ("samples_per_pixel", C.c_uint32),
        ("attenuation", C.c_int32), ("attenuation_offset", C.c_float),
        ("seed", C.c_uint32), ("flags", C.c_uint32),
        ("shard_index", C.c_uint32), ("shard_count", C.c_uint32),
        ("lens_radius", C.c_float), ("focus_distance", C.c_float), ("lens_samples", C.c_uint32),
        ("fog_density", C.c_float), ("fog_color", f3), ("reserved", C.c_uint32),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("rays_closest", C.c_uint64), ("rays_shadow", C.c_uint64), ("box_tests", C.c_uint64),
        ("tri_tests", C.c_uint64 * 4), ("sphere_tests", C.c_uint64 * 2), ("plane_tests", C.c_uint64),
        ("nodes_shaded", C.c_uint64), ("kernel_ms", C.c_double), ("total_ms", C.c_double),
        ("node_steps", C.c_uint64), ("node_step_lanes", C.c_uint64), ("node_step_idle", C.c_uint64),
        ("prim_steps", C.c_uint64), ("prim_step_lanes", C.c_uint64), ("prim_step_idle", C.c_uint64),
        ("queue_ms", C.c_double), ("last_warp_ms", C.c_double), ("warp_busy_ms", C.c_double),
    ]

    def as_dict(self):
        return {
            "rays_closest": int(self.rays_closest), "rays_shadow": int(self.rays_shadow),
            "box_tests": int(self.box_tests), "tri_tests": [int(x) for x in self.tri_tests],
            "sphere_tests": [int(x) for x in self.sphere_tests], "plane_tests": int(self.plane_tests),
            "nodes_shaded": int(self.nodes_shaded), "kernel_ms": float(self.kernel_ms),
            "total_ms": float(self.total_ms),
            "node_steps": int(self.node_steps), "node_step_lanes": int(self.node_step_lanes), "node_step_idle": int(self.node_step_idle),
            "prim_steps": int(self.prim_steps), "prim_step_lanes": int(self.prim_step_lanes), "prim_step_idle": int(self.prim_step_idle),
            "queue_ms": float(self.queue_ms), "last_warp_ms": float(self.last_warp_ms), "warp_busy_ms": float(self.warp_busy_ms),
        }


class FrameShare(C.Structure):
    _fields_ = [("work_counter", C.c_void_p), ("total_warps", C.c_uint32), ("reserved", C.c_uint32 * 5)]


class SceneInfo(C.Structure):
    _fields_ = [
        ("num_bounded", C.c_uint32), ("num_planes", C.c_uint32), ("num_emitters", C.c_uint32),
        ("total_lights", C.c_uint32), ("num_nodes", C.c_uint32), ("max_depth", C.c_uint32),
        ("device", C.c_uint32), ("sm_count", C.c_uint32),
        ("build_ms", C.c_double), ("upload_ms", C.c_double), ("device_bytes", C.c_uint64),
        ("built_on_device", C.c_uint32), ("reserved", C.c_uint32),
    ]


class PostOpts(C.Structure):
    _fields_ = [
        ("ops", C.c_uint32), ("brighten", C.c_float),
        ("dof_scale", C.c_float), ("dof_bias", C.c_float),
        ("dof_aperture", C.c_float), ("dof_focal_length", C.c_float), ("dof_plane_in_focus", C.c_float),
        ("mist_start", C.c_float), ("mist_depth", C.c_float), ("mist_falloff", C.c_uint32), ("mist_color", f3),
        ("reserved", C.c_uint32 * 3),
    ]


class PostStats(C.Structure):
    _fields_ = [
        ("kernel_ms", C.c_double), ("total_ms", C.c_double),
        ("dof_scale", C.c_float), ("dof_bias", C.c_float),
        ("dof_max_radius", C.c_uint32), ("launches", C.c_uint32),
    ]


# numpy views of the two record types (same memory layout)
OBJECT_DTYPE = np.dtype([
    ("type", "<i4"), ("material", "<i4"), ("num_lights", "<u4"), ("epsilon", "<f4"),
    ("a", "<f4", 3), ("b", "<f4", 3), ("c", "<f4", 3), ("n", "<f4", 3),
])
MATERIAL_DTYPE = np.dtype([
    ("ks", "<f4", 3), ("ka", "<f4", 3), ("kr", "<f4", 3), ("kt", "<f4", 3), ("ke", "<f4", 3),
    ("shininess", "<f4"), ("refractive_index", "<f4"),
    ("flags", "<u4"), ("tex_type", "<i4"), ("tex_func", "<i4"),
    ("color", "<f4", (2, 3)), ("scale", "<f4"), ("param1", "<f4"), ("param2", "<f4"),
    ("id", "<i4"), ("reserved", "<f4", 2),
])
assert OBJECT_DTYPE.itemsize == C.sizeof(Object) == 64
assert MATERIAL_DTYPE.itemsize == C.sizeof(Material) == 128


def algorithmic_flops(stats: "Stats") -> float:
    """FLOPs of the intersection tests of one frame, by the contract constants of
    SURVEY section 8(d): 3 per ray, 12 per box test, triangle 14/24/40/46 by exit
    stage, sphere 17/20, plane 12."""
    rays = stats.rays_closest + stats.rays_shadow
    tri = stats.tri_tests
    return (3.0 * rays + 12.0 * stats.box_tests
            + 14.0 * tri[0] + 24.0 * tri[1] + 40.0 * tri[2] + 46.0 * tri[3]
            + 17.0 * stats.sphere_tests[0] + 20.0 * stats.sphere_tests[1]
            + 12.0 * stats.plane_tests)
