/*
 * crt.h -- C ABI of the B200 trace-loop library (libcrt.so).
 *
 * This is the drop-in boundary for the per-pixel trace loop of
 * wojciech-graj/C-Raytracer.  The reference has no plugin / FFI interface:
 * the seam is `void render(void)` over file-scope globals
 * (reference src/raytracer/render.h:17-18, called from main.c:77-79), with the
 * two finer seams `get_closest_intersection()` and `is_light_blocked()`
 * (reference src/raytracer/render.c:48-49, :118-134).  Every entry point below
 * names the reference interface it replaces.
 *
 * Plain C, plain pointers and sizes; no C++/torch types.  All functions return
 * CRT_OK (0) or a negative error code; crt_last_error() gives the message.
 * No function calls exit(); the C host (src/raytracer) turns a non-zero return
 * into the reference's `error()` behaviour (reference src/core/error.h:21-26).
 *
 * There is NO CPU fallback behind this ABI: if no CUDA device is usable every
 * compute entry point fails with CRT_ERR_NO_DEVICE.
 */
#ifndef CRT_H
#define CRT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CRT_ABI_VERSION 2

enum crt_status {
	CRT_OK = 0,
	CRT_ERR_INVALID = -1,   /* bad argument / malformed scene description */
	CRT_ERR_NO_DEVICE = -2, /* no usable CUDA device */
	CRT_ERR_CUDA = -3,      /* CUDA runtime error, see crt_last_error() */
	CRT_ERR_NOMEM = -4,
	CRT_ERR_UNSUPPORTED = -5,
};

/* Object kinds: reference `enum ObjectType` (src/raytracer/object.h:18-24). */
enum crt_object_type {
	CRT_OBJ_SPHERE = 0,
	CRT_OBJ_TRIANGLE = 1,
	CRT_OBJ_PLANE = 2,
};

/* Texture kinds: the four `texture_get_color_*` (src/raytracer/material.c:152-200). */
enum crt_texture_type {
	CRT_TEX_UNIFORM = 0,
	CRT_TEX_CHECKERBOARD = 1,
	CRT_TEX_BRICK = 2,
	CRT_TEX_NOISY_PERIODIC = 3,
};

/* reference `enum PeriodicFunction` (src/raytracer/material.h:17-22) */
enum crt_periodic_func {
	CRT_FUNC_SIN = 0,
	CRT_FUNC_SAW = 1,
	CRT_FUNC_TRIANGLE = 2,
	CRT_FUNC_SQUARE = 3,
};

/* material flags, derived exactly as reference material.c:81-83 (|k| > 1e-6) */
#define CRT_MAT_REFLECTIVE 1u
#define CRT_MAT_TRANSPARENT 2u
#define CRT_MAT_EMITTANT 4u

/*
 * One material with its (single) procedural texture.
 * Replaces `struct Material` (material.h:31-44) + the texture structs
 * (material.c:27-53).  128 bytes.
 *
 * Texture parameter mapping:
 *   uniform        : color[0]
 *   checkerboard   : color[0], color[1], scale
 *   brick          : color[0], color[1], scale, param1 = mortar width
 *   noisy periodic : color[0] = color, color[1] = colour gradient,
 *                    scale = noise feature scale, param1 = noise scale,
 *                    param2 = frequency scale, tex_func = crt_periodic_func
 */
typedef struct crt_material {
	float ks[3];
	float ka[3];
	float kr[3];
	float kt[3];
	float ke[3];
	float shininess;
	float refractive_index;
	uint32_t flags;   /* CRT_MAT_* ; recomputed by crt_scene_create if 0xffffffff */
	int32_t tex_type; /* crt_texture_type */
	int32_t tex_func; /* crt_periodic_func */
	float color[2][3];
	float scale;
	float param1;
	float param2;
	int32_t id;       /* scene-file id, informational */
	float reserved[2];
} crt_material;

/*
 * One scene object, in SCENE ORDER (the order of the reference's `objects[]`,
 * object.h:79: JSON order, with each mesh expanded in place into its
 * triangles, scene.c:329-360 / object.c:542-587).  Order matters: it defines
 * emitter order (render.c:170), plane order and the RNG draw order.  64 bytes.
 *
 *   sphere   : a = centre, b[0] = radius                     (object.c:24-28)
 *   triangle : a = vertex 0, b = edge v1-v0, c = edge v2-v0,
 *              n = unit normal normalize(b x c)              (object.c:30-35, :327-334)
 *   plane    : a = unit normal, b[0] = d  (a.x = d)          (object.c:38-42, :457-468)
 *
 * epsilon is the RESOLVED value (auto rules object.c:235-236, :336-339,
 * :453-454 already applied by the loader).
 */
typedef struct crt_object {
	int32_t type;        /* crt_object_type */
	int32_t material;    /* index into crt_scene_desc.materials (not the JSON id) */
	uint32_t num_lights; /* shadow samples when this object is an emitter (scene.c:373) */
	float epsilon;
	float a[3];
	float b[3];
	float c[3];
	float n[3];
} crt_object;

/*
 * The scene as flat host arrays: everything `render()` reads from globals
 * (objects/num_objects object.h:79-86, materials material.h:59-60,
 * global_ambient_light_intensity render.h:20).  Caller-owned, only read during
 * crt_scene_create().
 */
typedef struct crt_scene_desc {
	uint32_t abi_version; /* CRT_ABI_VERSION */
	uint32_t num_objects;
	const crt_object *objects;
	uint32_t num_materials;
	const crt_material *materials;
	float ambient[3]; /* "AmbientLight", scene.c:104-105 */
	uint32_t build;   /* enum crt_build: where accel_init()'s tree is built */
} crt_scene_desc;

/* CRT_BUILD_HOST: binned-SAH tree on one host core, uploaded (fewest box tests per ray; 0.2-0.5 s for a
 * million triangles).  CRT_BUILD_DEVICE: Morton codes + sort + radix tree on the device in milliseconds
 * (accel.c:72-88, 227-264 as it was meant); more box tests per ray.  Same images. */
enum crt_build { CRT_BUILD_HOST = 0, CRT_BUILD_DEVICE = 1 };

enum crt_reflection_model { CRT_REFLECTION_PHONG = 0, CRT_REFLECTION_BLINN = 1 };
enum crt_gi_model { CRT_GI_AMBIENT = 0, CRT_GI_PATH = 1 };
enum crt_attenuation { CRT_ATTEN_NONE = 0, CRT_ATTEN_LINEAR = 1, CRT_ATTEN_SQUARE = 2 };

#define CRT_RENDER_COUNT_STATS 1u /* run the instrumented kernel (ray / test counters) */

/* Frame partition used for multi-GPU sharding (crt_render_opts.shard_*): the
 * frame is cut into CRT_TILE_W x CRT_TILE_H pixel tiles, numbered row-major over
 * ceil(W / CRT_TILE_W) columns; CRT_SHARD_CHUNK_TILES consecutive tiles form a
 * chunk; chunk k belongs to shard k % shard_count. */
#define CRT_TILE_W 8
#define CRT_TILE_H 4
#define CRT_SHARD_CHUNK_TILES 1

/*
 * Per-frame parameters: the image plane `struct Image` (image.h:17-26, set up
 * by image_init image.c:34-56), `camera.position` (camera.h:17-22) and the
 * render statics (render.c:52-59, parsed by render_init :61-116).
 */
typedef struct crt_render_opts {
	uint32_t width, height;
	float cam_pos[3];
	float corner[3]; /* image.corner */
	float vx[3];     /* image.vectors[X] */
	float vy[3];     /* image.vectors[Y] */
	uint32_t max_bounces;      /* -b, default 10 */
	float min_intensity_sqr;   /* (-a)^2, default 1e-4 */
	int32_t reflection_model;  /* -s */
	int32_t gi_model;          /* -g */
	uint32_t samples_per_pixel;/* -n */
	int32_t attenuation;       /* -l */
	float attenuation_offset;  /* -o, default 1 */
	uint32_t seed;             /* counter-based RNG seed (include/crt_rng.h) */
	uint32_t flags;            /* CRT_RENDER_* */
	/* Multi-GPU sharding: tile-chunk k is rendered iff k % shard_count ==
	 * shard_index.  Pixels of other shards are left zero.  (1-rank: 0 / 1.) */
	uint32_t shard_index, shard_count;
	/* In-tracer thin lens and fog.  EXTENSIONS: the reference has neither (its depth of
	 * field and mist are post-processes on the z-buffer, postproc.c:47-188 = crt_postprocess), so there is
	 * no parity oracle for them; all zero = the reference's pinhole camera in clear air, bit for bit.
	 *   lens_radius > 0: the primary ray of lens sample s starts at a point of the lens disk (spanned by
	 *     vx, vy around cam_pos) and passes through the point the pinhole ray meets the plane at
	 *     focus_distance in front of the camera; a pixel is the mean of lens_samples such rays
	 *     (sample 0 is the pinhole ray itself and writes the z-buffer).
	 *   fog_density > 0: every eye-path segment of length t (primary, reflected, refracted, GI) keeps
	 *     exp(-fog_density * t) of what it carries and gains (1 - exp(..)) of fog_color; a segment that
	 *     leaves the scene carries fog_color. */
	float lens_radius;
	float focus_distance;
	uint32_t lens_samples; /* 0 or 1: one ray per pixel */
	float fog_density;
	float fog_color[3];
	uint32_t reserved;
} crt_render_opts;

/* Counters of one crt_render call (all zero unless CRT_RENDER_COUNT_STATS). */
typedef struct crt_stats {
	uint64_t rays_closest; /* get_closest_intersection calls incl. inside-object shortcut */
	uint64_t rays_shadow;  /* is_light_blocked calls */
	uint64_t box_tests;    /* ray/AABB slab tests */
	uint64_t tri_tests[4]; /* Moeller-Trumbore tests by exit stage: parallel / u / v / complete */
	uint64_t sphere_tests[2]; /* miss / hit */
	uint64_t plane_tests;
	uint64_t nodes_shaded; /* cast_ray invocations that found a hit */
	double kernel_ms;      /* device time of the trace kernel(s), CUDA events */
	double total_ms;       /* wall time of the call incl. copies */
	/* walk efficiency of the direct-light loop (instrumented instance): warp-level NODE / PRIM steps taken,
	 * the lanes that took part in them, and the lanes that had no shadow ray at the time (waiting for a
	 * refill); 32 * steps - lanes - idle = lanes whose ray only had work of the other kind */
	uint64_t node_steps, node_step_lanes, node_step_idle;
	uint64_t prim_steps, prim_step_lanes, prim_step_idle;
	/* the persistent grid's time line (instrumented instance, device clock): from the first warp's start
	 * until a warp first found the work queue empty, until the last warp ended, and the mean life of a warp --
	 * last_warp_ms - queue_ms is the tail, last_warp_ms - warp_busy_ms the mean idle time of a warp slot */
	double queue_ms, last_warp_ms, warp_busy_ms;
} crt_stats;

typedef struct crt_scene crt_scene; /* opaque; owns all device memory */

/* Information about the built acceleration structure. */
typedef struct crt_scene_info {
	uint32_t num_bounded; /* spheres + triangles (BVH leaves' primitives) */
	uint32_t num_planes;
	uint32_t num_emitters;
	uint32_t total_lights; /* sum of num_lights over emitters */
	uint32_t num_nodes;    /* BVH nodes (two child boxes per 64-byte node) */
	uint32_t max_depth;
	uint32_t device;       /* CUDA device ordinal */
	uint32_t sm_count;
	double build_ms;       /* host BVH build */
	double upload_ms;
	uint64_t device_bytes;
	uint32_t built_on_device; /* 1: CRT_BUILD_DEVICE was used (build_ms is device + sync time) */
	uint32_t reserved;
} crt_scene_info;

const char *crt_last_error(void);
int crt_abi_version(void);
int crt_device_count(void);

/* Build the BVH, flatten and upload the scene to CUDA device `device`.
 * Replaces accel_init() (accel.c:266-315) plus the globals hand-off. */
int crt_scene_create(const crt_scene_desc *desc, int device, crt_scene **out);
void crt_scene_destroy(crt_scene *scene);
int crt_scene_get_info(const crt_scene *scene, crt_scene_info *info);
/* Diagnostics: the acceleration structure as the kernels read it -- info.num_nodes records of 128 bytes
 * (c-raytracer_b200/csrc/bvh_build.h, WideNode) and info.num_bounded primitive records of 64 bytes in tree
 * order (v0.xyz eps | e0.xyz info | e1.xyz object | normal.xyz -; a sphere: centre eps | radius - - info).
 * Copies the first nodes_bytes / prims_bytes bytes. */
int crt_scene_read_tree(crt_scene *scene, void *nodes, size_t nodes_bytes, void *prims, size_t prims_bytes);

/* Replaces render() (render.c:345-368).  raster = W*H*3 floats, z_buffer =
 * W*H floats, HOST memory, row-major top-left origin, fully overwritten
 * (zero where nothing was hit: reference relies on fresh zero pages,
 * image.c:45-46).  stats may be NULL. */
int crt_render(crt_scene *scene, const crt_render_opts *opts,
	float *raster, float *z_buffer, crt_stats *stats);

/* Same, but raster / z_buffer are DEVICE pointers on the scene's device and the
 * work is enqueued on `cuda_stream` (a cudaStream_t, may be NULL for the
 * default stream) without synchronising.  stats (may be NULL) is filled only
 * by crt_render_device_sync().  ONE frame in flight per crt_scene: the work
 * counter, pending-ray stacks and timing events are the scene's, so a second
 * crt_render_device() before crt_render_device_sync() is ordered AFTER the first
 * (its stream waits for the first frame's end event, whatever stream that ran on)
 * and the first frame's statistics are lost; render several frames concurrently
 * from several scenes. */
int crt_render_device(crt_scene *scene, const crt_render_opts *opts,
	float *d_raster, float *d_z_buffer, void *cuda_stream);
int crt_render_device_sync(crt_scene *scene, crt_stats *stats);

/* ---- one frame rendered by several devices (or processes) -------------------------
 * The reference parallelises inside render() itself (render.c:349-352: OpenMP threads over
 * rows, thread count from -m, system.c:54-74).  Here several scenes -- replicas of the same
 * description on different devices -- drain ONE work queue: every resident warp of every
 * participant pulls the next work unit (8 pixels; 2 pixels for the last part of the frame)
 * from `work_counter` with a system-scope atomic and writes its pixels straight into the
 * common frame.  work_counter, d_raster and d_z_buffer must be visible to every
 * participant: memory of one device mapped into the others by peer access
 * (cudaDeviceEnablePeerAccess over NVLink) or CUDA IPC; the owner zeroes work_counter
 * before the participants enqueue the frame.  Load balance is dynamic down to the last
 * unit; the frame does not depend on which device renders which unit (counter-based RNG,
 * fixed-order sums), so it is bit-identical to the single-device frame. */
typedef struct crt_frame_share {
	unsigned int *work_counter; /* one 32-bit counter, zero at the start of the frame */
	uint32_t total_warps;       /* sum of crt_scene_resident_warps() over the participants */
	uint32_t reserved[5];
} crt_frame_share;

/* resident warps of the render kernel on the scene's device (the persistent grid) */
int crt_scene_resident_warps(crt_scene *scene, const crt_render_opts *opts, uint32_t *warps);
/* crt_render_device() as one participant of a shared frame (share == NULL: alone) */
int crt_render_device_shared(crt_scene *scene, const crt_render_opts *opts,
	float *d_raster, float *d_z_buffer, void *cuda_stream, const crt_frame_share *share);

/* Page-locks a host range the caller owns (the reference's image.raster / image.z_buffer,
 * image.c:70-77, stay malloc'd): crt_render()'s device-to-host copies into a registered range run
 * at the link's rate instead of through the driver's bounce buffer.  Worth it when the same
 * buffers receive several frames (registering costs about as much as one pageable copy:
 * BASELINE.md section 6).  Unregister before free(). */
int crt_host_register(void *ptr, size_t bytes);
int crt_host_unregister(void *ptr);

/* Memory of one device made visible to other PROCESSES on the same node (one process per GPU,
 * torchrun): the owner allocates and passes `handle` to the others (any byte transport), who map
 * it with crt_shared_open() -- CUDA IPC, peer access over NVLink.  Used for the common frame and
 * the work counter of crt_frame_share. */
#define CRT_IPC_HANDLE_BYTES 64
int crt_shared_alloc(int device, size_t bytes, void **d_ptr, unsigned char handle[CRT_IPC_HANDLE_BYTES]);
int crt_shared_open(int device, const unsigned char handle[CRT_IPC_HANDLE_BYTES], void **d_ptr);
int crt_shared_release(int device, void *d_ptr, int owner);

/* The same within one process: the scene replicated on `num_devices` CUDA devices (the BVH is
 * built once), frame and work counter on devices[0], the others reach them by peer access.
 * Without peer access the devices render static shards (crt_render_opts.shard_*) that are
 * merged on the host.  A device may be listed twice (two replicas on one GPU). */
typedef struct crt_multi crt_multi;
int crt_multi_create(const crt_scene_desc *desc, const int *devices, int num_devices, crt_multi **out);
void crt_multi_destroy(crt_multi *multi);
int crt_multi_num_devices(const crt_multi *multi);
int crt_multi_is_dynamic(const crt_multi *multi); /* 1: shared queue over peer access, 0: static shards */
int crt_multi_get_info(const crt_multi *multi, int index, crt_scene_info *info);
/* replaces render(): HOST buffers as crt_render(); device_ms (may be NULL) receives the
 * render-kernel time of every device */
int crt_multi_render(crt_multi *multi, const crt_render_opts *opts,
	float *raster, float *z_buffer, crt_stats *stats, double *device_ms);

/* Ray-batch seams (tests, tools).  rays = n * 6 floats: direction xyz then
 * point xyz, the layout of `struct Ray` (object.h:27-30).
 * crt_trace_closest replaces get_closest_intersection() (render.c:118-124):
 *   object[i] = scene-order object index or -1, t[i] = distance (FLT_MAX on
 *   miss), normal = n*3 floats (may be NULL).
 * crt_trace_shadow replaces is_light_blocked() (render.c:126-134):
 *   light = n*3 floats in/out (multiplied by kt of every transparent occluder),
 *   emitter[i] = scene-order object index of the sampled emitter (or -1),
 *   blocked[i] = 1 if an opaque occluder lies within dist[i]. */
int crt_trace_closest(crt_scene *scene, const float *rays, size_t n,
	int32_t *object, float *t, float *normal);
int crt_trace_shadow(crt_scene *scene, const float *rays, const float *dist,
	const int32_t *emitter, size_t n, float *light, uint8_t *blocked);

/* Evaluate material `material`'s texture at n points (n*3 floats) -> n*3
 * colours.  Replaces Texture.get_color (material.h:27-29). */
int crt_texture_eval(crt_scene *scene, int32_t material, const float *points,
	size_t n, float *colors);

/* FP32 FMA micro-benchmark on the scene's device: measured TFLOP/s, the
 * denominator of the ray-test roofline. */
int crt_measure_fp32_peak(int device, double *tflops);

/* ---- post-processing of a rendered frame ------------------------
 * Replaces postprocess() (reference src/postprocess/postproc.c:36-92) and the
 * three operations behind it: brighten() :94-102, depth_of_field() :104-160,
 * mist() :162-188.  They consume exactly what crt_render() produces (raster
 * and z-buffer), in the reference's fixed order brighten -> depth of field ->
 * mist, each applied when its bit is set in `ops`. */
enum crt_post_op {
	CRT_POST_BRIGHTEN = 1,   /* -b <factor> */
	CRT_POST_DOF = 2,        /* --dof <scale> <bias> */
	CRT_POST_DOF_CAMERA = 4, /* --dof-camera <aperture> <focal length> <plane in focus>; ignored when CRT_POST_DOF is set (postproc.c:47-52) */
	CRT_POST_MIST = 8        /* --mist <start> <depth> <falloff> <r> <g> <b> */
};

enum crt_post_falloff { /* enum Falloff, postproc.c:26-30 */
	CRT_FALLOFF_QUADRATIC = 0,
	CRT_FALLOFF_LINEAR = 1,
	CRT_FALLOFF_INV_QUADRATIC = 2
};

typedef struct crt_post_opts {
	uint32_t ops; /* crt_post_op bits */
	float brighten;
	float dof_scale, dof_bias;
	float dof_aperture, dof_focal_length, dof_plane_in_focus;
	float mist_start, mist_depth;
	uint32_t mist_falloff; /* crt_post_falloff */
	float mist_color[3];
	uint32_t reserved[3];
} crt_post_opts;

typedef struct crt_post_stats {
	double kernel_ms;   /* all post-processing kernels, CUDA events */
	double total_ms;    /* including copies (crt_postprocess only) */
	float dof_scale, dof_bias; /* the values depth_of_field() ran with (derived for --dof-camera) */
	uint32_t dof_max_radius;   /* largest circle-of-confusion radius in pixels */
	uint32_t launches;         /* kernels launched */
} crt_post_stats;

/* raster = W*H*3 floats in/out, z_buffer = W*H floats in; HOST memory. */
int crt_postprocess(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *raster, const float *z_buffer, crt_post_stats *stats);
/* Same on DEVICE pointers of `device` (the buffers crt_render_device() filled),
 * enqueued on `cuda_stream`; synchronises before returning only when stats is
 * not NULL or CRT_POST_DOF_CAMERA needs the z range on the host. */
int crt_postprocess_device(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *d_raster, const float *d_z_buffer, void *cuda_stream, crt_post_stats *stats);

/* NO COUNTERPART IN THE REFERENCE: added for callers that
 * must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The depth-of-field path of crt_postprocess_device() -- CRT_POST_DOF must be set,
 * CRT_POST_BRIGHTEN / CRT_POST_MIST may be, CRT_POST_DOF_CAMERA is ignored as there -- with the
 * same kernels and bit-identical output, but its scratch (per-pixel weighted colours and radii,
 * per-segment and per-row radius maxima) lives in ONE fixed device array of
 * CRT_POST_NOALLOC_SCRATCH_BYTES inside the library, which every GPU that loads libcrt.so holds
 * until the process exits.  The call makes no cudaMalloc / cudaFree / event or stream creation and
 * keeps nothing per call; inputs and outputs are the caller's device pointers, as above.
 * Limits: width * height <= CRT_POST_NOALLOC_MAX_PIXELS (1920x1080 fits) and
 * 18 bytes per pixel + 2 per 32-pixel row segment + 4 per row <= CRT_POST_NOALLOC_SCRATCH_BYTES;
 * larger frames are refused with CRT_ERR_UNSUPPORTED and a message naming the limit.
 * The array is shared by all calls on a device and keeps the last call's contents (the call
 * clears the part that must start at zero): calls that overlap on the GPU -- different streams
 * with no ordering between them -- race on it; calls on one stream are safe.  Like the existing
 * path it waits for `cuda_stream` once (the largest radius is checked on the host). */
#define CRT_POST_NOALLOC_MAX_PIXELS (1u << 21)
#define CRT_POST_NOALLOC_SCRATCH_BYTES (37u << 20)
int crt_postprocess_device_noalloc(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *d_raster, const float *d_z_buffer, void *cuda_stream);

#ifdef __cplusplus
}
#endif

#endif /* CRT_H */
