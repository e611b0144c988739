/*
 * crt_api.cu -- implementation of the C ABI in include/crt.h on top of the
 * kernels in crt_kernels.cu.  No CPU fallback: without a CUDA device every
 * compute entry point fails with CRT_ERR_NO_DEVICE.
 */
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <vector>

#include "bvh_build.h"
#include "crt_bvh_device.cuh"
#include "crt_kernels.cuh"

using namespace crt;

namespace {

thread_local char g_error[512] = "";

constexpr uint32_t kMaxBounces = 4096;       /* pending-ray stack depth is max_bounces + 4 per lane */
constexpr uint32_t kMaxLights = 0x3fffffffu; /* light-sample indices and draw counters are 32-bit (2 draws per sample) */

int fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_error, sizeof g_error, fmt, ap);
	va_end(ap);
	return code;
}

#define CUDA_TRY(expr)                                                                          \
	do {                                                                                    \
		cudaError_t e_ = (expr);                                                        \
		if (e_ != cudaSuccess)                                                          \
			return fail(CRT_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); \
	} while (0)

double now_ms()
{
	using namespace std::chrono;
	return duration<double, std::milli>(steady_clock::now().time_since_epoch()).count();
}

float as_float(uint32_t u)
{
	float f;
	std::memcpy(&f, &u, 4);
	return f;
}

} // namespace

int crt::api_fail(int code, const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_error, sizeof g_error, fmt, ap);
	va_end(ap);
	return code;
}

struct crt_scene {
	int device = 0;
	int sm_count = 0;
	DScene d{};
	crt_scene_info info{};
	std::vector<void *> allocs;
	std::vector<int32_t> slot_object; /* host copy: prim slot -> scene object (for the ray-batch seams) */
	uint32_t num_materials = 0;
	uint32_t wide_depth = 0;
	uint32_t *deep = nullptr;  /* spill strips of the walk's inner stacks, only for trees deeper than kInnerStack */
	size_t deep_words = 0;
	/* per-frame scratch, grown on demand */
	float *row_pos = nullptr;
	size_t row_pos_floats = 0;
	float *pending = nullptr;
	size_t pending_floats = 0;
	unsigned int *tile_counter = nullptr;
	unsigned long long *counters = nullptr;
	float *d_raster = nullptr, *d_z = nullptr; /* staging for crt_render (host buffers) */
	size_t staging_pixels = 0;
	cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
	cudaStream_t last_stream = nullptr;
	bool frame_pending = false;
	bool frame_counted = false;
	uint32_t frame_warps = 0;
	size_t wide_bytes = 0;   /* size of d.wide: the window of the optional L2 persistence (CRT_L2_PERSIST_MB) */
	size_t persist_bytes = 0;
};

namespace {

template <typename T>
int upload(crt_scene *s, const std::vector<T> &host, const T **dev)
{
	*dev = nullptr;
	if (host.empty())
		return CRT_OK;
	void *p = nullptr;
	CUDA_TRY(cudaMalloc(&p, host.size() * sizeof(T)));
	s->allocs.push_back(p);
	s->info.device_bytes += host.size() * sizeof(T);
	CUDA_TRY(cudaMemcpy(p, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice));
	*dev = static_cast<const T *>(p);
	return CRT_OK;
}

constexpr float kMaxCoordinate = 1e12f;

int validate(const crt_scene_desc *desc)
{
	if (!desc)
		return fail(CRT_ERR_INVALID, "scene description is NULL");
	if (desc->abi_version != CRT_ABI_VERSION)
		return fail(CRT_ERR_INVALID, "ABI version mismatch: caller %u, library %d", desc->abi_version, CRT_ABI_VERSION);
	if (!desc->num_objects || !desc->objects)
		return fail(CRT_ERR_INVALID, "scene has no objects");
	if (!desc->num_materials || !desc->materials)
		return fail(CRT_ERR_INVALID, "scene has no materials");
	if (desc->num_materials > PRIM_MATERIAL_MASK)
		return fail(CRT_ERR_INVALID, "too many materials");
	for (uint32_t i = 0; i < desc->num_objects; i++) {
		const crt_object &o = desc->objects[i];
		if (o.type != CRT_OBJ_SPHERE && o.type != CRT_OBJ_TRIANGLE && o.type != CRT_OBJ_PLANE)
			return fail(CRT_ERR_INVALID, "object %u has unknown type %d", i, o.type);
		if (o.material < 0 || static_cast<uint32_t>(o.material) >= desc->num_materials)
			return fail(CRT_ERR_INVALID, "object %u references material index %d of %u", i, o.material, desc->num_materials);
		/* the tree builders form surface areas and Morton codes in FP32: coordinates must be finite and leave room for
		 * their squares (a = centre / vertex / normal, b, c = radius / edges / plane offset) */
		for (int k = 0; k < 3; k++) {
			const float v[3] = { o.a[k], o.b[k], o.c[k] };
			for (int j = 0; j < (o.type == CRT_OBJ_TRIANGLE ? 3 : (k == 0 ? 2 : 1)); j++)
				if (!(std::fabs(v[j]) <= kMaxCoordinate))
					return fail(CRT_ERR_UNSUPPORTED, "object %u has a coordinate that is not finite or beyond %g", i, (double)kMaxCoordinate);
		}
	}
	return CRT_OK;
}

uint32_t material_flags(const crt_material &m)
{
	if (m.flags != 0xffffffffu)
		return m.flags;
	auto mag = [](const float *k) { return std::sqrt(k[0] * k[0] + k[1] * k[1] + k[2] * k[2]); };
	return (mag(m.kr) > 1e-6f ? CRT_MAT_REFLECTIVE : 0u) | (mag(m.kt) > 1e-6f ? CRT_MAT_TRANSPARENT : 0u) |
		(mag(m.ke) > 1e-6f ? CRT_MAT_EMITTANT : 0u);
}

int ensure_device(int device)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
		cudaGetLastError();
		return fail(CRT_ERR_NO_DEVICE, "no CUDA device available (this library has no CPU fallback)");
	}
	if (device < 0 || device >= n)
		return fail(CRT_ERR_INVALID, "device %d out of range (%d devices)", device, n);
	CUDA_TRY(cudaSetDevice(device));
	return CRT_OK;
}

} // namespace

int crt::api_ensure_device(int device)
{
	return ensure_device(device);
}

extern "C" {

const char *crt_last_error(void)
{
	return g_error;
}

int crt_abi_version(void)
{
	return CRT_ABI_VERSION;
}

int crt_device_count(void)
{
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) {
		cudaGetLastError();
		return 0;
	}
	return n;
}

void crt_scene_destroy(crt_scene *s)
{
	if (!s)
		return;
	cudaSetDevice(s->device);
	for (void *p : s->allocs)
		cudaFree(p);
	cudaFree(s->row_pos);
	cudaFree(s->pending);
	cudaFree(s->deep);
	cudaFree(s->tile_counter);
	cudaFree(s->counters);
	cudaFree(s->d_raster);
	cudaFree(s->d_z);
	if (s->ev_start)
		cudaEventDestroy(s->ev_start);
	if (s->ev_stop)
		cudaEventDestroy(s->ev_stop);
	delete s;
}

/* The scene as the device wants it, built once on the host (accel_init(), accel.c:266-315) and
 * uploaded to one or several devices. */
struct HostScene {
	std::vector<float4> wide_nodes, prims, planes;
	std::vector<int32_t> obj_slot, slot_object;
	std::vector<DEmitter> emitters;
	std::vector<crt_material> mats;
	uint32_t num_objects = 0, num_bounded = 0, total_lights = 0, wide_nodes_count = 0, wide_depth = 0;
	float ambient[3] = { 0.f, 0.f, 0.f };
	double build_ms = 0.0;
	bool device_build = false; /* prims are in scene order, no tree: crt_bvh_device.cu builds it after the upload */
};

static int build_host_scene(const crt_scene_desc *desc, HostScene &h)
{
	int rc = validate(desc);
	if (rc)
		return rc;

	/* ---- materials ------------------------------------------------------ */
	h.mats.assign(desc->materials, desc->materials + desc->num_materials);
	for (auto &m : h.mats)
		m.flags = material_flags(m);
	const std::vector<crt_material> &mats = h.mats;

	/* ---- split objects: planes (linear list, object.c:168-197) / bounded --- */
	const uint32_t n = desc->num_objects;
	std::vector<uint32_t> bounded;
	h.obj_slot.assign(n, 0);
	uint32_t total_lights = 0;
	for (uint32_t i = 0; i < n; i++) {
		const crt_object &o = desc->objects[i];
		const crt_material &m = mats[o.material];
		const uint32_t info = static_cast<uint32_t>(o.material) | ((m.flags & CRT_MAT_TRANSPARENT) ? PRIM_TRANSPARENT : 0u);
		if (o.type == CRT_OBJ_PLANE) {
			if (m.flags & CRT_MAT_EMITTANT)
				return fail(CRT_ERR_INVALID, "Plane cannot be emittant");
			h.obj_slot[i] = ~static_cast<int32_t>(h.planes.size() / 2);
			h.planes.push_back(make_float4(o.a[0], o.a[1], o.a[2], o.b[0]));
			h.planes.push_back(make_float4(o.epsilon, as_float(info), as_float(i), 0.f));
			continue;
		}
		bounded.push_back(i);
		/* emitters: bounded objects with an emittant material, in scene order
		 * (scene.c:357-358); lights == 0 emits nothing onto others */
		if (m.flags & CRT_MAT_EMITTANT) {
			DEmitter e{};
			e.object = static_cast<int32_t>(i);
			e.num_lights = o.num_lights;
			e.first_light = total_lights;
			e.is_sphere = o.type == CRT_OBJ_SPHERE;
			if (e.is_sphere) {
				e.a = make_float4(o.a[0], o.a[1], o.a[2], o.b[0]);
			} else {
				e.a = make_float4(o.a[0], o.a[1], o.a[2], 0.f);
				e.b = make_float4(o.b[0], o.b[1], o.b[2], 0.f);
				e.c = make_float4(o.c[0], o.c[1], o.c[2], 0.f);
			}
			/* an emitter with no samples lights nothing: keep ke_n finite */
			const float inv = o.num_lights ? 1.f / static_cast<float>(o.num_lights) : 0.f;
			for (int k = 0; k < 3; k++)
				e.ke_n[k] = m.ke[k] * inv;
			if (static_cast<uint64_t>(total_lights) + o.num_lights > kMaxLights)
				return fail(CRT_ERR_INVALID, "more than %u light samples in the scene", kMaxLights);
			total_lights += o.num_lights;
			h.emitters.push_back(e);
		}
	}

	/* ---- per emitter: the box of all other bounded objects (DEmitter.rest_lo / rest_hi) ------- */
	if (!h.emitters.empty()) {
		struct Box {
			float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
			void add(const crt_object &o)
			{
				for (int a = 0; a < 3; a++) {
					float l, u;
					if (o.type == CRT_OBJ_SPHERE) {
						l = o.a[a] - std::fabs(o.b[0]);
						u = o.a[a] + std::fabs(o.b[0]);
					} else {
						const float p0 = o.a[a], p1 = o.a[a] + o.b[a], p2 = o.a[a] + o.c[a];
						l = std::fmin(p0, std::fmin(p1, p2));
						u = std::fmax(p0, std::fmax(p1, p2));
					}
					lo[a] = std::fmin(lo[a], l);
					hi[a] = std::fmax(hi[a], u);
				}
			}
		};
		/* rest(e) = (bounded objects before e) U (bounded objects after e): one pass each way */
		std::vector<Box> before(h.emitters.size()), after(h.emitters.size());
		Box run;
		size_t e = 0;
		for (size_t k = 0; k < bounded.size(); k++) {
			if (e < h.emitters.size() && h.emitters[e].object == static_cast<int32_t>(bounded[k]))
				before[e++] = run;
			run.add(desc->objects[bounded[k]]);
		}
		run = Box();
		e = h.emitters.size();
		for (size_t k = bounded.size(); k-- > 0;) {
			if (e > 0 && h.emitters[e - 1].object == static_cast<int32_t>(bounded[k]))
				after[--e] = run;
			run.add(desc->objects[bounded[k]]);
		}
		for (size_t i = 0; i < h.emitters.size(); i++) {
			DEmitter &em = h.emitters[i];
			for (int a = 0; a < 3; a++) {
				const float l = std::fmin(before[i].lo[a], after[i].lo[a]), u = std::fmax(before[i].hi[a], after[i].hi[a]);
				if (!(l <= u)) { /* nothing else is bounded (or a NaN coordinate: then never cull) */
					em.rest_lo[a] = l == l && u == u ? 1.f : -FLT_MAX;
					em.rest_hi[a] = l == l && u == u ? -1.f : FLT_MAX;
					continue;
				}
				/* padded like the tree's own boxes, and by the largest epsilon-free margin a hit test could need */
				const float pad = 1e-5f * std::fmax(1e-3f, std::fmax(std::fabs(l), std::fabs(u))) + 1e-5f * (u - l);
				em.rest_lo[a] = l - pad;
				em.rest_hi[a] = u + pad;
			}
			if (em.rest_lo[0] > em.rest_hi[0] || em.rest_lo[1] > em.rest_hi[1] || em.rest_lo[2] > em.rest_hi[2]) {
				em.rest_lo[0] = 1.f; /* the device tests axis 0 for "empty" */
				em.rest_hi[0] = -1.f;
			}
		}
	}

	/* ---- BVH over the bounded objects ------------------------------------- */
	const double t0 = now_ms();
	h.device_build = desc->build == CRT_BUILD_DEVICE;
	if (const char *e = std::getenv("CRT_BVH_BUILD")) /* tools / tests: override the caller's choice */
		h.device_build = std::strcmp(e, "device") == 0;
	if (bounded.size() > kWideMaxPrims)
		return fail(CRT_ERR_UNSUPPORTED, "scene too large for the device tree (%zu primitives)", bounded.size());
	std::vector<Aabb> boxes(h.device_build ? 0 : bounded.size());
	for (size_t k = 0; k < boxes.size(); k++) {
		const crt_object &o = desc->objects[bounded[k]];
		Aabb &b = boxes[k];
		for (int a = 0; a < 3; a++) {
			if (o.type == CRT_OBJ_SPHERE) {
				b.lo[a] = o.a[a] - o.b[0];
				b.hi[a] = o.a[a] + o.b[0];
			} else {
				const float p0 = o.a[a], p1 = o.a[a] + o.b[a], p2 = o.a[a] + o.c[a];
				b.lo[a] = std::fmin(p0, std::fmin(p1, p2));
				b.hi[a] = std::fmax(p0, std::fmax(p1, p2));
			}
			/* pad by a few ulps: the slab test must never cull a true hit */
			const float pad = 4e-7f * std::fmax(1e-3f, std::fmax(std::fabs(b.lo[a]), std::fabs(b.hi[a]))) + 1e-7f * (b.hi[a] - b.lo[a]);
			b.lo[a] -= pad;
			b.hi[a] += pad;
		}
	}
	BvhResult bvh;
	WideTree wide;
	if (!h.device_build) {
		build_bvh(boxes, bvh);
		/* the device tree: 8-wide collapse of the binary tree, one primitive per leaf slot, half-precision child boxes */
		build_wide(boxes, bvh, wide);
	}
	if (wide.max_depth > kWideMaxDepth || wide.nodes.size() > kWideMaxNodes || bounded.size() > kWideMaxPrims)
		return fail(CRT_ERR_UNSUPPORTED, "scene too large for the device tree (%zu nodes, %zu primitives, %u levels)",
			wide.nodes.size(), bounded.size(), wide.max_depth);
	h.build_ms = now_ms() - t0;

	/* ---- primitive records in the order of the wide tree -------------------- */
	h.prims.resize(4 * bounded.size());
	h.slot_object.resize(bounded.size());
	for (size_t slot = 0; slot < bounded.size(); slot++) {
		const uint32_t oi = bounded[h.device_build ? slot : wide.order[slot]];
		const crt_object &o = desc->objects[oi];
		const crt_material &m = mats[o.material];
		uint32_t info = static_cast<uint32_t>(o.material) | ((m.flags & CRT_MAT_TRANSPARENT) ? PRIM_TRANSPARENT : 0u);
		h.obj_slot[oi] = static_cast<int32_t>(slot);
		h.slot_object[slot] = static_cast<int32_t>(oi);
		if (o.type == CRT_OBJ_SPHERE) {
			info |= PRIM_SPHERE;
			h.prims[4 * slot + 0] = make_float4(o.a[0], o.a[1], o.a[2], o.epsilon);
			h.prims[4 * slot + 1] = make_float4(o.b[0], 0.f, 0.f, as_float(info));
			h.prims[4 * slot + 2] = make_float4(0.f, 0.f, 0.f, as_float(oi));
			h.prims[4 * slot + 3] = make_float4(0.f, 0.f, 0.f, 0.f);
		} else {
			h.prims[4 * slot + 0] = make_float4(o.a[0], o.a[1], o.a[2], o.epsilon);
			h.prims[4 * slot + 1] = make_float4(o.b[0], o.b[1], o.b[2], as_float(info));
			h.prims[4 * slot + 2] = make_float4(o.c[0], o.c[1], o.c[2], as_float(oi));
			h.prims[4 * slot + 3] = make_float4(o.n[0], o.n[1], o.n[2], 0.f);
		}
	}
	h.wide_nodes.resize(8 * wide.nodes.size());
	if (!wide.nodes.empty())
		std::memcpy(h.wide_nodes.data(), wide.nodes.data(), wide.nodes.size() * sizeof(WideNode));
	h.num_objects = n;
	h.num_bounded = static_cast<uint32_t>(bounded.size());
	h.total_lights = total_lights;
	h.wide_nodes_count = static_cast<uint32_t>(wide.nodes.size());
	h.wide_depth = wide.max_depth;
	std::memcpy(h.ambient, desc->ambient, sizeof h.ambient);
	return CRT_OK;
}

/* plain H2D copies of a built scene onto `device` */
static int upload_host_scene(const HostScene &h, int device, crt_scene **out)
{
	int rc = ensure_device(device);
	if (rc)
		return rc;
	const double t1 = now_ms();
	crt_scene *s = new crt_scene();
	s->device = device;
	s->num_materials = static_cast<uint32_t>(h.mats.size());
	s->slot_object = h.slot_object;
	cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device);
	rc = upload(s, h.wide_nodes, &s->d.wide);
	s->wide_bytes = h.wide_nodes.size() * sizeof(float4);
	if (!rc) rc = upload(s, h.prims, &s->d.prims);
	if (!rc) rc = upload(s, h.planes, &s->d.planes);
	if (!rc) rc = upload(s, h.obj_slot, &s->d.obj_slot);
	uint32_t nodes_count = h.wide_nodes_count, depth = h.wide_depth;
	double build_ms = h.build_ms;
	if (!rc && h.device_build && h.num_bounded) {
		/* Morton codes, sort, radix tree, boxes, 8-wide collapse and the permutation of the records: all here */
		const double tb = now_ms();
		float4 *sorted = nullptr;
		if (cudaMalloc(&sorted, h.prims.size() * sizeof(float4)) != cudaSuccess)
			rc = fail(CRT_ERR_CUDA, "cudaMalloc(primitive records) failed");
		DeviceTree tree;
		if (!rc) {
			s->allocs.push_back(sorted);
			rc = build_device_tree(s->d.prims, h.num_bounded, const_cast<int32_t *>(s->d.obj_slot), sorted, tree);
		}
		if (!rc) {
			s->allocs.push_back(tree.wide);
			s->d.wide = tree.wide;
			s->wide_bytes = static_cast<size_t>(tree.num_nodes) * sizeof(WideNode);
			s->info.device_bytes += s->wide_bytes;
			s->d.prims = sorted; /* the scene-order copy stays in allocs and goes with the scene */
			nodes_count = tree.num_nodes;
			depth = tree.max_depth;
			build_ms += now_ms() - tb;
			s->info.built_on_device = 1;
		}
	}
	if (!rc) rc = upload(s, h.emitters, &s->d.emitters);
	if (!rc) rc = upload(s, h.mats, &s->d.materials);
	if (!rc && cudaMalloc(&s->tile_counter, sizeof(unsigned int)) != cudaSuccess)
		rc = fail(CRT_ERR_CUDA, "cudaMalloc(tile counter) failed");
	if (!rc && cudaMalloc(&s->counters, kTotalCounters * sizeof(unsigned long long)) != cudaSuccess)
		rc = fail(CRT_ERR_CUDA, "cudaMalloc(counters) failed");
	if (!rc && (cudaEventCreate(&s->ev_start) != cudaSuccess || cudaEventCreate(&s->ev_stop) != cudaSuccess))
		rc = fail(CRT_ERR_CUDA, "cudaEventCreate failed");
	if (rc) {
		crt_scene_destroy(s);
		return rc;
	}
	s->d.num_nodes = nodes_count;
	s->wide_depth = depth;
	s->d.num_prims = h.num_bounded;
	s->d.num_planes = static_cast<uint32_t>(h.planes.size() / 2);
	s->d.num_emitters = static_cast<uint32_t>(h.emitters.size());
	s->d.num_objects = h.num_objects;
	s->d.total_lights = h.total_lights;
	std::memcpy(s->d.ambient, h.ambient, sizeof s->d.ambient);

	s->info.build_ms = build_ms;
	s->info.upload_ms = now_ms() - t1 - (build_ms - h.build_ms);
	s->info.num_bounded = s->d.num_prims;
	s->info.num_planes = s->d.num_planes;
	s->info.num_emitters = s->d.num_emitters;
	s->info.total_lights = h.total_lights;
	s->info.num_nodes = s->d.num_nodes;
	s->info.max_depth = depth;
	s->info.device = static_cast<uint32_t>(device);
	s->info.sm_count = static_cast<uint32_t>(s->sm_count);
	*out = s;
	return CRT_OK;
}

int crt_scene_create(const crt_scene_desc *desc, int device, crt_scene **out)
{
	if (!out)
		return fail(CRT_ERR_INVALID, "out is NULL");
	*out = nullptr;
	int rc = validate(desc);
	if (rc)
		return rc;
	rc = ensure_device(device);
	if (rc)
		return rc;
	HostScene h;
	rc = build_host_scene(desc, h);
	if (rc)
		return rc;
	return upload_host_scene(h, device, out);
}

int crt_scene_read_tree(crt_scene *s, void *nodes, size_t nodes_bytes, void *prims, size_t prims_bytes)
{
	if (!s || (nodes_bytes && !nodes) || (prims_bytes && !prims))
		return fail(CRT_ERR_INVALID, "NULL argument");
	const size_t have_nodes = static_cast<size_t>(s->d.num_nodes) * sizeof(WideNode), have_prims = static_cast<size_t>(s->d.num_prims) * 64u;
	if (nodes_bytes > have_nodes || prims_bytes > have_prims)
		return fail(CRT_ERR_INVALID, "the tree has %zu bytes of nodes and %zu bytes of primitive records", have_nodes, have_prims);
	CUDA_TRY(cudaSetDevice(s->device));
	if (nodes_bytes)
		CUDA_TRY(cudaMemcpy(nodes, s->d.wide, nodes_bytes, cudaMemcpyDeviceToHost));
	if (prims_bytes)
		CUDA_TRY(cudaMemcpy(prims, s->d.prims, prims_bytes, cudaMemcpyDeviceToHost));
	return CRT_OK;
}

int crt_scene_get_info(const crt_scene *s, crt_scene_info *info)
{
	if (!s || !info)
		return fail(CRT_ERR_INVALID, "NULL argument");
	*info = s->info;
	return CRT_OK;
}

int crt_render_device(crt_scene *s, const crt_render_opts *opts, float *d_raster, float *d_z, void *cuda_stream)
{
	return crt_render_device_shared(s, opts, d_raster, d_z, cuda_stream, nullptr);
}

/* lens or fog asked for: the kernel instance that has them (crt_kernels.cu, EXT) */
static bool uses_extensions(const crt_render_opts &o)
{
	return (o.lens_radius > 0.f && o.lens_samples > 1u) || o.fog_density > 0.f;
}

int crt_scene_resident_warps(crt_scene *s, const crt_render_opts *opts, uint32_t *warps)
{
	if (!s || !warps)
		return fail(CRT_ERR_INVALID, "NULL argument");
	CUDA_TRY(cudaSetDevice(s->device));
	const int per_sm = render_blocks_per_sm(opts && (opts->flags & CRT_RENDER_COUNT_STATS), opts && uses_extensions(*opts));
	if (per_sm <= 0)
		return fail(CRT_ERR_CUDA, "render kernel cannot be resident (%s)", cudaGetErrorString(cudaGetLastError()));
	*warps = static_cast<uint32_t>(per_sm * s->sm_count * kWarpsPerBlock);
	return CRT_OK;
}

int crt_render_device_shared(crt_scene *s, const crt_render_opts *opts, float *d_raster, float *d_z, void *cuda_stream,
	const crt_frame_share *share)
{
	if (!s || !opts || !d_raster || !d_z)
		return fail(CRT_ERR_INVALID, "NULL argument");
	if (share && (!share->work_counter || !share->total_warps || opts->shard_count != 1u))
		return fail(CRT_ERR_INVALID, "shared frame: needs a work counter, the total warp count and an unsharded frame");
	if (!opts->width || !opts->height)
		return fail(CRT_ERR_INVALID, "empty image");
	if (!opts->shard_count || opts->shard_index >= opts->shard_count)
		return fail(CRT_ERR_INVALID, "bad shard %u of %u", opts->shard_index, opts->shard_count);
	if (opts->gi_model == CRT_GI_PATH && opts->samples_per_pixel > 0x7fffffffu)
		return fail(CRT_ERR_INVALID, "too many samples per pixel");
	if (opts->max_bounces > kMaxBounces)
		return fail(CRT_ERR_INVALID, "max_bounces %u exceeds the supported %u", opts->max_bounces, kMaxBounces);
	if (!(opts->lens_radius >= 0.f) || !std::isfinite(opts->lens_radius) || opts->lens_samples > 65536u ||
		(opts->lens_radius > 0.f && opts->lens_samples > 1u && !(opts->focus_distance > 0.f && std::isfinite(opts->focus_distance))))
		return fail(CRT_ERR_INVALID, "thin lens: needs a finite radius >= 0, a focus distance > 0 and at most 65536 samples");
	if (!(opts->fog_density >= 0.f) || !std::isfinite(opts->fog_density))
		return fail(CRT_ERR_INVALID, "fog: needs a finite density >= 0");
	/* the per-frame scratch (work counter, pending stacks, contribution slots, events) belongs to
	 * the scene: one frame in flight per crt_scene (see the stream wait below) */
	CUDA_TRY(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	/* a frame enqueued earlier and not yet synchronised: this one starts after it, whatever the stream */
	if (s->frame_pending)
		CUDA_TRY(cudaStreamWaitEvent(stream, s->ev_stop, 0));
	const size_t pixels = static_cast<size_t>(opts->width) * opts->height;
	const bool count = (opts->flags & CRT_RENDER_COUNT_STATS) != 0;

	/* persistent grid: every resident warp slot of the device */
	const bool ext = uses_extensions(*opts);
	const int per_sm = render_blocks_per_sm(count, ext);
	if (per_sm <= 0)
		return fail(CRT_ERR_CUDA, "render kernel cannot be resident (%s)", cudaGetErrorString(cudaGetLastError()));
	const int grid = per_sm * s->sm_count;

	if (s->row_pos_floats < 3 * pixels) {
		cudaFree(s->row_pos);
		s->row_pos = nullptr;
		s->row_pos_floats = 0;
		CUDA_TRY(cudaMalloc(&s->row_pos, 3 * pixels * sizeof(float)));
		s->row_pos_floats = 3 * pixels;
	}
	RenderParams p{};
	p.pending_depth = opts->max_bounces + 4u;
	const size_t slots_floats = static_cast<size_t>(grid) * kWarpsPerBlock * 3 * kSlots;
	const size_t pending_floats = static_cast<size_t>(grid) * kWarpsPerBlock * p.pending_depth * kPendingWords * 32 + slots_floats;
	if (s->pending_floats < pending_floats) {
		cudaFree(s->pending);
		s->pending = nullptr;
		s->pending_floats = 0;
		CUDA_TRY(cudaMalloc(&s->pending, pending_floats * sizeof(float)));
		s->pending_floats = pending_floats;
	}

	if (s->wide_depth > static_cast<uint32_t>(kInnerStack)) {
		p.deep_depth = s->wide_depth - kInnerStack;
		const size_t words = static_cast<size_t>(grid) * kWarpsPerBlock * p.deep_depth * 32;
		if (s->deep_words < words) {
			cudaFree(s->deep);
			s->deep = nullptr;
			s->deep_words = 0;
			CUDA_TRY(cudaMalloc(&s->deep, words * sizeof(uint32_t)));
			s->deep_words = words;
		}
		p.deep = s->deep;
	}

	p.sc = s->d;
	p.o = *opts;
	p.raster = d_raster;
	p.z_buffer = d_z;
	p.row_pos = s->row_pos;
	p.tiles_x = (opts->width + kTileW - 1) / kTileW;
	p.tiles_y = (opts->height + kTileH - 1) / kTileH;
	const uint32_t tiles = p.tiles_x * p.tiles_y;
	const uint32_t chunks = (tiles + kChunkTiles - 1) / kChunkTiles;
	const uint32_t local_chunks = chunks / opts->shard_count + (opts->shard_index < chunks % opts->shard_count ? 1u : 0u);
	p.num_local_tiles = local_chunks * kChunkTiles;
	p.refill = kRefill;
	if (const char *e = std::getenv("CRT_REFILL"))
		p.refill = static_cast<uint32_t>(std::min(32, std::max(1, std::atoi(e))));
	/* work units (crt_kernels.cu): a few neighbouring pixels each -- bounds the tail of the persistent
	 * grid and keeps a warp's shading points close together (whole 8x4 tiles were never faster) */
	{
		/* a shared frame is drained by the resident warps of every participant */
		const uint32_t warps = share ? share->total_warps : static_cast<uint32_t>(grid) * kWarpsPerBlock;
		/* lanes per pixel: part of the frame's arithmetic (the order its GI samples are summed in),
		 * so it depends on the options alone -- never on the device count or the unit shape */
		uint32_t gi_lanes = CRT_GI_LANES;
		if (const char *e = std::getenv("CRT_GI_LANES"))
			gi_lanes = static_cast<uint32_t>(std::atoi(e));
		if (gi_lanes != 1u && gi_lanes != 2u && gi_lanes != 4u && gi_lanes != 8u && gi_lanes != 16u && gi_lanes != 32u)
			gi_lanes = CRT_GI_LANES;
		const bool split = opts->gi_model == CRT_GI_PATH && opts->samples_per_pixel > 1u && opts->max_bounces > 0u;
		p.gshift = 0;
		if (split)
			while ((1u << p.gshift) < gi_lanes)
				p.gshift++;
		const uint32_t tile_pixels = kTileW * kTileH;
		p.fine_pixels = std::min<uint32_t>(kTileW, tile_pixels >> p.gshift);
		/* units of s pixels while more than s * taper pixels per resident warp remain: a unit costs up to
		 * ~16 times the mean of its pixels, and whatever starts late must be small (crt_kernels.cu) */
		uint32_t taper = CRT_TAPER_PIXELS;
		if (const char *e = std::getenv("CRT_TAPER_PIXELS"))
			taper = static_cast<uint32_t>(std::max(0, std::atoi(e)));
		const uint32_t T = p.num_local_tiles;
		uint32_t first_tile = 0, first_unit = 0;
		for (int i = 0; i < kPhases; i++) {
			const uint32_t size = p.fine_pixels >> i;
			if (!size) {
				p.phase_unit[i] = 0xffffffffu;
				p.phase_tile[i] = T;
				continue;
			}
			p.phase_unit[i] = first_unit;
			p.phase_tile[i] = first_tile;
			/* tiles left when this phase ends */
			uint64_t left = size > 1u ? (static_cast<uint64_t>(size) * taper * warps + tile_pixels - 1) / tile_pixels : 0u;
			left = std::min<uint64_t>(left, T - first_tile);
			const uint32_t end_tile = T - static_cast<uint32_t>(left);
			first_unit += (end_tile - first_tile) * (tile_pixels / size);
			first_tile = end_tile;
		}
		p.num_units = first_unit;
	}
	p.tile_counter = share ? share->work_counter : s->tile_counter;
	p.system_counter = share ? 1u : 0u;
	p.counters = s->counters;
	p.pending = s->pending;
	p.slots = s->pending + (pending_floats - slots_floats);

	/* experiment (off by default): keep the shadow walk's nodes resident in a set-aside part of
	 * L2 -- the working set is 86 MB against 126 MB of L2 shared by two dies */
	if (const char *e = std::getenv("CRT_L2_PERSIST_MB")) {
		const size_t want = static_cast<size_t>(std::max(0, std::atoi(e))) << 20;
		if (want && s->wide_bytes) {
			if (s->persist_bytes != want) {
				CUDA_TRY(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want));
				s->persist_bytes = want;
			}
			cudaStreamAttrValue attr{};
			attr.accessPolicyWindow.base_ptr = const_cast<float4 *>(s->d.wide);
			attr.accessPolicyWindow.num_bytes = s->wide_bytes;
			attr.accessPolicyWindow.hitRatio = 1.f;
			attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
			attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
			CUDA_TRY(cudaStreamSetAttribute(stream, cudaStreamAttributeAccessPolicyWindow, &attr));
		}
	}

	if (!share)
		CUDA_TRY(cudaMemsetAsync(s->tile_counter, 0, sizeof(unsigned int), stream));
	CUDA_TRY(cudaMemsetAsync(s->counters, 0, kTotalCounters * sizeof(unsigned long long), stream));
	/* pixels of other shards (and misses) are zero: the reference relies on fresh zero pages */
	if (!share) { /* the participants of a shared frame write every pixel between them */
		CUDA_TRY(cudaMemsetAsync(d_raster, 0, 3 * pixels * sizeof(float), stream));
		CUDA_TRY(cudaMemsetAsync(d_z, 0, pixels * sizeof(float), stream));
	}
	CUDA_TRY(cudaEventRecord(s->ev_start, stream));
	launch_raygen_rows(*opts, s->row_pos, stream);
	launch_render(p, grid, count, ext, stream);
	CUDA_TRY(cudaEventRecord(s->ev_stop, stream));
	CUDA_TRY(cudaGetLastError());
	s->last_stream = stream;
	s->frame_pending = true;
	s->frame_counted = count;
	s->frame_warps = static_cast<uint32_t>(grid) * kWarpsPerBlock;
	return CRT_OK;
}

int crt_render_device_sync(crt_scene *s, crt_stats *stats)
{
	if (!s)
		return fail(CRT_ERR_INVALID, "NULL argument");
	if (!s->frame_pending)
		return fail(CRT_ERR_INVALID, "no frame in flight");
	CUDA_TRY(cudaSetDevice(s->device));
	CUDA_TRY(cudaEventSynchronize(s->ev_stop));
	s->frame_pending = false;
	if (stats) {
		std::memset(stats, 0, sizeof *stats);
		float ms = 0.f;
		CUDA_TRY(cudaEventElapsedTime(&ms, s->ev_start, s->ev_stop));
		stats->kernel_ms = ms;
		if (s->frame_counted) {
			unsigned long long c[kTotalCounters];
			CUDA_TRY(cudaMemcpy(c, s->counters, sizeof c, cudaMemcpyDeviceToHost));
			stats->rays_closest = c[0];
			stats->rays_shadow = c[1];
			stats->box_tests = c[2];
			stats->tri_tests[0] = c[3];
			stats->tri_tests[1] = c[4];
			stats->tri_tests[2] = c[5];
			stats->tri_tests[3] = c[6];
			stats->sphere_tests[0] = c[7];
			stats->sphere_tests[1] = c[8];
			stats->plane_tests = c[9];
			stats->nodes_shaded = c[10];
			stats->node_steps = c[11];
			stats->node_step_lanes = c[12];
			stats->node_step_idle = c[13];
			stats->prim_steps = c[14];
			stats->prim_step_lanes = c[15];
			stats->prim_step_idle = c[16];
			if (c[kTimeStart]) {
				const unsigned long long first = ~c[kTimeStart], dry = ~c[kTimeDry];
				stats->queue_ms = static_cast<double>(dry - first) * 1e-6;
				stats->last_warp_ms = static_cast<double>(c[kTimeEnd] - first) * 1e-6;
				stats->warp_busy_ms = static_cast<double>(c[kTimeBusy]) * 1e-6 / (s->frame_warps ? s->frame_warps : 1u);
			}
		}
	}
	return CRT_OK;
}

int crt_render(crt_scene *s, const crt_render_opts *opts, float *raster, float *z_buffer, crt_stats *stats)
{
	if (!s || !opts || !raster || !z_buffer)
		return fail(CRT_ERR_INVALID, "NULL argument");
	const double t0 = now_ms();
	CUDA_TRY(cudaSetDevice(s->device));
	const size_t pixels = static_cast<size_t>(opts->width) * opts->height;
	if (s->staging_pixels < pixels) {
		cudaFree(s->d_raster);
		cudaFree(s->d_z);
		s->d_raster = s->d_z = nullptr;
		s->staging_pixels = 0;
		CUDA_TRY(cudaMalloc(&s->d_raster, 3 * pixels * sizeof(float)));
		CUDA_TRY(cudaMalloc(&s->d_z, pixels * sizeof(float)));
		s->staging_pixels = pixels;
	}
	int rc = crt_render_device(s, opts, s->d_raster, s->d_z, nullptr);
	if (rc)
		return rc;
	rc = crt_render_device_sync(s, stats);
	if (rc)
		return rc;
	CUDA_TRY(cudaMemcpy(raster, s->d_raster, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost));
	CUDA_TRY(cudaMemcpy(z_buffer, s->d_z, pixels * sizeof(float), cudaMemcpyDeviceToHost));
	if (stats)
		stats->total_ms = now_ms() - t0;
	return CRT_OK;
}

/* ---- several devices, one frame (crt_multi) -------------------------------------- */

struct crt_multi {
	std::vector<crt_scene *> scenes;
	std::vector<cudaStream_t> streams;
	bool dynamic = false;          /* shared work queue + frame on scenes[0]'s device over peer access */
	unsigned int *work_counter = nullptr; /* on scenes[0]'s device */
	float *h_raster = nullptr, *h_z = nullptr; /* static mode: pinned staging, one frame per device */
	size_t h_pixels = 0;
};

void crt_multi_destroy(crt_multi *m)
{
	if (!m)
		return;
	for (size_t k = 0; k < m->scenes.size(); k++) {
		if (m->scenes[k]) {
			cudaSetDevice(m->scenes[k]->device);
			if (k < m->streams.size() && m->streams[k])
				cudaStreamDestroy(m->streams[k]);
		}
	}
	if (!m->scenes.empty() && m->scenes[0]) {
		cudaSetDevice(m->scenes[0]->device);
		cudaFree(m->work_counter);
	}
	cudaFreeHost(m->h_raster);
	cudaFreeHost(m->h_z);
	for (crt_scene *sc : m->scenes)
		crt_scene_destroy(sc);
	delete m;
}

int crt_multi_create(const crt_scene_desc *desc, const int *devices, int num_devices, crt_multi **out)
{
	if (!out)
		return fail(CRT_ERR_INVALID, "out is NULL");
	*out = nullptr;
	if (!devices || num_devices < 1 || num_devices > 64)
		return fail(CRT_ERR_INVALID, "bad device list (%d devices)", num_devices);
	int rc = validate(desc);
	if (rc)
		return rc;
	for (int k = 0; k < num_devices; k++) {
		rc = ensure_device(devices[k]);
		if (rc)
			return rc;
	}
	HostScene h;
	rc = build_host_scene(desc, h); /* once: every device gets the same tree */
	if (rc)
		return rc;
	crt_multi *m = new crt_multi();
	m->scenes.assign(num_devices, nullptr);
	m->streams.assign(num_devices, nullptr);
	for (int k = 0; k < num_devices && !rc; k++) {
		rc = upload_host_scene(h, devices[k], &m->scenes[k]);
		if (!rc && cudaStreamCreateWithFlags(&m->streams[k], cudaStreamNonBlocking) != cudaSuccess)
			rc = fail(CRT_ERR_CUDA, "cudaStreamCreate failed on device %d", devices[k]);
	}
	/* peer access of every other device to devices[0] (frame + work counter live there) */
	bool peers = !rc;
	for (int k = 1; k < num_devices && peers; k++) {
		if (devices[k] == devices[0])
			continue;
		int can = 0;
		if (cudaDeviceCanAccessPeer(&can, devices[k], devices[0]) != cudaSuccess || !can) {
			cudaGetLastError();
			peers = false;
			break;
		}
		cudaSetDevice(devices[k]);
		const cudaError_t e = cudaDeviceEnablePeerAccess(devices[0], 0);
		if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
			peers = false;
		cudaGetLastError();
	}
	if (const char *e = std::getenv("CRT_MULTI_STATIC")) /* force the static-shard path (tests) */
		if (std::atoi(e))
			peers = false;
	m->dynamic = peers;
	if (!rc && m->dynamic) {
		cudaSetDevice(devices[0]);
		if (cudaMalloc(&m->work_counter, sizeof(unsigned int)) != cudaSuccess)
			rc = fail(CRT_ERR_CUDA, "cudaMalloc(work counter) failed");
	}
	if (rc) {
		crt_multi_destroy(m);
		return rc;
	}
	*out = m;
	return CRT_OK;
}

int crt_multi_num_devices(const crt_multi *m)
{
	return m ? static_cast<int>(m->scenes.size()) : 0;
}

int crt_multi_is_dynamic(const crt_multi *m)
{
	return m && m->dynamic ? 1 : 0;
}

int crt_multi_get_info(const crt_multi *m, int index, crt_scene_info *info)
{
	if (!m || !info || index < 0 || index >= static_cast<int>(m->scenes.size()))
		return fail(CRT_ERR_INVALID, "bad argument");
	*info = m->scenes[index]->info;
	return CRT_OK;
}

static void add_stats(crt_stats &acc, const crt_stats &st)
{
	acc.rays_closest += st.rays_closest;
	acc.rays_shadow += st.rays_shadow;
	acc.box_tests += st.box_tests;
	for (int i = 0; i < 4; i++)
		acc.tri_tests[i] += st.tri_tests[i];
	acc.sphere_tests[0] += st.sphere_tests[0];
	acc.sphere_tests[1] += st.sphere_tests[1];
	acc.plane_tests += st.plane_tests;
	acc.nodes_shaded += st.nodes_shaded;
	acc.node_steps += st.node_steps;
	acc.node_step_lanes += st.node_step_lanes;
	acc.node_step_idle += st.node_step_idle;
	acc.prim_steps += st.prim_steps;
	acc.prim_step_lanes += st.prim_step_lanes;
	acc.prim_step_idle += st.prim_step_idle;
	acc.queue_ms = std::max(acc.queue_ms, st.queue_ms);
	acc.last_warp_ms = std::max(acc.last_warp_ms, st.last_warp_ms);
	acc.warp_busy_ms = std::max(acc.warp_busy_ms, st.warp_busy_ms);
	acc.kernel_ms = std::fmax(acc.kernel_ms, st.kernel_ms);
}

int crt_multi_render(crt_multi *m, const crt_render_opts *opts, float *raster, float *z_buffer, crt_stats *stats, double *device_ms)
{
	if (!m || !opts || !raster || !z_buffer)
		return fail(CRT_ERR_INVALID, "NULL argument");
	if (opts->shard_count != 1u || opts->shard_index != 0u)
		return fail(CRT_ERR_INVALID, "crt_multi_render shards the frame itself: pass shard 0 of 1");
	const double t0 = now_ms();
	const size_t n = m->scenes.size();
	const size_t pixels = static_cast<size_t>(opts->width) * opts->height;
	crt_stats total;
	std::memset(&total, 0, sizeof total);
	int rc = CRT_OK;

	if (m->dynamic) {
		/* frame + counter on scenes[0]'s device; everybody pulls units from the one counter */
		crt_scene *s0 = m->scenes[0];
		CUDA_TRY(cudaSetDevice(s0->device));
		if (s0->staging_pixels < pixels) {
			cudaFree(s0->d_raster);
			cudaFree(s0->d_z);
			s0->d_raster = s0->d_z = nullptr;
			s0->staging_pixels = 0;
			CUDA_TRY(cudaMalloc(&s0->d_raster, 3 * pixels * sizeof(float)));
			CUDA_TRY(cudaMalloc(&s0->d_z, pixels * sizeof(float)));
			s0->staging_pixels = pixels;
		}
		crt_frame_share share;
		std::memset(&share, 0, sizeof share);
		share.work_counter = m->work_counter;
		for (size_t k = 0; k < n; k++) {
			uint32_t w = 0;
			rc = crt_scene_resident_warps(m->scenes[k], opts, &w);
			if (rc)
				return rc;
			share.total_warps += w;
		}
		CUDA_TRY(cudaSetDevice(s0->device));
		CUDA_TRY(cudaMemset(m->work_counter, 0, sizeof(unsigned int))); /* synchronous: visible before any participant starts */
		for (size_t k = 0; k < n && !rc; k++)
			rc = crt_render_device_shared(m->scenes[k], opts, s0->d_raster, s0->d_z, m->streams[k], &share);
		for (size_t k = 0; k < n; k++) { /* wait for every frame that was enqueued, also after an error */
			if (!m->scenes[k]->frame_pending)
				continue;
			crt_stats st;
			const int r2 = crt_render_device_sync(m->scenes[k], &st);
			if (r2 && !rc)
				rc = r2;
			if (!r2) {
				add_stats(total, st);
				if (device_ms)
					device_ms[k] = st.kernel_ms;
			}
		}
		if (rc)
			return rc;
		CUDA_TRY(cudaSetDevice(s0->device));
		CUDA_TRY(cudaMemcpy(raster, s0->d_raster, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost));
		CUDA_TRY(cudaMemcpy(z_buffer, s0->d_z, pixels * sizeof(float), cudaMemcpyDeviceToHost));
	} else {
		/* static shards (tile k -> device k mod n, crt_render_opts.shard_*), merged on the host:
		 * shards are disjoint and zero elsewhere, so the sum is exact */
		if (m->h_pixels < pixels * n) {
			cudaFreeHost(m->h_raster);
			cudaFreeHost(m->h_z);
			m->h_raster = m->h_z = nullptr;
			m->h_pixels = 0;
			CUDA_TRY(cudaHostAlloc(&m->h_raster, n * 3 * pixels * sizeof(float), cudaHostAllocPortable));
			CUDA_TRY(cudaHostAlloc(&m->h_z, n * pixels * sizeof(float), cudaHostAllocPortable));
			m->h_pixels = pixels * n;
		}
		for (size_t k = 0; k < n && !rc; k++) {
			crt_scene *sk = m->scenes[k];
			CUDA_TRY(cudaSetDevice(sk->device));
			if (sk->staging_pixels < pixels) {
				cudaFree(sk->d_raster);
				cudaFree(sk->d_z);
				sk->d_raster = sk->d_z = nullptr;
				sk->staging_pixels = 0;
				CUDA_TRY(cudaMalloc(&sk->d_raster, 3 * pixels * sizeof(float)));
				CUDA_TRY(cudaMalloc(&sk->d_z, pixels * sizeof(float)));
				sk->staging_pixels = pixels;
			}
			crt_render_opts o = *opts;
			o.shard_index = static_cast<uint32_t>(k);
			o.shard_count = static_cast<uint32_t>(n);
			rc = crt_render_device_shared(sk, &o, sk->d_raster, sk->d_z, m->streams[k], nullptr);
			if (!rc) {
				CUDA_TRY(cudaMemcpyAsync(m->h_raster + k * 3 * pixels, sk->d_raster, 3 * pixels * sizeof(float), cudaMemcpyDeviceToHost, m->streams[k]));
				CUDA_TRY(cudaMemcpyAsync(m->h_z + k * pixels, sk->d_z, pixels * sizeof(float), cudaMemcpyDeviceToHost, m->streams[k]));
			}
		}
		for (size_t k = 0; k < n; k++) {
			if (!m->scenes[k]->frame_pending)
				continue;
			crt_stats st;
			const int r2 = crt_render_device_sync(m->scenes[k], &st);
			if (r2 && !rc)
				rc = r2;
			if (!r2) {
				add_stats(total, st);
				if (device_ms)
					device_ms[k] = st.kernel_ms;
			}
			cudaSetDevice(m->scenes[k]->device);
			if (cudaStreamSynchronize(m->streams[k]) != cudaSuccess && !rc)
				rc = fail(CRT_ERR_CUDA, "copy of shard %zu failed: %s", k, cudaGetErrorString(cudaGetLastError()));
		}
		if (rc)
			return rc;
		std::memcpy(raster, m->h_raster, 3 * pixels * sizeof(float));
		std::memcpy(z_buffer, m->h_z, pixels * sizeof(float));
		for (size_t k = 1; k < n; k++) {
			const float *r = m->h_raster + k * 3 * pixels, *z = m->h_z + k * pixels;
			for (size_t i = 0; i < 3 * pixels; i++)
				raster[i] += r[i];
			for (size_t i = 0; i < pixels; i++)
				z_buffer[i] += z[i];
		}
	}
	if (stats) {
		*stats = total;
		stats->total_ms = now_ms() - t0;
	}
	return CRT_OK;
}

/* ---- memory shared between processes (one process per GPU) ------------------------ */

int crt_shared_alloc(int device, size_t bytes, void **d_ptr, unsigned char handle[CRT_IPC_HANDLE_BYTES])
{
	if (!d_ptr || !handle || !bytes)
		return fail(CRT_ERR_INVALID, "NULL argument");
	static_assert(sizeof(cudaIpcMemHandle_t) <= CRT_IPC_HANDLE_BYTES, "IPC handle does not fit");
	int rc = ensure_device(device);
	if (rc)
		return rc;
	void *p = nullptr;
	CUDA_TRY(cudaMalloc(&p, bytes));
	cudaIpcMemHandle_t hnd;
	const cudaError_t e = cudaIpcGetMemHandle(&hnd, p);
	if (e != cudaSuccess) {
		cudaFree(p);
		return fail(CRT_ERR_CUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
	}
	std::memset(handle, 0, CRT_IPC_HANDLE_BYTES);
	std::memcpy(handle, &hnd, sizeof hnd);
	*d_ptr = p;
	return CRT_OK;
}

int crt_shared_open(int device, const unsigned char handle[CRT_IPC_HANDLE_BYTES], void **d_ptr)
{
	if (!d_ptr || !handle)
		return fail(CRT_ERR_INVALID, "NULL argument");
	int rc = ensure_device(device);
	if (rc)
		return rc;
	cudaIpcMemHandle_t hnd;
	std::memcpy(&hnd, handle, sizeof hnd);
	void *p = nullptr;
	CUDA_TRY(cudaIpcOpenMemHandle(&p, hnd, cudaIpcMemLazyEnablePeerAccess));
	*d_ptr = p;
	return CRT_OK;
}

int crt_shared_release(int device, void *d_ptr, int owner)
{
	if (!d_ptr)
		return CRT_OK;
	int rc = ensure_device(device);
	if (rc)
		return rc;
	if (owner)
		CUDA_TRY(cudaFree(d_ptr));
	else
		CUDA_TRY(cudaIpcCloseMemHandle(d_ptr));
	return CRT_OK;
}

/* ---- the caller's frame buffers, page-locked ------------------------------------ */

int crt_host_register(void *ptr, size_t bytes)
{
	if (!ptr || !bytes)
		return fail(CRT_ERR_INVALID, "NULL or empty host range");
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
		cudaGetLastError();
		return fail(CRT_ERR_NO_DEVICE, "no CUDA device available (this library has no CPU fallback)");
	}
	CUDA_TRY(cudaHostRegister(ptr, bytes, cudaHostRegisterPortable));
	return CRT_OK;
}

int crt_host_unregister(void *ptr)
{
	if (!ptr)
		return CRT_OK;
	CUDA_TRY(cudaHostUnregister(ptr));
	return CRT_OK;
}

/* ---- ray-batch seams ---------------------------------------------------------- */

namespace {
struct DeviceBuffer {
	void *p = nullptr;
	~DeviceBuffer() { cudaFree(p); }
	cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
};
} // namespace

int crt_trace_closest(crt_scene *s, const float *rays, size_t n, int32_t *object, float *t, float *normal)
{
	if (!s || (n && (!rays || !object || !t)))
		return fail(CRT_ERR_INVALID, "NULL argument");
	CUDA_TRY(cudaSetDevice(s->device));
	DeviceBuffer d_rays, d_obj, d_t, d_n;
	CUDA_TRY(d_rays.alloc(n * 6 * sizeof(float)));
	CUDA_TRY(d_obj.alloc(n * sizeof(int32_t)));
	CUDA_TRY(d_t.alloc(n * sizeof(float)));
	CUDA_TRY(d_n.alloc(n * 3 * sizeof(float)));
	CUDA_TRY(cudaMemcpy(d_rays.p, rays, n * 6 * sizeof(float), cudaMemcpyHostToDevice));
	launch_trace_closest(s->d, static_cast<float *>(d_rays.p), n, static_cast<int32_t *>(d_obj.p), static_cast<float *>(d_t.p),
		static_cast<float *>(d_n.p), nullptr);
	CUDA_TRY(cudaGetLastError());
	CUDA_TRY(cudaMemcpy(object, d_obj.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost));
	CUDA_TRY(cudaMemcpy(t, d_t.p, n * sizeof(float), cudaMemcpyDeviceToHost));
	if (normal)
		CUDA_TRY(cudaMemcpy(normal, d_n.p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	return CRT_OK;
}

int crt_trace_shadow(crt_scene *s, const float *rays, const float *dist, const int32_t *emitter, size_t n, float *light, uint8_t *blocked)
{
	if (!s || (n && (!rays || !dist || !light || !blocked)))
		return fail(CRT_ERR_INVALID, "NULL argument");
	CUDA_TRY(cudaSetDevice(s->device));
	DeviceBuffer d_rays, d_dist, d_em, d_light, d_blocked;
	CUDA_TRY(d_rays.alloc(n * 6 * sizeof(float)));
	CUDA_TRY(d_dist.alloc(n * sizeof(float)));
	CUDA_TRY(d_em.alloc(n * sizeof(int32_t)));
	CUDA_TRY(d_light.alloc(n * 3 * sizeof(float)));
	CUDA_TRY(d_blocked.alloc(n));
	CUDA_TRY(cudaMemcpy(d_rays.p, rays, n * 6 * sizeof(float), cudaMemcpyHostToDevice));
	CUDA_TRY(cudaMemcpy(d_dist.p, dist, n * sizeof(float), cudaMemcpyHostToDevice));
	if (emitter)
		CUDA_TRY(cudaMemcpy(d_em.p, emitter, n * sizeof(int32_t), cudaMemcpyHostToDevice));
	CUDA_TRY(cudaMemcpy(d_light.p, light, n * 3 * sizeof(float), cudaMemcpyHostToDevice));
	launch_trace_shadow(s->d, static_cast<float *>(d_rays.p), static_cast<float *>(d_dist.p),
		emitter ? static_cast<int32_t *>(d_em.p) : nullptr, n, static_cast<float *>(d_light.p), static_cast<uint8_t *>(d_blocked.p), nullptr);
	CUDA_TRY(cudaGetLastError());
	CUDA_TRY(cudaMemcpy(light, d_light.p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	CUDA_TRY(cudaMemcpy(blocked, d_blocked.p, n, cudaMemcpyDeviceToHost));
	return CRT_OK;
}

int crt_texture_eval(crt_scene *s, int32_t material, const float *points, size_t n, float *colors)
{
	if (!s || (n && (!points || !colors)))
		return fail(CRT_ERR_INVALID, "NULL argument");
	if (material < 0 || static_cast<uint32_t>(material) >= s->num_materials)
		return fail(CRT_ERR_INVALID, "bad material index %d (scene has %u materials)", material, s->num_materials);
	CUDA_TRY(cudaSetDevice(s->device));
	DeviceBuffer d_p, d_c;
	CUDA_TRY(d_p.alloc(n * 3 * sizeof(float)));
	CUDA_TRY(d_c.alloc(n * 3 * sizeof(float)));
	CUDA_TRY(cudaMemcpy(d_p.p, points, n * 3 * sizeof(float), cudaMemcpyHostToDevice));
	launch_texture_eval(s->d, material, static_cast<float *>(d_p.p), n, static_cast<float *>(d_c.p), nullptr);
	CUDA_TRY(cudaGetLastError());
	CUDA_TRY(cudaMemcpy(colors, d_c.p, n * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	return CRT_OK;
}

int crt_measure_fp32_peak(int device, double *tflops)
{
	if (!tflops)
		return fail(CRT_ERR_INVALID, "NULL argument");
	int rc = ensure_device(device);
	if (rc)
		return rc;
	int sms = 0;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
	DeviceBuffer sink;
	CUDA_TRY(sink.alloc(sizeof(float)));
	cudaEvent_t a, b;
	CUDA_TRY(cudaEventCreate(&a));
	CUDA_TRY(cudaEventCreate(&b));
	const int blocks = sms * 8, threads = 256, iters = 4096;
	double best = 0.0;
	for (int rep = 0; rep < 5; rep++) {
		CUDA_TRY(cudaEventRecord(a));
		const double flops = launch_fma_peak(blocks, threads, iters, static_cast<float *>(sink.p), nullptr);
		CUDA_TRY(cudaEventRecord(b));
		CUDA_TRY(cudaEventSynchronize(b));
		float ms = 0.f;
		CUDA_TRY(cudaEventElapsedTime(&ms, a, b));
		if (rep > 0 && ms > 0.f)
			best = std::fmax(best, flops / (ms * 1e-3) / 1e12);
	}
	cudaEventDestroy(a);
	cudaEventDestroy(b);
	*tflops = best;
	return CRT_OK;
}

} // extern "C"
