/*
 * crt_device.cuh -- device data layout and device functions of the trace loop
 * (sm_100a).  Everything here is FP32 CUDA-core work: there is no dense
 * contraction on this path, so no tensor cores.
 *
 * Device layout (all arrays in HBM, read through the read-only path; the
 * working set of every configuration fits the 126 MB L2):
 *   wide    : WideNode[ ], 128 B, four 256-bit loads per visit (bvh_build.h)
 *   prims   : 4 x float4 (64 B) per bounded primitive, in the order of the wide tree ("slot")
 *               triangle: (v0.xyz, eps) (e0.xyz, info) (e1.xyz, object) (unit normal, 0)
 *               sphere  : (c.xyz,  eps) (r, 0, 0, info) (0, 0, 0, object) (0, 0, 0, 0)
 *             info = material index | PRIM_TRANSPARENT | PRIM_SPHERE,
 *             object = scene-order object index (both as raw int bits);
 *             the normal is read by closest-hit shading only
 *   planes  : 2 x float4 per plane: (n.xyz, d) (eps, info, object, 0)
 *   obj_slot: int per scene object: slot >= 0, or ~plane_index
 *   emitters: DEmitter per emitter, scene order
 *   materials: crt_material (128 B) as in the ABI
 */
#ifndef CRT_DEVICE_CUH
#define CRT_DEVICE_CUH

#include <cfloat>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "crt.h"
#include "crt_rng.h"

namespace crt {

constexpr uint32_t PRIM_SPHERE = 0x80000000u;
constexpr uint32_t PRIM_TRANSPARENT = 0x40000000u;
constexpr uint32_t PRIM_MATERIAL_MASK = 0x00ffffffu;

constexpr float kWideGridF = 1023.f;     /* bvh_build.h kWideGrid */
constexpr float kWideTauScale = 2048.f;  /* half-precision crossings are in units of 2048 grid steps: the cube spans [0, 1/2] */
constexpr float kWideSlopeMax = 48.f;    /* clamp of |idir| / 2048: beyond it the ray moves < 1/48 unit inside the cube; h |i| <= 516 * 48 stays finite */
constexpr float kPi = 3.1415927f;       /* reference src/core/type.h:30 */

struct DEmitter {
	float4 a;      /* sphere: centre.xyz, radius ; triangle: v0 */
	float4 b;      /* triangle: e0 */
	float4 c;      /* triangle: e1 */
	float ke_n[3]; /* ke / num_lights (render.c:175) */
	int32_t object;
	uint32_t num_lights;
	uint32_t first_light; /* prefix sum of num_lights over preceding emitters */
	int32_t is_sphere;
	/* box of every OTHER bounded object (padded; lo > hi: there is none).  The sampled emitter never occludes its own
	 * shadow rays (accel.c:367), so a ray that misses this box cannot be blocked or tinted by the tree and does not walk it */
	float rest_lo[3], rest_hi[3];
};

struct DScene {
	const float4 *wide;  /* 8-wide nodes, 8 float4 each (bvh_build.h WideNode) */
	const float4 *prims;
	const float4 *planes;
	const int32_t *obj_slot;
	const DEmitter *emitters;
	const crt_material *materials;
	uint32_t num_nodes, num_prims, num_planes, num_emitters, num_objects, total_lights;
	float ambient[3];
};

/* per-lane test counters of the instrumented kernel instance */
struct Counters {
	uint32_t closest, shadow, box, tri0, tri1, tri2, tri3, sph0, sph1, plane, shaded;
	/* walk efficiency of the direct-light loop, counted by lane 0 per warp-level step: steps of each kind, lanes
	 * that took them, and why the others did not (no ray: waiting for a refill; other: only work of the other kind) */
	uint32_t node_steps, node_lanes, node_noray, prim_steps, prim_lanes, prim_noray;
};

template <bool COUNT>
struct Cnt {
	__device__ __forceinline__ static void add(uint32_t &c, uint32_t v = 1u)
	{
		if (COUNT)
			c += v;
	}
};

struct Ray {
	float ox, oy, oz;
	float dx, dy, dz;
};

/* ---- small vector helpers ---------------------------------------------- */

__device__ __forceinline__ float dot3(float ax, float ay, float az, float bx, float by, float bz)
{
	return ax * bx + ay * by + az * bz;
}

__device__ __forceinline__ float4 ldg4(const float4 *p) { return __ldg(p); }


/* ---- primitive tests ------------------------------------------------------
 * Same formulation, comparisons and epsilon rules as the reference, so that
 * hits/misses agree except within rounding of the thresholds. */

/* Exactly-rounded FP32 helpers.  The three primitive tests below are written
 * with them (no FMA contraction, the reference's operation order), so that a
 * hit / miss decision and the hit distance are bit-identical to the reference's
 * scalar C for the same ray: grazing sphere hits and triangle-edge hits are
 * ill-conditioned, and a 1-ulp difference there flips whole pixels.  The box
 * tests stay fused (they only need to be conservative). */
__device__ __forceinline__ float xmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float xadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float xsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float xdot(float ax, float ay, float az, float bx, float by, float bz)
{
	return xadd(xadd(xmul(ax, bx), xmul(ay, by)), xmul(az, bz));
}

/* Moeller-Trumbore, two-sided: reference object.c:422-441 */
template <bool COUNT>
__device__ __forceinline__ bool hit_triangle(const float4 p0, const float4 p1, const float4 p2, const Ray &r, float &t, Counters &cn)
{
	const float eps = p0.w;
	/* h = d x e1 */
	const float hx = xsub(xmul(r.dy, p2.z), xmul(r.dz, p2.y));
	const float hy = xsub(xmul(r.dz, p2.x), xmul(r.dx, p2.z));
	const float hz = xsub(xmul(r.dx, p2.y), xmul(r.dy, p2.x));
	float a = xdot(p1.x, p1.y, p1.z, hx, hy, hz);
	asm volatile("" : "+f"(a)); /* pin a: it was recomputed for the division */
	if (a < eps && a > -eps) {
		Cnt<COUNT>::add(cn.tri0);
		return false;
	}
	const float sx = xsub(r.ox, p0.x), sy = xsub(r.oy, p0.y), sz = xsub(r.oz, p0.z);
	const float du = xdot(sx, sy, sz, hx, hy, hz);
	/* u = fl(fl(1 / a) * du).  Over half of all tests end at u < 0 or u > 1, most of
	 * them far from the thresholds, where the outcome follows from du and a
	 * without the (16-instruction, exactly rounded) division: opposite signs give
	 * u < 0 (the bounds keep the product from underflowing to -0, which would
	 * pass), and |du| > |a| (1 + 8 ulp) gives u > 1 whatever the two roundings do. */
	const float dus = a < 0.f ? -du : du, aa = fabsf(a);
	if ((dus < -1e-28f && aa < 1e8f) || dus > aa * 1.000001f) {
		Cnt<COUNT>::add(cn.tri1);
		return false;
	}
	const float f = __fdiv_rn(1.f, a);
	float u = xmul(f, du);
	asm volatile("" : "+f"(u)); /* pin u: under register pressure it was recomputed (h, s.h, the product) for u + v */
	if (u < 0.f || u > 1.f) {
		Cnt<COUNT>::add(cn.tri1);
		return false;
	}
	/* q = s x e0 */
	const float qx = xsub(xmul(sy, p1.z), xmul(sz, p1.y));
	const float qy = xsub(xmul(sz, p1.x), xmul(sx, p1.z));
	const float qz = xsub(xmul(sx, p1.y), xmul(sy, p1.x));
	const float v = xmul(f, xdot(r.dx, r.dy, r.dz, qx, qy, qz));
	if (v < 0.f || xadd(u, v) > 1.f) {
		Cnt<COUNT>::add(cn.tri2);
		return false;
	}
	Cnt<COUNT>::add(cn.tri3);
	t = xmul(f, xdot(p2.x, p2.y, p2.z, qx, qy, qz));
	return t > eps;
}

/* unit-direction quadratic: reference object.c:306-321 */
template <bool COUNT>
__device__ __forceinline__ bool hit_sphere(const float4 p0, const float radius, const Ray &r, float &t, Counters &cn)
{
	const float rx = xsub(r.ox, p0.x), ry = xsub(r.oy, p0.y), rz = xsub(r.oz, p0.z);
	const float b = -xdot(r.dx, r.dy, r.dz, rx, ry, rz);
	const float c = xsub(xdot(rx, ry, rz, rx, ry, rz), xmul(radius, radius));
	const float det = xsub(xmul(b, b), c);
	if (det < 0.f) {
		Cnt<COUNT>::add(cn.sph0);
		return false;
	}
	Cnt<COUNT>::add(cn.sph1);
	const float sq = __fsqrt_rn(det);
	t = xsub(b, sq);
	if (t > p0.w)
		return true;
	t = xadd(b, sq);
	return t > p0.w;
}

/* reference object.c:473-498; a = n.d is returned for the normal flip */
template <bool COUNT>
__device__ __forceinline__ bool hit_plane(const float4 pl, const float eps, const Ray &r, float &t, float &a, Counters &cn)
{
	Cnt<COUNT>::add(cn.plane);
	a = xdot(pl.x, pl.y, pl.z, r.dx, r.dy, r.dz);
	if (fabsf(a) < eps)
		return false;
	t = __fdiv_rn(xsub(pl.w, xdot(pl.x, pl.y, pl.z, r.ox, r.oy, r.oz)), a);
	return t > eps;
}

template <bool COUNT>
__device__ __forceinline__ bool hit_prim(const DScene &sc, int slot, const Ray &r, float &t, uint32_t &info, int &object, Counters &cn)
{
	const float4 *rec = sc.prims + 4 * (size_t)slot;
	const float4 p0 = ldg4(rec), p1 = ldg4(rec + 1), p2 = ldg4(rec + 2);
	info = __float_as_uint(p1.w);
	object = __float_as_int(p2.w);
	if (info & PRIM_SPHERE)
		return hit_sphere<COUNT>(p0, p1.x, r, t, cn);
	return hit_triangle<COUNT>(p0, p1, p2, r, t, cn);
}

/* ---- BVH traversal ---------------------------------------------------------
 *
 * Both walks (closest hit and the any-hit shadow query) read the same tree: 8-wide
 * nodes of 128 bytes (bvh_build.h WideNode), four 256-bit loads per visit.  A node
 * visit tests its eight child boxes in HALF2 arithmetic, two children per
 * instruction:
 *
 *   - the ray is first clipped, in FP32, against the node's grid cube [0, 1023]^3
 *     and re-based to its entry point p (so every quantity of the child test is at
 *     most ~1023 grid units, where half precision resolves 1/2 unit), with the ray
 *     parameter rescaled so that the dominant axis advances one grid unit per unit
 *     (tau' in [0, 1023] inside the cube);
 *   - per axis and child pair, with tau'' = tau' / 2048 (the cube spans [0, 1/2]):
 *       a = c - p;  q = a * i;  tn = q - h |i|;  tf = q + h |i|
 *     (HADD2, HMUL2, two HFMA2; i = idir / 2048 clamped to |i| <= 48), c = box
 *     centre, h = half extent padded by kWideSlack = 3.5 grid units, which bounds
 *     every rounding above: 1/4 unit for p, 1/4 for c - p, 2^-10 relative for the
 *     product and the FMA on magnitudes <= 1535 units, leaving 3/4 unit for the FP32
 *     clip.  With the clamp no product overflows (|a i| <= 1100 * 48, h |i| <= 516 * 48),
 *     so tn / tf are finite or a correctly signed infinity, never inf - inf; a slope
 *     beyond the clamp moves the ray by less than 1/48 of a unit inside the cube,
 *     far inside the slack, so evaluating it with the clamped slope stays
 *     conservative.  When the FP32 clip itself cannot guarantee its 3/4 unit
 *     (origins ~10^6 grid units from the node) the visit reports every child as
 *     hit -- slower, never wrong;
 *   - tmin = max(tn.x, tn.y, tn.z, 0), tmax = min(tf.x, tf.y, tf.z, limit) with the
 *     three-input half2 min/max, one HSET2 per child pair.
 *
 * A visit yields an 8-bit mask of hit slots; leaf slots and inner slots are
 * contiguous in primitive / node index (bvh_build.h), so the walk state is a pair
 * of (base << 8 | mask) words per kind and a short stack of such words. */

constexpr int kInnerStack = 12;  /* inner-group entries per lane kept in shared memory; deeper ones spill (DeepStack) */
#ifndef CRT_LEAF_STACK
#define CRT_LEAF_STACK 8
#endif
constexpr int kLeafStack = CRT_LEAF_STACK;    /* leaf-group entries per lane ([slot][lane]); a node visit adds at most one */
constexpr int kDeepStack = 64;   /* >= kWideMaxDepth */

/* ray constants of the wide walk (beside the ray itself) */
struct WideRay {
	float inx, iny, inz;    /* maxd / d: grid units of the dominant axis per grid unit of this axis (|.| >= 1) */
	float maxd, inv_maxd;   /* max |d| and its reciprocal: tau = t * maxd / scale */
	uint32_t hx, hy, hz;    /* half2 (i, i) of (maxd / d) / 2048, clamped to +-kWideSlopeMax */
};

__device__ __forceinline__ uint32_t h2_bits(__half2 v) { return *reinterpret_cast<uint32_t *>(&v); }
__device__ __forceinline__ __half2 h2_from(uint32_t v) { return *reinterpret_cast<__half2 *>(&v); }

__device__ __forceinline__ uint32_t pack_half2(float lo, float hi)
{
	return h2_bits(__floats2half2_rn(lo, hi));
}

__device__ __forceinline__ float safe_idir(float d, float maxd)
{
	/* maxd / d with a floor on |d|: finite everywhere, huge for an axis the ray is parallel to */
	const float a = fabsf(d) > 1e-18f ? d : copysignf(1e-18f, d);
	return maxd / a;
}

__device__ __forceinline__ WideRay make_wide_ray(const Ray &r)
{
	WideRay w;
	w.maxd = fmaxf(fmaxf(fabsf(r.dx), fabsf(r.dy)), fabsf(r.dz));
	w.inv_maxd = 1.f / w.maxd;
	w.inx = safe_idir(r.dx, w.maxd);
	w.iny = safe_idir(r.dy, w.maxd);
	w.inz = safe_idir(r.dz, w.maxd);
	const float big = kWideSlopeMax, sc = 1.f / kWideTauScale;
	const float cx = fminf(fmaxf(w.inx * sc, -big), big), cy = fminf(fmaxf(w.iny * sc, -big), big), cz = fminf(fmaxf(w.inz * sc, -big), big);
	w.hx = pack_half2(cx, cx); w.hy = pack_half2(cy, cy); w.hz = pack_half2(cz, cz);
	return w;
}

#ifndef CRT_REST_BOX
#define CRT_REST_BOX 1 /* 0: every shadow ray walks the tree from the root (experiments) */
#endif

/* Can the segment [0, dist] of the ray be shown to miss the box [lo, hi]?  FP32 slab test in units of tau = t * maxd
 * with the ray's safe reciprocals; margins of 1e-5 relative (the products carry a few ulp) on a box the host padded.
 * Written so that a NaN anywhere answers false (the ray then walks the tree as before). */
__device__ __forceinline__ bool ray_misses_box(const float *lo, const float *hi, const Ray &r, const WideRay &w, float dist)
{
	const float lx = __ldg(lo), ly = __ldg(lo + 1), lz = __ldg(lo + 2), ux = __ldg(hi), uy = __ldg(hi + 1), uz = __ldg(hi + 2);
	if (lx > ux)
		return true; /* no other bounded object */
	const float ax = (lx - r.ox) * w.inx, bx = (ux - r.ox) * w.inx;
	const float ay = (ly - r.oy) * w.iny, by = (uy - r.oy) * w.iny;
	const float az = (lz - r.oz) * w.inz, bz = (uz - r.oz) * w.inz;
	const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
	const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
	const float lim = dist * w.maxd;
	const float slack = fmaf(fabsf(tn) + fabsf(tf), 1e-5f, 1e-30f);
	return tn - tf > slack || tf < -slack || tn > fmaf(fabsf(lim), 1e-5f, lim) + slack;
}

struct alignas(32) Vec8 {
	uint32_t v[8];
};

__device__ __forceinline__ Vec8 ldg8(const float4 *p)
{
	Vec8 r; /* one 256-bit load through the read-only path (LDG.E.ENL2.256.CONSTANT) */
	asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
		: "=r"(r.v[0]), "=r"(r.v[1]), "=r"(r.v[2]), "=r"(r.v[3]), "=r"(r.v[4]), "=r"(r.v[5]), "=r"(r.v[6]), "=r"(r.v[7])
		: "l"(p));
	return r;
}

struct WideHits {
	uint32_t inner; /* (child_base << 8) | mask of hit inner slots */
	uint32_t leaf;  /* (prim_base << 8) | mask of hit leaf slots */
};

/* One visit of wide node `node`: which of its children may the segment [0, limit] of the ray meet? */
template <bool COUNT>
__device__ __forceinline__ WideHits wide_node_test(const DScene &sc, const Ray &r, const WideRay &w, float limit, int node, Counters &cn)
{
	const float4 *n = sc.wide + 8 * (size_t)node;
	const Vec8 head = ldg8(n), vx = ldg8(n + 2), vy = ldg8(n + 4), vz = ldg8(n + 6);
	const uint32_t meta = head.v[3];
	const float inv_s = __uint_as_float((254u - (meta & 0xffu)) << 23); /* 1 / scale, exact */
	/* ray in grid coordinates, clipped against the grid cube */
	const float olx = (r.ox - __uint_as_float(head.v[0])) * inv_s;
	const float oly = (r.oy - __uint_as_float(head.v[1])) * inv_s;
	const float olz = (r.oz - __uint_as_float(head.v[2])) * inv_s;
	const float ax = -olx * w.inx, ay = -oly * w.iny, az = -olz * w.inz;                 /* tau of the planes g = 0 */
	const float bx = fmaf(kWideGridF, w.inx, ax), by = fmaf(kWideGridF, w.iny, ay), bz = fmaf(kWideGridF, w.inz, az); /* g = 1023 */
	float tin = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.f));
	const float tout = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
	tin = fminf(tin, 1e30f);
	const float tind = tin * w.inv_maxd;
	const float px = fmaf(tind, r.dx, olx), py = fmaf(tind, r.dy, oly), pz = fmaf(tind, r.dz, olz);
	const float lim = fmaf(limit * w.maxd, inv_s, -tin); /* the ray's limit in tau', relative to the entry point */
	/* FP32 error of p and lim: a few ulp of the magnitudes they were formed from */
	const float mag = (fabsf(olx) + fabsf(oly)) + (fabsf(olz) + tin);
	const uint32_t p2x = pack_half2(px, px), p2y = pack_half2(py, py), p2z = pack_half2(pz, pz);
	/* + 2^-8 relative + 1 unit: above the true limit after the conversion (an overflow becomes +infinity) */
	const float lim_up = fmaf(fabsf(lim), 0.00390625f, lim + 1.f) * (1.f / kWideTauScale);
	const uint32_t l2 = pack_half2(lim_up, lim_up);
	uint32_t hit = 0;
	const __half2 nax = h2_from(w.hx | 0x80008000u), nay = h2_from(w.hy | 0x80008000u), naz = h2_from(w.hz | 0x80008000u); /* -|idir| */
#pragma unroll
	for (int k = 0; k < 4; k++) {
		const __half2 cx = h2_from(vx.v[k]), hx = h2_from(vx.v[4 + k]);
		const __half2 cy = h2_from(vy.v[k]), hy = h2_from(vy.v[4 + k]);
		const __half2 cz = h2_from(vz.v[k]), hz = h2_from(vz.v[4 + k]);
		const __half2 qx = __hmul2(__hsub2(cx, h2_from(p2x)), h2_from(w.hx));
		const __half2 qy = __hmul2(__hsub2(cy, h2_from(p2y)), h2_from(w.hy));
		const __half2 qz = __hmul2(__hsub2(cz, h2_from(p2z)), h2_from(w.hz));
		const __half2 tnx = __hfma2(hx, nax, qx), tfx = __hfma2(hx, __hneg2(nax), qx);
		const __half2 tny = __hfma2(hy, nay, qy), tfy = __hfma2(hy, __hneg2(nay), qy);
		const __half2 tnz = __hfma2(hz, naz, qz), tfz = __hfma2(hz, __hneg2(naz), qz);
		const __half2 tmin = __hmax2(__hmax2(__hmax2(tnx, tny), tnz), h2_from(0u));
		const __half2 tmax = __hmin2(__hmin2(__hmin2(tfx, tfy), tfz), h2_from(l2));
		const uint32_t m = __hle2_mask(tmin, tmax);
		hit |= m & (0x00020001u << (2 * k)); /* slot 2k -> bit 2k, slot 2k + 1 -> bit 17 + 2k */
	}
	hit = (hit | (hit >> 16)) & 0xffu;
	const uint32_t valid = ((meta >> 8) | (meta >> 16)) & 0xffu;
	/* precision guard: beyond ~10^6 grid units the clip cannot place p to a fraction of a unit */
	if (!(mag <= 1048576.f))
		hit = valid;
	/* the ray misses the grid cube (the parent's test is conservative), or the cube lies beyond the limit */
	if (tin > fmaf(fabsf(tout) + tin, 1e-6f, tout) || lim < -1.f - mag * 1e-6f)
		hit = 0;
	hit &= valid;
	if (COUNT)
		cn.box += (uint32_t)__popc(valid);
	WideHits res;
	res.leaf = ((uint32_t)head.v[5] << 8) | (hit & (meta >> 8) & 0xffu);
	res.inner = ((uint32_t)head.v[4] << 8) | (hit & (meta >> 16) & 0xffu);
	return res;
}

struct Hit {
	float t;
	int object; /* scene-order object index, -1 = none */
	int slot;   /* prim slot, or ~plane index */
};

/* lowest set bit of the 8-bit mask of a group word, removed from it */
__device__ __forceinline__ int group_take(uint32_t &g)
{
	const int j = __ffs((int)(g & 0xffu)) - 1;
	g &= g - 1u;
	return j;
}

__device__ __forceinline__ int group_base(uint32_t g) { return (int)g >> 8; }

/* Nearest hit over planes then the BVH: get_closest_intersection()
 * (reference render.c:118-124, object.c:168-181, accel.c:322-353).
 * Strict '<' like the reference; on an exact tie between bounded primitives
 * the lower scene index wins (the reference's winner depends on its Morton
 * order).  Depth-first over the wide tree with the current nearest
 * distance as the limit of every node test; the primitives of a node are tested
 * before its inner children are entered. */
template <bool COUNT, class STACKS>
__device__ __forceinline__ Hit trace_closest(const DScene &sc, const Ray &r, const STACKS &st, Counters &cn)
{
	Hit best;
	best.t = FLT_MAX;
	best.object = -1;
	best.slot = 0;
	for (uint32_t k = 0; k < sc.num_planes; k++) {
		const float4 pl = ldg4(sc.planes + 2 * k);
		const float4 px = ldg4(sc.planes + 2 * k + 1);
		float t, a;
		if (hit_plane<COUNT>(pl, px.x, r, t, a, cn) && t < best.t) {
			best.t = t;
			best.object = __float_as_int(px.z);
			best.slot = ~static_cast<int>(k);
		}
	}
	if (sc.num_nodes == 0)
		return best;

	const WideRay w = make_wide_ray(r);
	bool bounded_found = false;
	int sp = 0;
	uint32_t cur = 1u; /* group (base 0, slot 0): the root */
	while (cur & 0xffu) {
		const int node = group_base(cur) + group_take(cur);
		WideHits h = wide_node_test<COUNT>(sc, r, w, fminf(best.t, 1e30f), node, cn);
		while (h.leaf & 0xffu) {
			const int slot = group_base(h.leaf) + group_take(h.leaf);
			float t;
			uint32_t info;
			int object;
			if (hit_prim<COUNT>(sc, slot, r, t, info, object, cn)) {
				if (t < best.t || (t == best.t && bounded_found && object < best.object)) {
					best.t = t;
					best.object = object;
					best.slot = slot;
					bounded_found = true;
				}
			}
		}
		if (h.inner & 0xffu) {
			if (cur & 0xffu)
				st.put_inner(sp++, cur);
			cur = h.inner;
		} else if (!(cur & 0xffu) && sp > 0) {
			cur = st.get_inner(--sp);
		}
	}
	return best;
}

/* Any-hit with transmittance: is_light_blocked() (reference render.c:126-134,
 * object.c:183-197, accel.c:360-387).  Blocked at the first opaque occluder
 * closer than dist; every transparent occluder multiplies the light by its
 * kt, once per primitive; the sampled emitter never occludes
 * (accel.c:367).  It runs in resumable pieces, so that the render kernel's
 * lanes can be refilled and scheduled step by step:
 * shadow_planes() is the linear plane scan, the ShadowWalk steps below walk
 * the wide tree; trace_shadow_wide() strings them together for
 * one ray per thread (the crt_trace_shadow seam). */
template <bool COUNT>
__device__ __forceinline__ bool shadow_planes(const DScene &sc, const Ray &r, float dist, float &lr, float &lg, float &lb, Counters &cn)
{
	for (uint32_t k = 0; k < sc.num_planes; k++) {
		const float4 pl = ldg4(sc.planes + 2 * k);
		const float4 px = ldg4(sc.planes + 2 * k + 1);
		float t, a;
		if (hit_plane<COUNT>(pl, px.x, r, t, a, cn) && t < dist) {
			const uint32_t info = __float_as_uint(px.y);
			if (!(info & PRIM_TRANSPARENT))
				return true;
			const crt_material *m = sc.materials + (info & PRIM_MATERIAL_MASK);
			lr *= __ldg(&m->kt[0]);
			lg *= __ldg(&m->kt[1]);
			lb *= __ldg(&m->kt[2]);
		}
	}
	return false;
}

/* Any-hit order is free (every occluder within dist counts, accel.c:360-387),
 * so the walk is split into two kinds of unit steps that a lane can take in
 * any interleaving: a NODE step (one wide node: eight slab tests; the hit leaf
 * slots become one leaf group, the hit inner slots one inner group) and a PRIM
 * step (one primitive of the current leaf group).  The render kernel lets the
 * warp vote each iteration for the kind more lanes are ready for, instead of a
 * while-while loop in which lanes at a leaf wait for the longest descent of the
 * warp.  The current group of each kind lives in a register; only a group that
 * is superseded while slots remain goes to the lane's stack of that kind. */
struct ShadowWalk {
	uint32_t inner; /* current inner group, mask 0: none */
	uint32_t leaf;  /* current leaf group, mask 0: none */
	int isp, lsp;   /* entries on the inner / leaf stack */
};

__device__ __forceinline__ void shadow_walk_reset(ShadowWalk &w)
{
	w.inner = 0;
	w.leaf = 0;
	w.isp = 0;
	w.lsp = 0;
}

__device__ __forceinline__ bool shadow_walk_busy(const ShadowWalk &w)
{
	return ((w.inner | w.leaf) & 0xffu) != 0u;
}

__device__ __forceinline__ bool shadow_can_inner(const ShadowWalk &w)
{
	return (w.inner & 0xffu) != 0u && w.lsp < kLeafStack;
}

__device__ __forceinline__ bool shadow_can_prim(const ShadowWalk &w)
{
	return (w.leaf & 0xffu) != 0u;
}

/* The lane's two stacks.  In the render kernel they live in shared memory as
 * [slot][lane] words; the lane's base is kept as a 32-bit shared address in a
 * register the compiler cannot rematerialise (re-deriving it from %tid at every
 * push cost 15 instructions).  Inner entries beyond kInnerStack (trees deeper than
 * that are rare) spill to a per-lane strip of global scratch.  The
 * one-ray-per-thread seam kernels use private arrays. */
struct SharedStacks {
	uint32_t base;  /* shared-space byte address of inner slot 0 of this lane; leaf slots follow the inner ones */
	uint32_t *deep; /* this lane's spill strip (stride 32 words), may be null when the tree is shallow */
	__device__ __forceinline__ SharedStacks(const uint32_t *lane_slot0, uint32_t *deep_strip) : deep(deep_strip)
	{
		uint32_t b = static_cast<uint32_t>(__cvta_generic_to_shared(lane_slot0));
		asm volatile("mov.u32 %0, %1;" : "=r"(base) : "r"(b));
	}
	__device__ __forceinline__ void put_inner(int slot, uint32_t v) const
	{
		if (slot < kInnerStack)
			asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + (uint32_t)slot * 128u), "r"(v) : "memory");
		else
			deep[(slot - kInnerStack) * 32] = v;
	}
	__device__ __forceinline__ uint32_t get_inner(int slot) const
	{
		uint32_t v;
		if (slot < kInnerStack)
			asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + (uint32_t)slot * 128u) : "memory");
		else
			v = deep[(slot - kInnerStack) * 32];
		return v;
	}
	__device__ __forceinline__ void put_leaf(int slot, uint32_t v) const
	{
		asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + (uint32_t)(kInnerStack + slot) * 128u), "r"(v) : "memory");
	}
	__device__ __forceinline__ uint32_t get_leaf(int slot) const
	{
		uint32_t v;
		asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + (uint32_t)(kInnerStack + slot) * 128u) : "memory");
		return v;
	}
};

struct PrivateStacks {
	uint32_t *inner, *leaf;
	__device__ __forceinline__ void put_inner(int slot, uint32_t v) const { inner[slot] = v; }
	__device__ __forceinline__ uint32_t get_inner(int slot) const { return inner[slot]; }
	__device__ __forceinline__ void put_leaf(int slot, uint32_t v) const { leaf[slot] = v; }
	__device__ __forceinline__ uint32_t get_leaf(int slot) const { return leaf[slot]; }
};

template <bool COUNT, class STACKS>
__device__ __forceinline__ void shadow_inner_step(const DScene &sc, const Ray &r, const WideRay &rw, float dist, ShadowWalk &w,
	const STACKS &st, Counters &cn)
{
	const int node = group_base(w.inner) + group_take(w.inner);
	const WideHits h = wide_node_test<COUNT>(sc, r, rw, dist, node, cn);
	if (h.inner & 0xffu) {
		if (w.inner & 0xffu)
			st.put_inner(w.isp++, w.inner);
		w.inner = h.inner;
	} else if (!(w.inner & 0xffu) && w.isp > 0) {
		w.inner = st.get_inner(--w.isp);
	}
	if (h.leaf & 0xffu) {
		if (w.leaf & 0xffu)
			st.put_leaf(w.lsp++, w.leaf);
		w.leaf = h.leaf;
	}
}

template <bool COUNT, class STACKS>
__device__ __forceinline__ void shadow_prim_step(const DScene &sc, const Ray &r, float dist, int emitter_object, ShadowWalk &w,
	const STACKS &st, bool &blocked, float &lr, float &lg, float &lb, Counters &cn)
{
	const int slot = group_base(w.leaf) + group_take(w.leaf);
	if (!(w.leaf & 0xffu) && w.lsp > 0)
		w.leaf = st.get_leaf(--w.lsp);
	float t;
	uint32_t info;
	int object;
	const bool hit = hit_prim<COUNT>(sc, slot, r, t, info, object, cn);
	if (hit && t < dist && object != emitter_object) {
		if (!(info & PRIM_TRANSPARENT)) {
			blocked = true;
			shadow_walk_reset(w); /* first opaque occluder ends the walk (accel.c:369-371) */
			return;
		}
		const crt_material *m = sc.materials + (info & PRIM_MATERIAL_MASK);
		lr *= __ldg(&m->kt[0]);
		lg *= __ldg(&m->kt[1]);
		lb *= __ldg(&m->kt[2]);
	}
}

/* is_light_blocked() through the production walk, one ray per thread (the
 * crt_trace_shadow seam): planes, then node / primitive unit steps until done. */
template <bool COUNT>
__device__ __forceinline__ bool trace_shadow_wide(const DScene &sc, const Ray &r, float dist, int emitter_object,
	float &lr, float &lg, float &lb, Counters &cn)
{
	if (shadow_planes<COUNT>(sc, r, dist, lr, lg, lb, cn))
		return true;
	if (sc.num_nodes == 0)
		return false;
	uint32_t inner[kDeepStack], leaf[kLeafStack];
	const PrivateStacks st = { inner, leaf };
	ShadowWalk w;
	shadow_walk_reset(w);
	w.inner = 1u; /* the root */
	const WideRay rw = make_wide_ray(r);
	bool blocked = false;
	while (shadow_walk_busy(w)) {
		if (shadow_can_inner(w))
			shadow_inner_step<COUNT>(sc, r, rw, dist, w, st, cn);
		else
			shadow_prim_step<COUNT>(sc, r, dist, emitter_object, w, st, blocked, lr, lg, lb, cn);
	}
	return blocked;
}

/* get_intersection of one given object (the inside-object shortcut,
 * reference render.c:143-144) and the geometric normal of a hit
 * (sphere outward object.c:258-261, triangle from winding
 * object.c:361, plane facing the ray object.c:481-484). */
template <bool COUNT>
__device__ __forceinline__ bool hit_object(const DScene &sc, int object, const Ray &r, Hit &h, Counters &cn)
{
	const int slot = __ldg(sc.obj_slot + object);
	h.object = object;
	h.slot = slot;
	if (slot < 0) {
		const int k = ~slot;
		float a;
		return hit_plane<COUNT>(ldg4(sc.planes + 2 * k), ldg4(sc.planes + 2 * k + 1).x, r, h.t, a, cn);
	}
	uint32_t info;
	int obj;
	return hit_prim<COUNT>(sc, slot, r, h.t, info, obj, cn);
}

__device__ __forceinline__ void hit_normal(const DScene &sc, const Hit &h, const Ray &r, float &nx, float &ny, float &nz,
	uint32_t &material, float &eps)
{
	if (h.slot < 0) {
		const int k = ~h.slot;
		const float4 pl = ldg4(sc.planes + 2 * k);
		const float4 px = ldg4(sc.planes + 2 * k + 1);
		const float a = xdot(pl.x, pl.y, pl.z, r.dx, r.dy, r.dz);
		const float s = signbit(a) ? 1.f : -1.f;
		nx = pl.x * s;
		ny = pl.y * s;
		nz = pl.z * s;
		material = __float_as_uint(px.y) & PRIM_MATERIAL_MASK;
		eps = px.x;
		return;
	}
	const float4 p0 = ldg4(sc.prims + 4 * (size_t)h.slot);
	const float4 p1 = ldg4(sc.prims + 4 * (size_t)h.slot + 1);
	const uint32_t info = __float_as_uint(p1.w);
	material = info & PRIM_MATERIAL_MASK;
	eps = p0.w;
	if (info & PRIM_SPHERE) {
		const float inv = __fdiv_rn(1.f, p1.x);
		nx = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dx, h.t), r.ox), p0.x), inv);
		ny = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dy, h.t), r.oy), p0.y), inv);
		nz = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dz, h.t), r.oz), p0.z), inv);
	} else {
		const float4 n = ldg4(sc.prims + 4 * (size_t)h.slot + 3);
		nx = n.x;
		ny = n.y;
		nz = n.z;
	}
}

/* ---- procedural textures: reference material.c:152-200 ------------------- */

__device__ const uint8_t kPerm[256] = {
	151,160,137,91,90,15,131,13,201,95,96,53,194,233,7,225,140,36,103,30,69,142,8,99,37,240,21,10,23,190,6,148,
	247,120,234,75,0,26,197,62,94,252,219,203,117,35,11,32,57,177,33,88,237,149,56,87,174,20,125,136,171,168,68,175,
	74,165,71,134,139,48,27,166,77,146,158,231,83,111,229,122,60,211,133,230,220,105,92,41,55,46,245,40,244,102,143,54,
	65,25,63,161,1,216,80,73,209,76,132,187,208,89,18,169,200,196,135,130,116,188,159,86,164,100,109,198,173,186,3,64,
	52,217,226,250,124,123,5,202,38,147,118,126,255,82,85,212,207,206,59,227,47,16,58,17,182,189,28,42,223,183,170,213,
	119,248,152,2,44,154,163,70,221,153,101,155,167,43,172,9,129,22,39,253,19,98,108,110,79,113,224,232,178,185,112,104,
	218,246,97,228,251,34,242,193,238,210,144,12,191,179,162,241,81,51,145,235,249,14,239,107,49,192,214,31,181,199,106,157,
	184,84,204,176,115,121,50,45,127,4,150,254,138,236,205,93,222,114,67,29,24,72,243,141,128,195,78,66,215,61,156,180
};

__device__ __forceinline__ int perm_at(int i) { return kPerm[i & 255]; }

__device__ __forceinline__ float simplex_corner(int gi, float x, float y, float z)
{
	/* no FMA contraction here: the falloff sign test must match the CPU */
	float t = __fsub_rn(__fsub_rn(__fsub_rn(0.6f, __fmul_rn(x, x)), __fmul_rn(y, y)), __fmul_rn(z, z));
	if (t < 0.f)
		return 0.f;
	t = __fmul_rn(t, t);
	const int h = gi & 15;
	const float u = h < 8 ? x : y;
	const float v = h < 4 ? y : (h == 12 || h == 14) ? x : z;
	const float g = __fadd_rn((h & 1) ? -u : u, (h & 2) ? -v : v);
	return __fmul_rn(__fmul_rn(t, t), g);
}

/* 3-D simplex noise: reference lib/SimplexNoise/SimplexNoise.c:117-212 */
__device__ inline float simplex_noise(float x, float y, float z)
{
	const float F3 = 1.0f / 3.0f, G3 = 1.0f / 6.0f;
	const float s = __fmul_rn(__fadd_rn(__fadd_rn(x, y), z), F3);
	const int i = (int)floorf(__fadd_rn(x, s)), j = (int)floorf(__fadd_rn(y, s)), k = (int)floorf(__fadd_rn(z, s));
	const float t = __fmul_rn((float)(i + j + k), G3);
	const float x0 = __fsub_rn(x, __fsub_rn((float)i, t));
	const float y0 = __fsub_rn(y, __fsub_rn((float)j, t));
	const float z0 = __fsub_rn(z, __fsub_rn((float)k, t));
	int i1, j1, k1, i2, j2, k2;
	if (x0 >= y0) {
		if (y0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
		else if (x0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 0; k2 = 1; }
		else { i1 = 0; j1 = 0; k1 = 1; i2 = 1; j2 = 0; k2 = 1; }
	} else {
		if (y0 < z0) { i1 = 0; j1 = 0; k1 = 1; i2 = 0; j2 = 1; k2 = 1; }
		else if (x0 < z0) { i1 = 0; j1 = 1; k1 = 0; i2 = 0; j2 = 1; k2 = 1; }
		else { i1 = 0; j1 = 1; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
	}
	const float G32 = __fmul_rn(2.0f, G3), G33 = __fmul_rn(3.0f, G3);
	const float x1 = __fadd_rn(__fsub_rn(x0, (float)i1), G3), y1 = __fadd_rn(__fsub_rn(y0, (float)j1), G3), z1 = __fadd_rn(__fsub_rn(z0, (float)k1), G3);
	const float x2 = __fadd_rn(__fsub_rn(x0, (float)i2), G32), y2 = __fadd_rn(__fsub_rn(y0, (float)j2), G32), z2 = __fadd_rn(__fsub_rn(z0, (float)k2), G32);
	const float x3 = __fadd_rn(__fsub_rn(x0, 1.0f), G33), y3 = __fadd_rn(__fsub_rn(y0, 1.0f), G33), z3 = __fadd_rn(__fsub_rn(z0, 1.0f), G33);
	const int gi0 = perm_at(i + perm_at(j + perm_at(k)));
	const int gi1 = perm_at(i + i1 + perm_at(j + j1 + perm_at(k + k1)));
	const int gi2 = perm_at(i + i2 + perm_at(j + j2 + perm_at(k + k2)));
	const int gi3 = perm_at(i + 1 + perm_at(j + 1 + perm_at(k + 1)));
	const float n0 = simplex_corner(gi0, x0, y0, z0);
	const float n1 = simplex_corner(gi1, x1, y1, z1);
	const float n2 = simplex_corner(gi2, x2, y2, z2);
	const float n3 = simplex_corner(gi3, x3, y3, z3);
	return __fmul_rn(32.0f, __fadd_rn(__fadd_rn(__fadd_rn(n0, n1), n2), n3));
}

/* (uint32_t) of a possibly negative float the way gcc/x86-64 evaluates it:
 * truncate to a 64-bit integer, keep the low 32 bits.  A plain CUDA
 * float->unsigned conversion saturates negatives to 0. */
__device__ __forceinline__ uint32_t wrap_u32(float f)
{
	return static_cast<uint32_t>(static_cast<long long>(f));
}

__device__ inline void texture_color(const crt_material *m, float px, float py, float pz, float &r, float &g, float &b)
{
	const int type = __ldg(&m->tex_type);
	int pick = 0;
	if (type == CRT_TEX_UNIFORM) {
		pick = 0;
	} else if (type == CRT_TEX_CHECKERBOARD) {
		const float s = __ldg(&m->scale);
		pick = (wrap_u32(__fmul_rn(px, s)) + wrap_u32(__fmul_rn(py, s)) + wrap_u32(__fmul_rn(pz, s))) & 1u;
	} else if (type == CRT_TEX_BRICK) {
		const float s = __ldg(&m->scale), mw = __ldg(&m->param1);
		const float sx = __fmul_rn(px, s);
		float sy = __fmul_rn(py, s);
		const uint32_t parity = wrap_u32(sx) & 1u;
		sy = __fsub_rn(sy, __fmul_rn((float)parity, .5f));
		pick = (__fsub_rn(sx, floorf(sx)) < mw) || (__fsub_rn(sy, floorf(sy)) < mw);
	} else {
		const float s = __ldg(&m->scale);
		const float noise = simplex_noise(__fmul_rn(px, s), __fmul_rn(py, s), __fmul_rn(pz, s));
		const float angle = __fmul_rn(__fadd_rn(px, __fmul_rn(noise, __ldg(&m->param1))), __ldg(&m->param2));
		float w;
		switch (__ldg(&m->tex_func)) {
		case CRT_FUNC_SIN: w = __fmul_rn(__fadd_rn(1.f, sinf(angle)), .5f); break;
		case CRT_FUNC_SAW: w = __fsub_rn(angle, floorf(angle)); break;
		case CRT_FUNC_TRIANGLE: w = fabsf(__fmul_rn(2.f, __fsub_rn(__fsub_rn(angle, floorf(angle)), .5f))); break;
		default: w = signbit(sinf(angle)) ? 0.f : 1.f; break;
		}
		r = __fadd_rn(__fmul_rn(__ldg(&m->color[1][0]), w), __ldg(&m->color[0][0]));
		g = __fadd_rn(__fmul_rn(__ldg(&m->color[1][1]), w), __ldg(&m->color[0][1]));
		b = __fadd_rn(__fmul_rn(__ldg(&m->color[1][2]), w), __ldg(&m->color[0][2]));
		return;
	}
	r = __ldg(&m->color[pick][0]);
	g = __ldg(&m->color[pick][1]);
	b = __ldg(&m->color[pick][2]);
}

/* ---- light sampling: reference object.c:293-304, :403-419 ---------------- */

__device__ __forceinline__ void light_point(const DEmitter &e, float px, float py, float pz, uint32_t node, uint32_t draw,
	float &lx, float &ly, float &lz)
{
	const float u0 = crt_rand(node, draw), u1 = crt_rand(node, draw + 1u);
	if (e.is_sphere) {
		const float inclination = u0 * 2.f * kPi;
		const float azimuth = u1 * 2.f * kPi;
		float si, ci, sa, ca;
		sincosf(inclination, &si, &ci);
		sincosf(azimuth, &sa, &ca);
		const float rad = e.a.w;
		float dx = rad * ca * si, dy = rad * sa * si, dz = rad * ci;
		/* negated unless exactly perpendicular to (centre - point): SURVEY Q7 */
		const float d = dot3(e.a.x - px, e.a.y - py, e.a.z - pz, dx, dy, dz);
		if (d != 0.f) {
			dx = -dx;
			dy = -dy;
			dz = -dz;
		}
		lx = e.a.x + dx;
		ly = e.a.y + dy;
		lz = e.a.z + dz;
	} else {
		float p = u0, q = u1;
		if (p + q > 1.f) {
			p = 1.f - p;
			q = 1.f - q;
		}
		lx = e.a.x + e.b.x * p + e.c.x * q;
		ly = e.a.y + e.b.y * p + e.c.y * q;
		lz = e.a.z + e.b.z * p + e.c.z * q;
	}
}

} // namespace crt

#endif /* CRT_DEVICE_CUH */
