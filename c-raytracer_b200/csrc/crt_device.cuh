/*
 * crt_device.cuh -- device data layout and device functions of the trace loop
 * (sm_100a).  Everything here is FP32 CUDA-core work: there is no dense
 * contraction on this path, so no tensor cores.
 *
 * Device layout (all arrays in HBM, read through the read-only path; the
 * working set of every configuration fits the 126 MB L2):
 *   nodes   : BvhNode[ ], 64 B, four float4 loads per visit (bvh_build.h)
 *   prims   : 3 x float4 per bounded primitive, in BVH leaf order ("slot")
 *               triangle: (v0.xyz, eps) (e0.xyz, info) (e1.xyz, object)
 *               sphere  : (c.xyz,  eps) (r, 0, 0, info) (0, 0, 0, object)
 *             info = material index | PRIM_TRANSPARENT | PRIM_SPHERE,
 *             object = scene-order object index (both as raw int bits)
 *   normals : float4 per slot, unit triangle normal (closest-hit shading only)
 *   planes  : 2 x float4 per plane: (n.xyz, d) (eps, info, object, 0)
 *   obj_slot: int per scene object: slot >= 0, or ~plane_index
 *   emitters: DEmitter per emitter, scene order
 *   materials: crt_material (128 B) as in the ABI
 */
#ifndef CRT_DEVICE_CUH
#define CRT_DEVICE_CUH

#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

#include "crt.h"
#include "crt_rng.h"

namespace crt {

constexpr uint32_t PRIM_SPHERE = 0x80000000u;
constexpr uint32_t PRIM_TRANSPARENT = 0x40000000u;
constexpr uint32_t PRIM_MATERIAL_MASK = 0x00ffffffu;

constexpr int kStackSize = 128;         /* >= binary tree depth (closest hit) and 7 * wide tree depth + 1 (shadow walk): checked at scene creation */
constexpr int kStackSentinel = 0x7fffffff;
constexpr int kLeafBits = 3;
#ifndef CRT_WIDE_NODE_STRIDE
#define CRT_WIDE_NODE_STRIDE 6             /* float4 per quantised node in memory: 6 = packed (96 B), 8 = one 128-byte line each */
#endif
constexpr int kWideStride = CRT_WIDE_NODE_STRIDE;
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
	int32_t pad;
};

struct DScene {
	const float4 *nodes;
	const float4 *wide;  /* quantised 8-wide nodes of the shadow walk, 6 float4 each (bvh_build.h QuantNode) */
	const float4 *prims;
	const float4 *normals;
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
	const float4 p0 = ldg4(sc.prims + 3 * slot);
	const float4 p1 = ldg4(sc.prims + 3 * slot + 1);
	const float4 p2 = ldg4(sc.prims + 3 * slot + 2);
	info = __float_as_uint(p1.w);
	object = __float_as_int(p2.w);
	if (info & PRIM_SPHERE)
		return hit_sphere<COUNT>(p0, p1.x, r, t, cn);
	return hit_triangle<COUNT>(p0, p1, p2, r, t, cn);
}

/* ---- BVH traversal --------------------------------------------------------- */

struct RayBox {
	float idx, idy, idz; /* 1 / d */
	float ox, oy, oz;    /* o / d, so that t = plane * id - o */
};

__device__ __forceinline__ RayBox make_raybox(const Ray &r)
{
	RayBox b;
	b.idx = 1.f / r.dx;
	b.idy = 1.f / r.dy;
	b.idz = 1.f / r.dz;
	b.ox = r.ox * b.idx;
	b.oy = r.oy * b.idy;
	b.oz = r.oz * b.idz;
	return b;
}

/* Slab test of both child boxes of a node against [0, tlimit].
 * fminf/fmaxf drop NaNs (0 * inf for an axis-parallel ray starting on a slab
 * plane), which keeps the interval -- the reference's comparisons behave the
 * same way (accel.c:112-158).  The reference's extra `tmax > node epsilon`
 * cull is deliberately not reproduced. */
__device__ __forceinline__ void test_children(const float4 xy0, const float4 xy1, const float4 z01, const Ray &r,
	const RayBox &rb, float tlimit, bool &hit0, bool &hit1, float &tmin0, float &tmin1)
{
	/* (plane - o) * id, written as in the reference so grazing cases round alike */
	const float c0lox = (xy0.x - r.ox) * rb.idx, c0hix = (xy0.y - r.ox) * rb.idx;
	const float c0loy = (xy0.z - r.oy) * rb.idy, c0hiy = (xy0.w - r.oy) * rb.idy;
	const float c0loz = (z01.x - r.oz) * rb.idz, c0hiz = (z01.y - r.oz) * rb.idz;
	const float c1lox = (xy1.x - r.ox) * rb.idx, c1hix = (xy1.y - r.ox) * rb.idx;
	const float c1loy = (xy1.z - r.oy) * rb.idy, c1hiy = (xy1.w - r.oy) * rb.idy;
	const float c1loz = (z01.z - r.oz) * rb.idz, c1hiz = (z01.w - r.oz) * rb.idz;
	tmin0 = fmaxf(fmaxf(fminf(c0lox, c0hix), fminf(c0loy, c0hiy)), fmaxf(fminf(c0loz, c0hiz), 0.f));
	const float tmax0 = fminf(fminf(fmaxf(c0lox, c0hix), fmaxf(c0loy, c0hiy)), fminf(fmaxf(c0loz, c0hiz), tlimit));
	tmin1 = fmaxf(fmaxf(fminf(c1lox, c1hix), fminf(c1loy, c1hiy)), fmaxf(fminf(c1loz, c1hiz), 0.f));
	const float tmax1 = fminf(fminf(fmaxf(c1lox, c1hix), fmaxf(c1loy, c1hiy)), fminf(fmaxf(c1loz, c1hiz), tlimit));
	hit0 = tmin0 <= tmax0;
	hit1 = tmin1 <= tmax1;
}

struct Hit {
	float t;
	int object; /* scene-order object index, -1 = none */
	int slot;   /* prim slot, or ~plane index */
};

/* Nearest hit over planes then the BVH: get_closest_intersection()
 * (reference render.c:118-124, object.c:168-181, accel.c:322-353).
 * Strict '<' like the reference; on an exact tie between bounded primitives
 * the lower scene index wins (the reference's winner depends on its Morton
 * order). */
template <bool COUNT>
__device__ __forceinline__ Hit trace_closest(const DScene &sc, const Ray &r, int *stack, Counters &cn)
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

	const RayBox rb = make_raybox(r);
	int sp = 0;
	stack[sp++] = kStackSentinel;
	int node = 0;
	bool bounded_found = false;
	while (node != kStackSentinel) {
		while (node >= 0 && node != kStackSentinel) {
			const float4 *n = sc.nodes + 4 * node;
			const float4 xy0 = ldg4(n), xy1 = ldg4(n + 1), z01 = ldg4(n + 2);
			const float4 ch = ldg4(n + 3);
			bool h0, h1;
			float t0, t1;
			test_children(xy0, xy1, z01, r, rb, best.t, h0, h1, t0, t1);
			Cnt<COUNT>::add(cn.box, 2u);
			const int c0 = __float_as_int(ch.x), c1 = __float_as_int(ch.y);
			if (!h0 && !h1) {
				node = stack[--sp];
			} else if (h0 && h1) {
				const bool swap = t1 < t0;
				node = swap ? c1 : c0;
				stack[sp++] = swap ? c0 : c1;
			} else {
				node = h0 ? c0 : c1;
			}
		}
		if (node < 0) {
			const int leaf = ~node;
			const int first = leaf >> kLeafBits, count = (leaf & ((1 << kLeafBits) - 1)) + 1;
			for (int i = 0; i < count; i++) {
				float t;
				uint32_t info;
				int object;
				if (hit_prim<COUNT>(sc, first + i, r, t, info, object, cn)) {
					if (t < best.t || (t == best.t && bounded_found && object < best.object)) {
						best.t = t;
						best.object = object;
						best.slot = first + i;
						bounded_found = true;
					}
				}
			}
			node = stack[--sp];
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
 * the quantised 8-wide tree; trace_shadow_wide() strings them together for
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
 * any interleaving: an INNER step (one node: two slab tests, hit leaves go to
 * the lane's leaf queue, hit inner children to its stack) and a PRIM step (one
 * primitive of the queued leaves).  The render kernel lets the warp vote each
 * iteration for the kind more lanes are ready for, instead of a while-while
 * loop in which lanes at a leaf wait for the longest descent of the warp. */
constexpr int kLeafQueue = 16; /* queued leaf codes per lane ([slot][lane]); a node step adds up to 8 */

struct ShadowWalk {
	int node;       /* wide node to visit next, kStackSentinel: none */
	int top;        /* top of the stack, kept in a register: a pop's load is only needed by the NEXT pop */
	int sp;         /* entries in stack[] below top */
	int leaf_first; /* primitive slot being tested next */
	int leaf_left;  /* primitives left in the current leaf */
	int qhead, qcount;
};

__device__ __forceinline__ void shadow_walk_reset(ShadowWalk &w)
{
	w.node = kStackSentinel;
	w.top = kStackSentinel;
	w.sp = 0;
	w.leaf_first = 0;
	w.leaf_left = 0;
	w.qhead = 0;
	w.qcount = 0;
}

__device__ __forceinline__ bool shadow_walk_busy(const ShadowWalk &w)
{
	return w.node != kStackSentinel || w.leaf_left > 0 || w.qcount > 0;
}

__device__ __forceinline__ bool shadow_can_inner(const ShadowWalk &w)
{
	return w.node != kStackSentinel && w.qcount <= kLeafQueue - 8;
}

__device__ __forceinline__ bool shadow_can_prim(const ShadowWalk &w)
{
	return w.leaf_left > 0 || w.qcount > 0;
}

/* The per-ray leaf queue, kLeafQueue entries.  In the render kernel it lives in
 * shared memory as [slot][lane]; the lane's base is kept as a 32-bit shared
 * address in a register the compiler cannot rematerialise (re-deriving it from
 * %tid at every push cost 15 instructions, a fifth of the node step).  The
 * one-ray-per-thread seam kernel uses a private array. */
struct SharedLeafQueue {
	uint32_t base; /* shared-space byte address of slot 0 of this lane */
	__device__ __forceinline__ explicit SharedLeafQueue(const int *lane_slot0)
	{
		uint32_t b = static_cast<uint32_t>(__cvta_generic_to_shared(lane_slot0));
		asm volatile("mov.u32 %0, %1;" : "=r"(base) : "r"(b));
	}
	__device__ __forceinline__ void put(int slot, int v) const
	{
		asm volatile("st.shared.b32 [%0], %1;" ::"r"(base + (uint32_t)slot * 128u), "r"(v) : "memory");
	}
	__device__ __forceinline__ int get(int slot) const
	{
		int v;
		asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(base + (uint32_t)slot * 128u) : "memory");
		return v;
	}
};

struct PrivateLeafQueue {
	int *slots;
	__device__ __forceinline__ void put(int slot, int v) const { slots[slot] = v; }
	__device__ __forceinline__ int get(int slot) const { return slots[slot]; }
};

/* 1/d with a floor on |d|: keeps the quantised slab arithmetic finite (no 0 * inf) and
 * still puts both planes of an axis at -+huge for an axis-parallel ray */
__device__ __forceinline__ float safe_rcp(float d)
{
	return 1.f / (fabsf(d) > 1e-30f ? d : copysignf(1e-30f, d));
}

/* ray constants of the wide walk: idir, and per-axis byte selectors are derived per node */
struct RayWide {
	float idx, idy, idz;
};

/* 2^23 + byte k of word, without a conversion instruction: the byte goes into the mantissa of 2^23 */
__device__ __forceinline__ float byte_to_float(uint32_t word, uint32_t selector)
{
	return __uint_as_float(__byte_perm(word, 0x4b000000u, selector));
}

constexpr float kSlabUlps = 9.5367431640625e-7f; /* 16 * 2^-24 */

/* One visit of a quantised 8-wide node (bvh_build.h QuantNode): six vector loads,
 * eight slab tests; the child pushes are predicated, not branched (lanes hit
 * different children). */
template <bool COUNT, class QUEUE>
__device__ __forceinline__ void shadow_inner_step(const DScene &sc, const Ray &r, const RayWide &rw, float dist, ShadowWalk &w,
	int *stack, const QUEUE &q, Counters &cn)
{
	const float4 *n = sc.wide + kWideStride * (size_t)w.node;
	const float4 w0 = ldg4(n), w1 = ldg4(n + 1), w2 = ldg4(n + 2), w3 = ldg4(n + 3), w4 = ldg4(n + 4), w5 = ldg4(n + 5);
	const uint32_t eb = __float_as_uint(w0.w);
	const float adjx = __uint_as_float((eb & 0xffu) << 23) * rw.idx;
	const float adjy = __uint_as_float(((eb >> 8) & 0xffu) << 23) * rw.idy;
	const float adjz = __uint_as_float(((eb >> 16) & 0xffu) << 23) * rw.idz;
	const float orgx = (w0.x - r.ox) * rw.idx, orgy = (w0.y - r.oy) * rw.idy, orgz = (w0.z - r.oz) * rw.idz;
	/* The walk must never skip a child whose primitive the exact test would accept
	 * (rays through shared edges and vertices, occluders within an ulp of the light
	 * distance).  The plane distances q * adj + org carry a rounding error of at most
	 * 6 ulp of (255 |adj| + |org|) against the true crossing (reciprocal, the two
	 * products, the sums), so the near planes are pulled in and the far planes
	 * pushed out by 16 ulp of (256 |adj| + |org|), per node and axis. */
	const float ex = fmaf(fabsf(adjx), 256.f * kSlabUlps, fabsf(orgx) * kSlabUlps);
	const float ey = fmaf(fabsf(adjy), 256.f * kSlabUlps, fabsf(orgy) * kSlabUlps);
	const float ez = fmaf(fabsf(adjz), 256.f * kSlabUlps, fabsf(orgz) * kSlabUlps);
	/* byte -> float costs one PRMT (2^23 + q in the mantissa); the 2^23 is taken
	 * off in the constant term instead of per plane.  That term then has the
	 * granularity of one quantisation step (ulp of 2^23 |adj|), so it is rounded
	 * towards the outside of the slab. */
	const float onx = __fmaf_rd(-8388608.f, adjx, orgx - ex), ofx = __fmaf_ru(-8388608.f, adjx, orgx + ex);
	const float ony = __fmaf_rd(-8388608.f, adjy, orgy - ey), ofy = __fmaf_ru(-8388608.f, adjy, orgy + ey);
	const float onz = __fmaf_rd(-8388608.f, adjz, orgz - ez), ofz = __fmaf_ru(-8388608.f, adjz, orgz + ez);
	/* near / far planes by the sign of the direction */
	const bool nx = rw.idx < 0.f, ny = rw.idy < 0.f, nz = rw.idz < 0.f;
	const uint32_t nearx[2] = { __float_as_uint(nx ? w1.z : w1.x), __float_as_uint(nx ? w1.w : w1.y) };
	const uint32_t farx[2] = { __float_as_uint(nx ? w1.x : w1.z), __float_as_uint(nx ? w1.y : w1.w) };
	const uint32_t neary[2] = { __float_as_uint(ny ? w2.z : w2.x), __float_as_uint(ny ? w2.w : w2.y) };
	const uint32_t fary[2] = { __float_as_uint(ny ? w2.x : w2.z), __float_as_uint(ny ? w2.y : w2.w) };
	const uint32_t nearz[2] = { __float_as_uint(nz ? w3.z : w3.x), __float_as_uint(nz ? w3.w : w3.y) };
	const uint32_t farz[2] = { __float_as_uint(nz ? w3.x : w3.z), __float_as_uint(nz ? w3.y : w3.w) };
	const int code[8] = { __float_as_int(w4.x), __float_as_int(w4.y), __float_as_int(w4.z), __float_as_int(w4.w),
		__float_as_int(w5.x), __float_as_int(w5.y), __float_as_int(w5.z), __float_as_int(w5.w) };
#pragma unroll
	for (int i = 0; i < 8; i++) {
		const uint32_t sel = 0x7540u + (uint32_t)(i & 3); /* bytes: (b_i, 0x00, 0x00, 0x4b) */
		const float tnx = fmaf(byte_to_float(nearx[i >> 2], sel), adjx, onx);
		const float tny = fmaf(byte_to_float(neary[i >> 2], sel), adjy, ony);
		const float tnz = fmaf(byte_to_float(nearz[i >> 2], sel), adjz, onz);
		const float tfx = fmaf(byte_to_float(farx[i >> 2], sel), adjx, ofx);
		const float tfy = fmaf(byte_to_float(fary[i >> 2], sel), adjy, ofy);
		const float tfz = fmaf(byte_to_float(farz[i >> 2], sel), adjz, ofz);
		const float tmin = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.f));
		const float tmax = fminf(fminf(tfx, tfy), fminf(tfz, dist));
		if (code[i] != kStackSentinel)
			Cnt<COUNT>::add(cn.box);
		/* an empty child has near 255 / far 0 planes on every axis (and the sentinel code) */
		const bool hit = tmin <= tmax && code[i] != kStackSentinel;
		const bool to_stack = hit && code[i] >= 0, to_queue = hit && code[i] < 0;
		if (to_stack)
			stack[w.sp] = w.top;
		if (to_queue)
			q.put((w.qhead + w.qcount) & (kLeafQueue - 1), code[i]);
		w.sp += to_stack ? 1 : 0;
		w.top = to_stack ? code[i] : w.top;
		w.qcount += to_queue ? 1 : 0;
	}
	w.node = w.top;
	w.top = w.sp > 0 ? stack[--w.sp] : kStackSentinel;
}

template <bool COUNT, class QUEUE>
__device__ __forceinline__ void shadow_prim_step(const DScene &sc, const Ray &r, float dist, int emitter_object, ShadowWalk &w,
	const QUEUE &q, bool &blocked, float &lr, float &lg, float &lb, Counters &cn)
{
	if (w.leaf_left == 0) {
		const int leaf = ~q.get(w.qhead);
		w.qhead = (w.qhead + 1) & (kLeafQueue - 1);
		w.qcount--;
		w.leaf_first = leaf >> kLeafBits;
		w.leaf_left = (leaf & ((1 << kLeafBits) - 1)) + 1;
	}
	float t;
	uint32_t info;
	int object;
	const bool hit = hit_prim<COUNT>(sc, w.leaf_first, r, t, info, object, cn);
	w.leaf_first++;
	w.leaf_left--;
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
	int stack[kStackSize];
	int slots[kLeafQueue];
	const PrivateLeafQueue q = { slots };
	ShadowWalk w;
	shadow_walk_reset(w);
	w.node = 0;
	const RayWide rw = { safe_rcp(r.dx), safe_rcp(r.dy), safe_rcp(r.dz) };
	bool blocked = false;
	while (shadow_walk_busy(w)) {
		if (shadow_can_inner(w))
			shadow_inner_step<COUNT>(sc, r, rw, dist, w, stack, q, cn);
		else
			shadow_prim_step<COUNT>(sc, r, dist, emitter_object, w, q, blocked, lr, lg, lb, cn);
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
	const float4 p0 = ldg4(sc.prims + 3 * h.slot);
	const float4 p1 = ldg4(sc.prims + 3 * h.slot + 1);
	const uint32_t info = __float_as_uint(p1.w);
	material = info & PRIM_MATERIAL_MASK;
	eps = p0.w;
	if (info & PRIM_SPHERE) {
		const float inv = __fdiv_rn(1.f, p1.x);
		nx = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dx, h.t), r.ox), p0.x), inv);
		ny = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dy, h.t), r.oy), p0.y), inv);
		nz = __fmul_rn(__fsub_rn(__fadd_rn(__fmul_rn(r.dz, h.t), r.oz), p0.z), inv);
	} else {
		const float4 n = ldg4(sc.normals + h.slot);
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
