/*
 * crt_oracle.c -- TEST INFRASTRUCTURE (oracle).  Not part of the product.
 *
 * CPU restatement, in plain C, of the reference's per-pixel trace loop, over
 * the same flat scene description (include/crt.h) the CUDA library consumes.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library, and only as the checker.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_ref.py compares this file
 * against the unmodified reference compiled in place (oracle/_ref/ref_engine,
 * see oracle/Makefile) whenever /root/reference is present, and against the
 * golden images/ray batches that build produced (tests/golden/) everywhere
 * else.  Floating-point operations are written in the reference's order and
 * the file is compiled with -O2 -ffp-contract=off like the numerics build of
 * the reference, so images agree bit for bit except where two surfaces are
 * hit at exactly the same distance or a hit lies closer than a
 * sibling's epsilon in the reference's BVH.
 *
 * Each function cites the reference code it follows (paths relative to
 * /root/reference).
 */
#define _GNU_SOURCE
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "crt.h"
#include "crt_rng.h"

#define OR_PI 3.1415927f /* src/core/type.h:30 */

typedef float v3[3];

/* ---- vector helpers: src/core/calc.c ----------------------------------- */

static inline float sqr(float v) { return v * v; }
static inline float dot3(const float *a, const float *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline float magsqr3(const float *a) { return sqr(a[0]) + sqr(a[1]) + sqr(a[2]); }
static inline float mag3(const float *a) { return sqrtf(magsqr3(a)); }
static inline void mul3s(const float *a, float s, float *r) { r[0] = a[0] * s; r[1] = a[1] * s; r[2] = a[2] * s; }
static inline void mul3v(const float *a, const float *b, float *r) { r[0] = a[0] * b[0]; r[1] = a[1] * b[1]; r[2] = a[2] * b[2]; }
static inline void add3v(const float *a, const float *b, float *r) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
static inline void sub3v(const float *a, const float *b, float *r) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
static inline void assign3(float *d, const float *s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
static inline void norm3(float *a) { mul3s(a, 1.f / mag3(a), a); }
static inline void cross3(const float *a, const float *b, float *r)
{
	const float x = a[1] * b[2] - a[2] * b[1];
	const float y = a[2] * b[0] - a[0] * b[2];
	const float z = a[0] * b[1] - a[1] * b[0];
	r[0] = x; r[1] = y; r[2] = z;
}

/* ---- scene ------------------------------------------------------------- */

struct or_node { /* own BVH: median split, results are topology independent */
	float lo[3], hi[3];
	int32_t left, right; /* children, or left = -1 for a leaf */
	uint32_t first, count; /* leaf range in prim[] */
};

typedef struct oracle_scene {
	uint32_t num_objects, num_materials;
	crt_object *objects;
	crt_material *materials;
	float ambient[3];
	uint32_t num_planes, num_bounded, num_emitters;
	uint32_t *planes;   /* object indices, scene order */
	uint32_t *prim;     /* bounded object indices, BVH order */
	uint32_t *emitters; /* object indices, scene order */
	struct or_node *nodes;
	uint32_t num_nodes;
	int brute_force;
} oracle_scene;

struct or_counters {
	uint64_t closest, shadow, shaded;
};

/* ---- simplex noise: lib/SimplexNoise/SimplexNoise.c:74-212 ------------- */

static const uint8_t PERM[256] = {
	151,160,137,91,90,15,131,13,201,95,96,53,194,233,7,225,140,36,103,30,69,142,8,99,37,240,21,10,23,190,6,148,
	247,120,234,75,0,26,197,62,94,252,219,203,117,35,11,32,57,177,33,88,237,149,56,87,174,20,125,136,171,168,68,175,
	74,165,71,134,139,48,27,166,77,146,158,231,83,111,229,122,60,211,133,230,220,105,92,41,55,46,245,40,244,102,143,54,
	65,25,63,161,1,216,80,73,209,76,132,187,208,89,18,169,200,196,135,130,116,188,159,86,164,100,109,198,173,186,3,64,
	52,217,226,250,124,123,5,202,38,147,118,126,255,82,85,212,207,206,59,227,47,16,58,17,182,189,28,42,223,183,170,213,
	119,248,152,2,44,154,163,70,221,153,101,155,167,43,172,9,129,22,39,253,19,98,108,110,79,113,224,232,178,185,112,104,
	218,246,97,228,251,34,242,193,238,210,144,12,191,179,162,241,81,51,145,235,249,14,239,107,49,192,214,31,181,199,106,157,
	184,84,204,176,115,121,50,45,127,4,150,254,138,236,205,93,222,114,67,29,24,72,243,141,128,195,78,66,215,61,156,180
};

static inline int perm_at(int i) { return PERM[(uint8_t)i]; } /* int -> uint8 wrap, SURVEY Q23 */

static float corner_gradient(int hash, float x, float y, float z) /* grad(), :100-106 */
{
	const int h = hash & 15;
	const float u = h < 8 ? x : y;
	const float v = h < 4 ? y : (h == 12 || h == 14) ? x : z;
	return ((h & 1) ? -u : u) + ((h & 2) ? -v : v);
}

static float corner_term(int gi, float x, float y, float z)
{
	float t = 0.6f - x * x - y * y - z * z;
	if (t < 0)
		return 0.0f;
	t *= t;
	return t * t * corner_gradient(gi, x, y, z);
}

float oracle_simplex_noise(float x, float y, float z)
{
	const float F3 = 1.0f / 3.0f, G3 = 1.0f / 6.0f;
	const float s = (x + y + z) * F3;
	const int i = (int)floorf(x + s), j = (int)floorf(y + s), k = (int)floorf(z + s);
	const float t = (float)(i + j + k) * G3;
	const float x0 = x - ((float)i - t), y0 = y - ((float)j - t), z0 = z - ((float)k - t);

	int i1, j1, k1, i2, j2, k2; /* second and third simplex corner, by rank order of x0,y0,z0 */
	if (x0 >= y0) {
		if (y0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
		else if (x0 >= z0) { i1 = 1; j1 = 0; k1 = 0; i2 = 1; j2 = 0; k2 = 1; }
		else { i1 = 0; j1 = 0; k1 = 1; i2 = 1; j2 = 0; k2 = 1; }
	} else {
		if (y0 < z0) { i1 = 0; j1 = 0; k1 = 1; i2 = 0; j2 = 1; k2 = 1; }
		else if (x0 < z0) { i1 = 0; j1 = 1; k1 = 0; i2 = 0; j2 = 1; k2 = 1; }
		else { i1 = 0; j1 = 1; k1 = 0; i2 = 1; j2 = 1; k2 = 0; }
	}
	const float x1 = x0 - (float)i1 + G3, y1 = y0 - (float)j1 + G3, z1 = z0 - (float)k1 + G3;
	const float x2 = x0 - (float)i2 + 2.0f * G3, y2 = y0 - (float)j2 + 2.0f * G3, z2 = z0 - (float)k2 + 2.0f * G3;
	const float x3 = x0 - 1.0f + 3.0f * G3, y3 = y0 - 1.0f + 3.0f * G3, z3 = z0 - 1.0f + 3.0f * G3;

	const int gi0 = perm_at(i + perm_at(j + perm_at(k)));
	const int gi1 = perm_at(i + i1 + perm_at(j + j1 + perm_at(k + k1)));
	const int gi2 = perm_at(i + i2 + perm_at(j + j2 + perm_at(k + k2)));
	const int gi3 = perm_at(i + 1 + perm_at(j + 1 + perm_at(k + 1)));

	const float n0 = corner_term(gi0, x0, y0, z0);
	const float n1 = corner_term(gi1, x1, y1, z1);
	const float n2 = corner_term(gi2, x2, y2, z2);
	const float n3 = corner_term(gi3, x3, y3, z3);
	return 32.0f * (n0 + n1 + n2 + n3);
}

/* ---- textures: src/raytracer/material.c:152-200 ------------------------ */

/* (uint32_t) of a possibly negative float as gcc/x86-64 evaluates it: convert
 * to a 64-bit integer, keep the low 32 bits */
static inline uint32_t wrap_u32(float f) { return (uint32_t)(int64_t)f; }

static void texture_color(const crt_material *m, const float *p, float *color)
{
	switch (m->tex_type) {
	case CRT_TEX_UNIFORM:
		assign3(color, m->color[0]);
		break;
	case CRT_TEX_CHECKERBOARD: {
		v3 sp;
		mul3s(p, m->scale, sp);
		const uint32_t parity = (wrap_u32(sp[0]) + wrap_u32(sp[1]) + wrap_u32(sp[2])) % 2u;
		assign3(color, m->color[parity]);
	} break;
	case CRT_TEX_BRICK: {
		v3 sp;
		mul3s(p, m->scale, sp);
		const uint32_t parity = wrap_u32(sp[0]) % 2u;
		sp[1] -= (float)parity * .5f;
		const int mortar = (sp[0] - floorf(sp[0]) < m->param1) || (sp[1] - floorf(sp[1]) < m->param1);
		assign3(color, m->color[mortar]);
	} break;
	default: { /* noisy periodic */
		v3 sp;
		mul3s(p, m->scale, sp);
		const float angle = (p[0] + oracle_simplex_noise(sp[0], sp[1], sp[2]) * m->param1) * m->param2;
		float w;
		switch (m->tex_func) {
		case CRT_FUNC_SIN: w = (1.f + sinf(angle)) * .5f; break;
		case CRT_FUNC_SAW: w = angle - floorf(angle); break;
		case CRT_FUNC_TRIANGLE: w = fabsf(2.f * (angle - floorf(angle) - .5f)); break;
		default: w = (float)!signbit(sinf(angle)); break;
		}
		mul3s(m->color[1], w, color);
		add3v(color, m->color[0], color);
	} break;
	}
}

/* ---- primitives: src/raytracer/object.c -------------------------------- */

static int hit_sphere(const crt_object *o, const float *org, const float *dir, float *t) /* :306-321 */
{
	v3 rel;
	sub3v(org, o->a, rel);
	const float b = -dot3(dir, rel);
	const float c = dot3(rel, rel) - sqr(o->b[0]);
	const float det = sqr(b) - c;
	if (det < 0)
		return 0;
	const float sq = sqrtf(det);
	*t = b - sq;
	if (*t > o->epsilon)
		return 1;
	*t = b + sq;
	return *t > o->epsilon;
}

static int hit_triangle(const crt_object *o, const float *org, const float *dir, float *t) /* :422-441 */
{
	v3 h, s, q;
	cross3(dir, o->c, h);
	const float a = dot3(o->b, h);
	if (a < o->epsilon && a > -o->epsilon)
		return 0;
	const float f = 1.f / a;
	sub3v(org, o->a, s);
	const float u = f * dot3(s, h);
	if (u < 0.f || u > 1.f)
		return 0;
	cross3(s, o->b, q);
	const float v = f * dot3(dir, q);
	if (v < 0.f || u + v > 1.f)
		return 0;
	*t = f * dot3(o->c, q);
	return *t > o->epsilon;
}

static int hit_plane(const crt_object *o, const float *org, const float *dir, float *t, float *a_out) /* :473-498 */
{
	const float a = dot3(o->a, dir);
	if (fabsf(a) < o->epsilon)
		return 0;
	*t = (o->b[0] - dot3(o->a, org)) / dot3(o->a, dir);
	*a_out = a;
	return *t > o->epsilon;
}

/* get_intersection of any object kind, with the normal rules of SURVEY Q4:
 * sphere outward (:254-265), triangle from winding (:356-365), plane facing
 * the ray (:481-484) */
static int object_intersect(const crt_object *o, const float *org, const float *dir, float *t, float *normal)
{
	switch (o->type) {
	case CRT_OBJ_SPHERE:
		if (!hit_sphere(o, org, dir, t))
			return 0;
		mul3s(dir, *t, normal);
		add3v(normal, org, normal);
		sub3v(normal, o->a, normal);
		mul3s(normal, 1.f / o->b[0], normal);
		return 1;
	case CRT_OBJ_TRIANGLE:
		if (!hit_triangle(o, org, dir, t))
			return 0;
		assign3(normal, o->n);
		return 1;
	default: {
		float a;
		if (!hit_plane(o, org, dir, t, &a))
			return 0;
		if (signbit(a))
			assign3(normal, o->a);
		else
			mul3s(o->a, -1.f, normal);
		return 1;
	}
	}
}

static int bounded_hit(const crt_object *o, const float *org, const float *dir, float *t)
{
	return o->type == CRT_OBJ_SPHERE ? hit_sphere(o, org, dir, t) : hit_triangle(o, org, dir, t);
}

/* ---- own BVH (the reference's tree is implementation-defined, SURVEY F8;
 * closest-hit and transmittance results do not depend on the topology) ----- */

static void object_bounds(const crt_object *o, float *lo, float *hi)
{
	if (o->type == CRT_OBJ_SPHERE) {
		for (int k = 0; k < 3; k++) {
			lo[k] = o->a[k] - o->b[0];
			hi[k] = o->a[k] + o->b[0];
		}
	} else {
		for (int k = 0; k < 3; k++) {
			const float p0 = o->a[k], p1 = o->a[k] + o->b[k], p2 = o->a[k] + o->c[k];
			lo[k] = fminf(p0, fminf(p1, p2));
			hi[k] = fmaxf(p0, fmaxf(p1, p2));
		}
	}
	for (int k = 0; k < 3; k++) { /* pad: the slab test must never cull a true hit */
		const float pad = 1e-5f * fmaxf(1.f, fmaxf(fabsf(lo[k]), fabsf(hi[k])));
		lo[k] -= pad;
		hi[k] += pad;
	}
}

struct build_ctx {
	oracle_scene *sc;
	float *centroid; /* 3 per object index */
	int axis;
};

static struct build_ctx *g_sort_ctx; /* qsort has no context argument; builds are serialised */

static int cmp_centroid(const void *pa, const void *pb)
{
	const uint32_t a = *(const uint32_t *)pa, b = *(const uint32_t *)pb;
	const float ca = g_sort_ctx->centroid[3 * a + g_sort_ctx->axis], cb = g_sort_ctx->centroid[3 * b + g_sort_ctx->axis];
	return (ca > cb) - (ca < cb);
}

static int32_t build_node(struct build_ctx *bc, uint32_t first, uint32_t count)
{
	oracle_scene *sc = bc->sc;
	const int32_t id = (int32_t)sc->num_nodes++;
	struct or_node *n = &sc->nodes[id];
	float clo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, chi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
	for (int k = 0; k < 3; k++) {
		n->lo[k] = FLT_MAX;
		n->hi[k] = -FLT_MAX;
	}
	for (uint32_t i = first; i < first + count; i++) {
		float lo[3], hi[3];
		object_bounds(&sc->objects[sc->prim[i]], lo, hi);
		for (int k = 0; k < 3; k++) {
			n->lo[k] = fminf(n->lo[k], lo[k]);
			n->hi[k] = fmaxf(n->hi[k], hi[k]);
			const float c = bc->centroid[3 * sc->prim[i] + k];
			clo[k] = fminf(clo[k], c);
			chi[k] = fmaxf(chi[k], c);
		}
	}
	n->first = first;
	n->count = count;
	n->left = n->right = -1;
	if (count <= 2)
		return id;
	int axis = 0;
	if (chi[1] - clo[1] > chi[axis] - clo[axis])
		axis = 1;
	if (chi[2] - clo[2] > chi[axis] - clo[axis])
		axis = 2;
	bc->axis = axis;
	g_sort_ctx = bc;
	qsort(sc->prim + first, count, sizeof(uint32_t), cmp_centroid);
	const uint32_t half = count / 2;
	const int32_t l = build_node(bc, first, half);
	const int32_t r = build_node(bc, first + half, count - half);
	n = &sc->nodes[id]; /* nodes[] is preallocated, pointer stays valid; re-read for clarity */
	n->left = l;
	n->right = r;
	return id;
}

static int slab(const struct or_node *n, const float *org, const float *inv, float tmax_limit)
{
	float t0 = 0.f, t1 = tmax_limit;
	for (int k = 0; k < 3; k++) {
		float a = (n->lo[k] - org[k]) * inv[k], b = (n->hi[k] - org[k]) * inv[k];
		if (a > b) { const float s = a; a = b; b = s; }
		/* NaN (0 * inf) compares false: keeps the interval, like the reference's test */
		if (a > t0) t0 = a;
		if (b < t1) t1 = b;
	}
	return t0 <= t1;
}

/* ---- the two queries: src/raytracer/render.c:118-134 -------------------- */

/* closest hit over planes (object.c:168-181) then bounded objects
 * (accel.c:322-353).  Strict "<" as the reference; on an exact tie between
 * bounded objects the lower scene index wins (the reference's winner depends
 * on its Morton order). */
static void closest_hit(const oracle_scene *sc, const float *org, const float *dir, int32_t *obj, float *normal, float *best)
{
	for (uint32_t i = 0; i < sc->num_planes; i++) {
		float t;
		v3 n;
		if (object_intersect(&sc->objects[sc->planes[i]], org, dir, &t, n) && t < *best) {
			*best = t;
			*obj = (int32_t)sc->planes[i];
			assign3(normal, n);
		}
	}
	int32_t bounded = -1;
	float bt = *best;
	if (sc->brute_force || !sc->num_nodes) {
		for (uint32_t i = 0; i < sc->num_bounded; i++) {
			const uint32_t oi = sc->prim[i];
			float t;
			if (bounded_hit(&sc->objects[oi], org, dir, &t) && (t < bt || (t == bt && bounded >= 0 && (int32_t)oi < bounded))) {
				bt = t;
				bounded = (int32_t)oi;
			}
		}
	} else {
		const float inv[3] = { 1.f / dir[0], 1.f / dir[1], 1.f / dir[2] };
		int32_t stack[128];
		int sp = 0;
		stack[sp++] = 0;
		while (sp) {
			const struct or_node *n = &sc->nodes[stack[--sp]];
			if (!slab(n, org, inv, bt))
				continue;
			if (n->left < 0) {
				for (uint32_t i = n->first; i < n->first + n->count; i++) {
					const uint32_t oi = sc->prim[i];
					float t;
					if (bounded_hit(&sc->objects[oi], org, dir, &t) && (t < bt || (t == bt && bounded >= 0 && (int32_t)oi < bounded))) {
						bt = t;
						bounded = (int32_t)oi;
					}
				}
			} else {
				stack[sp++] = n->left;
				stack[sp++] = n->right;
			}
		}
	}
	if (bounded >= 0) {
		float t;
		object_intersect(&sc->objects[bounded], org, dir, &t, normal);
		*best = t;
		*obj = bounded;
	}
}

static int occluder(const oracle_scene *sc, const crt_object *o, float *light)
{
	const crt_material *m = &sc->materials[o->material];
	if (m->flags & CRT_MAT_TRANSPARENT) {
		mul3v(light, m->kt, light);
		return 0;
	}
	return 1;
}

/* any-hit with transmittance: planes (object.c:183-197) then bounded objects
 * (accel.c:360-387); the sampled emitter itself never occludes (accel.c:367) */
static int light_blocked(const oracle_scene *sc, const float *org, const float *dir, float dist, float *light, int32_t emitter)
{
	for (uint32_t i = 0; i < sc->num_planes; i++) {
		const crt_object *o = &sc->objects[sc->planes[i]];
		float t, a;
		if (hit_plane(o, org, dir, &t, &a) && t < dist && occluder(sc, o, light))
			return 1;
	}
	if (sc->brute_force || !sc->num_nodes) {
		for (uint32_t i = 0; i < sc->num_bounded; i++) {
			const uint32_t oi = sc->prim[i];
			float t;
			if ((int32_t)oi != emitter && bounded_hit(&sc->objects[oi], org, dir, &t) && t < dist && occluder(sc, &sc->objects[oi], light))
				return 1;
		}
		return 0;
	}
	const float inv[3] = { 1.f / dir[0], 1.f / dir[1], 1.f / dir[2] };
	int32_t stack[128];
	int sp = 0;
	stack[sp++] = 0;
	while (sp) {
		const struct or_node *n = &sc->nodes[stack[--sp]];
		if (!slab(n, org, inv, dist))
			continue;
		if (n->left < 0) {
			for (uint32_t i = n->first; i < n->first + n->count; i++) {
				const uint32_t oi = sc->prim[i];
				float t;
				if ((int32_t)oi != emitter && bounded_hit(&sc->objects[oi], org, dir, &t) && t < dist && occluder(sc, &sc->objects[oi], light))
					return 1;
			}
		} else {
			stack[sp++] = n->left;
			stack[sp++] = n->right;
		}
	}
	return 0;
}

/* ---- light sampling: src/raytracer/object.c:293-304, :403-419 ---------- */

static void light_point(const crt_object *e, const float *point, uint32_t node, uint32_t *draw, float *out)
{
	if (e->type == CRT_OBJ_SPHERE) {
		v3 to_centre;
		sub3v(e->a, point, to_centre);
		const float inclination = crt_rand(node, (*draw)++) * 2.f * OR_PI;
		const float azimuth = crt_rand(node, (*draw)++) * 2.f * OR_PI;
		const float r = e->b[0];
		v3 d = { r * cosf(azimuth) * sinf(inclination), r * sinf(azimuth) * sinf(inclination), r * cosf(inclination) };
		if (dot3(to_centre, d)) /* non-zero, i.e. practically always */
			mul3s(d, -1.f, d);
		add3v(e->a, d, out);
	} else {
		float p = crt_rand(node, (*draw)++);
		float q = crt_rand(node, (*draw)++);
		if (p + q > 1.f) {
			p = 1.f - p;
			q = 1.f - q;
		}
		/* vertices[1]-vertices[0] and vertices[2]-vertices[0] are exactly the stored edges */
		for (int k = 0; k < 3; k++)
			out[k] = e->a[k] + e->b[k] * p + e->c[k] * q;
	}
}

/* ---- cast_ray: src/raytracer/render.c:136-343 -------------------------- */

struct or_frame_ctx {
	const oracle_scene *sc;
	const crt_render_opts *o;
	struct or_counters *cnt;
};

static float cast_ray(const struct or_frame_ctx *cx, const float *org, const float *dir, const float *kr,
	float *color, uint32_t remaining, int32_t inside, uint32_t node)
{
	const oracle_scene *sc = cx->sc;
	const crt_render_opts *o = cx->o;
	int32_t obj = -1;
	v3 normal;
	float t;
	uint32_t draw = 0, children = 0;

	cx->cnt->closest++;
	if (inside >= 0 && object_intersect(&sc->objects[inside], org, dir, &t, normal)) {
		obj = inside; /* inside-object shortcut, :143-144 */
	} else {
		t = FLT_MAX;
		closest_hit(sc, org, dir, &obj, normal, &t);
	}
	if (obj < 0)
		return 0.f;
	cx->cnt->shaded++;

	v3 P, out_dir;
	mul3s(dir, t, P);
	add3v(P, org, P);

	const crt_object *object = &sc->objects[obj];
	const crt_material *m = &sc->materials[object->material];
	v3 oc;
	assign3(oc, m->ke);

	const float b = dot3(normal, dir);
	const int outside = signbit(b) != 0;

	for (uint32_t ei = 0; ei < sc->num_emitters; ei++) { /* :170-229 */
		const uint32_t eidx = sc->emitters[ei];
		if ((int32_t)eidx == obj)
			continue;
		const crt_object *e = &sc->objects[eidx];
		v3 base;
		mul3s(sc->materials[e->material].ke, 1.f / (float)e->num_lights, base);
		for (uint32_t j = 0; j < e->num_lights; j++) {
			v3 lp, li, ld;
			light_point(e, P, node, &draw, lp);
			assign3(li, base);
			sub3v(lp, P, ld);
			const float ldist = mag3(ld);
			mul3s(ld, 1.f / ldist, ld);
			const float a = dot3(ld, normal);
			if (!outside)
				continue;
			cx->cnt->shadow++;
			if (light_blocked(sc, P, ld, ldist, li, (int32_t)eidx))
				continue;
			v3 dvec;
			sub3v(lp, P, dvec);
			if (o->attenuation == CRT_ATTEN_LINEAR)
				mul3s(li, 1.f / (o->attenuation_offset + mag3(dvec)), li);
			else if (o->attenuation == CRT_ATTEN_SQUARE)
				mul3s(li, 1.f / (o->attenuation_offset + magsqr3(dvec)), li);

			v3 diffuse;
			texture_color(m, P, diffuse);
			mul3v(diffuse, li, diffuse);
			mul3s(diffuse, fmaxf(0.f, a), diffuse);

			v3 refl;
			float smul;
			if (o->reflection_model == CRT_REFLECTION_BLINN) { /* :215-220, SURVEY Q27 */
				mul3s(ld, -1.f, refl);
				add3v(refl, dir, refl);
				norm3(refl);
				smul = -dot3(normal, refl);
			} else {
				mul3s(normal, 2 * a, refl);
				sub3v(refl, ld, refl);
				smul = -dot3(refl, dir);
			}
			v3 spec;
			mul3v(m->ks, li, spec);
			mul3s(spec, fmaxf(0.f, powf(smul, m->shininess)), spec); /* raw base, SURVEY Q11 */
			for (int k = 0; k < 3; k++)
				oc[k] = oc[k] + diffuse[k] + spec[k];
		}
	}

	if (o->gi_model == CRT_GI_AMBIENT) { /* :233-237 */
		v3 amb;
		mul3v(m->ka, sc->ambient, amb);
		add3v(oc, amb, oc);
	} else if (remaining && outside) { /* :239-288, SURVEY Q12 */
		float R[3][3];
		if (normal[1] - object->epsilon < -1.f) {
			const float flip[3][3] = { { 1.f, 0.f, 0.f }, { 0.f, -1.f, 0.f }, { 0.f, 0.f, -1.f } };
			memcpy(R, flip, sizeof R);
		} else {
			const v3 up = { 0.f, 1.f, 0.f };
			const float mul = 1.f / (1.f + dot3(up, normal));
			R[0][0] = 1.f - sqr(normal[0]) * mul;
			R[0][1] = normal[0];
			R[0][2] = -normal[0] * normal[2] * mul;
			R[1][0] = -normal[0];
			R[1][1] = 1.f - (sqr(normal[0]) + sqr(normal[2])) * mul;
			R[1][2] = -normal[2];
			R[2][0] = -normal[0] * normal[2] * mul;
			R[2][1] = normal[2];
			R[2][2] = 1.f - sqr(normal[2]) * mul;
		}
		v3 delta = { 1.f, 1.f, 1.f };
		uint32_t samples = 1;
		if (remaining == o->max_bounces) {
			samples = o->samples_per_pixel;
			mul3s(delta, 1.f / (float)samples, delta);
		}
		for (uint32_t i = 0; i < samples; i++) {
			const float inclination = acosf(crt_rand(node, draw++) * 2.f - 1.f);
			const float azimuth = crt_rand(node, draw++) * OR_PI;
			const v3 s = { 1 * cosf(azimuth) * sinf(inclination), 1 * sinf(azimuth) * sinf(inclination), 1 * cosf(inclination) };
			out_dir[0] = dot3(R[0], s);
			out_dir[1] = dot3(R[1], s);
			out_dir[2] = dot3(R[2], s);
			v3 weight;
			mul3s(delta, dot3(normal, out_dir), weight);
			cast_ray(cx, P, out_dir, weight, oc, 0, -1, crt_node_child(node, children++));
		}
	}

	mul3v(oc, kr, oc); /* :291-302, SURVEY Q10 */
	if (o->attenuation == CRT_ATTEN_LINEAR)
		mul3s(oc, 1.f / (o->attenuation_offset + t), oc);
	else if (o->attenuation == CRT_ATTEN_SQUARE)
		mul3s(oc, 1.f / sqr(o->attenuation_offset + t), oc);
	add3v(color, oc, color);

	if (!remaining)
		return 0.f; /* :304-305, SURVEY Q3 */

	if (inside != obj && (m->flags & CRT_MAT_REFLECTIVE)) { /* :308-317, SURVEY Q13, Q16 */
		v3 rkr;
		mul3v(kr, m->kr, rkr);
		if (o->min_intensity_sqr < magsqr3(rkr)) {
			mul3s(normal, 2 * b, out_dir);
			sub3v(dir, out_dir, out_dir);
			cast_ray(cx, P, out_dir, rkr, color, remaining - 1, -1, crt_node_child(node, children++));
		}
	}

	if (m->flags & CRT_MAT_TRANSPARENT) { /* :320-340, SURVEY Q15 */
		v3 tkt;
		mul3v(kr, m->kt, tkt);
		if (o->min_intensity_sqr < magsqr3(tkt)) {
			const float incident = acosf(fabsf(b));
			const float ratio = outside ? 1.f / m->refractive_index : m->refractive_index;
			const float refracted = asinf(sinf(incident) * ratio);
			const float turn = refracted - incident;
			v3 c, f, g, h;
			cross3(dir, normal, c);
			norm3(c);
			if (!outside)
				mul3s(c, -1.f, c);
			cross3(c, dir, f);
			mul3s(dir, cosf(turn), g);
			mul3s(f, sinf(turn), h);
			add3v(g, h, out_dir);
			norm3(out_dir);
			cast_ray(cx, P, out_dir, tkt, color, remaining - 1, obj, crt_node_child(node, children++));
		}
	}
	return t;
}

/* ---- public entry points ------------------------------------------------ */

void oracle_scene_destroy(oracle_scene *sc)
{
	if (!sc)
		return;
	free(sc->objects);
	free(sc->materials);
	free(sc->planes);
	free(sc->prim);
	free(sc->emitters);
	free(sc->nodes);
	free(sc);
}

oracle_scene *oracle_scene_create(const crt_scene_desc *desc, int brute_force)
{
	oracle_scene *sc = calloc(1, sizeof *sc);
	const uint32_t n = desc->num_objects;
	sc->num_objects = n;
	sc->num_materials = desc->num_materials;
	sc->objects = malloc(sizeof(crt_object) * (n ? n : 1));
	sc->materials = malloc(sizeof(crt_material) * (desc->num_materials ? desc->num_materials : 1));
	memcpy(sc->objects, desc->objects, sizeof(crt_object) * n);
	memcpy(sc->materials, desc->materials, sizeof(crt_material) * desc->num_materials);
	memcpy(sc->ambient, desc->ambient, sizeof sc->ambient);
	sc->planes = malloc(sizeof(uint32_t) * (n ? n : 1));
	sc->prim = malloc(sizeof(uint32_t) * (n ? n : 1));
	sc->emitters = malloc(sizeof(uint32_t) * (n ? n : 1));
	sc->brute_force = brute_force;
	float *centroid = malloc(sizeof(float) * 3 * (n ? n : 1));
	for (uint32_t i = 0; i < n; i++) {
		const crt_object *o = &sc->objects[i];
		if (o->type == CRT_OBJ_PLANE) {
			sc->planes[sc->num_planes++] = i;
		} else {
			float lo[3], hi[3];
			object_bounds(o, lo, hi);
			for (int k = 0; k < 3; k++)
				centroid[3 * i + k] = 0.5f * (lo[k] + hi[k]);
			sc->prim[sc->num_bounded++] = i;
			/* emitters: bounded objects with an emittant material, scene order
			 * (scene.c:357-358; planes cannot emit, object.c:452) */
			if (sc->materials[o->material].flags & CRT_MAT_EMITTANT)
				sc->emitters[sc->num_emitters++] = i;
		}
	}
	if (sc->num_bounded && !brute_force) {
		sc->nodes = malloc(sizeof(struct or_node) * 2 * sc->num_bounded);
		struct build_ctx bc = { sc, centroid, 0 };
		build_node(&bc, 0, sc->num_bounded);
	}
	free(centroid);
	return sc;
}

/* render(): src/raytracer/render.c:345-368, including the running FP32 sum of
 * the pixel position along a row and the (col+1) offset */
int oracle_render(const oracle_scene *sc, const crt_render_opts *o, float *raster, float *z_buffer, crt_stats *stats)
{
	const size_t pixels = (size_t)o->width * o->height;
	memset(raster, 0, pixels * 3 * sizeof(float));
	memset(z_buffer, 0, pixels * sizeof(float));
	uint64_t n_closest = 0, n_shadow = 0, n_shaded = 0;
	const v3 kr = { 1.f, 1.f, 1.f };
	const uint32_t shards = o->shard_count ? o->shard_count : 1u;
	const uint32_t tiles_x = (o->width + CRT_TILE_W - 1) / CRT_TILE_W;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_closest, n_shadow, n_shaded)
	for (uint32_t row = 0; row < o->height; row++) {
		struct or_counters cnt = { 0, 0, 0 };
		const struct or_frame_ctx cx = { sc, o, &cnt };
		v3 pos, dir;
		mul3s(o->vy, (float)row, pos);
		add3v(pos, o->corner, pos);
		uint32_t pixel = o->width * row;
		for (uint32_t col = 0; col < o->width; col++, pixel++) {
			add3v(pos, o->vx, pos);
			if (shards > 1) {
				/* frame partition of include/crt.h: tile -> chunk -> shard */
				const uint32_t tile = (row / CRT_TILE_H) * tiles_x + col / CRT_TILE_W;
				if ((tile / CRT_SHARD_CHUNK_TILES) % shards != o->shard_index)
					continue;
			}
			sub3v(pos, o->cam_pos, dir);
			norm3(dir);
			z_buffer[pixel] = cast_ray(&cx, o->cam_pos, dir, kr, raster + 3 * (size_t)pixel, o->max_bounces, -1,
				crt_node_root(o->seed, pixel));
		}
		n_closest += cnt.closest;
		n_shadow += cnt.shadow;
		n_shaded += cnt.shaded;
	}
	if (stats) {
		memset(stats, 0, sizeof *stats);
		stats->rays_closest = n_closest;
		stats->rays_shadow = n_shadow;
		stats->nodes_shaded = n_shaded;
	}
	return 0;
}

/* same conventions as crt_trace_closest / crt_trace_shadow / crt_texture_eval (include/crt.h) */
int oracle_trace_closest(const oracle_scene *sc, const float *rays, size_t n, int32_t *object, float *t, float *normal)
{
	for (size_t i = 0; i < n; i++) {
		int32_t obj = -1;
		float best = FLT_MAX;
		v3 nrm = { 0.f, 0.f, 0.f };
		closest_hit(sc, rays + 6 * i + 3, rays + 6 * i, &obj, nrm, &best);
		object[i] = obj;
		t[i] = best;
		if (normal)
			assign3(normal + 3 * i, nrm);
	}
	return 0;
}

int oracle_trace_shadow(const oracle_scene *sc, const float *rays, const float *dist, const int32_t *emitter,
	size_t n, float *light, uint8_t *blocked)
{
	for (size_t i = 0; i < n; i++)
		blocked[i] = (uint8_t)light_blocked(sc, rays + 6 * i + 3, rays + 6 * i, dist[i], light + 3 * i, emitter ? emitter[i] : -1);
	return 0;
}

int oracle_texture_eval(const oracle_scene *sc, int32_t material, const float *points, size_t n, float *colors)
{
	if (material < 0 || (uint32_t)material >= sc->num_materials)
		return -1;
	for (size_t i = 0; i < n; i++)
		texture_color(&sc->materials[material], points + 3 * i, colors + 3 * i);
	return 0;
}

/* primitive known-answer seams: kind 0 triangle, 1 sphere (see tests/test_kat.py) */
int oracle_primitive_hit(int kind, const crt_object *o, const float *org, const float *dir, float *t)
{
	*t = 0.f;
	return kind == 0 ? hit_triangle(o, org, dir, t) : hit_sphere(o, org, dir, t);
}
