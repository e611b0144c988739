/* node_test_emu -- the device's node test (c-raytracer_b200/csrc/crt_device.cuh, wide_node_test) restated on the
 * CPU with its half-precision operations emulated bit for bit (every HADD2 / HMUL2 / HFMA2 / float -> half conversion = the exact result
 * rounded once to binary16, round to nearest even), run on the host builder's tree:
 *     node_test_emu <boxes> <rays per kind> <seed>
 * For every (ray, node, slot) tried: if the segment of the ray meets the TRUE box of what lies below the slot
 * (exact test in double precision), the emulated test must report the slot.  A miss there is a primitive the
 * walk would never see.  Kinds of rays: random; with one or two zero direction components; aimed at the
 * faces, edges and corners of the slot's box from near and from far; starting inside the node.
 * Prints the looseness too (slots reported / slots truly met).  Exits 1 on the first false negative. */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "bvh_build.h"

using namespace crt;

/* ---- binary16 ------------------------------------------------------------------ */
static double round_half(double v)
{
	if (std::isnan(v))
		return v;
	const double a = std::fabs(v);
	if (a == 0.0)
		return v;
	if (std::isinf(a))
		return v;
	int e;
	std::frexp(a, &e); /* a = m * 2^e, m in [0.5, 1) */
	int ulp_exp = e - 11;          /* 11 significant bits */
	if (ulp_exp < -24)
		ulp_exp = -24;         /* subnormal spacing 2^-24 */
	const double ulp = std::ldexp(1.0, ulp_exp);
	const double q = a / ulp;      /* exact: a power-of-two scaling */
	double r = std::nearbyint(q);  /* default rounding mode: to nearest even */
	double out = r * ulp;
	if (out >= 65520.0)
		out = INFINITY;
	return std::copysign(out, v);
}
static double h_add(double a, double b) { return round_half(a + b); }
static double h_mul(double a, double b) { return round_half(a * b); }
static double h_fma(double a, double b, double c) { return round_half(a * b + c); } /* exact in double for the magnitudes here */
static double h_max(double a, double b) { return a > b ? a : b; }
static double h_min(double a, double b) { return a < b ? a : b; }
static double f2h(float f) { return round_half((double)f); }

/* ---- the device code, line by line ------------------------------------------------ */
struct EmuRay {
	float ox, oy, oz, dx, dy, dz;
	float inx, iny, inz, maxd, inv_maxd;
	double hx, hy, hz;
};

static float safe_idir(float d, float maxd)
{
	const float a = std::fabs(d) > 1e-18f ? d : std::copysign(1e-18f, d);
	return maxd / a;
}

static void make_wide_ray(EmuRay &w)
{
	w.maxd = std::fmax(std::fmax(std::fabs(w.dx), std::fabs(w.dy)), std::fabs(w.dz));
	w.inv_maxd = 1.f / w.maxd;
	w.inx = safe_idir(w.dx, w.maxd);
	w.iny = safe_idir(w.dy, w.maxd);
	w.inz = safe_idir(w.dz, w.maxd);
	const float big = 48.f, sc = 1.f / 2048.f;
	w.hx = f2h(std::fmin(std::fmax(w.inx * sc, -big), big));
	w.hy = f2h(std::fmin(std::fmax(w.iny * sc, -big), big));
	w.hz = f2h(std::fmin(std::fmax(w.inz * sc, -big), big));
}

static uint32_t emu_node_test(const WideNode &n, const EmuRay &w, float limit)
{
	const float grid = kWideGrid;
	const uint32_t meta = n.meta;
	const float inv_s = std::ldexp(1.f, 127 - (int)(meta & 0xffu));
	const float olx = (w.ox - n.origin[0]) * inv_s, oly = (w.oy - n.origin[1]) * inv_s, olz = (w.oz - n.origin[2]) * inv_s;
	const float ax = -olx * w.inx, ay = -oly * w.iny, az = -olz * w.inz;
	const float bx = std::fmaf(grid, w.inx, ax), by = std::fmaf(grid, w.iny, ay), bz = std::fmaf(grid, w.inz, az);
	float tin = std::fmax(std::fmax(std::fmin(ax, bx), std::fmin(ay, by)), std::fmax(std::fmin(az, bz), 0.f));
	const float tout = std::fmin(std::fmin(std::fmax(ax, bx), std::fmax(ay, by)), std::fmax(az, bz));
	tin = std::fmin(tin, 1e30f);
	const float tind = tin * w.inv_maxd;
	const float p[3] = { std::fmaf(tind, w.dx, olx), std::fmaf(tind, w.dy, oly), std::fmaf(tind, w.dz, olz) };
	const float lim = std::fmaf(limit * w.maxd, inv_s, -tin);
	const float mag = (std::fabs(olx) + std::fabs(oly)) + (std::fabs(olz) + tin);
	const double i[3] = { w.hx, w.hy, w.hz };
	double ph[3], na[3];
	for (int k = 0; k < 3; k++) {
		ph[k] = f2h(p[k]);            /* pack_half2(p, p) */
		na[k] = -std::fabs(i[k]);     /* h2 | 0x8000: -|i| */
	}
	const float lim_up = std::fmaf(std::fabs(lim), 0.00390625f, lim + 1.f) * (1.f / 2048.f);
	const double l2 = f2h(lim_up);
	uint32_t hit = 0;
	for (int j = 0; j < 8; j++) {
		double tn[3], tf[3];
		for (int k = 0; k < 3; k++) {
			const double c = half_value(n.plane[k][0][j]), h = half_value(n.plane[k][1][j]);
			const double q = h_mul(h_add(c, -ph[k]), i[k]);
			tn[k] = h_fma(h, na[k], q);
			tf[k] = h_fma(h, -na[k], q);
		}
		const double tmin = h_max(h_max(h_max(tn[0], tn[1]), tn[2]), 0.0);
		const double tmax = h_min(h_min(h_min(tf[0], tf[1]), tf[2]), l2);
		if (tmin <= tmax)
			hit |= 1u << j;
	}
	const uint32_t valid = ((meta >> 8) | (meta >> 16)) & 0xffu;
	if (!(mag <= 1048576.f))
		hit = valid;
	if (tin > std::fmaf(std::fabs(tout) + tin, 1e-6f, tout) || lim < -1.f - mag * 1e-6f)
		hit = 0;
	return hit & valid;
}

/* ---- exact reference: does the segment o + t d, t in [0, limit], meet the box? ----- */
static bool exact_hit(const Aabb &b, const EmuRay &r, double limit)
{
	double t0 = 0.0, t1 = limit;
	const double o[3] = { r.ox, r.oy, r.oz }, d[3] = { r.dx, r.dy, r.dz };
	for (int k = 0; k < 3; k++) {
		if (d[k] == 0.0) {
			if (o[k] < b.lo[k] || o[k] > b.hi[k])
				return false;
			continue;
		}
		double a = (b.lo[k] - o[k]) / d[k], c = (b.hi[k] - o[k]) / d[k];
		if (a > c)
			std::swap(a, c);
		t0 = std::max(t0, a);
		t1 = std::min(t1, c);
	}
	return t0 <= t1;
}

int main(int argc, char **argv)
{
	const uint32_t n = argc > 1 ? (uint32_t)atoi(argv[1]) : 20000;
	const int per_kind = argc > 2 ? atoi(argv[2]) : 20000;
	std::mt19937 rng(argc > 3 ? (uint32_t)atoi(argv[3]) : 1);
	std::uniform_real_distribution<float> pos(-10.f, 10.f), size(0.f, 0.05f), uni(0.f, 1.f);
	std::vector<Aabb> boxes(n);
	for (uint32_t i = 0; i < n; i++)
		for (int k = 0; k < 3; k++) {
			float c = pos(rng), s = size(rng);
			if (i % 5 == 0 && k == (int)(i % 3))
				s = 0.f; /* flat boxes */
			boxes[i].lo[k] = c - s;
			boxes[i].hi[k] = c + s;
		}
	BvhResult bvh;
	build_bvh(boxes, bvh);
	WideTree w;
	build_wide(boxes, bvh, w);
	/* what lies below every node */
	std::vector<Aabb> below(w.nodes.size());
	auto used_slots = [&](const WideNode &wn) { return (uint32_t)__builtin_popcount(((wn.meta >> 8) | (wn.meta >> 16)) & 0xffu); };
	auto slot_content = [&](size_t v, uint32_t j) {
		const WideNode &wn = w.nodes[v];
		return ((wn.meta >> 8) >> j) & 1u ? boxes[w.order[wn.prim_base + (int)j]] : below[wn.child_base + (int)j];
	};
	for (size_t v = w.nodes.size(); v-- > 0;) {
		Aabb m;
		for (int k = 0; k < 3; k++) {
			m.lo[k] = INFINITY;
			m.hi[k] = -INFINITY;
		}
		const uint32_t used = used_slots(w.nodes[v]);
		for (uint32_t j = 0; j < used; j++) {
			const Aabb c = slot_content(v, j);
			for (int k = 0; k < 3; k++) {
				m.lo[k] = std::min(m.lo[k], c.lo[k]);
				m.hi[k] = std::max(m.hi[k], c.hi[k]);
			}
		}
		below[v] = m;
	}
	unsigned long long tried = 0, true_hits = 0, reported = 0;
	const char *kinds[] = { "random", "axis-parallel", "faces / edges / corners, near", "the same from far away", "origin inside the node" };
	for (int kind = 0; kind < 5; kind++) {
		for (int it = 0; it < per_kind; it++) {
			const size_t v = rng() % w.nodes.size();
			const WideNode &wn = w.nodes[v];
			const uint32_t used = used_slots(wn);
			const uint32_t j = rng() % used;
			const Aabb target = slot_content(v, j);
			const Aabb &node_box = below[v];
			const float ext = std::max({ node_box.hi[0] - node_box.lo[0], node_box.hi[1] - node_box.lo[1], node_box.hi[2] - node_box.lo[2], 1e-6f });
			EmuRay r;
			float tgt[3], org[3];
			for (int k = 0; k < 3; k++) {
				const float u = uni(rng);
				tgt[k] = target.lo[k] + u * (target.hi[k] - target.lo[k]);
				if (kind == 2 || kind == 3) { /* on a face, an edge or a corner of the slot's box */
					const int pick = rng() % 3;
					if (pick == 0)
						tgt[k] = target.lo[k];
					if (pick == 1)
						tgt[k] = target.hi[k];
				}
			}
			const float far = kind == 3 ? ext * std::pow(10.f, 1.f + 4.f * uni(rng)) : ext * (0.5f + 4.f * uni(rng));
			float d[3] = { uni(rng) - 0.5f, uni(rng) - 0.5f, uni(rng) - 0.5f };
			if (kind == 1) {
				d[rng() % 3] = 0.f;
				if (rng() & 1u)
					d[rng() % 3] = 0.f;
				if (d[0] == 0.f && d[1] == 0.f && d[2] == 0.f)
					d[rng() % 3] = 1.f;
			}
			const float len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
			for (int k = 0; k < 3; k++) {
				d[k] /= len;
				org[k] = kind == 4 ? node_box.lo[k] + uni(rng) * (node_box.hi[k] - node_box.lo[k]) : tgt[k] - d[k] * far;
			}
			if (kind == 4) { /* aim at the target from inside */
				float e[3] = { tgt[0] - org[0], tgt[1] - org[1], tgt[2] - org[2] };
				const float l = std::sqrt(e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
				if (l > 0.f)
					for (int k = 0; k < 3; k++)
						d[k] = e[k] / l;
			}
			r.ox = org[0]; r.oy = org[1]; r.oz = org[2];
			r.dx = d[0]; r.dy = d[1]; r.dz = d[2];
			make_wide_ray(r);
			const float limit = (rng() & 3u) ? 1e30f : far * (0.5f + uni(rng)); /* some segments end near the target */
			const uint32_t got = emu_node_test(wn, r, limit);
			for (uint32_t s = 0; s < used; s++) {
				const bool truth = exact_hit(slot_content(v, s), r, (double)limit);
				tried++;
				true_hits += truth;
				reported += (got >> s) & 1u;
				if (truth && !((got >> s) & 1u)) {
					printf("FALSE NEGATIVE: kind '%s', node %zu slot %u, ray o (%.9g %.9g %.9g) d (%.9g %.9g %.9g) limit %g\n", kinds[kind], v, s,
						r.ox, r.oy, r.oz, r.dx, r.dy, r.dz, limit);
					return 1;
				}
			}
		}
	}
	printf("ok: %llu (ray, slot) pairs, %llu truly met, %llu reported (%.3f x), no false negative; %zu nodes over %u boxes\n", tried, true_hits,
		reported, (double)reported / (double)std::max(1ull, true_hits), w.nodes.size(), n);
	return 0;
}
