/* bvh_check -- invariants of the host tree builders (c-raytracer_b200/csrc/bvh_build.cpp),
 * run by tests/test_bvh_build.py on seeded boxes:
 *   bvh_check <n> <seed> <kind>      kind 0: uniform cloud, 1: flat axis-aligned sheets, 2: many coincident boxes
 * Exits 0 and prints a summary line, or prints the violated invariant and exits 1. */
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "bvh_build.h"

using namespace crt;

#define CHECK(cond, ...)                      \
	do {                                  \
		if (!(cond)) {                \
			printf(__VA_ARGS__);  \
			printf("\n");         \
			return 1;             \
		}                             \
	} while (0)

static bool contains(const float *lo, const float *hi, const Aabb &b)
{
	for (int k = 0; k < 3; k++)
		if (b.lo[k] < lo[k] || b.hi[k] > hi[k])
			return false;
	return true;
}

static float plane(float origin, uint32_t exponent, uint8_t q)
{
	const uint32_t bits = exponent << 23;
	float scale;
	memcpy(&scale, &bits, 4);
	return origin + (float)q * scale; /* the device evaluates fl(q * scale + origin) too */
}

int main(int argc, char **argv)
{
	if (argc < 4)
		return 2;
	const uint32_t n = (uint32_t)atoi(argv[1]);
	std::mt19937 rng((uint32_t)atoi(argv[2]));
	const int kind = atoi(argv[3]);
	std::uniform_real_distribution<float> pos(-10.f, 10.f), size(0.f, 0.05f);
	std::vector<Aabb> boxes(n);
	for (uint32_t i = 0; i < n; i++) {
		for (int k = 0; k < 3; k++) {
			float c = pos(rng), s = size(rng);
			if (kind == 1 && k == (int)(i % 3))
				s = 0.f, c = std::floor(c); /* zero thickness, on a lattice */
			if (kind == 2 && i % 4 != 0)
				c = 1.5f, s = 0.01f; /* three quarters of the boxes coincide */
			boxes[i].lo[k] = c - s;
			boxes[i].hi[k] = c + s;
		}
	}

	setenv("CRT_BVH_THREADS", "1", 1);
	BvhResult a;
	build_bvh(boxes, a);
	for (const char *threads : { "2", "5", "16" }) {
		setenv("CRT_BVH_THREADS", threads, 1);
		BvhResult b;
		build_bvh(boxes, b);
		CHECK(a.nodes.size() == b.nodes.size() && a.order == b.order && a.max_depth == b.max_depth &&
				(a.nodes.empty() || !memcmp(a.nodes.data(), b.nodes.data(), a.nodes.size() * sizeof(BvhNode))),
			"tree built with %s threads differs from the single-threaded tree", threads);
	}
	if (n == 0) {
		CHECK(a.nodes.empty() && a.order.empty(), "empty input must give an empty tree");
		printf("ok n 0\n");
		return 0;
	}

	/* order is a permutation */
	std::vector<uint32_t> seen(n, 0);
	CHECK(a.order.size() == n, "order has %zu entries for %u boxes", a.order.size(), n);
	for (uint32_t p : a.order) {
		CHECK(p < n, "order entry %u out of range", p);
		seen[p]++;
	}
	for (uint32_t i = 0; i < n; i++)
		CHECK(seen[i] == 1, "box %u appears %u times in order", i, seen[i]);

	/* binary tree: every slot in exactly one leaf, child boxes contain their primitives, depth bound */
	std::vector<uint32_t> covered(n, 0);
	struct Item { int32_t node; uint32_t depth; };
	std::vector<Item> stack = { { 0, 1 } };
	uint32_t depth_seen = 0, leaves = 0;
	while (!stack.empty()) {
		const Item it = stack.back();
		stack.pop_back();
		CHECK(it.node >= 0 && (size_t)it.node < a.nodes.size(), "node index %d out of range", it.node);
		const BvhNode &nd = a.nodes[it.node];
		depth_seen = std::max(depth_seen, it.depth);
		for (int c = 0; c < 2; c++) {
			const float *xy = c == 0 ? nd.xy0 : nd.xy1;
			const float lo[3] = { xy[0], xy[2], nd.z01[2 * c] }, hi[3] = { xy[1], xy[3], nd.z01[2 * c + 1] };
			if (nd.child[c] >= 0) {
				stack.push_back({ nd.child[c], it.depth + 1 });
				continue;
			}
			if (lo[0] >= 1e30f)
				continue; /* the absent child of a one-leaf scene */
			const uint32_t leaf = (uint32_t)~nd.child[c];
			const uint32_t first = leaf >> kLeafCountBits, count = (leaf & ((1u << kLeafCountBits) - 1u)) + 1u;
			CHECK(count <= (uint32_t)kMaxLeafPrims && first + count <= n, "leaf %u+%u malformed", first, count);
			leaves++;
			for (uint32_t s = first; s < first + count; s++) {
				covered[s]++;
				CHECK(contains(lo, hi, boxes[a.order[s]]), "leaf box of node %d does not contain primitive slot %u", it.node, s);
			}
		}
	}
	for (uint32_t s = 0; s < n; s++)
		CHECK(covered[s] == 1, "primitive slot %u is in %u leaves (the shadow query tints once per primitive)", s, covered[s]);
	CHECK(depth_seen <= a.max_depth && a.max_depth <= 128, "depth %u (reported %u) exceeds the device stack", depth_seen, a.max_depth);

	/* 8-wide collapse and quantisation: same coverage, quantised boxes contain the exact ones */
	WideResult w;
	build_wide(a, w, 0);
	std::vector<QuantNode> q;
	quantize_wide(w, q);
	CHECK(q.size() == w.nodes.size() && !w.nodes.empty(), "no wide nodes");
	std::fill(covered.begin(), covered.end(), 0u);
	std::vector<int32_t> wstack = { 0 };
	size_t visited = 0;
	while (!wstack.empty()) {
		const int32_t id = wstack.back();
		wstack.pop_back();
		CHECK(id >= 0 && (size_t)id < w.nodes.size(), "wide node index %d out of range", id);
		visited++;
		const WideNode &wn = w.nodes[id];
		const QuantNode &qn = q[id];
		for (int j = 0; j < 8; j++) {
			int32_t code;
			memcpy(&code, &wn.b[j][2], 4);
			CHECK(code == qn.code[j], "codes of wide node %d differ", id);
			if (code == kWideEmpty) {
				for (int k = 0; k < 3; k++)
					CHECK(qn.q[k][0][j] == 255 && qn.q[k][1][j] == 0, "empty child %d of node %d is not inverted", j, id);
				continue;
			}
			Aabb exact;
			exact.lo[0] = wn.a[j][0]; exact.lo[1] = wn.a[j][1]; exact.lo[2] = wn.a[j][2];
			exact.hi[0] = wn.a[j][3]; exact.hi[1] = wn.b[j][0]; exact.hi[2] = wn.b[j][1];
			float lo[3], hi[3];
			for (int k = 0; k < 3; k++) {
				const uint32_t e = (qn.exponents >> (8 * k)) & 0xffu;
				lo[k] = plane(qn.origin[k], e, qn.q[k][0][j]);
				hi[k] = plane(qn.origin[k], e, qn.q[k][1][j]);
			}
			CHECK(contains(lo, hi, exact), "quantised box of child %d of wide node %d is not conservative", j, id);
			if (code >= 0) {
				wstack.push_back(code);
				continue;
			}
			const uint32_t leaf = (uint32_t)~code;
			const uint32_t first = leaf >> kLeafCountBits, count = (leaf & ((1u << kLeafCountBits) - 1u)) + 1u;
			for (uint32_t s = first; s < first + count; s++) {
				CHECK(s < n, "wide leaf slot %u out of range", s);
				covered[s]++;
				CHECK(contains(exact.lo, exact.hi, boxes[a.order[s]]), "wide leaf box does not contain slot %u", s);
			}
		}
	}
	CHECK(visited == w.nodes.size(), "%zu of %zu wide nodes reachable", visited, w.nodes.size());
	for (uint32_t s = 0; s < n; s++)
		CHECK(covered[s] == 1, "primitive slot %u is in %u wide leaves", s, covered[s]);
	printf("ok n %u nodes %zu leaves %u depth %u wide %zu depth %u\n", n, a.nodes.size(), leaves, a.max_depth, w.nodes.size(), w.max_depth);
	return 0;
}
