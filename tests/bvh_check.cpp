/* bvh_check -- invariants of the host tree builders (c-raytracer_b200/csrc/bvh_build.cpp),
 * run by tests/test_bvh_build.py on seeded boxes:
 *   bvh_check <n> <seed> <kind>      kind 0: uniform cloud, 1: flat axis-aligned sheets, 2: many coincident boxes,
 *                                    3: exponential chain (deep tree), 4: the same up to 10^30 (surfaces overflow FP32)
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
		if (kind == 3 || kind == 4) {
			/* exponential chain: SAH peels the largest box off at every level, the wide tree gets deeper than the
			 * walk's shared-memory stack (kind 3, coordinates <= 10^6); kind 4 grows until the FP32 surfaces overflow */
			const double c = 1e-6 * std::pow(kind == 3 ? 1.2 : 2.0, (double)(i % (kind == 3 ? 150u : 120u))), r = 0.2 * c;
			for (int k = 0; k < 3; k++) {
				boxes[i].lo[k] = (float)(c - r);
				boxes[i].hi[k] = (float)(c + r);
			}
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

	/* 8-wide device tree: every primitive in exactly one leaf slot, slots of a kind contiguous (leaves,
	 * then inner children), child nodes and primitives of a node contiguous, and the stored boxes
	 * -- centre +- (half extent - slack) on the node's grid -- contain the exact ones */
	WideTree w;
	build_wide(boxes, a, w);
	CHECK(!w.nodes.empty() && w.order.size() == n, "no wide nodes / order has %zu entries", w.order.size());
	std::fill(seen.begin(), seen.end(), 0u);
	for (uint32_t p : w.order) {
		CHECK(p < n, "wide order entry %u out of range", p);
		seen[p]++;
	}
	for (uint32_t i = 0; i < n; i++)
		CHECK(seen[i] == 1, "box %u appears %u times in the wide order", i, seen[i]);
	std::fill(covered.begin(), covered.end(), 0u);
	struct WItem { int32_t node; uint32_t depth; Aabb bound; bool bounded; };
	std::vector<WItem> wstack;
	wstack.push_back({ 0, 1, Aabb(), false });
	size_t visited = 0;
	uint32_t wdepth = 0;
	while (!wstack.empty()) {
		const WItem it = wstack.back();
		wstack.pop_back();
		CHECK(it.node >= 0 && (size_t)it.node < w.nodes.size(), "wide node index %d out of range", it.node);
		visited++;
		wdepth = std::max(wdepth, it.depth);
		const WideNode &wn = w.nodes[it.node];
		const uint32_t lmask = (wn.meta >> 8) & 0xffu, imask = (wn.meta >> 16) & 0xffu;
		const int nl = __builtin_popcount(lmask), ni = __builtin_popcount(imask);
		CHECK(lmask == (1u << nl) - 1u && imask == (((1u << ni) - 1u) << nl) && nl + ni >= 1 && nl + ni <= 8,
			"slots of wide node %d are not leaves-then-inner (masks %02x %02x)", it.node, lmask, imask);
		const int e = (int)(wn.meta & 0xffu) - 127;
		CHECK(e >= -40 && e <= (kind == 4 ? 127 : 60), "grid exponent %d of node %d out of range", e, it.node);
		const double scale = std::ldexp(1.0, e);
		for (int j = 0; j < 8; j++) {
			const bool used = ((lmask | imask) >> j) & 1u;
			Aabb q;
			for (int k = 0; k < 3; k++) {
				const float c = half_value(wn.plane[k][0][j]), h = half_value(wn.plane[k][1][j]);
				if (!used) {
					CHECK(std::isinf(h) && h < 0.f, "empty slot %d of node %d is not inverted", j, it.node);
					continue;
				}
				CHECK(c >= 0.f && c <= kWideGrid && h >= kWideSlack && std::floor(2.f * c) == 2.f * c && std::floor(2.f * h) == 2.f * h,
					"slot %d of node %d: centre %g / half extent %g off the half-unit grid", j, it.node, c, h);
				CHECK(c - (h - kWideSlack) >= 0.f && c + (h - kWideSlack) <= kWideGrid, "slot %d of node %d leaves the grid cube", j, it.node);
				q.lo[k] = (float)((double)wn.origin[k] + ((double)c - ((double)h - kWideSlack)) * scale);
				q.hi[k] = (float)((double)wn.origin[k] + ((double)c + ((double)h - kWideSlack)) * scale);
				/* the float conversion above may round inwards by an ulp: compare in double */
				const double lo = (double)wn.origin[k] + ((double)c - ((double)h - kWideSlack)) * scale;
				const double hi = (double)wn.origin[k] + ((double)c + ((double)h - kWideSlack)) * scale;
				q.lo[k] = (float)lo > lo ? std::nextafter((float)lo, -INFINITY) : (float)lo;
				q.hi[k] = (float)hi < hi ? std::nextafter((float)hi, INFINITY) : (float)hi;
			}
			if (!used)
				continue;
			if (it.bounded)
				for (int k = 0; k < 3; k++)
					CHECK(q.lo[k] >= it.bound.lo[k] - 8.0 * scale && q.hi[k] <= it.bound.hi[k] + 8.0 * scale,
						"child box %d of node %d sticks out of the node's own box", j, it.node);
			if ((imask >> j) & 1u) {
				const int32_t child = wn.child_base + j;
				CHECK(child > it.node && (size_t)child < w.nodes.size() && w.nodes[child].parent == it.node,
					"inner slot %d of node %d -> %d: bad child index or parent link", j, it.node, child);
				wstack.push_back({ child, it.depth + 1, q, true });
				continue;
			}
			const int32_t s = wn.prim_base + j;
			CHECK(s >= 0 && (uint32_t)s < n, "leaf slot %d of node %d -> primitive slot %d out of range", j, it.node, s);
			covered[s]++;
			CHECK(contains(q.lo, q.hi, boxes[w.order[s]]), "stored box of leaf slot %d of node %d does not contain primitive slot %d", j, it.node, s);
		}
		/* an inner child's stored box must contain everything below it: checked through `bound` one level down,
		 * and transitively for the primitives (each node's leaf boxes contain their primitives) */
	}
	CHECK(visited == w.nodes.size(), "%zu of %zu wide nodes reachable", visited, w.nodes.size());
	for (uint32_t s = 0; s < n; s++)
		CHECK(covered[s] == 1, "primitive slot %u is in %u wide leaf slots", s, covered[s]);
	CHECK(wdepth == w.max_depth && w.max_depth <= kWideMaxDepth, "wide depth %u (reported %u) exceeds the device stacks", wdepth, w.max_depth);
	printf("ok n %u nodes %zu leaves %u depth %u wide %zu depth %u cost %.2f\n", n, a.nodes.size(), leaves, a.max_depth, w.nodes.size(), w.max_depth, w.cost);
	return 0;
}
