/* bvh_stats -- offline statistics of the two trees built from a binary STL
 * (fill of the 8-wide nodes, surface-area estimate of visits per random ray).
 *   g++ -O2 -std=c++17 -Ic-raytracer_b200/csrc tools/bvh_stats.cpp c-raytracer_b200/csrc/bvh_build.cpp -o build/bvh_stats
 *   build/bvh_stats meshes/dragon.stl */
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "bvh_build.h"

using namespace crt;

static float half_area(const float *lo, const float *hi)
{
	const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
	return x * y + y * z + z * x;
}

int main(int argc, char **argv)
{
	if (argc < 2)
		return 1;
	FILE *f = fopen(argv[1], "rb");
	if (!f)
		return 1;
	unsigned char head[84];
	if (fread(head, 1, 84, f) != 84)
		return 1;
	uint32_t n;
	memcpy(&n, head + 80, 4);
	std::vector<Aabb> boxes(n);
	for (uint32_t i = 0; i < n; i++) {
		unsigned char rec[50];
		if (fread(rec, 1, 50, f) != 50)
			return 1;
		float v[12];
		memcpy(v, rec, 48);
		for (int k = 0; k < 3; k++) {
			boxes[i].lo[k] = std::min({ v[3 + k], v[6 + k], v[9 + k] });
			boxes[i].hi[k] = std::max({ v[3 + k], v[6 + k], v[9 + k] });
		}
	}
	fclose(f);
	BvhResult bvh;
	build_bvh(boxes, bvh);
	WideTree wide;
	build_wide(boxes, bvh, wide);
	/* per node: its own box is the parent's stored box of it; the surface-area estimate uses the stored
	 * (quantised, padded) boxes, i.e. what the device walk actually tests */
	size_t hist[9] = { 0 }, leaf_children = 0, inner_children = 0;
	double visits = 0, filled = 0, prim_tests = 0;
	std::vector<double> area(wide.nodes.size(), 0.0);
	auto stored_area = [&](const WideNode &w, int j) {
		double ext[3];
		const double scale = std::ldexp(1.0, (int)(w.meta & 0xffu) - 127);
		for (int k = 0; k < 3; k++)
			ext[k] = 2.0 * half_value(w.plane[k][1][j]) * scale;
		return ext[0] * ext[1] + ext[1] * ext[2] + ext[2] * ext[0];
	};
	{
		float lo[3] = { 1e30f, 1e30f, 1e30f }, hi[3] = { -1e30f, -1e30f, -1e30f };
		for (const Aabb &b : boxes)
			for (int k = 0; k < 3; k++) {
				lo[k] = std::min(lo[k], b.lo[k]);
				hi[k] = std::max(hi[k], b.hi[k]);
			}
		area[0] = half_area(lo, hi);
	}
	for (size_t i = 0; i < wide.nodes.size(); i++) {
		const WideNode &w = wide.nodes[i];
		const uint32_t lmask = (w.meta >> 8) & 0xffu, imask = (w.meta >> 16) & 0xffu;
		hist[__builtin_popcount(lmask | imask)]++;
		const double p = area[i] / area[0]; /* chance that a random line through the root box meets this node */
		visits += p;
		for (int j = 0; j < 8; j++) {
			if (!(((lmask | imask) >> j) & 1u))
				continue;
			filled += p;
			const double a = stored_area(w, j);
			if ((lmask >> j) & 1u) {
				leaf_children++;
				prim_tests += a / area[0];
			} else {
				inner_children++;
				area[w.child_base + j] = a; /* children come after their parent */
			}
		}
	}
	printf("prims %u  bvh2 nodes %zu depth %u | wide nodes %zu depth %u, modelled cost %.2f\n", n, bvh.nodes.size(), bvh.max_depth,
		wide.nodes.size(), wide.max_depth, wide.cost);
	printf("children per wide node:");
	for (int k = 0; k <= 8; k++)
		printf(" %d:%zu", k, hist[k]);
	printf("\nmean fill %.2f  leaf slots %zu  inner slots %zu\n",
		(double)(leaf_children + inner_children) / wide.nodes.size(), leaf_children, inner_children);
	printf("surface-area estimate per random line (stored boxes): node visits %.1f, filled slots %.1f, primitive tests %.1f\n", visits, filled, prim_tests);
	return 0;
}
