/* bvh_stats -- offline statistics of the two trees built from a binary STL
 * (fill of the 8-wide nodes, surface-area estimate of visits per random ray).
 *   g++ -O2 -std=c++17 -Ic-raytracer_b200/csrc tools/bvh_stats.cpp c-raytracer_b200/csrc/bvh_build.cpp -o build/bvh_stats
 *   build/bvh_stats meshes/dragon.stl [merge_prims] */
#include <algorithm>
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
	WideResult wide;
	build_wide(bvh, wide, argc > 2 ? atoi(argv[2]) : 0);
	size_t hist[9] = { 0 }, leaf_children = 0, inner_children = 0, leaf_prims = 0;
	double visits = 0, slots = 0, filled = 0, prim_tests = 0;
	float root_area = 0;
	std::vector<float> area(wide.nodes.size(), 0.f);
	/* node area = area of the union of its children; the root first */
	for (size_t i = 0; i < wide.nodes.size(); i++) {
		const WideNode &w = wide.nodes[i];
		float lo[3] = { 1e30f, 1e30f, 1e30f }, hi[3] = { -1e30f, -1e30f, -1e30f };
		int kids = 0;
		for (int j = 0; j < 8; j++) {
			int32_t code;
			memcpy(&code, &w.b[j][2], 4);
			if (code == kWideEmpty)
				continue;
			kids++;
			const float clo[3] = { w.a[j][0], w.a[j][1], w.a[j][2] }, chi[3] = { w.a[j][3], w.b[j][0], w.b[j][1] };
			for (int k = 0; k < 3; k++) {
				lo[k] = std::min(lo[k], clo[k]);
				hi[k] = std::max(hi[k], chi[k]);
			}
		}
		hist[kids]++;
		area[i] = half_area(lo, hi);
		if (i == 0)
			root_area = area[0];
	}
	for (size_t i = 0; i < wide.nodes.size(); i++) {
		const WideNode &w = wide.nodes[i];
		const double p = area[i] / root_area; /* chance that a random line through the root box meets this node */
		visits += p;
		slots += 8 * p;
		for (int j = 0; j < 8; j++) {
			int32_t code;
			memcpy(&code, &w.b[j][2], 4);
			if (code == kWideEmpty)
				continue;
			filled += p;
			if (code < 0) {
				const uint32_t cnt = ((~code) & 7) + 1;
				leaf_children++;
				leaf_prims += cnt;
				const float clo[3] = { w.a[j][0], w.a[j][1], w.a[j][2] }, chi[3] = { w.a[j][3], w.b[j][0], w.b[j][1] };
				prim_tests += cnt * half_area(clo, chi) / root_area;
			} else {
				inner_children++;
			}
		}
	}
	printf("prims %u  bvh2 nodes %zu depth %u | wide nodes %zu depth %u\n", n, bvh.nodes.size(), bvh.max_depth, wide.nodes.size(), wide.max_depth);
	printf("children per wide node:");
	for (int k = 0; k <= 8; k++)
		printf(" %d:%zu", k, hist[k]);
	printf("\nmean fill %.2f  leaf children %zu (%.2f prims each)  inner children %zu\n",
		(double)(leaf_children + inner_children) / wide.nodes.size(), leaf_children, (double)leaf_prims / leaf_children, inner_children);
	printf("surface-area estimate per random line: node visits %.1f, filled slots %.1f (of %.1f), prim tests %.1f\n", visits, filled, slots, prim_tests);
	return 0;
}
