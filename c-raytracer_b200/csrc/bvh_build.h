/*
 * bvh_build.h -- host-side BVH construction for the device traversal kernels.
 *
 * Replaces the reference's accel_init() (reference src/raytracer/accel.c:266-315):
 * that builder's Morton normalisation and comparator are flawed and its tree
 * is implementation-defined, while closest-hit and transmittance
 * results do not depend on the topology -- so this is an independent binned
 * SAH builder producing the 64-byte two-child-boxes-per-node layout the
 * kernels read with four float4 loads.
 *
 * Every primitive lives in exactly ONE leaf (no spatial splits): the shadow
 * query multiplies the light by kt once per transparent primitive hit
 * (accel.c:369-373), so a duplicated reference would tint twice.
 */
#ifndef CRT_BVH_BUILD_H
#define CRT_BVH_BUILD_H

#include <cstdint>
#include <vector>

namespace crt {

struct Aabb {
	float lo[3], hi[3];
};

/* 64 bytes, read as float4 x 4 on the device:
 *   xy0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
 *   xy1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
 *   z01 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
 *   child[0], child[1] : >= 0 inner node index,
 *                        <  0 leaf: ~((first << 3) | (count - 1)), count in 1..8
 *   pad[2]
 * An absent child (single-primitive scene) is an empty leaf with an inverted box. */
struct alignas(16) BvhNode {
	float xy0[4];
	float xy1[4];
	float z01[4];
	int32_t child[2];
	int32_t pad[2];
};

static_assert(sizeof(BvhNode) == 64, "BvhNode must be 64 bytes");

constexpr int kMaxLeafPrims = 4;
constexpr int kLeafCountBits = 3;

struct BvhResult {
	std::vector<BvhNode> nodes;   /* nodes[0] is the root; empty if there are no primitives */
	std::vector<uint32_t> order;  /* order[slot] = index into the input boxes */
	uint32_t max_depth = 0;
};

/* boxes: one per bounded primitive (already padded by the caller). */
void build_bvh(const std::vector<Aabb> &boxes, BvhResult &out);

} // namespace crt

#endif
