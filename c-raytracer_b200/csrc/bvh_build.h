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

/* 8-wide tree of the shadow (any-hit) walk, collapsed from the binary tree (same
 * primitive order).  Intermediate, full-precision form: child j has
 * a[j] = (lo.x, lo.y, lo.z, hi.x), b[j] = (hi.y, hi.z, code, 0).  code (int
 * bits): >= 0 wide node index; < 0 leaf ~((first << 3) | (count - 1)), a
 * contiguous primitive range; kWideEmpty: no child. */
struct alignas(16) WideNode {
	float a[8][4];
	float b[8][4];
};

static_assert(sizeof(WideNode) == 256, "WideNode must be 256 bytes");

constexpr int32_t kWideEmpty = 0x7fffffff;
constexpr uint32_t kWideLeafPrims = 8;

struct WideResult {
	std::vector<WideNode> nodes; /* nodes[0] is the root; empty if there are no primitives */
	uint32_t max_depth = 0;
};

/* Collapses the binary tree into the 8-wide tree (same primitive order). */
/* merge_prims: binary subtrees with at most this many primitives (<= 8) become one leaf;
 * 0 keeps exactly the binary tree's leaves. */
void build_wide(const BvhResult &bvh, WideResult &out, uint32_t merge_prims);

/* The device form, 96 bytes = 6 x float4 (two visits per 3 sectors of a line):
 *   origin.xyz, exponents (one IEEE exponent byte per axis: scale = 2^(e-127))
 *   q[axis][0 = lo, 1 = hi][child] : 8-bit planes, plane = origin + q * scale,
 *                                    rounded outwards; empty child: lo 255, hi 0
 *   code[child]
 * One visit tests eight boxes from six vector loads (the binary node: two boxes
 * from four), and the tree is a third as deep: fewer dependent load rounds and
 * fewer L1 wavefronts per ray. */
struct alignas(16) QuantNode {
	float origin[3];
	uint32_t exponents;
	uint8_t q[3][2][8];
	int32_t code[8];
};

static_assert(sizeof(QuantNode) == 96, "QuantNode must be 96 bytes");

void quantize_wide(const WideResult &wide, std::vector<QuantNode> &out);


} // namespace crt


#endif
