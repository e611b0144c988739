/*
 * bvh_build.h -- host-side BVH construction for the device traversal kernels.
 *
 * Replaces the reference's accel_init() (reference src/raytracer/accel.c:266-315):
 * that builder's Morton normalisation and comparator are flawed and its tree
 * is implementation-defined, while closest-hit and transmittance
 * results do not depend on the topology -- so this is an independent binned
 * SAH builder (binary tree, one primitive per leaf) whose tree is collapsed into
 * the 8-wide, 128-byte-per-node form both device walks read.
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

/* Binary tree (host only, the input of the 8-wide collapse), 64 bytes:
 *   xy0 = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
 *   xy1 = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
 *   z01 = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
 *   child[0], child[1] : >= 0 inner node index,
 *                        <  0 leaf: ~((first << 3) | (count - 1)), count == 1
 * An absent child (single-primitive scene) is an empty leaf with a far-away box. */
struct alignas(16) BvhNode {
	float xy0[4];
	float xy1[4];
	float z01[4];
	int32_t child[2];
	int32_t pad[2];
};

static_assert(sizeof(BvhNode) == 64, "BvhNode must be 64 bytes");

constexpr int kMaxLeafPrims = 1; /* a leaf of the binary tree is one primitive: the wide nodes box every primitive on its own */
constexpr int kLeafCountBits = 3;

struct BvhResult {
	std::vector<BvhNode> nodes;   /* nodes[0] is the root; empty if there are no primitives */
	std::vector<uint32_t> order;  /* order[slot] = index into the input boxes */
	uint32_t max_depth = 0;
};

/* boxes: one per bounded primitive (already padded by the caller). */
void build_bvh(const std::vector<Aabb> &boxes, BvhResult &out);

/* ---- the device tree: 8-wide nodes of 128 bytes ---------------------------------
 *
 * A node has up to eight children in SLOTS 0..7: first its leaf children (each ONE
 * primitive), then its inner children, then empty slots.
 *   primitive of leaf slot j  = prim_base  + j      (primitive slots of a node are contiguous)
 *   node of inner slot j      = child_base + j      (child nodes of a node are contiguous;
 *                                                    child_base = first child - number of leaf slots, may be negative)
 * so a traversal stack entry is (base << 8 | 8-bit mask of slots still to visit) and a
 * node visit pushes at most one entry per kind -- no per-child codes are read.
 *
 * Child boxes live on the node's own grid: origin + g * 2^(exponent - 127), g in
 * [0, 1023], the same power-of-two scale on the three axes.  Per axis and child the
 * node stores, as IEEE half floats, the box CENTRE c = (lo_q + hi_q) / 2 and the
 * PADDED HALF EXTENT h = (hi_q - lo_q) / 2 + kWideSlack (lo_q rounded down, hi_q up;
 * all multiples of 1/2 below 1024, exact in half precision).  With centre and half
 * extent the slab test needs no near / far selection by the sign of the ray
 * direction: t = (c - p) * idir -+ h * |idir|.  The device evaluates that in
 * half2 arithmetic, two children per instruction, on a ray re-based to its entry
 * point p into the node's grid cube; kWideSlack covers every rounding of that
 * evaluation (crt_device.cuh, wide_node_test).  An empty slot has h = -infinity.
 *
 * Memory layout = four 32-byte vectors (one 256-bit load each):
 *   [0] origin.xyz, meta, child_base, prim_base, parent, -
 *   [1] cx[8] hx[8]    [2] cy[8] hy[8]    [3] cz[8] hz[8]     (half bits)
 * meta: bits 0-7 biased exponent of the grid scale, 8-15 mask of leaf slots,
 *       16-23 mask of inner slots, 24-25 axis the slots of each kind are sorted along. */
constexpr float kWideGrid = 1023.f;
constexpr float kWideSlack = 3.5f;
constexpr uint32_t kWideMaxDepth = 64;       /* deeper trees are refused (device stacks) */
constexpr uint32_t kWideMaxNodes = 1u << 23; /* stack entries hold a signed 24-bit base */
constexpr uint32_t kWideMaxPrims = 1u << 24;

struct alignas(32) WideNode {
	float origin[3];
	uint32_t meta;
	int32_t child_base;
	int32_t prim_base;
	int32_t parent; /* wide node index of the parent, -1 for the root (diagnostics) */
	uint32_t pad;
	uint16_t plane[3][2][8]; /* [axis][0 = centre, 1 = padded half extent][slot], IEEE half bits */
};

static_assert(sizeof(WideNode) == 128, "WideNode must be 128 bytes");

struct WideTree {
	std::vector<WideNode> nodes;  /* nodes[0] is the root; empty if there are no primitives */
	std::vector<uint32_t> order;  /* order[primitive slot] = index into the input boxes */
	uint32_t max_depth = 0;       /* nodes on the longest root-to-leaf path */
	double cost = 0.0;            /* modelled cost per unit of root surface: sum of node surfaces + prim_cost * leaf surfaces */
};

/* Collapses the binary tree into the device tree.  Which binary nodes become wide nodes is
 * chosen by dynamic programming so that the summed surface of the wide nodes (the expected
 * number of node visits of a random ray) plus prim_cost times the summed surface of the
 * primitive boxes is minimal. */
void build_wide(const std::vector<Aabb> &boxes, const BvhResult &bvh, WideTree &out);

/* half <-> float on the host (round to nearest even; the builder only stores exactly representable values) */
uint16_t half_bits(float f);
float half_value(uint16_t h);

} // namespace crt

#endif
