/*
 * bvh_build.cpp -- binned-SAH BVH builder (host).  See bvh_build.h.
 */
#include "bvh_build.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace crt {

namespace {

constexpr int kBins = 16;
constexpr float kTraversalCost = 1.2f; /* relative to one primitive test */
constexpr uint32_t kMedianDepth = 40;  /* below this depth: median splits only, bounds the device stack */

/* A subtree set aside while the top of the tree is built: it is built later (by
 * a worker thread) into its own node vector and spliced in. */
struct Deferred {
	uint32_t first, count, depth;
	std::vector<BvhNode> nodes; /* local pre-order indices */
	int32_t root = 0;           /* local node index, or encoded leaf */
	Aabb bounds;
	uint32_t max_depth = 0;
};

constexpr int32_t kDeferredBase = 0x40000000; /* child references >= this are indices into Builder::deferred */

struct Builder {
	const std::vector<Aabb> &boxes;
	const std::vector<float> &centroid; /* 3 per box */
	std::vector<uint32_t> &order;
	std::vector<BvhNode> &nodes;
	uint32_t max_depth = 0;
	/* ranges of at most this many primitives are deferred (0: build everything here) */
	uint32_t defer_below = 0;
	std::vector<Deferred> *deferred = nullptr;

	Builder(const std::vector<Aabb> &b, const std::vector<float> &c, std::vector<uint32_t> &o, std::vector<BvhNode> &n)
		: boxes(b), centroid(c), order(o), nodes(n) {}
};

inline void box_reset(Aabb &b)
{
	for (int k = 0; k < 3; k++) {
		b.lo[k] = FLT_MAX;
		b.hi[k] = -FLT_MAX;
	}
}

inline void box_grow(Aabb &b, const Aabb &o)
{
	for (int k = 0; k < 3; k++) {
		b.lo[k] = std::min(b.lo[k], o.lo[k]);
		b.hi[k] = std::max(b.hi[k], o.hi[k]);
	}
}

inline float half_area(const Aabb &b)
{
	const float dx = b.hi[0] - b.lo[0], dy = b.hi[1] - b.lo[1], dz = b.hi[2] - b.lo[2];
	if (dx < 0.f || dy < 0.f || dz < 0.f)
		return 0.f;
	return dx * dy + dy * dz + dz * dx;
}

inline int32_t encode_leaf(uint32_t first, uint32_t count)
{
	return ~static_cast<int32_t>((first << kLeafCountBits) | (count - 1u));
}

void store_child(BvhNode &n, int which, int32_t ref, const Aabb &b)
{
	float *xy = which == 0 ? n.xy0 : n.xy1;
	xy[0] = b.lo[0];
	xy[1] = b.hi[0];
	xy[2] = b.lo[1];
	xy[3] = b.hi[1];
	n.z01[2 * which + 0] = b.lo[2];
	n.z01[2 * which + 1] = b.hi[2];
	n.child[which] = ref;
}

/* Builds the subtree over order[first, first+count); returns the child
 * reference (node index or encoded leaf) and its bounds. */
int32_t build_range(Builder &B, uint32_t first, uint32_t count, uint32_t depth, Aabb &bounds)
{
	B.max_depth = std::max(B.max_depth, depth);
	box_reset(bounds);
	Aabb cbox;
	box_reset(cbox);
	for (uint32_t i = first; i < first + count; i++) {
		const uint32_t p = B.order[i];
		box_grow(bounds, B.boxes[p]);
		for (int k = 0; k < 3; k++) {
			cbox.lo[k] = std::min(cbox.lo[k], B.centroid[3 * p + k]);
			cbox.hi[k] = std::max(cbox.hi[k], B.centroid[3 * p + k]);
		}
	}
	if (count == 1)
		return encode_leaf(first, 1);
	if (B.deferred && count <= B.defer_below) {
		B.deferred->emplace_back();
		Deferred &d = B.deferred->back();
		d.first = first;
		d.count = count;
		d.depth = depth;
		return kDeferredBase + static_cast<int32_t>(B.deferred->size() - 1);
	}

	/* --- choose a split ------------------------------------------------ */
	int best_axis = -1, best_bin = -1;
	float best_cost = FLT_MAX;
	const float parent_area = half_area(bounds);
	if (depth < kMedianDepth && parent_area > 0.f) {
		/* one pass over the (permuted, cache-unfriendly) range fills the bins of all three axes */
		Aabb bin_box[3][kBins];
		uint32_t bin_count[3][kBins];
		float scale[3];
		bool usable[3];
		for (int axis = 0; axis < 3; axis++) {
			const float extent = cbox.hi[axis] - cbox.lo[axis];
			usable[axis] = extent > 0.f;
			scale[axis] = usable[axis] ? kBins / extent : 0.f;
			for (int b = 0; b < kBins; b++) {
				box_reset(bin_box[axis][b]);
				bin_count[axis][b] = 0;
			}
		}
		for (uint32_t i = first; i < first + count; i++) {
			const uint32_t p = B.order[i];
			const Aabb &pb = B.boxes[p];
			for (int axis = 0; axis < 3; axis++) {
				if (!usable[axis])
					continue;
				int b = static_cast<int>((B.centroid[3 * p + axis] - cbox.lo[axis]) * scale[axis]);
				b = std::min(std::max(b, 0), kBins - 1);
				bin_count[axis][b]++;
				box_grow(bin_box[axis][b], pb);
			}
		}
		for (int axis = 0; axis < 3; axis++) {
			if (!usable[axis])
				continue;
			float right_area[kBins];
			uint32_t right_count[kBins];
			Aabb acc;
			box_reset(acc);
			uint32_t n = 0;
			for (int b = kBins - 1; b > 0; b--) {
				box_grow(acc, bin_box[axis][b]);
				n += bin_count[axis][b];
				right_area[b] = half_area(acc);
				right_count[b] = n;
			}
			box_reset(acc);
			n = 0;
			for (int b = 0; b < kBins - 1; b++) {
				box_grow(acc, bin_box[axis][b]);
				n += bin_count[axis][b];
				if (n == 0 || right_count[b + 1] == 0)
					continue;
				const float cost = half_area(acc) * n + right_area[b + 1] * right_count[b + 1];
				if (cost < best_cost) {
					best_cost = cost;
					best_axis = axis;
					best_bin = b;
				}
			}
		}
	}

	if (best_axis >= 0 && count <= static_cast<uint32_t>(kMaxLeafPrims)) {
		const float split_cost = kTraversalCost + best_cost / parent_area;
		if (split_cost >= static_cast<float>(count))
			return encode_leaf(first, count);
	}

	uint32_t mid;
	if (best_axis >= 0) {
		const float lo = cbox.lo[best_axis];
		const float scale = kBins / (cbox.hi[best_axis] - lo);
		uint32_t *begin = B.order.data() + first;
		uint32_t *split = std::partition(begin, begin + count, [&](uint32_t p) {
			int b = static_cast<int>((B.centroid[3 * p + best_axis] - lo) * scale);
			b = std::min(std::max(b, 0), kBins - 1);
			return b <= best_bin;
		});
		mid = static_cast<uint32_t>(split - begin);
	} else {
		mid = 0;
	}
	if (mid == 0 || mid == count) {
		/* binning failed (coincident centroids) or forced: median split on the widest axis */
		if (count <= static_cast<uint32_t>(kMaxLeafPrims) && best_axis < 0 && depth < kMedianDepth)
			return encode_leaf(first, count);
		int axis = 0;
		for (int k = 1; k < 3; k++)
			if (cbox.hi[k] - cbox.lo[k] > cbox.hi[axis] - cbox.lo[axis])
				axis = k;
		mid = count / 2;
		uint32_t *begin = B.order.data() + first;
		std::nth_element(begin, begin + mid, begin + count, [&](uint32_t a, uint32_t b) {
			const float ca = B.centroid[3 * a + axis], cb = B.centroid[3 * b + axis];
			return ca < cb || (ca == cb && a < b);
		});
		if (count <= static_cast<uint32_t>(kMaxLeafPrims) && depth >= kMedianDepth)
			return encode_leaf(first, count);
	}

	const int32_t id = static_cast<int32_t>(B.nodes.size());
	B.nodes.emplace_back();
	std::memset(&B.nodes[id], 0, sizeof(BvhNode));
	Aabb lb, rb;
	const int32_t l = build_range(B, first, mid, depth + 1, lb);
	const int32_t r = build_range(B, first + mid, count - mid, depth + 1, rb);
	store_child(B.nodes[id], 0, l, lb);
	store_child(B.nodes[id], 1, r, rb);
	return id;
}

/* Copies the top tree and the deferred subtrees into `out` in pre-order (the
 * numbering of a single-threaded build) and returns the new reference of `ref`.
 * Child bounds were already known when the top was built. */
int32_t splice(const std::vector<BvhNode> &top, const std::vector<Deferred> &deferred, int32_t ref, std::vector<BvhNode> &out)
{
	if (ref < 0)
		return ref;
	if (ref >= kDeferredBase) {
		const Deferred &d = deferred[ref - kDeferredBase];
		if (d.root < 0)
			return d.root;
		const int32_t base = static_cast<int32_t>(out.size());
		for (BvhNode n : d.nodes) {
			for (int c = 0; c < 2; c++)
				if (n.child[c] >= 0)
					n.child[c] += base;
			out.push_back(n);
		}
		return base + d.root;
	}
	const int32_t id = static_cast<int32_t>(out.size());
	out.push_back(top[ref]);
	for (int c = 0; c < 2; c++) {
		const int32_t r = splice(top, deferred, top[ref].child[c], out);
		out[id].child[c] = r;
	}
	return id;
}

constexpr uint32_t kForkAbove = 16384; /* smaller ranges are built by the thread that has them */

struct ForkState {
	uint32_t max_threads;
	std::atomic<uint32_t> running{ 1 };
};

/* Builds the subtree of d.first / d.count into d.nodes (local pre-order).  A
 * large range is split once by the calling thread; its halves are then built
 * concurrently when a thread is free, one after the other otherwise. */
void build_subtree(const std::vector<Aabb> &boxes, const std::vector<float> &centroid, std::vector<uint32_t> &order, Deferred &d, ForkState &fork)
{
	if (fork.max_threads <= 1 || d.count < kForkAbove) {
		Builder B(boxes, centroid, order, d.nodes);
		d.root = build_range(B, d.first, d.count, d.depth, d.bounds);
		d.max_depth = B.max_depth;
		return;
	}
	/* split this range once: both halves come back deferred */
	std::vector<BvhNode> top;
	std::vector<Deferred> halves;
	halves.reserve(2);
	Builder B(boxes, centroid, order, top);
	B.defer_below = d.count - 1;
	B.deferred = &halves;
	const int32_t root = build_range(B, d.first, d.count, d.depth, d.bounds);
	d.max_depth = B.max_depth;
	if (halves.size() == 2 && fork.running.fetch_add(1) < fork.max_threads) {
		std::thread other([&]() {
			build_subtree(boxes, centroid, order, halves[1], fork);
			fork.running.fetch_sub(1);
		});
		build_subtree(boxes, centroid, order, halves[0], fork);
		other.join();
	} else {
		if (halves.size() == 2)
			fork.running.fetch_sub(1);
		for (Deferred &h : halves)
			build_subtree(boxes, centroid, order, h, fork);
	}
	for (const Deferred &h : halves)
		d.max_depth = std::max(d.max_depth, h.max_depth);
	d.root = splice(top, halves, root, d.nodes);
}

} // namespace

void build_bvh(const std::vector<Aabb> &boxes, BvhResult &out)
{
	out.nodes.clear();
	out.order.clear();
	out.max_depth = 0;
	const uint32_t n = static_cast<uint32_t>(boxes.size());
	if (n == 0)
		return;
	out.order.resize(n);
	for (uint32_t i = 0; i < n; i++)
		out.order[i] = i;
	std::vector<float> centroid(3 * static_cast<size_t>(n));
	for (uint32_t i = 0; i < n; i++)
		for (int k = 0; k < 3; k++)
			centroid[3 * i + k] = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
	/* Fork-join over the top of the tree: a range is split by one thread, its two
	 * halves are built concurrently while threads are free (they partition
	 * disjoint ranges of `order`), and the results are laid out in the pre-order
	 * a single-threaded build produces, so the tree does not depend on the
	 * thread count or on the schedule. */
	uint32_t threads = std::max(1u, std::thread::hardware_concurrency());
	if (const char *e = std::getenv("CRT_BVH_THREADS"))
		threads = static_cast<uint32_t>(std::max(1, std::atoi(e)));
	threads = std::min(threads, 64u);
	Deferred whole;
	whole.first = 0;
	whole.count = n;
	whole.depth = 1;
	ForkState fork;
	fork.max_threads = threads;
	build_subtree(boxes, centroid, out.order, whole, fork);
	out.nodes = std::move(whole.nodes);
	const Aabb root_box = whole.bounds;
	const int32_t root = whole.root;
	if (root < 0) {
		/* the whole scene fits one leaf: wrap it so that node 0 is an inner node */
		BvhNode node;
		std::memset(&node, 0, sizeof node);
		/* a far-away point box: the slab test misses it for every ray (an inverted
		 * box would pass a min/max slab test) */
		Aabb empty;
		for (int k = 0; k < 3; k++) {
			empty.lo[k] = 1e30f;
			empty.hi[k] = 1e30f;
		}
		store_child(node, 0, root, root_box);
		store_child(node, 1, encode_leaf(0, 1), empty);
		out.nodes.push_back(node);
	}
	out.max_depth = whole.max_depth;
}

/* ---- 8-wide collapse ------------------------------------------------------ */

uint16_t half_bits(float f)
{
	uint32_t x;
	std::memcpy(&x, &f, 4);
	const uint32_t sign = (x >> 16) & 0x8000u;
	const int32_t exp = static_cast<int32_t>((x >> 23) & 0xffu) - 127 + 15;
	uint32_t man = x & 0x7fffffu;
	if (((x >> 23) & 0xffu) == 0xffu)
		return static_cast<uint16_t>(sign | 0x7c00u | (man ? 0x200u : 0u));
	if (exp >= 31)
		return static_cast<uint16_t>(sign | 0x7c00u);
	if (exp <= 0) {
		if (exp < -10)
			return static_cast<uint16_t>(sign);
		man |= 0x800000u;
		const int shift = 14 - exp;
		uint32_t h = man >> shift;
		const uint32_t rem = man & ((1u << shift) - 1u), half = 1u << (shift - 1);
		if (rem > half || (rem == half && (h & 1u)))
			h++;
		return static_cast<uint16_t>(sign | h);
	}
	uint32_t h = (static_cast<uint32_t>(exp) << 10) | (man >> 13);
	const uint32_t rem = man & 0x1fffu;
	if (rem > 0x1000u || (rem == 0x1000u && (h & 1u)))
		h++;
	return static_cast<uint16_t>(sign | h);
}

float half_value(uint16_t h)
{
	const uint32_t sign = (h & 0x8000u) << 16, exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
	float f;
	if (exp == 0) {
		f = std::ldexp(static_cast<float>(man), -24);
	} else if (exp == 31) {
		uint32_t x = 0x7f800000u | (man << 13);
		std::memcpy(&f, &x, 4);
	} else {
		f = std::ldexp(static_cast<float>(man | 0x400u), static_cast<int>(exp) - 25);
	}
	uint32_t x;
	std::memcpy(&x, &f, 4);
	x |= sign;
	std::memcpy(&f, &x, 4);
	return f;
}

namespace {

struct WideRef {
	int32_t ref; /* binary node index, or encoded leaf */
	Aabb box;
};

struct WideBuilder {
	const BvhResult &bvh;
	WideTree &out;
	/* optimal collapse (see plan_collapse): per binary node, cost[i - 1] of covering its subtree with at
	 * most i wide-tree children, and the split that achieves it */
	std::vector<float> cost;    /* 7 per node */
	std::vector<uint8_t> split; /* 8 per node: [0] = children given to the left side when the node is a wide
	                               node (8 slots); [i - 1], i = 2..7: left share of i slots, 0 = use i - 1 */
	float prim_cost = 0.6f;     /* one primitive test relative to one visit of a wide node */
	uint32_t next_prim = 0;
};

Aabb child_box(const BvhNode &n, int which)
{
	const float *xy = which == 0 ? n.xy0 : n.xy1;
	Aabb b;
	b.lo[0] = xy[0];
	b.hi[0] = xy[1];
	b.lo[1] = xy[2];
	b.hi[1] = xy[3];
	b.lo[2] = n.z01[2 * which + 0];
	b.hi[2] = n.z01[2 * which + 1];
	return b;
}

bool absent(const Aabb &b)
{
	return b.lo[0] >= 1e30f; /* the second child of a one-primitive scene */
}

/* Which binary nodes become 8-wide nodes?  Every visit of a wide node costs the
 * same eight slab tests, so the expected cost of a walk is the sum over wide
 * nodes of the chance that a ray meets them (their surface) plus, for every
 * primitive, prim_cost times the surface of its box.  That sum is minimised
 * exactly by dynamic programming over the binary tree (Ylitie, Karras, Laine
 * 2017, "Efficient incoherent ray traversal on GPUs through compressed wide
 * BVHs", section 4.2):
 *   C(n, 1) = area(n) + min over k of  T(left, k) + T(right, 8 - k)
 *   C(n, i) = min( C(n, i - 1), min over k of  T(left, k) + T(right, i - k) ),  i = 2..7
 *   T(x, i) = area(x) * prim_cost for a leaf (one primitive), C(x, i) for an inner node. */
float forest_cost(const WideBuilder &W, int32_t ref, const Aabb &box, int slots)
{
	if (ref < 0)
		return absent(box) ? 0.f : half_area(box) * W.prim_cost;
	return W.cost[7 * static_cast<size_t>(ref) + (slots - 1)];
}

void plan_collapse(WideBuilder &W)
{
	const size_t n = W.bvh.nodes.size();
	W.cost.assign(7 * n, 0.f);
	W.split.assign(8 * n, 0);
	/* children have larger indices than their parent (pre-order): go backwards */
	for (size_t v = n; v-- > 0;) {
		const BvhNode &node = W.bvh.nodes[v];
		const int32_t ref[2] = { node.child[0], node.child[1] };
		const Aabb cbox[2] = { child_box(node, 0), child_box(node, 1) };
		Aabb box;
		box_reset(box);
		for (int c = 0; c < 2; c++)
			if (!absent(cbox[c]))
				box_grow(box, cbox[c]);
		float *cost = &W.cost[7 * v];
		uint8_t *split = &W.split[8 * v];
		float best = FLT_MAX;
		split[0] = 4; /* any share 1..7 is a valid collapse: kept when no cost below compares smaller (surfaces that overflow, NaN boxes) */
		for (int k = 1; k <= 7; k++) {
			const float c = forest_cost(W, ref[0], cbox[0], k) + forest_cost(W, ref[1], cbox[1], 8 - k);
			if (c < best) {
				best = c;
				split[0] = static_cast<uint8_t>(k);
			}
		}
		cost[0] = half_area(box) + best;
		for (int i = 2; i <= 7; i++) {
			cost[i - 1] = cost[i - 2];
			split[i - 1] = 0;
			for (int k = 1; k < i; k++) {
				const float c = forest_cost(W, ref[0], cbox[0], k) + forest_cost(W, ref[1], cbox[1], i - k);
				if (c < cost[i - 1]) {
					cost[i - 1] = c;
					split[i - 1] = static_cast<uint8_t>(k);
				}
			}
		}
	}
}

/* the wide-tree children that cover child `which` of binary node `node` with at most `slots` of them */
void expand(const WideBuilder &W, int32_t node, int which, int slots, std::vector<WideRef> &kids)
{
	WideRef r;
	r.ref = W.bvh.nodes[node].child[which];
	r.box = child_box(W.bvh.nodes[node], which);
	if (r.ref < 0) {
		if (!absent(r.box))
			kids.push_back(r);
		return;
	}
	while (slots > 1 && W.split[8 * static_cast<size_t>(r.ref) + (slots - 1)] == 0)
		slots--;
	if (slots == 1) {
		kids.push_back(r); /* a wide node of its own */
		return;
	}
	const int k = W.split[8 * static_cast<size_t>(r.ref) + (slots - 1)];
	expand(W, r.ref, 0, k, kids);
	expand(W, r.ref, 1, slots - k, kids);
}

/* Grid of a node: origin = lower corner of the union of the child boxes, scale = the
 * smallest power of two with extent <= (kWideGrid - 1) * scale on every axis.  Child
 * planes are rounded outwards to grid lines in exact (double) arithmetic. */
void quantize_node(WideNode &n, const std::vector<WideRef> &kids, const std::vector<int> &slot_of)
{
	Aabb box;
	box_reset(box);
	for (const WideRef &k : kids)
		box_grow(box, k.box);
	double extent = 0.0;
	for (int a = 0; a < 3; a++) {
		n.origin[a] = box.lo[a];
		extent = std::max(extent, static_cast<double>(box.hi[a]) - static_cast<double>(box.lo[a]));
	}
	int e = -40;
	if (extent > 0.0) {
		int ex;
		std::frexp(extent / (kWideGrid - 1.0), &ex); /* extent / 1022 = m * 2^ex, m in [0.5, 1) => 2^ex >= extent / 1022 */
		e = std::min(std::max(ex, -40), 40);
	}
	double scale = std::ldexp(1.0, e);
	/* a scale clamped from above cannot hold the node: widen until it does (boxes that large are not built in practice;
	 * the exponent byte ends at 2^127, which also ends the loop for an infinite extent) */
	while (extent > (kWideGrid - 1.0) * scale && e < 127) {
		e++;
		scale = std::ldexp(1.0, e);
	}
	n.meta = (n.meta & ~0xffu) | static_cast<uint32_t>(e + 127);
	const uint16_t empty_c = half_bits(0.f), empty_h = 0xfc00u; /* -infinity: the slab of an empty slot is inverted on every axis */
	for (int a = 0; a < 3; a++)
		for (int j = 0; j < 8; j++) {
			n.plane[a][0][j] = empty_c;
			n.plane[a][1][j] = empty_h;
		}
	for (size_t k = 0; k < kids.size(); k++) {
		const int j = slot_of[k];
		for (int a = 0; a < 3; a++) {
			const double o = n.origin[a];
			double lo = std::floor((static_cast<double>(kids[k].box.lo[a]) - o) / scale);
			double hi = std::ceil((static_cast<double>(kids[k].box.hi[a]) - o) / scale);
			lo = std::min(std::max(lo, 0.0), static_cast<double>(kWideGrid));
			hi = std::min(std::max(hi, lo), static_cast<double>(kWideGrid));
			const float c = static_cast<float>(0.5 * (lo + hi)), h = static_cast<float>(0.5 * (hi - lo)) + kWideSlack;
			n.plane[a][0][j] = half_bits(c);
			n.plane[a][1][j] = half_bits(h);
		}
	}
}

void emit_node(WideBuilder &W, int32_t binary, int32_t id, int32_t parent, uint32_t depth)
{
	W.out.max_depth = std::max(W.out.max_depth, depth);
	std::vector<WideRef> kids;
	const int k = W.split[8 * static_cast<size_t>(binary)];
	expand(W, binary, 0, k, kids);
	expand(W, binary, 1, 8 - k, kids);
	/* slots: leaves first, then inner children; each kind sorted along the longest axis of the node
	 * (the closest-hit walk takes the slots of a kind in that order, or in reverse for a ray going the
	 * other way, so that near children tend to come first) */
	Aabb box;
	box_reset(box);
	for (const WideRef &r : kids)
		box_grow(box, r.box);
	int axis = 0;
	for (int a = 1; a < 3; a++)
		if (box.hi[a] - box.lo[a] > box.hi[axis] - box.lo[axis])
			axis = a;
	std::vector<int> idx(kids.size());
	for (size_t i = 0; i < kids.size(); i++)
		idx[i] = static_cast<int>(i);
	std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) {
		const bool lx = kids[x].ref < 0, ly = kids[y].ref < 0;
		if (lx != ly)
			return lx;
		return kids[x].box.lo[axis] + kids[x].box.hi[axis] < kids[y].box.lo[axis] + kids[y].box.hi[axis];
	});
	std::vector<WideRef> sorted(kids.size());
	std::vector<int> slot_of(kids.size());
	uint32_t leaves = 0;
	for (size_t i = 0; i < kids.size(); i++) {
		sorted[i] = kids[idx[i]];
		slot_of[i] = static_cast<int>(i);
		if (sorted[i].ref < 0)
			leaves++;
	}
	const uint32_t inner = static_cast<uint32_t>(sorted.size()) - leaves;
	const int32_t first_child = static_cast<int32_t>(W.out.nodes.size());
	W.out.nodes.resize(W.out.nodes.size() + inner);
	WideNode n;
	std::memset(&n, 0, sizeof n);
	n.meta = (((1u << leaves) - 1u) << 8) | ((((1u << inner) - 1u) << leaves) << 16) | (static_cast<uint32_t>(axis) << 24);
	n.child_base = first_child - static_cast<int32_t>(leaves);
	n.prim_base = static_cast<int32_t>(W.next_prim);
	n.parent = parent;
	for (uint32_t j = 0; j < leaves; j++) {
		const uint32_t first = static_cast<uint32_t>(~sorted[j].ref) >> kLeafCountBits;
		W.out.order[W.next_prim++] = W.bvh.order[first];
		W.out.cost += static_cast<double>(W.prim_cost) * half_area(sorted[j].box);
	}
	W.out.cost += half_area(box);
	quantize_node(n, sorted, slot_of);
	W.out.nodes[id] = n;
	for (uint32_t j = 0; j < inner; j++)
		emit_node(W, sorted[leaves + j].ref, first_child + static_cast<int32_t>(j), id, depth + 1);
}

} // namespace

void build_wide(const std::vector<Aabb> &boxes, const BvhResult &bvh, WideTree &out)
{
	out.nodes.clear();
	out.order.clear();
	out.max_depth = 0;
	out.cost = 0.0;
	if (bvh.nodes.empty())
		return;
	WideBuilder W{ bvh, out, {}, {} };
	if (const char *e = std::getenv("CRT_WIDE_PRIM_COST"))
		W.prim_cost = static_cast<float>(std::atof(e));
	plan_collapse(W);
	out.order.assign(boxes.size(), 0);
	out.nodes.reserve(bvh.nodes.size() / 3 + 1);
	out.nodes.emplace_back();
	emit_node(W, 0, 0, -1, 1);
	const BvhNode &root = bvh.nodes[0];
	Aabb rb;
	box_reset(rb);
	for (int c = 0; c < 2; c++)
		if (!absent(child_box(root, c)))
			box_grow(rb, child_box(root, c));
	const float ra = half_area(rb);
	out.cost = ra > 0.f ? out.cost / ra : 0.0;
}

} // namespace crt
