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

namespace {

struct WideRef {
	int32_t ref; /* binary node index, or encoded leaf */
	Aabb box;
	uint32_t first, count;
};

struct WideBuilder {
	const BvhResult &bvh;
	WideResult &out;
	std::vector<uint32_t> first, count; /* primitive range of every binary node */
	uint32_t merge_prims = 0;            /* binary subtrees with at most this many primitives become one leaf */
	/* optimal collapse (see plan_collapse): per binary node, cost[i - 1] of covering its subtree with at
	 * most i wide-tree children, and the split that achieves it */
	std::vector<float> cost;             /* 7 per node */
	std::vector<uint8_t> split;          /* 8 per node: [0] = children given to the left side when the node is a wide
	                                        node (8 slots); [i - 1], i = 2..7: left share of i slots, 0 = use i - 1 */
	bool optimal = false;
	bool order_by_area = false;
	float prim_cost = 0.35f;             /* one primitive test of a leaf relative to one visit of a wide node */
};

constexpr uint8_t kPlanLeaf = 255;          /* split[0]: the whole subtree becomes one leaf */

void decode_leaf(int32_t ref, uint32_t &first, uint32_t &count)
{
	const uint32_t leaf = static_cast<uint32_t>(~ref);
	first = leaf >> kLeafCountBits;
	count = (leaf & ((1u << kLeafCountBits) - 1u)) + 1u;
}

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

void ranges(WideBuilder &W, int32_t node)
{
	uint32_t lo = 0xffffffffu, total = 0;
	for (int c = 0; c < 2; c++) {
		const int32_t ref = W.bvh.nodes[node].child[c];
		uint32_t f, n;
		if (ref < 0) {
			decode_leaf(ref, f, n);
			if (child_box(W.bvh.nodes[node], c).lo[0] >= 1e30f)
				continue; /* the absent child of a one-leaf scene */
		} else {
			ranges(W, ref);
			f = W.first[ref];
			n = W.count[ref];
		}
		lo = std::min(lo, f);
		total += n;
	}
	W.first[node] = lo;
	W.count[node] = total;
}

WideRef make_ref(WideBuilder &W, int32_t node, int which)
{
	WideRef r;
	r.ref = W.bvh.nodes[node].child[which];
	r.box = child_box(W.bvh.nodes[node], which);
	if (r.ref < 0) {
		decode_leaf(r.ref, r.first, r.count);
	} else {
		r.first = W.first[r.ref];
		r.count = W.count[r.ref];
	}
	return r;
}

/* Which binary nodes become 8-wide nodes?  Opening the child with the largest
 * surface until eight are found (the loop in wide_node) leaves many
 * bottom-level nodes with two or three children.  Every visit of a wide node
 * costs the same eight slab tests, so the expected cost of the shadow walk is
 * the sum over wide nodes of the chance that a ray meets them (their surface):
 * that sum is minimised exactly by dynamic programming over the binary tree
 * (Ylitie, Karras, Laine 2017, "Efficient incoherent ray traversal on GPUs
 * through compressed wide BVHs", section 4.2), with a leaf of P primitives
 * costing prim_cost * P per ray that meets its box; a binary subtree of at
 * most merge_prims primitives may become one leaf where that is cheaper:
 *   C(n, 1) = min( area(n) + min over k of  T(left, k) + T(right, 8 - k),  area(n) * P(n) * prim_cost )
 *   C(n, i) = min( C(n, i - 1), min over k of  T(left, k) + T(right, i - k) ),  i = 2..7
 *   T(x, i) = area(x) * P(x) * prim_cost for a binary leaf, C(x, i) for an inner node. */
float forest_cost(const WideBuilder &W, int32_t ref, uint32_t count, const Aabb &box, int slots)
{
	if (ref < 0)
		return half_area(box) * static_cast<float>(count) * W.prim_cost;
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
		int32_t ref[2];
		uint32_t cnt[2];
		Aabb cbox[2];
		Aabb box;
		for (int k = 0; k < 3; k++) {
			box.lo[k] = FLT_MAX;
			box.hi[k] = -FLT_MAX;
		}
		for (int c = 0; c < 2; c++) {
			ref[c] = node.child[c];
			if (ref[c] < 0) {
				uint32_t f;
				decode_leaf(ref[c], f, cnt[c]);
			} else {
				cnt[c] = W.count[ref[c]];
			}
			const Aabb b = child_box(node, c);
			cbox[c] = b;
			if (b.lo[0] >= 1e30f) {
				cnt[c] = 0; /* the absent child of a one-leaf scene */
				continue;
			}
			for (int k = 0; k < 3; k++) {
				box.lo[k] = std::min(box.lo[k], b.lo[k]);
				box.hi[k] = std::max(box.hi[k], b.hi[k]);
			}
		}
		float *cost = &W.cost[7 * v];
		uint8_t *split = &W.split[8 * v];
		float best = FLT_MAX;
		for (int k = 1; k <= 7; k++) {
			const float c = forest_cost(W, ref[0], cnt[0], cbox[0], k) + forest_cost(W, ref[1], cnt[1], cbox[1], 8 - k);
			if (c < best) {
				best = c;
				split[0] = static_cast<uint8_t>(k);
			}
		}
		cost[0] = half_area(box) + best;
		const uint32_t prims = cnt[0] + cnt[1];
		if (prims <= W.merge_prims) {
			const float as_leaf = half_area(box) * static_cast<float>(prims) * W.prim_cost;
			if (as_leaf < cost[0]) {
				cost[0] = as_leaf;
				split[0] = kPlanLeaf;
			}
		}
		for (int i = 2; i <= 7; i++) {
			cost[i - 1] = cost[i - 2];
			split[i - 1] = 0;
			for (int k = 1; k < i; k++) {
				const float c = forest_cost(W, ref[0], cnt[0], cbox[0], k) + forest_cost(W, ref[1], cnt[1], cbox[1], i - k);
				if (c < cost[i - 1]) {
					cost[i - 1] = c;
					split[i - 1] = static_cast<uint8_t>(k);
				}
			}
		}
	}
}

/* the wide-tree children that cover child `which` of binary node `node` with at most `slots` of them */
void expand(WideBuilder &W, int32_t node, int which, int slots, std::vector<WideRef> &kids)
{
	const WideRef r = make_ref(W, node, which);
	if (r.ref < 0 && r.box.lo[0] >= 1e30f)
		return; /* the absent child of a one-leaf scene */
	if (r.ref < 0) {
		kids.push_back(r);
		return;
	}
	while (slots > 1 && W.split[8 * static_cast<size_t>(r.ref) + (slots - 1)] == 0)
		slots--;
	if (slots == 1) {
		kids.push_back(r); /* a wide node of its own, or -- split[0] == kPlanLeaf -- one merged leaf */
		return;
	}
	const int k = W.split[8 * static_cast<size_t>(r.ref) + (slots - 1)];
	expand(W, r.ref, 0, k, kids);
	expand(W, r.ref, 1, slots - k, kids);
}

int32_t wide_node(WideBuilder &W, int32_t node, uint32_t depth)
{
	W.out.max_depth = std::max(W.out.max_depth, depth);
	const int32_t id = static_cast<int32_t>(W.out.nodes.size());
	W.out.nodes.emplace_back();
	std::vector<WideRef> kids;
	if (W.optimal) {
		const int k = W.split[8 * static_cast<size_t>(node)];
		expand(W, node, 0, k, kids);
		expand(W, node, 1, 8 - k, kids);
	} else {
		for (int c = 0; c < 2; c++) {
			const WideRef r = make_ref(W, node, c);
			if (!(r.ref < 0 && r.box.lo[0] >= 1e30f))
				kids.push_back(r);
		}
		/* open the binary child with the largest surface that is too big for a wide leaf */
		while (kids.size() < 8) {
			int best = -1;
			float best_area = -1.f;
			for (size_t i = 0; i < kids.size(); i++)
				if (kids[i].ref >= 0 && kids[i].count > W.merge_prims && half_area(kids[i].box) > best_area) {
					best_area = half_area(kids[i].box);
					best = static_cast<int>(i);
				}
			if (best < 0)
				break;
			const int32_t open = kids[best].ref;
			kids[best] = make_ref(W, open, 0);
			kids.push_back(make_ref(W, open, 1));
		}
	}
	/* experiment (CRT_WIDE_ORDER=area, off by default): children are pushed in slot order and
	 * popped in reverse, so the largest boxes last = visited first, which finds an opaque
	 * occluder sooner; any-hit results do not depend on the order */
	if (W.order_by_area)
		std::stable_sort(kids.begin(), kids.end(), [](const WideRef &a, const WideRef &b) { return half_area(a.box) < half_area(b.box); });
	for (int j = 0; j < 8; j++) {
		int32_t code = kWideEmpty;
		Aabb box;
		for (int k = 0; k < 3; k++)
			box.lo[k] = box.hi[k] = 1e30f;
		if (j < static_cast<int>(kids.size())) {
			const WideRef &r = kids[j];
			box = r.box;
			if (r.ref < 0 || r.count <= W.merge_prims)
				code = encode_leaf(r.first, r.count); /* a whole small subtree becomes one leaf */
			else
				code = wide_node(W, r.ref, depth + 1);
		}
		WideNode &n = W.out.nodes[id]; /* re-fetch: the vector may have grown */
		n.a[j][0] = box.lo[0];
		n.a[j][1] = box.lo[1];
		n.a[j][2] = box.lo[2];
		n.a[j][3] = box.hi[0];
		n.b[j][0] = box.hi[1];
		n.b[j][1] = box.hi[2];
		std::memcpy(&n.b[j][2], &code, sizeof code);
		n.b[j][3] = 0.f;
	}
	return id;
}

} // namespace

void build_wide(const BvhResult &bvh, WideResult &out, uint32_t merge_prims)
{
	out.nodes.clear();
	out.max_depth = 0;
	if (bvh.nodes.empty())
		return;
	WideBuilder W{ bvh, out, {}, {}, std::min(merge_prims, kWideLeafPrims) };
	W.first.assign(bvh.nodes.size(), 0);
	W.count.assign(bvh.nodes.size(), 0);
	ranges(W, 0);
	const char *mode = std::getenv("CRT_WIDE_COLLAPSE"); /* "greedy": largest surface first (the round-1 trees) */
	W.optimal = !(mode && !std::strcmp(mode, "greedy"));
	if (const char *e = std::getenv("CRT_WIDE_ORDER"))
		W.order_by_area = !std::strcmp(e, "area");
	if (const char *e = std::getenv("CRT_WIDE_PRIM_COST"))
		W.prim_cost = static_cast<float>(std::atof(e));
	if (W.optimal)
		plan_collapse(W);
	out.nodes.reserve(bvh.nodes.size() / 2 + 1);
	wide_node(W, 0, 1);
}


/* ---- quantisation of the 8-wide nodes ----------------------------------------- */

namespace {

/* smallest power of two s with extent <= 255 * s, as an IEEE exponent byte */
uint32_t scale_exponent(float extent)
{
	if (!(extent > 0.f))
		return 1u; /* 2^-126: all planes quantise to 0 */
	int e;
	std::frexp(extent / 255.f, &e); /* extent / 255 = m * 2^e, m in [0.5, 1) => 2^e >= extent / 255 */
	const int biased = std::min(std::max(e + 127, 1), 254);
	return static_cast<uint32_t>(biased);
}

float exponent_scale(uint32_t biased)
{
	const uint32_t bits = biased << 23;
	float f;
	std::memcpy(&f, &bits, 4);
	return f;
}

} // namespace

void quantize_wide(const WideResult &wide, std::vector<QuantNode> &out)
{
	out.resize(wide.nodes.size());
	for (size_t n = 0; n < wide.nodes.size(); n++) {
		const WideNode &w = wide.nodes[n];
		QuantNode &qn = out[n];
		std::memset(&qn, 0, sizeof qn);
		float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
		int32_t code[8];
		float clo[8][3], chi[8][3];
		for (int j = 0; j < 8; j++) {
			std::memcpy(&code[j], &w.b[j][2], 4);
			clo[j][0] = w.a[j][0]; clo[j][1] = w.a[j][1]; clo[j][2] = w.a[j][2];
			chi[j][0] = w.a[j][3]; chi[j][1] = w.b[j][0]; chi[j][2] = w.b[j][1];
			if (code[j] == kWideEmpty)
				continue;
			for (int k = 0; k < 3; k++) {
				lo[k] = std::min(lo[k], clo[j][k]);
				hi[k] = std::max(hi[k], chi[j][k]);
			}
		}
		uint32_t ebits = 0;
		float scale[3];
		for (int k = 0; k < 3; k++) {
			if (lo[k] > hi[k])
				lo[k] = hi[k] = 0.f;
			qn.origin[k] = lo[k];
			uint32_t e = scale_exponent(hi[k] - lo[k]);
			/* origin + 255 * scale must reach hi in FP32 arithmetic */
			while (e < 254u && qn.origin[k] + 255.f * exponent_scale(e) < hi[k])
				e++;
			scale[k] = exponent_scale(e);
			ebits |= e << (8 * k);
		}
		qn.exponents = ebits;
		for (int j = 0; j < 8; j++) {
			qn.code[j] = code[j];
			for (int k = 0; k < 3; k++) {
				uint8_t ql = 255, qh = 0; /* empty child: inverted, never hit */
				if (code[j] != kWideEmpty) {
					int a = static_cast<int>(std::floor((clo[j][k] - qn.origin[k]) / scale[k]));
					a = std::min(std::max(a, 0), 255);
					while (a > 0 && qn.origin[k] + static_cast<float>(a) * scale[k] > clo[j][k])
						a--;
					int b = static_cast<int>(std::ceil((chi[j][k] - qn.origin[k]) / scale[k]));
					b = std::min(std::max(b, 0), 255);
					while (b < 255 && qn.origin[k] + static_cast<float>(b) * scale[k] < chi[j][k])
						b++;
					ql = static_cast<uint8_t>(a);
					qh = static_cast<uint8_t>(b);
				}
				qn.q[k][0][j] = ql;
				qn.q[k][1][j] = qh;
			}
		}
	}
}

} // namespace crt
