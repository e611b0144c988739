/*
 * crt_bvh_device.cu -- the acceleration structure built ON THE DEVICE.
 *
 * Replaces accel_init() (reference src/raytracer/accel.c:266-315) the way the reference meant
 * it: Morton codes of the primitive centres (accel.c:72-88; here 21 bits per axis, normalised
 * over the box of the centres on every axis -- the reference normalises with one wrong extent,
 * SURVEY F8), a sort, and the radix tree over the sorted codes that accel.c:227-264 builds by
 * recursion (Karras 2012 builds every inner node independently instead).  Then, still on the
 * device: boxes bottom-up, the 8-wide collapse (each wide node keeps opening its largest child
 * until eight slots are used), the half-precision child boxes of bvh_build.h and the primitive
 * records permuted into tree order.  Nothing but counters crosses PCIe.
 *
 * The result has the layout and the invariants of the host builder's tree (bvh_build.h;
 * tests/bvh_check.cpp checks both), not its quality: no SAH, so a ray tests more boxes.
 * It is deterministic (stable sort, scans instead of atomics for every index), so every
 * device of a multi-device job builds the same tree.
 */
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cuda_fp16.h>

#include <algorithm>
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <type_traits>

#include "bvh_build.h"
#include "crt_bvh_device.cuh"

namespace crt {
namespace {

constexpr int kThreads = 256;

struct Box {
	float lo[3], hi[3];
};

__device__ __forceinline__ float area_of(const Box &b)
{
	const float x = b.hi[0] - b.lo[0], y = b.hi[1] - b.lo[1], z = b.hi[2] - b.lo[2];
	return x * y + y * z + z * x;
}

/* monotone float <-> uint map, for atomicMin / atomicMax on floats */
__device__ __forceinline__ uint32_t ordered(float f)
{
	const uint32_t u = __float_as_uint(f);
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float unordered(uint32_t u)
{
	return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

/* ---- 1. primitive boxes (the padding of crt_api.cu build_host_scene) and the box of their centres --- */
__global__ void prim_boxes_kernel(const float4 *prims, uint32_t n, Box *boxes, uint32_t *centre_bounds)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	const float4 a = prims[4 * (size_t)i], b = prims[4 * (size_t)i + 1], c = prims[4 * (size_t)i + 2];
	const bool sphere = (__float_as_uint(b.w) & PRIM_SPHERE) != 0u;
	const float pa[3] = { a.x, a.y, a.z }, pb[3] = { b.x, b.y, b.z }, pc[3] = { c.x, c.y, c.z };
	Box box;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		float lo, hi;
		if (sphere) {
			lo = __fsub_rn(pa[k], b.x);
			hi = __fadd_rn(pa[k], b.x);
		} else {
			const float p1 = __fadd_rn(pa[k], pb[k]), p2 = __fadd_rn(pa[k], pc[k]);
			lo = fminf(pa[k], fminf(p1, p2));
			hi = fmaxf(pa[k], fmaxf(p1, p2));
		}
		const float pad = __fadd_rn(__fmul_rn(4e-7f, fmaxf(1e-3f, fmaxf(fabsf(lo), fabsf(hi)))), __fmul_rn(1e-7f, __fsub_rn(hi, lo)));
		box.lo[k] = __fsub_rn(lo, pad);
		box.hi[k] = __fadd_rn(hi, pad);
		const float centre = 0.5f * (box.lo[k] + box.hi[k]);
		atomicMin(centre_bounds + k, ordered(centre));
		atomicMax(centre_bounds + 3 + k, ordered(centre));
	}
	boxes[i] = box;
}

__device__ __forceinline__ uint64_t spread21(uint64_t v)
{
	v &= 0x1fffffull;
	v = (v | (v << 32)) & 0x1f00000000ffffull;
	v = (v | (v << 16)) & 0x1f0000ff0000ffull;
	v = (v | (v << 8)) & 0x100f00f00f00f00full;
	v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
	v = (v | (v << 2)) & 0x1249249249249249ull;
	return v;
}

/* ---- 2. Morton codes (accel.c:72-88 with every axis normalised by its own extent) ---------------- */
__global__ void morton_kernel(const Box *boxes, uint32_t n, const uint32_t *centre_bounds, uint64_t *keys, uint32_t *ids)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	uint64_t key = 0;
#pragma unroll
	for (int k = 0; k < 3; k++) {
		const float lo = unordered(centre_bounds[k]), hi = unordered(centre_bounds[3 + k]);
		const float centre = 0.5f * (boxes[i].lo[k] + boxes[i].hi[k]);
		const float ext = hi - lo;
		float u = ext > 0.f ? (centre - lo) / ext : 0.f;
		u = fminf(fmaxf(u, 0.f), 1.f);
		const uint64_t q = (uint64_t)fminf(u * 2097152.f, 2097151.f);
		key |= spread21(q) << (2 - k);
	}
	keys[i] = key;
	ids[i] = i;
}

/* ---- 3. the radix tree over the sorted codes (Karras 2012; accel.c:227-264 by recursion) ----------
 * nodes 0 .. n-2 are inner (0 is the root), node n-1+k is the leaf of sorted position k */
__device__ __forceinline__ int delta(const uint64_t *keys, int n, int i, int j)
{
	if (j < 0 || j >= n)
		return -1;
	const uint64_t x = keys[i] ^ keys[j];
	return x ? __clzll((long long)x) : 64 + __clz(i ^ j);
}

__global__ void radix_tree_kernel(const uint64_t *keys, int n, int2 *children, int *parent)
{
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n - 1)
		return;
	const int d = delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1) >= 0 ? 1 : -1;
	const int dmin = delta(keys, n, i, i - d);
	int lmax = 2;
	while (delta(keys, n, i, i + lmax * d) > dmin)
		lmax <<= 1;
	int l = 0;
	for (int t = lmax >> 1; t >= 1; t >>= 1)
		if (delta(keys, n, i, i + (l + t) * d) > dmin)
			l += t;
	const int j = i + l * d;
	const int dnode = delta(keys, n, i, j);
	int s = 0, t = l;
	do {
		t = (t + 1) >> 1;
		if (delta(keys, n, i, i + (s + t) * d) > dnode)
			s += t;
	} while (t > 1);
	const int gamma = i + s * d + min(d, 0);
	const int left = min(i, j) == gamma ? n - 1 + gamma : gamma;
	const int right = max(i, j) == gamma + 1 ? n - 1 + gamma + 1 : gamma + 1;
	children[i] = make_int2(left, right);
	parent[left] = i;
	parent[right] = i;
	if (i == 0)
		parent[0] = -1;
}

/* ---- 4. boxes bottom-up: the second thread to reach an inner node has both children ----------------- */
__global__ void refit_kernel(const Box *prim_boxes, const uint32_t *ids, int n, const int2 *children, const int *parent,
	Box *node_boxes, unsigned int *visits)
{
	const int k = blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n)
		return;
	node_boxes[n - 1 + k] = prim_boxes[ids[k]];
	__threadfence();
	int cur = parent[n - 1 + k];
	while (cur >= 0) {
		if (atomicAdd(visits + cur, 1u) == 0u)
			return;
		const int2 c = children[cur];
		const volatile Box *a = node_boxes + c.x, *b = node_boxes + c.y;
		Box box;
#pragma unroll
		for (int x = 0; x < 3; x++) {
			box.lo[x] = fminf(a->lo[x], b->lo[x]);
			box.hi[x] = fmaxf(a->hi[x], b->hi[x]);
		}
		node_boxes[cur] = box;
		__threadfence();
		cur = parent[cur];
	}
}

/* ---- 5. the 8-wide collapse, one level of the wide tree per launch -------------------------------- */
struct WideItem {
	int binary;     /* node of the radix tree this wide node stands for */
	int parent;     /* wide node index, -1 for the root */
	int slots[8];   /* radix-tree nodes of the slots, leaves first, each kind sorted along `axis`; -1 = empty */
	int leaves, inner, axis;
};

/* opens the child with the largest box until eight slots are used or only leaves remain */
__global__ void collapse_plan_kernel(WideItem *items, uint32_t count, int n, const int2 *children, const Box *node_boxes,
	uint32_t *leaf_counts, uint32_t *inner_counts)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	WideItem it = items[i];
	int slot[8], used = 0;
	if (it.binary >= n - 1) {
		slot[used++] = it.binary; /* a scene of one primitive: the root holds it */
	} else {
		const int2 c = children[it.binary];
		slot[used++] = c.x;
		slot[used++] = c.y;
	}
	while (used < 8) {
		int best = -1;
		float best_area = -1.f;
		for (int k = 0; k < used; k++)
			if (slot[k] < n - 1) {
				const float a = area_of(node_boxes[slot[k]]);
				if (a > best_area) {
					best_area = a;
					best = k;
				}
			}
		if (best < 0)
			break;
		const int2 c = children[slot[best]];
		slot[best] = c.x;
		slot[used++] = c.y;
	}
	/* longest axis of the union of the slots */
	Box u = node_boxes[slot[0]];
	for (int k = 1; k < used; k++)
		for (int x = 0; x < 3; x++) {
			u.lo[x] = fminf(u.lo[x], node_boxes[slot[k]].lo[x]);
			u.hi[x] = fmaxf(u.hi[x], node_boxes[slot[k]].hi[x]);
		}
	int axis = 0;
	for (int x = 1; x < 3; x++)
		if (u.hi[x] - u.lo[x] > u.hi[axis] - u.lo[axis])
			axis = x;
	/* leaves first, each kind by centre along the axis (insertion sort, stable) */
	for (int a = 1; a < used; a++) {
		const int v = slot[a];
		const bool vl = v >= n - 1;
		const float vc = node_boxes[v].lo[axis] + node_boxes[v].hi[axis];
		int b = a - 1;
		while (b >= 0) {
			const bool bl = slot[b] >= n - 1;
			const float bc = node_boxes[slot[b]].lo[axis] + node_boxes[slot[b]].hi[axis];
			const bool before = vl != bl ? vl : vc < bc;
			if (!before)
				break;
			slot[b + 1] = slot[b];
			b--;
		}
		slot[b + 1] = v;
	}
	it.leaves = it.inner = 0;
	for (int k = 0; k < 8; k++) {
		it.slots[k] = k < used ? slot[k] : -1;
		if (k < used) {
			if (slot[k] >= n - 1)
				it.leaves++;
			else
				it.inner++;
		}
	}
	it.axis = axis;
	items[i] = it;
	leaf_counts[i] = (uint32_t)it.leaves;
	inner_counts[i] = (uint32_t)it.inner;
}

/* half bits of a value that is exactly representable (multiples of 1/2 below 1024 + the slack) */
__device__ __forceinline__ uint16_t half_bits_dev(float f)
{
	return __half_as_ushort(__float2half_rn(f));
}

/* writes the wide nodes of one level, the next level's items and the primitive order (bvh_build.cpp:
 * emit_node + quantize_node, same arithmetic) */
__global__ void collapse_emit_kernel(const WideItem *items, uint32_t count, uint32_t level_first, int n, const Box *node_boxes,
	const uint32_t *ids, const uint32_t *leaf_offsets, const uint32_t *inner_offsets, uint32_t prim_first, uint32_t next_first,
	WideNode *nodes, uint32_t *order, WideItem *next_items)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= count)
		return;
	const WideItem it = items[i];
	const uint32_t id = level_first + i;
	const uint32_t prim_base = prim_first + leaf_offsets[i], first_child = next_first + inner_offsets[i];
	WideNode node;
	node.child_base = (int32_t)first_child - it.leaves;
	node.prim_base = (int32_t)prim_base;
	node.parent = it.parent;
	node.pad = 0;
	Box u = node_boxes[it.slots[0]];
	for (int k = 1; k < it.leaves + it.inner; k++)
		for (int x = 0; x < 3; x++) {
			u.lo[x] = fminf(u.lo[x], node_boxes[it.slots[k]].lo[x]);
			u.hi[x] = fmaxf(u.hi[x], node_boxes[it.slots[k]].hi[x]);
		}
	double extent = 0.0;
	for (int x = 0; x < 3; x++) {
		node.origin[x] = u.lo[x];
		extent = fmax(extent, (double)u.hi[x] - (double)u.lo[x]);
	}
	int e = -40;
	if (extent > 0.0) {
		int ex;
		frexp(extent / ((double)kWideGrid - 1.0), &ex);
		e = min(max(ex, -40), 40);
	}
	double scale = ldexp(1.0, e);
	while (extent > ((double)kWideGrid - 1.0) * scale) {
		e++;
		scale = ldexp(1.0, e);
	}
	node.meta = (uint32_t)(e + 127) | (((1u << it.leaves) - 1u) << 8) | ((((1u << it.inner) - 1u) << it.leaves) << 16) | ((uint32_t)it.axis << 24);
	for (int j = 0; j < 8; j++) {
		const int ref = it.slots[j];
		for (int x = 0; x < 3; x++) {
			if (ref < 0) {
				node.plane[x][0][j] = 0;        /* centre 0 */
				node.plane[x][1][j] = 0xfc00u;  /* half extent -infinity: never hit */
				continue;
			}
			const double o = node.origin[x];
			double lo = floor(((double)node_boxes[ref].lo[x] - o) / scale);
			double hi = ceil(((double)node_boxes[ref].hi[x] - o) / scale);
			lo = fmin(fmax(lo, 0.0), (double)kWideGrid);
			hi = fmin(fmax(hi, lo), (double)kWideGrid);
			node.plane[x][0][j] = half_bits_dev((float)(0.5 * (lo + hi)));
			node.plane[x][1][j] = half_bits_dev((float)(0.5 * (hi - lo)) + kWideSlack);
		}
	}
	nodes[id] = node;
	for (int j = 0; j < it.leaves; j++)
		order[prim_base + j] = ids[it.slots[j] - (n - 1)];
	for (int j = 0; j < it.inner; j++) {
		WideItem c;
		c.binary = it.slots[it.leaves + j];
		c.parent = (int)id;
		for (int k = 0; k < 8; k++)
			c.slots[k] = -1;
		c.leaves = c.inner = c.axis = 0;
		next_items[inner_offsets[i] + j] = c;
	}
}

/* ---- 6. primitive records into tree order ------------------------------------------------------- */
__global__ void permute_prims_kernel(const float4 *in, const uint32_t *order, uint32_t n, float4 *out, int32_t *obj_slot)
{
	const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= n)
		return;
	const uint32_t src = order[slot];
	const float4 c = in[4 * (size_t)src + 2];
#pragma unroll
	for (int k = 0; k < 4; k++)
		out[4 * (size_t)slot + k] = k == 2 ? c : in[4 * (size_t)src + k];
	obj_slot[__float_as_uint(c.w)] = (int32_t)slot;
}

struct Scratch {
	void *p = nullptr;
	~Scratch() { cudaFree(p); }
	template <typename T> cudaError_t alloc(T **out, size_t count)
	{
		const cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
		*out = static_cast<T *>(p);
		return e;
	}
};

inline unsigned blocks_for(size_t n)
{
	return static_cast<unsigned>((n + kThreads - 1) / kThreads);
}

} // namespace

#define BVH_TRY(call) \
	do { \
		const cudaError_t err__ = (call); \
		if (err__ != cudaSuccess) \
			return api_fail(CRT_ERR_CUDA, "device BVH build: %s: %s", #call, cudaGetErrorString(err__)); \
	} while (0)

int build_device_tree(const float4 *d_prims_in, uint32_t n, int32_t *d_obj_slot, float4 *d_prims_out, DeviceTree &out)
{
	out = DeviceTree();
	if (n == 0)
		return CRT_OK;
	cudaStream_t stream = nullptr;
	/* CRT_BVH_TIMING=1: milliseconds per stage on stderr (each stage is synchronised for it) */
	const bool timing = std::getenv("CRT_BVH_TIMING") != nullptr;
	auto wall = [] {
		timespec ts;
		clock_gettime(CLOCK_MONOTONIC, &ts);
		return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
	};
	double t_prev = wall();
	auto stage = [&](const char *name) {
		if (!timing)
			return;
		cudaStreamSynchronize(stream);
		const double t = wall();
		fprintf(stderr, "crt device BVH: %-28s %8.3f ms\n", name, t - t_prev);
		t_prev = t;
	};
	/* one allocation for all the scratch (a dozen cudaMalloc / cudaFree pairs cost more than the build) */
	const size_t max_items = std::max<size_t>(n, 1);
	/* the wide tree has at most n - 1 nodes (every wide node consumes an inner radix node), at least 1 */
	const size_t max_nodes = std::max<size_t>(n > 1 ? n - 1 : 1, 1);
	size_t sort_bytes = 0, scan_bytes = 0;
	BVH_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (uint64_t *)nullptr, (uint64_t *)nullptr, (uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 63, stream));
	BVH_TRY(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr, (int)(max_items + 1), stream));
	Box *boxes = nullptr, *node_boxes = nullptr;
	uint32_t *bounds = nullptr, *ids = nullptr, *ids2 = nullptr, *order = nullptr, *lc = nullptr, *ic = nullptr, *lo = nullptr, *io = nullptr;
	uint64_t *keys = nullptr, *keys2 = nullptr;
	int2 *children = nullptr;
	int *parent = nullptr;
	unsigned int *visits = nullptr;
	unsigned char *sort_tmp = nullptr, *scan_tmp = nullptr;
	WideItem *items[2] = { nullptr, nullptr };
	WideNode *nodes = nullptr;
	Scratch arena;
	unsigned char *base = nullptr;
	size_t total = 0;
	for (int pass = 0; pass < 2; pass++) {
		size_t off = 0;
		auto take = [&](auto **ptr, size_t count) {
			using T = std::remove_pointer_t<std::remove_pointer_t<decltype(ptr)>>;
			if (pass)
				*ptr = reinterpret_cast<T *>(base + off);
			off += (count * sizeof(T) + 255) / 256 * 256;
		};
		take(&boxes, n);
		take(&bounds, 6);
		take(&keys, n);
		take(&keys2, n);
		take(&ids, n);
		take(&ids2, n);
		take(&children, n);
		take(&parent, 2 * (size_t)n);
		take(&node_boxes, 2 * (size_t)n);
		take(&visits, n);
		take(&order, n);
		take(&sort_tmp, sort_bytes);
		take(&scan_tmp, scan_bytes);
		take(&items[0], max_items);
		take(&items[1], max_items);
		take(&lc, max_items + 1);
		take(&ic, max_items + 1);
		take(&lo, max_items + 1);
		take(&io, max_items + 1);
		take(&nodes, max_nodes);
		if (!pass) {
			total = off;
			BVH_TRY(arena.alloc(&base, total));
		}
	}
	stage("scratch allocation");
	/* 1, 2: boxes, centre bounds, codes */
	const uint32_t init_bounds[6] = { 0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u };
	BVH_TRY(cudaMemcpyAsync(bounds, init_bounds, sizeof init_bounds, cudaMemcpyHostToDevice, stream));
	prim_boxes_kernel<<<blocks_for(n), kThreads, 0, stream>>>(d_prims_in, n, boxes, bounds);
	morton_kernel<<<blocks_for(n), kThreads, 0, stream>>>(boxes, n, bounds, keys, ids);
	stage("boxes + Morton codes");
	/* sort (stable: equal codes keep the scene order) */
	BVH_TRY(cub::DeviceRadixSort::SortPairs(sort_tmp, sort_bytes, keys, keys2, ids, ids2, (int)n, 0, 63, stream));
	stage("radix sort");
	/* 3, 4: radix tree and its boxes */
	BVH_TRY(cudaMemsetAsync(visits, 0, n * sizeof(unsigned int), stream));
	BVH_TRY(cudaMemsetAsync(parent, 0xff, 2 * (size_t)n * sizeof(int), stream));
	if (n > 1)
		radix_tree_kernel<<<blocks_for(n - 1), kThreads, 0, stream>>>(keys2, (int)n, children, parent);
	if (timing)
		stage("radix tree");
	refit_kernel<<<blocks_for(n), kThreads, 0, stream>>>(boxes, ids2, (int)n, children, parent, node_boxes, visits);
	BVH_TRY(cudaGetLastError());
	stage("boxes bottom-up");

	/* 5: wide levels.  A level has at most as many nodes as there are inner radix nodes + 1. */
	WideItem root;
	root.binary = n > 1 ? 0 : 0; /* n == 1: node 0 is the only leaf (n - 1 + 0) */
	root.parent = -1;
	for (int k = 0; k < 8; k++)
		root.slots[k] = -1;
	root.leaves = root.inner = root.axis = 0;
	BVH_TRY(cudaMemcpyAsync(items[0], &root, sizeof root, cudaMemcpyHostToDevice, stream));
	uint32_t count = 1, level_first = 0, prim_first = 0, depth = 0;
	int cur = 0;
	while (count) {
		depth++;
		if (depth > kWideMaxDepth)
			return api_fail(CRT_ERR_UNSUPPORTED, "device BVH build: tree deeper than %u levels", kWideMaxDepth);
		collapse_plan_kernel<<<blocks_for(count), kThreads, 0, stream>>>(items[cur], count, (int)n, children, node_boxes, lc, ic);
		/* one element more than the level: the scans' last entries are the totals */
		BVH_TRY(cudaMemsetAsync(lc + count, 0, sizeof(uint32_t), stream));
		BVH_TRY(cudaMemsetAsync(ic + count, 0, sizeof(uint32_t), stream));
		BVH_TRY(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, lc, lo, (int)(count + 1), stream));
		BVH_TRY(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, ic, io, (int)(count + 1), stream));
		uint32_t totals[2];
		BVH_TRY(cudaMemcpyAsync(&totals[0], lo + count, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
		BVH_TRY(cudaMemcpyAsync(&totals[1], io + count, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
		BVH_TRY(cudaStreamSynchronize(stream));
		const uint32_t next_first = level_first + count;
		if ((size_t)next_first + totals[1] > max_nodes || (size_t)prim_first + totals[0] > n)
			return api_fail(CRT_ERR_CUDA, "device BVH build: inconsistent level (%u nodes, %u primitives)", next_first + totals[1], prim_first + totals[0]);
		collapse_emit_kernel<<<blocks_for(count), kThreads, 0, stream>>>(items[cur], count, level_first, (int)n, node_boxes, ids2, lo, io,
			prim_first, next_first, nodes, order, items[cur ^ 1]);
		level_first = next_first;
		prim_first += totals[0];
		count = totals[1];
		cur ^= 1;
	}
	stage("8-wide collapse");
	if (prim_first != n)
		return api_fail(CRT_ERR_CUDA, "device BVH build: %u of %u primitives placed", prim_first, n);
	/* 6 */
	permute_prims_kernel<<<blocks_for(n), kThreads, 0, stream>>>(d_prims_in, order, n, d_prims_out, d_obj_slot);
	BVH_TRY(cudaGetLastError());
	BVH_TRY(cudaStreamSynchronize(stream));
	stage("primitive records");
	out.num_nodes = level_first;
	out.max_depth = depth;
	if (out.num_nodes > kWideMaxNodes || n > kWideMaxPrims)
		return api_fail(CRT_ERR_UNSUPPORTED, "scene too large for the device tree (%u nodes, %u primitives)", out.num_nodes, n);
	/* the nodes in a buffer of their own size (the scratch one was sized for the worst case) */
	float4 *wide = nullptr;
	BVH_TRY(cudaMalloc(&wide, (size_t)out.num_nodes * sizeof(WideNode)));
	const cudaError_t e = cudaMemcpy(wide, nodes, (size_t)out.num_nodes * sizeof(WideNode), cudaMemcpyDeviceToDevice);
	if (e != cudaSuccess) {
		cudaFree(wide);
		return api_fail(CRT_ERR_CUDA, "device BVH build: copy of the nodes: %s", cudaGetErrorString(e));
	}
	out.wide = wide;
	return CRT_OK;
}

} // namespace crt
