/*
 * crt_post.cu -- the post-processing operations on the frame the trace loop
 * produced (raster + z-buffer), on the device: crt_postprocess[_device]() of
 * include/crt.h.  Replaces reference src/postprocess/postproc.c:36-188.
 *
 * brighten() and mist() are per-pixel maps (HBM-bound: 28 B per pixel).
 * depth_of_field() is a forward scatter in the reference: every source pixel
 * adds colour * alpha to all pixels of its circle of confusion that are not in
 * front of it, in source order, and the sums are normalised at the end.  Here
 * it is a gather: one thread per TARGET pixel walks the candidate sources in
 * ascending source index -- the order in which the reference's serial loop
 * reaches that target -- so every per-pixel float sum is formed in the same
 * order and the result is bit-identical to the (IEEE-compiled) reference,
 * with no atomics.  Candidates are pruned by the maximum radius of each source
 * row and of each 32-pixel row segment.
 *
 * All arithmetic the reference rounds once is written with __fmul_rn /
 * __fadd_rn / __fdiv_rn (no FMA contraction).
 */
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <mutex>

#include "crt_kernels.cuh"

using namespace crt;

namespace {

#define POST_TRY(expr)                                                                              \
	do {                                                                                        \
		cudaError_t e_ = (expr);                                                            \
		if (e_ != cudaSuccess)                                                              \
			return api_fail(CRT_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e_)); \
	} while (0)

constexpr int kSeg = 32;           /* source pixels per row segment (one warp of the prepare kernel) */
constexpr int kRows = 8;           /* rows per block */
constexpr int kMaxRadius = 32767;  /* radius * radius must fit the reference's int (postproc.c:115) */

struct PointOps {
	uint32_t brighten, mist, falloff;
	float factor;
	float start, inv_depth;
	float cr, cg, cb;
};

/* mist(), postproc.c:162-188; clamp() is calc.c:164-168 */
__device__ __forceinline__ void mist_pixel(const PointOps &o, float z, float &r, float &g, float &b)
{
	const float x = __fmul_rn(__fsub_rn(z, o.start), o.inv_depth);
	const float lo = x < 0.f ? 0.f : x;
	const float factor = lo > 1.f ? 1.f : lo;
	float opacity;
	if (o.falloff == CRT_FALLOFF_LINEAR)
		opacity = factor;
	else if (o.falloff == CRT_FALLOFF_QUADRATIC)
		opacity = __fmul_rn(factor, factor);
	else
		opacity = __fsqrt_rn(factor);
	const float keep = __fsub_rn(1.f, opacity);
	r = __fadd_rn(__fmul_rn(r, keep), __fmul_rn(o.cr, opacity));
	g = __fadd_rn(__fmul_rn(g, keep), __fmul_rn(o.cg, opacity));
	b = __fadd_rn(__fmul_rn(b, keep), __fmul_rn(o.cb, opacity));
}

__device__ __forceinline__ void point_pixel(const PointOps &o, float z, float &r, float &g, float &b)
{
	if (o.brighten) { /* brighten(), postproc.c:94-102 */
		r = __fmul_rn(r, o.factor);
		g = __fmul_rn(g, o.factor);
		b = __fmul_rn(b, o.factor);
	}
	if (o.mist)
		mist_pixel(o, z, r, g, b);
}

/* brighten and / or mist in one pass, four pixels per thread: three 16-byte
 * loads of colour, one of depth, three stores */
__global__ void __launch_bounds__(256) pointwise4_kernel(float4 *raster, const float4 *z, size_t quads, PointOps o)
{
	const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= quads)
		return;
	float4 a = raster[3 * i], b = raster[3 * i + 1], c = raster[3 * i + 2];
	const float4 d = o.mist ? __ldg(z + i) : make_float4(0.f, 0.f, 0.f, 0.f);
	point_pixel(o, d.x, a.x, a.y, a.z);
	point_pixel(o, d.y, a.w, b.x, b.y);
	point_pixel(o, d.z, b.z, b.w, c.x);
	point_pixel(o, d.w, c.y, c.z, c.w);
	raster[3 * i] = a;
	raster[3 * i + 1] = b;
	raster[3 * i + 2] = c;
}

/* the tail pixels, and buffers that are not 16-byte aligned */
__global__ void __launch_bounds__(256) pointwise1_kernel(float *raster, const float *z, size_t first, size_t pixels, PointOps o)
{
	const size_t i = first + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= pixels)
		return;
	float r = raster[3 * i], g = raster[3 * i + 1], b = raster[3 * i + 2];
	point_pixel(o, o.mist ? z[i] : 0.f, r, g, b);
	raster[3 * i] = r;
	raster[3 * i + 1] = g;
	raster[3 * i + 2] = b;
}

/* order-preserving float <-> uint for atomicMin / atomicMax */
__device__ __forceinline__ uint32_t ordered(float f)
{
	const uint32_t u = __float_as_uint(f);
	return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__host__ float unordered(uint32_t u)
{
	const uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
	float f;
	memcpy(&f, &v, 4);
	return f;
}

/* --dof-camera: the z range, postproc.c:53-61 (strict comparisons drop NaNs) */
__global__ void __launch_bounds__(256) z_range_kernel(const float *z, size_t pixels, uint32_t *range)
{
	float lo = FLT_MAX, hi = FLT_MIN;
	for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < pixels; i += (size_t)gridDim.x * blockDim.x) {
		const float v = __ldg(z + i);
		if (v < lo)
			lo = v;
		if (v > hi)
			hi = v;
	}
	for (int d = 16; d > 0; d >>= 1) {
		lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
		hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
	}
	if ((threadIdx.x & 31) == 0) {
		atomicMin(range, ordered(lo));
		atomicMax(range + 1, ordered(hi));
	}
}

struct DofParams {
	uint32_t w, h, nseg;
	float scale, bias;
	uint32_t brighten;
	float factor;
};

/* Per source pixel (postproc.c:112-119): radius of the circle of confusion,
 * alpha = min(1 / radius^2, 1), colour * alpha; and the largest radius per row
 * segment, per row and overall.  One warp = one 32-pixel row segment. */
__global__ void __launch_bounds__(kSeg *kRows) dof_prepare_kernel(const float *raster, const float *z, DofParams p, float4 *src,
	uint16_t *radius, uint16_t *seg_rmax, int *row_rmax, int *frame_rmax)
{
	const uint32_t x = blockIdx.x * kSeg + threadIdx.x, y = blockIdx.y * kRows + threadIdx.y;
	if (y >= p.h)
		return;
	int r = 0;
	if (x < p.w) {
		const size_t i = (size_t)y * p.w + x;
		const float depth = z[i];
		const float coc = fabsf(__fadd_rn(__fmul_rn(depth, p.scale), p.bias));
		const float half = __fmul_rn(coc, .5f);
		r = half < (float)kMaxRadius ? (int)half : kMaxRadius; /* NaN -> kMaxRadius; the host rejects the frame */
		const float alpha = fminf(__fdiv_rn(1.f, (float)(r * r)), 1.f);
		float cr = raster[3 * i], cg = raster[3 * i + 1], cb = raster[3 * i + 2];
		if (p.brighten) {
			cr = __fmul_rn(cr, p.factor);
			cg = __fmul_rn(cg, p.factor);
			cb = __fmul_rn(cb, p.factor);
		}
		src[i] = make_float4(__fmul_rn(cr, alpha), __fmul_rn(cg, alpha), __fmul_rn(cb, alpha), alpha);
		radius[i] = (uint16_t)r;
	}
	int m = r;
	for (int d = 16; d > 0; d >>= 1)
		m = max(m, __shfl_xor_sync(0xffffffffu, m, d));
	if (threadIdx.x == 0) {
		seg_rmax[(size_t)y * p.nseg + blockIdx.x] = (uint16_t)m;
		/* test before the atomic: 65 k warps of a 1080p frame would otherwise serialise on one word */
		if (m > *(volatile int *)(row_rmax + y))
			atomicMax(row_rmax + y, m);
		if (m > *(volatile int *)frame_rmax)
			atomicMax(frame_rmax, m);
	}
}

/* One thread per target pixel: the sources whose disc covers it
 * (postproc.c:121-151: |dx| <= radius, |dy| <= (int)sqrtf(radius^2 - dx^2),
 * clipped to the frame) and that are not behind it (:137, :148), in ascending
 * source index; then the normalisation of :156-157, and mist() if requested. */
__global__ void __launch_bounds__(kSeg *kRows) dof_gather_kernel(float *raster, const float *z, DofParams p, const float4 *src,
	const uint16_t *radius, const uint16_t *seg_rmax, const int *row_rmax, const int *frame_rmax, PointOps o)
{
	const int tx = blockIdx.x * kSeg + threadIdx.x, ty = blockIdx.y * kRows + threadIdx.y;
	if (tx >= (int)p.w || ty >= (int)p.h)
		return;
	const int R = *frame_rmax;
	const size_t t = (size_t)ty * p.w + tx;
	const float zt = z[t];
	float ar = 0.f, ag = 0.f, ab = 0.f, aa = 0.f;
	const int y0 = max(0, ty - R), y1 = min((int)p.h - 1, ty + R);
	const int sg0 = max(0, tx - R) / kSeg, sg1 = min((int)p.w - 1, tx + R) / kSeg;
	for (int sy = y0; sy <= y1; sy++) {
		const int dy = abs(ty - sy);
		if (__ldg(row_rmax + sy) < dy)
			continue;
		for (int sg = sg0; sg <= sg1; sg++) {
			const int s0 = sg * kSeg;
			const int reach = max(dy, max(s0 - tx, tx - (s0 + kSeg - 1)));
			const int rm = __ldg(seg_rmax + (size_t)sy * p.nseg + sg);
			if (rm < reach)
				continue;
			/* no source of this segment reaches further than rm pixels */
			const int x0 = max(s0, tx - rm), x1 = min(min(s0 + kSeg, (int)p.w), tx + rm + 1);
			const size_t row = (size_t)sy * p.w;
			for (int sx = x0; sx < x1; sx++) {
				const int r = __ldg(radius + row + sx), dx = abs(tx - sx);
				if (dx > r || dy > r)
					continue;
				const int hh = (int)__fsqrt_rn((float)(r * r - dx * dx));
				if (dy > hh || !(__ldg(z + row + sx) <= zt))
					continue;
				const float4 s = __ldg(src + row + sx);
				ar = __fadd_rn(ar, s.x);
				ag = __fadd_rn(ag, s.y);
				ab = __fadd_rn(ab, s.z);
				aa = __fadd_rn(aa, s.w);
			}
		}
	}
	const float inv = __fdiv_rn(1.f, aa);
	float cr = __fmul_rn(ar, inv), cg = __fmul_rn(ag, inv), cb = __fmul_rn(ab, inv);
	if (o.mist)
		mist_pixel(o, zt, cr, cg, cb);
	raster[3 * t] = cr;
	raster[3 * t + 1] = cg;
	raster[3 * t + 2] = cb;
}

struct Scratch {
	void *p = nullptr;
	~Scratch() { cudaFree(p); }
};

/* a timing event that is released on every return path */
struct Event {
	cudaEvent_t e = nullptr;
	~Event()
	{
		if (e)
			cudaEventDestroy(e);
	}
};

/* Scratch of the depth-of-field kernels, kept per device between calls (a
 * cudaMalloc / cudaFree pair per frame costs more than the kernels).  Calls
 * that reuse it are ordered on the device by an event: the next call's stream
 * waits for the previous call's last kernel. */
struct Workspace {
	void *p = nullptr;
	size_t bytes = 0;
	cudaEvent_t last_use = nullptr;
};
constexpr int kMaxDevices = 64;
Workspace g_workspace[kMaxDevices];
std::mutex g_workspace_mutex;

int workspace_acquire(int device, size_t bytes, cudaStream_t stream, void **out)
{
	Workspace &w = g_workspace[device];
	if (!w.last_use)
		POST_TRY(cudaEventCreateWithFlags(&w.last_use, cudaEventDisableTiming));
	if (w.bytes < bytes) {
		if (w.p) {
			POST_TRY(cudaEventSynchronize(w.last_use));
			POST_TRY(cudaFree(w.p));
			w.p = nullptr;
			w.bytes = 0;
		}
		POST_TRY(cudaMalloc(&w.p, bytes));
		w.bytes = bytes;
	} else {
		POST_TRY(cudaStreamWaitEvent(stream, w.last_use, 0));
	}
	*out = w.p;
	return CRT_OK;
}

size_t align_up(size_t v)
{
	return (v + 255) & ~(size_t)255;
}

int run_pointwise(float *d_raster, const float *d_z, size_t pixels, const PointOps &o, cudaStream_t stream, uint32_t &launches)
{
	const bool aligned = ((uintptr_t)d_raster % 16 == 0) && ((uintptr_t)d_z % 16 == 0);
	const size_t quads = aligned ? pixels / 4 : 0;
	if (quads) {
		pointwise4_kernel<<<(unsigned)((quads + 255) / 256), 256, 0, stream>>>(reinterpret_cast<float4 *>(d_raster),
			reinterpret_cast<const float4 *>(d_z), quads, o);
		launches++;
	}
	const size_t rest = pixels - 4 * quads;
	if (rest) {
		pointwise1_kernel<<<(unsigned)((rest + 255) / 256), 256, 0, stream>>>(d_raster, d_z, 4 * quads, pixels, o);
		launches++;
	}
	POST_TRY(cudaGetLastError());
	return CRT_OK;
}

} // namespace

/* ---- crt_postprocess_device_noalloc: the depth-of-field scratch as a fixed device array -----------
 * Every device that loads the library holds these CRT_POST_NOALLOC_SCRATCH_BYTES until the process
 * exits.  Host code must not pass the array's name to a kernel (on the host it is only a stand-in
 * for the symbol): noalloc_scratch() asks the runtime for its device address, once per device. */
static __device__ __align__(256) unsigned char g_post_noalloc_scratch[CRT_POST_NOALLOC_SCRATCH_BYTES];

namespace {

void *g_noalloc_ptr[kMaxDevices];
std::mutex g_noalloc_mutex;

/* `device` is current (api_ensure_device) */
int noalloc_scratch(int device, void **out)
{
	std::lock_guard<std::mutex> lock(g_noalloc_mutex);
	if (!g_noalloc_ptr[device]) {
		void *p = nullptr;
		POST_TRY(cudaGetSymbolAddress(&p, g_post_noalloc_scratch));
		if (!p)
			return api_fail(CRT_ERR_CUDA, "cudaGetSymbolAddress gave no address for the static scratch on device %d", device);
		g_noalloc_ptr[device] = p;
	}
	*out = g_noalloc_ptr[device];
	return CRT_OK;
}

} // namespace

extern "C" {

int crt_postprocess_device(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *d_raster, const float *d_z, void *cuda_stream, crt_post_stats *stats)
{
	if (!opts || !d_raster || !d_z)
		return api_fail(CRT_ERR_INVALID, "NULL argument");
	if (width == 0 || height == 0)
		return api_fail(CRT_ERR_INVALID, "empty frame %ux%u", width, height);
	if (opts->ops & ~(uint32_t)(CRT_POST_BRIGHTEN | CRT_POST_DOF | CRT_POST_DOF_CAMERA | CRT_POST_MIST))
		return api_fail(CRT_ERR_INVALID, "unknown post-processing operation bits 0x%x", opts->ops);
	if ((opts->ops & CRT_POST_MIST) && opts->mist_falloff > CRT_FALLOFF_INV_QUADRATIC)
		return api_fail(CRT_ERR_INVALID, "unrecognized falloff type %u", opts->mist_falloff);
	int rc = api_ensure_device(device);
	if (rc)
		return rc;
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const size_t pixels = (size_t)width * height;
	crt_post_stats st{};
	Event ev0, ev1;
	if (stats) {
		POST_TRY(cudaEventCreate(&ev0.e));
		POST_TRY(cudaEventCreate(&ev1.e));
		POST_TRY(cudaEventRecord(ev0.e, stream));
	}

	PointOps before{}, after{};
	before.brighten = (opts->ops & CRT_POST_BRIGHTEN) ? 1u : 0u;
	before.factor = opts->brighten;
	after.mist = (opts->ops & CRT_POST_MIST) ? 1u : 0u;
	after.falloff = opts->mist_falloff;
	after.start = opts->mist_start;
	after.inv_depth = 1.f / opts->mist_depth;
	after.cr = opts->mist_color[0];
	after.cg = opts->mist_color[1];
	after.cb = opts->mist_color[2];

	const bool dof = (opts->ops & (CRT_POST_DOF | CRT_POST_DOF_CAMERA)) != 0;
	if (!dof) {
		PointOps both = after;
		both.brighten = before.brighten;
		both.factor = before.factor;
		if (both.brighten || both.mist) {
			rc = run_pointwise(d_raster, d_z, pixels, both, stream, st.launches);
			if (rc)
				return rc;
		}
	} else {
		float scale = opts->dof_scale, bias = opts->dof_bias;
		if (!(opts->ops & CRT_POST_DOF)) {
			/* --dof-camera, postproc.c:49-65 */
			Scratch range;
			POST_TRY(cudaMalloc(&range.p, 2 * sizeof(uint32_t)));
			const uint32_t init[2] = { 0xffffffffu, 0u };
			uint32_t host_range[2];
			POST_TRY(cudaMemcpyAsync(range.p, init, sizeof init, cudaMemcpyHostToDevice, stream));
			z_range_kernel<<<1184, 256, 0, stream>>>(d_z, pixels, static_cast<uint32_t *>(range.p));
			st.launches++;
			POST_TRY(cudaMemcpyAsync(host_range, range.p, sizeof host_range, cudaMemcpyDeviceToHost, stream));
			POST_TRY(cudaStreamSynchronize(stream));
			const float z_min = unordered(host_range[0]), z_max = unordered(host_range[1]);
			const float aperture = opts->dof_aperture, focal_length = opts->dof_focal_length, plane_in_focus = opts->dof_plane_in_focus;
			scale = (aperture * focal_length * plane_in_focus * (z_max - z_min)) / ((plane_in_focus - focal_length) * z_min * z_max);
			bias = (aperture * focal_length * (z_min - plane_in_focus)) / ((plane_in_focus * focal_length) * z_min);
			if (!std::isfinite(scale) || !std::isfinite(bias))
				return api_fail(CRT_ERR_INVALID, "--dof-camera: scale %g / bias %g from the z range [%g, %g] (a frame with sky pixels has z_min = 0)",
					(double)scale, (double)bias, (double)z_min, (double)z_max);
		}
		st.dof_scale = scale;
		st.dof_bias = bias;
		DofParams p{};
		p.w = width;
		p.h = height;
		p.nseg = (width + kSeg - 1) / kSeg;
		p.scale = scale;
		p.bias = bias;
		p.brighten = before.brighten;
		p.factor = before.factor;
		if (device >= kMaxDevices)
			return api_fail(CRT_ERR_UNSUPPORTED, "device ordinal %d", device);
		std::lock_guard<std::mutex> lock(g_workspace_mutex);
		const size_t src_bytes = align_up(pixels * sizeof(float4)), radius_bytes = align_up(pixels * sizeof(uint16_t));
		const size_t seg_bytes = align_up((size_t)p.nseg * height * sizeof(uint16_t)), rows_bytes = align_up(((size_t)height + 1) * sizeof(int));
		void *base = nullptr;
		rc = workspace_acquire(device, src_bytes + radius_bytes + seg_bytes + rows_bytes, stream, &base);
		if (rc)
			return rc;
		char *cursor = static_cast<char *>(base);
		float4 *src = reinterpret_cast<float4 *>(cursor);
		cursor += src_bytes;
		uint16_t *radius = reinterpret_cast<uint16_t *>(cursor);
		cursor += radius_bytes;
		uint16_t *seg = reinterpret_cast<uint16_t *>(cursor);
		cursor += seg_bytes;
		int *row_rmax = reinterpret_cast<int *>(cursor), *frame_rmax = row_rmax + height;
		POST_TRY(cudaMemsetAsync(row_rmax, 0, ((size_t)height + 1) * sizeof(int), stream));
		const dim3 block(kSeg, kRows), grid(p.nseg, (height + kRows - 1) / kRows);
		dof_prepare_kernel<<<grid, block, 0, stream>>>(d_raster, d_z, p, src, radius, seg, row_rmax, frame_rmax);
		st.launches++;
		int rmax = 0;
		POST_TRY(cudaMemcpyAsync(&rmax, frame_rmax, sizeof rmax, cudaMemcpyDeviceToHost, stream));
		POST_TRY(cudaStreamSynchronize(stream));
		if (rmax >= kMaxRadius)
			return api_fail(CRT_ERR_UNSUPPORTED, "depth of field: a circle of confusion of %d pixels or more (scale %g, bias %g)", kMaxRadius,
				(double)scale, (double)bias);
		st.dof_max_radius = (uint32_t)rmax;
		dof_gather_kernel<<<grid, block, 0, stream>>>(d_raster, d_z, p, src, radius, seg, row_rmax, frame_rmax, after);
		st.launches++;
		POST_TRY(cudaGetLastError());
		POST_TRY(cudaEventRecord(g_workspace[device].last_use, stream));
	}
	if (stats) {
		float ms = 0.f;
		POST_TRY(cudaEventRecord(ev1.e, stream));
		POST_TRY(cudaEventSynchronize(ev1.e));
		POST_TRY(cudaEventElapsedTime(&ms, ev0.e, ev1.e));
		st.kernel_ms = ms;
		st.total_ms = ms;
		*stats = st;
	}
	return CRT_OK;
}

int crt_postprocess(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *raster, const float *z_buffer, crt_post_stats *stats)
{
	if (!opts || !raster || !z_buffer)
		return api_fail(CRT_ERR_INVALID, "NULL argument");
	if (width == 0 || height == 0)
		return api_fail(CRT_ERR_INVALID, "empty frame %ux%u", width, height);
	int rc = api_ensure_device(device);
	if (rc)
		return rc;
	const size_t pixels = (size_t)width * height;
	Event ev0, ev1;
	POST_TRY(cudaEventCreate(&ev0.e));
	POST_TRY(cudaEventCreate(&ev1.e));
	POST_TRY(cudaEventRecord(ev0.e, nullptr));
	Scratch d_raster, d_z;
	POST_TRY(cudaMalloc(&d_raster.p, pixels * 3 * sizeof(float)));
	POST_TRY(cudaMalloc(&d_z.p, pixels * sizeof(float)));
	POST_TRY(cudaMemcpy(d_raster.p, raster, pixels * 3 * sizeof(float), cudaMemcpyHostToDevice));
	POST_TRY(cudaMemcpy(d_z.p, z_buffer, pixels * sizeof(float), cudaMemcpyHostToDevice));
	crt_post_stats st{};
	rc = crt_postprocess_device(device, opts, width, height, static_cast<float *>(d_raster.p), static_cast<const float *>(d_z.p), nullptr, &st);
	if (rc)
		return rc;
	POST_TRY(cudaMemcpy(raster, d_raster.p, pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	float ms = 0.f;
	POST_TRY(cudaEventRecord(ev1.e, nullptr));
	POST_TRY(cudaEventSynchronize(ev1.e));
	POST_TRY(cudaEventElapsedTime(&ms, ev0.e, ev1.e));
	st.total_ms = ms;
	if (stats)
		*stats = st;
	return CRT_OK;
}


/* The depth-of-field path of crt_postprocess_device() (same kernels, same order, same bits) with its
 * scratch in g_post_noalloc_scratch: no cudaMalloc / cudaFree / event creation, nothing kept per call.
 * The array is shared by every call on the device and keeps the last call's contents; the per-row and
 * per-frame radius maxima are the only part that must start at zero and are cleared on `stream` first.
 * Calls that overlap on the GPU (different streams without an ordering between them) race on it. */
int crt_postprocess_device_noalloc(int device, const crt_post_opts *opts, uint32_t width, uint32_t height,
	float *d_raster, const float *d_z, void *cuda_stream)
{
	if (!opts || !d_raster || !d_z)
		return api_fail(CRT_ERR_INVALID, "NULL argument");
	if (width == 0 || height == 0)
		return api_fail(CRT_ERR_INVALID, "empty frame %ux%u", width, height);
	if (opts->ops & ~(uint32_t)(CRT_POST_BRIGHTEN | CRT_POST_DOF | CRT_POST_DOF_CAMERA | CRT_POST_MIST))
		return api_fail(CRT_ERR_INVALID, "unknown post-processing operation bits 0x%x", opts->ops);
	if (!(opts->ops & CRT_POST_DOF))
		return api_fail(CRT_ERR_INVALID, "crt_postprocess_device_noalloc is the depth-of-field path: CRT_POST_DOF (explicit scale and bias) "
			"must be set; the other operations need no scratch, use crt_postprocess_device");
	if ((opts->ops & CRT_POST_MIST) && opts->mist_falloff > CRT_FALLOFF_INV_QUADRATIC)
		return api_fail(CRT_ERR_INVALID, "unrecognized falloff type %u", opts->mist_falloff);
	if (device < 0 || device >= kMaxDevices)
		return api_fail(CRT_ERR_UNSUPPORTED, "device ordinal %d", device);
	const uint64_t pixels = (uint64_t)width * height;
	const uint64_t nseg = ((uint64_t)width + kSeg - 1) / kSeg;
	if (pixels > CRT_POST_NOALLOC_MAX_PIXELS)
		return api_fail(CRT_ERR_UNSUPPORTED, "frame %ux%u = %llu pixels exceeds the static scratch's limit of %u pixels (CRT_POST_NOALLOC_MAX_PIXELS); "
			"use crt_postprocess_device", width, height, (unsigned long long)pixels, (unsigned)CRT_POST_NOALLOC_MAX_PIXELS);
	const size_t src_bytes = align_up((size_t)pixels * sizeof(float4)), radius_bytes = align_up((size_t)pixels * sizeof(uint16_t));
	const size_t seg_bytes = align_up((size_t)(nseg * height) * sizeof(uint16_t)), rows_bytes = align_up(((size_t)height + 1) * sizeof(int));
	const size_t need = src_bytes + radius_bytes + seg_bytes + rows_bytes;
	if (need > (size_t)CRT_POST_NOALLOC_SCRATCH_BYTES)
		return api_fail(CRT_ERR_UNSUPPORTED, "frame %ux%u needs %zu bytes of scratch, the static array has %zu (CRT_POST_NOALLOC_SCRATCH_BYTES); "
			"use crt_postprocess_device", width, height, need, (size_t)CRT_POST_NOALLOC_SCRATCH_BYTES);
	int rc = api_ensure_device(device);
	if (rc)
		return rc;
	void *base = nullptr;
	rc = noalloc_scratch(device, &base);
	if (rc)
		return rc;
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);

	PointOps after{};
	after.mist = (opts->ops & CRT_POST_MIST) ? 1u : 0u;
	after.falloff = opts->mist_falloff;
	after.start = opts->mist_start;
	after.inv_depth = 1.f / opts->mist_depth;
	after.cr = opts->mist_color[0];
	after.cg = opts->mist_color[1];
	after.cb = opts->mist_color[2];
	DofParams p{};
	p.w = width;
	p.h = height;
	p.nseg = (uint32_t)nseg;
	p.scale = opts->dof_scale;
	p.bias = opts->dof_bias;
	p.brighten = (opts->ops & CRT_POST_BRIGHTEN) ? 1u : 0u;
	p.factor = opts->brighten;

	char *cursor = static_cast<char *>(base);
	float4 *src = reinterpret_cast<float4 *>(cursor);
	cursor += src_bytes;
	uint16_t *radius = reinterpret_cast<uint16_t *>(cursor);
	cursor += radius_bytes;
	uint16_t *seg = reinterpret_cast<uint16_t *>(cursor);
	cursor += seg_bytes;
	int *row_rmax = reinterpret_cast<int *>(cursor), *frame_rmax = row_rmax + height;
	POST_TRY(cudaMemsetAsync(row_rmax, 0, ((size_t)height + 1) * sizeof(int), stream));
	const dim3 block(kSeg, kRows), grid(p.nseg, (height + kRows - 1) / kRows);
	dof_prepare_kernel<<<grid, block, 0, stream>>>(d_raster, d_z, p, src, radius, seg, row_rmax, frame_rmax);
	POST_TRY(cudaGetLastError());
	int rmax = 0;
	POST_TRY(cudaMemcpyAsync(&rmax, frame_rmax, sizeof rmax, cudaMemcpyDeviceToHost, stream));
	POST_TRY(cudaStreamSynchronize(stream));
	if (rmax >= kMaxRadius)
		return api_fail(CRT_ERR_UNSUPPORTED, "depth of field: a circle of confusion of %d pixels or more (scale %g, bias %g)", kMaxRadius,
			(double)p.scale, (double)p.bias);
	dof_gather_kernel<<<grid, block, 0, stream>>>(d_raster, d_z, p, src, radius, seg, row_rmax, frame_rmax, after);
	POST_TRY(cudaGetLastError());
	return CRT_OK;
}

} // extern "C"
