/*
 * crt_kernels.cu -- the sm_100a kernels of the trace loop.
 *
 * render_kernel replaces the body of the reference's render()/cast_ray()
 * (reference src/raytracer/render.c:136-368).  Design:
 *
 *   - persistent warps; each warp pulls 8x4-pixel tiles from a global atomic
 *     counter (dynamic load balance; tiles cost 0 .. thousands of shadow rays);
 *   - lane = pixel for everything that is one ray per cast_ray() call
 *     (closest hit, shading set-up, spawning reflection / refraction / GI);
 *     the reference's recursion becomes a per-lane stack of pending-ray records;
 *   - the >99 % of rays that are shadow rays are processed
 *     warp-cooperatively: the (shading point x light sample) index space of
 *     the warp's current shading points is flattened into a work list the 32
 *     lanes pull from (dynamic refill), the any-hit walk is a sequence of
 *     node / primitive unit steps scheduled by warp vote, and the per-point
 *     sums are formed in a fixed order from per-ray slots (deterministic
 *     image, no float atomics).
 */
#include "crt_kernels.cuh"

namespace crt {

namespace {

constexpr unsigned kFull = 0xffffffffu;

/* shading-point entry fields kept in shared memory, [field][lane] per warp */
enum EntryField {
	E_PX, E_PY, E_PZ,    /* hit point */
	E_NX, E_NY, E_NZ,    /* normal */
	E_DX, E_DY, E_DZ,    /* incoming ray direction */
	E_TR, E_TG, E_TB,    /* texture colour at the hit point (hoisted) */
	E_KSR, E_KSG, E_KSB, /* ks */
	E_SHIN,
	E_OBJ,               /* hit object (int bits) */
	E_NODE,              /* RNG node id (uint bits) */
	E_SKIP,              /* index of the emitter that IS the hit object, or -1 (int bits) */
	E_FIELDS
};

struct WarpShared {
	float entry[E_FIELDS][32];
};

/* pending-ray record kinds, stored in the top bits of word 9 */
constexpr uint32_t REC_GI = 0x80000000u;

struct Pending {
	float ox, oy, oz; /* ray origin             | GI: hit point P */
	float dx, dy, dz; /* ray direction          | GI: normal N */
	float kr, kg, kb; /* throughput             | GI: parent weight W */
	uint32_t remaining; /* remaining bounces    | GI: REC_GI | next sample k */
	int32_t inside;     /* inside object or -1  | GI: sample count n */
	uint32_t node;      /* RNG node id          | GI: parent's node id */
	uint32_t aux;       /*                      | GI: first draw index | flip << 31 */
};

__device__ __forceinline__ float *pending_slot(const RenderParams &p, uint32_t warp_global, uint32_t level, uint32_t lane)
{
	return p.pending + ((size_t)warp_global * p.pending_depth + level) * (kPendingWords * 32) + lane;
}

__device__ __forceinline__ void pending_store(float *s, const Pending &r)
{
	s[0 * 32] = r.ox; s[1 * 32] = r.oy; s[2 * 32] = r.oz;
	s[3 * 32] = r.dx; s[4 * 32] = r.dy; s[5 * 32] = r.dz;
	s[6 * 32] = r.kr; s[7 * 32] = r.kg; s[8 * 32] = r.kb;
	s[9 * 32] = __uint_as_float(r.remaining);
	s[10 * 32] = __int_as_float(r.inside);
	s[11 * 32] = __uint_as_float(r.node);
	s[12 * 32] = __uint_as_float(r.aux);
}

__device__ __forceinline__ void pending_load(const float *s, Pending &r)
{
	r.ox = s[0 * 32]; r.oy = s[1 * 32]; r.oz = s[2 * 32];
	r.dx = s[3 * 32]; r.dy = s[4 * 32]; r.dz = s[5 * 32];
	r.kr = s[6 * 32]; r.kg = s[7 * 32]; r.kb = s[8 * 32];
	r.remaining = __float_as_uint(s[9 * 32]);
	r.inside = __float_as_int(s[10 * 32]);
	r.node = __float_as_uint(s[11 * 32]);
	r.aux = __float_as_uint(s[12 * 32]);
}

/* position of the n-th (0-based) set bit of mask */
__device__ __forceinline__ uint32_t nth_set_bit(uint32_t mask, uint32_t n)
{
	uint32_t pos = 0;
#pragma unroll
	for (uint32_t w = 16; w; w >>= 1) {
		const uint32_t c = (uint32_t)__popc(mask & (((1u << w) - 1u) << pos));
		if (n >= c) {
			n -= c;
			pos += w;
		}
	}
	return pos;
}

/* which emitter and which of its samples light index li is; false: that emitter is the hit object itself (no ray) */
__device__ __forceinline__ bool light_of_index(const DScene &sc, uint32_t li, int skip_e, uint32_t &ei, uint32_t &l)
{
	ei = 0;
	while (ei + 1u < sc.num_emitters && li >= sc.emitters[ei + 1u].first_light)
		ei++;
	/* draw index counts only the emitters that are sampled, in order */
	l = li;
	if (skip_e >= 0 && (int)ei > skip_e)
		l -= sc.emitters[skip_e].num_lights;
	return (int)ei != skip_e;
}

/* eye-segment attenuation applied to every cast_ray result (render.c:292-301) */
__device__ __forceinline__ float eye_attenuation(const crt_render_opts &o, float t)
{
	if (o.attenuation == CRT_ATTEN_LINEAR)
		return 1.f / (o.attenuation_offset + t);
	if (o.attenuation == CRT_ATTEN_SQUARE) {
		const float s = o.attenuation_offset + t;
		return 1.f / (s * s);
	}
	return 1.f;
}

/* k-th GI direction of a shading point: rotation taking +Y to N and a uniform
 * hemisphere sample weighted by cos (render.c:240-284) */
__device__ __forceinline__ void gi_direction(float nx, float ny, float nz, bool flip, float u0, float u1,
	float &dx, float &dy, float &dz)
{
	const float inclination = acosf(u0 * 2.f - 1.f);
	const float azimuth = u1 * kPi;
	float si, ci, sa, ca;
	sincosf(inclination, &si, &ci);
	sincosf(azimuth, &sa, &ca);
	const float sx = ca * si, sy = sa * si, sz = ci;
	if (flip) {
		dx = sx;
		dy = -sy;
		dz = -sz;
		return;
	}
	const float mul = 1.f / (1.f + ny);
	dx = (1.f - nx * nx * mul) * sx + nx * sy + (-nx * nz * mul) * sz;
	dy = -nx * sx + (1.f - (nx * nx + nz * nz) * mul) * sy - nz * sz;
	dz = (-nx * nz * mul) * sx + nz * sy + (1.f - nz * nz * mul) * sz;
}

/* Thin lens: moves the primary ray's origin to a point of the lens disk (uniform, two draws of the
 * ray's own stream far above those of the lights and the GI samples) and aims it at the point where
 * the pinhole ray crosses the plane in focus.  vx x vy is the viewing direction (camera.c:30-32). */
__device__ __forceinline__ void lens_ray(const crt_render_opts &o, Pending &r)
{
	float fx = o.vx[1] * o.vy[2] - o.vx[2] * o.vy[1], fy = o.vx[2] * o.vy[0] - o.vx[0] * o.vy[2], fz = o.vx[0] * o.vy[1] - o.vx[1] * o.vy[0];
	const float finv = rsqrtf(fx * fx + fy * fy + fz * fz);
	fx *= finv; fy *= finv; fz *= finv;
	float along = r.dx * fx + r.dy * fy + r.dz * fz;
	if (!(fabsf(along) > 1e-6f))
		return;
	const float ft = o.focus_distance / along;
	const float px = r.ox + r.dx * ft, py = r.oy + r.dy * ft, pz = r.oz + r.dz * ft;
	const float xinv = rsqrtf(o.vx[0] * o.vx[0] + o.vx[1] * o.vx[1] + o.vx[2] * o.vx[2]);
	const float yinv = rsqrtf(o.vy[0] * o.vy[0] + o.vy[1] * o.vy[1] + o.vy[2] * o.vy[2]);
	const float rad = o.lens_radius * sqrtf(crt_rand(r.node, 0xfffffff0u)), phi = 6.2831853f * crt_rand(r.node, 0xfffffff1u);
	float sn, cs;
	sincosf(phi, &sn, &cs);
	const float a = rad * cs * xinv, b = rad * sn * yinv;
	r.ox += a * o.vx[0] + b * o.vy[0];
	r.oy += a * o.vx[1] + b * o.vy[1];
	r.oz += a * o.vx[2] + b * o.vy[2];
	const float dx = px - r.ox, dy = py - r.oy, dz = pz - r.oz;
	const float dinv = rsqrtf(dx * dx + dy * dy + dz * dz);
	r.dx = dx * dinv; r.dy = dy * dinv; r.dz = dz * dinv;
}

template <bool COUNT>
__device__ __forceinline__ void flush_counters(const RenderParams &p, Counters &cn)
{
	if (!COUNT)
		return;
	uint32_t *c = reinterpret_cast<uint32_t *>(&cn);
	for (int i = 0; i < kNumCounters; i++) {
		unsigned long long v = c[i];
		for (int off = 16; off; off >>= 1)
			v += __shfl_down_sync(kFull, v, off);
		if ((threadIdx.x & 31) == 0 && v)
			atomicAdd(p.counters + i, v);
		c[i] = 0;
	}
}

__device__ __forceinline__ unsigned long long device_clock_ns()
{
	unsigned long long t;
	asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
	return t;
}

/* EXT: the instance with the in-tracer lens and fog (crt_render_opts.lens_* / fog_*); the frames of the
 * reference's own options run the instance without a trace of them */
template <bool COUNT, bool EXT>
__global__ void __launch_bounds__(kBlockThreads, CRT_MIN_BLOCKS) render_kernel(const RenderParams p)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	WarpShared *shared = reinterpret_cast<WarpShared *>(smem_raw);
	uint32_t *wstacks = reinterpret_cast<uint32_t *>(smem_raw + sizeof(WarpShared) * kWarpsPerBlock);
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t warp_local = threadIdx.x >> 5;
	const uint32_t warp_global = blockIdx.x * kWarpsPerBlock + warp_local;
	WarpShared &ws = shared[warp_local];
	const DScene &sc = p.sc;
	const crt_render_opts &o = p.o;
	Counters cn = {};
	/* the lane's traversal stacks (closest-hit walk, then the shadow walks): shared memory, [slot][lane] */
	const SharedStacks stacks(wstacks + warp_local * (kLaneStackWords * 32) + lane,
		p.deep ? p.deep + (size_t)warp_global * p.deep_depth * 32 + lane : nullptr);
	const unsigned long long warp_t0 = COUNT ? device_clock_ns() : 0ull;

	for (;;) {
		/* ---- next unit of this shard -------------------------------------- */
		/* work unit: p.fine_pixels neighbouring pixels of an 8x4 tile, then -- towards the end of
		 * the queue -- 4, 2 and 1 of them: eight pixels through the glass mesh are 20 to 100 ms of
		 * one warp, 16 times the mean, and a unit that starts late outlives everything handed
		 * out after it (1/8 of a 4K frame measured a 22 ms tail on 104 ms with 8-pixel units up
		 * to the last 64 pixels per warp; crt_api.cu sizes the phases).
		 * A pixel is traced by 1 << p.gshift neighbouring lanes: with path-traced GI the lanes
		 * of a pixel share its GI samples (sample k goes to lane k mod G; lane 0 of the pixel
		 * also owns the direct light and the specular chain), their partial sums are added in
		 * lane order at the end.  G is a property of the frame, not of the unit, so a pixel's
		 * value does not depend on the unit it is traced in or on the devices that took part. */
		uint32_t unit = 0;
		if (lane == 0)
			unit = p.system_counter ? atomicAdd_system(p.tile_counter, 1u) : atomicAdd(p.tile_counter, 1u);
		unit = __shfl_sync(kFull, unit, 0);
		if (unit >= p.num_units) {
			if (COUNT && lane == 0) { /* crt_stats.queue_ms / warp_busy_ms / last_warp_ms */
				const unsigned long long t = device_clock_ns();
				atomicMax(p.counters + kTimeStart, ~warp_t0);
				atomicMax(p.counters + kTimeDry, ~t);
				atomicMax(p.counters + kTimeEnd, t);
				atomicAdd(p.counters + kTimeBusy, t - warp_t0);
			}
			break;
		}
		/* lane's place among the lanes of its pixel: recomputed where it is needed, not kept in a register */
		const auto lane_group = [&]() { return lane & ((1u << p.gshift) - 1u); };
		/* which phase of the queue: units shrink from fine_pixels to one pixel towards its end */
		uint32_t ph = 0;
#pragma unroll
		for (uint32_t i = 1; i < (uint32_t)kPhases; i++)
			ph += unit >= p.phase_unit[i] ? 1u : 0u;
		const uint32_t upx = p.fine_pixels >> ph, per_tile = (uint32_t)(kTileW * kTileH) / upx;
		const uint32_t v = unit - p.phase_unit[ph];
		const uint32_t local_tile = p.phase_tile[ph] + v / per_tile;
		const uint32_t pix_lane = (v % per_tile) * upx + (lane >> p.gshift);
		const bool lane_on = (lane >> p.gshift) < upx;
		const uint32_t chunk = (local_tile / kChunkTiles) * o.shard_count + o.shard_index;
		const uint32_t tile = chunk * kChunkTiles + local_tile % kChunkTiles;
		const uint32_t tx = tile % p.tiles_x, ty = tile / p.tiles_x;
		const uint32_t px = tx * kTileW + (pix_lane & 7u), py = ty * kTileH + (pix_lane >> 3);
		const bool in_image = lane_on && ty < p.tiles_y && px < o.width && py < o.height;
		const uint32_t pixel = py * o.width + px;

		float acc_r = 0.f, acc_g = 0.f, acc_b = 0.f, z_value = 0.f;
		uint32_t sp = 0;
		/* thin lens (crt_render_opts.lens_*, an extension): a pixel is the mean of lens_n primary rays, traced one
		 * after the other; lens sample 0 is the reference's pinhole ray */
		const uint32_t lens_n = EXT && o.lens_radius > 0.f && o.lens_samples > 1u ? o.lens_samples : 1u;
		for (uint32_t ls = 0; ls < lens_n; ls++) {
		bool primary = true;

		if (in_image) {
			/* primary ray: reference render.c:352-363 (position table holds the running sum) */
			Pending r;
			r.ox = o.cam_pos[0]; r.oy = o.cam_pos[1]; r.oz = o.cam_pos[2];
			const float vx = __fsub_rn(p.row_pos[3 * (size_t)pixel + 0], r.ox);
			const float vy = __fsub_rn(p.row_pos[3 * (size_t)pixel + 1], r.oy);
			const float vz = __fsub_rn(p.row_pos[3 * (size_t)pixel + 2], r.oz);
			const float len2 = __fadd_rn(__fadd_rn(__fmul_rn(vx, vx), __fmul_rn(vy, vy)), __fmul_rn(vz, vz));
			const float inv = __fdiv_rn(1.f, __fsqrt_rn(len2));
			r.dx = __fmul_rn(vx, inv); r.dy = __fmul_rn(vy, inv); r.dz = __fmul_rn(vz, inv);
			r.kr = r.kg = r.kb = 1.f;
			r.remaining = o.max_bounces;
			r.inside = -1;
			r.node = crt_node_root(EXT ? o.seed + 0x9e3779b9u * ls : o.seed, pixel);
			r.aux = 0;
			if (EXT && ls > 0u)
				lens_ray(o, r);
			pending_store(pending_slot(p, warp_global, sp++, lane), r);
		}

		/* ---- one iteration = one cast_ray() level for every lane ---------- */
		for (;;) {
			bool have = false;
			Pending r;
			if (sp > 0) {
				float *slot = pending_slot(p, warp_global, sp - 1, lane);
				pending_load(slot, r);
				have = true;
				if (r.remaining & REC_GI) {
					/* GI generator: emit sample k, keep the record while samples remain */
					const uint32_t k = r.remaining & ~REC_GI, n = (uint32_t)r.inside;
					const uint32_t draw = (r.aux & 0x7fffffffu) + 2u * k;
					float dx, dy, dz;
					gi_direction(r.dx, r.dy, r.dz, (r.aux >> 31) != 0u, crt_rand(r.node, draw), crt_rand(r.node, draw + 1u), dx, dy, dz);
					const float w = (1.f / (float)n) * dot3(r.dx, r.dy, r.dz, dx, dy, dz);
					/* the lanes of a pixel take the samples of the primary hit in turn (deeper generators have n = 1) */
					const uint32_t kn = k + (1u << p.gshift);
					if (kn < n)
						slot[9 * 32] = __uint_as_float(REC_GI | kn);
					else
						sp--;
					r.dx = dx; r.dy = dy; r.dz = dz;
					r.kr *= w; r.kg *= w; r.kb *= w;
					r.node = crt_node_child(r.node, k);
					r.remaining = 0;
					r.inside = -1;
				} else {
					sp--;
				}
			}
			if (!__any_sync(kFull, have))
				break;

			/* ---- closest hit (render.c:143-151) ---------------------------- */
			Ray ray = { r.ox, r.oy, r.oz, r.dx, r.dy, r.dz };
			Hit hit;
			hit.object = -1;
			hit.t = 0.f;
			hit.slot = 0;
			/* a pixel's other lanes take the primary hit from its first lane */
			{
			const bool ghost = primary && lane_group() != 0u;
			if (have && !ghost) {
				Cnt<COUNT>::add(cn.closest);
				/* a NaN direction (total internal reflection) hits nothing */
				if (ray.dx == ray.dx && ray.dy == ray.dy && ray.dz == ray.dz) {
					bool found = false;
					if (r.inside >= 0)
						found = hit_object<COUNT>(sc, r.inside, ray, hit, cn);
					if (!found)
						hit = trace_closest<COUNT>(sc, ray, stacks, cn);
				}
			}
			}
			if (primary && p.gshift) { /* warp-uniform */
				const uint32_t lead = lane - lane_group();
				hit.t = __shfl_sync(kFull, hit.t, lead);
				hit.object = __shfl_sync(kFull, hit.object, lead);
				hit.slot = __shfl_sync(kFull, hit.slot, lead);
			}
			const bool is_hit = have && hit.object >= 0;
			if (EXT && o.fog_density > 0.f && have && !is_hit && ray.dx == ray.dx && !(primary && lane_group() != 0u)) {
				/* a segment that leaves the scene ends in fog */
				acc_r += o.fog_color[0] * r.kr; acc_g += o.fog_color[1] * r.kg; acc_b += o.fog_color[2] * r.kb;
			}

			/* ---- shading set-up (render.c:155-168) ------------------------- */
			float nx = 0.f, ny = 0.f, nz = 0.f, bdot = 0.f, eps = 0.f;
			float wr = 0.f, wg = 0.f, wb = 0.f;
			float hx = 0.f, hy = 0.f, hz = 0.f;
			uint32_t mat_index = 0;
			bool outside = false, lit = false;
			int skip = -1;
			if (is_hit) {
				const bool ghost = primary && lane_group() != 0u;
				if (!ghost)
					Cnt<COUNT>::add(cn.shaded);
				hit_normal(sc, hit, ray, nx, ny, nz, mat_index, eps);
				hx = __fadd_rn(__fmul_rn(ray.dx, hit.t), ray.ox);
				hy = __fadd_rn(__fmul_rn(ray.dy, hit.t), ray.oy);
				hz = __fadd_rn(__fmul_rn(ray.dz, hit.t), ray.oz);
				const crt_material *m = sc.materials + mat_index;
				bdot = __fadd_rn(__fadd_rn(__fmul_rn(nx, ray.dx), __fmul_rn(ny, ray.dy)), __fmul_rn(nz, ray.dz));
				outside = signbit(bdot);
				const float att = eye_attenuation(o, hit.t);
				wr = r.kr * att; wg = r.kg * att; wb = r.kb * att;
				if (EXT && o.fog_density > 0.f) { /* in-tracer fog (extension): the segment keeps T of what it carries */
					const float T = __expf(-o.fog_density * hit.t);
					if (!ghost) {
						acc_r += o.fog_color[0] * r.kr * (1.f - T);
						acc_g += o.fog_color[1] * r.kg * (1.f - T);
						acc_b += o.fog_color[2] * r.kb * (1.f - T);
					}
					wr *= T; wg *= T; wb *= T;
				}
				/* emission + ambient term (render.c:164, :233-237) */
				float cr = __ldg(&m->ke[0]), cg = __ldg(&m->ke[1]), cb = __ldg(&m->ke[2]);
				if (o.gi_model == CRT_GI_AMBIENT) {
					cr += __ldg(&m->ka[0]) * sc.ambient[0];
					cg += __ldg(&m->ka[1]) * sc.ambient[1];
					cb += __ldg(&m->ka[2]) * sc.ambient[2];
				}
				if (!ghost) {
					acc_r += cr * wr; acc_g += cg * wg; acc_b += cb * wb;
				}
				if (primary && (!EXT || ls == 0u) && !ghost && o.max_bounces > 0u)
					z_value = hit.t; /* render.c:304-305, :342 */
				/* is the hit object itself an emitter? it is skipped as a light (render.c:172) */
				uint32_t lights = sc.total_lights;
				for (uint32_t e = 0; e < sc.num_emitters; e++)
					if (sc.emitters[e].object == hit.object) {
						skip = (int)e;
						lights -= sc.emitters[e].num_lights;
					}
				lit = outside && lights > 0u && !ghost;
				if (lit) {
					float tr, tg, tb;
					texture_color(m, hx, hy, hz, tr, tg, tb);
					ws.entry[E_PX][lane] = hx; ws.entry[E_PY][lane] = hy; ws.entry[E_PZ][lane] = hz;
					ws.entry[E_NX][lane] = nx; ws.entry[E_NY][lane] = ny; ws.entry[E_NZ][lane] = nz;
					ws.entry[E_DX][lane] = ray.dx; ws.entry[E_DY][lane] = ray.dy; ws.entry[E_DZ][lane] = ray.dz;
					ws.entry[E_TR][lane] = tr; ws.entry[E_TG][lane] = tg; ws.entry[E_TB][lane] = tb;
					ws.entry[E_KSR][lane] = __ldg(&m->ks[0]);
					ws.entry[E_KSG][lane] = __ldg(&m->ks[1]);
					ws.entry[E_KSB][lane] = __ldg(&m->ks[2]);
					ws.entry[E_SHIN][lane] = __ldg(&m->shininess);
					ws.entry[E_OBJ][lane] = __int_as_float(hit.object);
					ws.entry[E_NODE][lane] = __uint_as_float(r.node);
					ws.entry[E_SKIP][lane] = __int_as_float(skip);
				}
			}
			const unsigned lit_mask = __ballot_sync(kFull, lit);
			__syncwarp();

			/* ---- direct light (render.c:170-229) --------------------------------
			 * The (lit shading point x light sample) index space of the warp is cut
			 * into chunks of <= kSlots shadow rays.  Inside a chunk the lanes pull
			 * ray indices from a warp counter: a lane whose any-hit walk ends takes
			 * the next ray as soon as fewer than kRefill lanes are still walking, so
			 * long walks (through the glass mesh) do not idle the lanes of short
			 * ones.  The walk itself is split into unit steps the warp votes on
			 * (crt_device.cuh, ShadowWalk).  Every ray writes its contribution to its
			 * own slot and the slots of a point are summed in a fixed order
			 * afterwards: the image does not depend on which lane traced which ray,
			 * so frames are bit-reproducible whatever the schedule or GPU count. --- */
			float sum_r = 0.f, sum_g = 0.f, sum_b = 0.f;
			const uint32_t npts = (uint32_t)__popc(lit_mask);
			const uint32_t L = sc.total_lights;
			if (npts > 0u && L > 0u) {
				/* contribution slots: this warp's private scratch in global memory (stays in L2) */
				float *slot_r = p.slots + (size_t)warp_global * (3 * kSlots), *slot_g = slot_r + kSlots, *slot_b = slot_g + kSlots;
				const uint32_t pts_per_chunk = (uint32_t)kSlots / L; /* 0: a point spans several chunks */
				uint32_t p0 = 0, s0 = 0;
				while (p0 < npts) {
					uint32_t cpts, cL;
					if (pts_per_chunk) {
						cpts = min(pts_per_chunk, npts - p0);
						cL = L;
					} else {
						cpts = 1u;
						cL = min((uint32_t)kSlots, L - s0);
					}
					const uint32_t nwork = cpts * cL;
					uint32_t next = 0;
					/* per-lane shadow-ray state */
					bool have = false;
					uint32_t w = 0, e_lane = 0;
					Ray sr = { 0.f, 0.f, 0.f, 0.f, 0.f, 1.f };
					WideRay rw = make_wide_ray(sr);
					float dist = 0.f, d2 = 0.f, ir = 0.f, ig = 0.f, ib = 0.f;
					int em_object = -1;
					bool blocked = false;
					ShadowWalk sw;
					shadow_walk_reset(sw);
					for (;;) {
						/* -- service: finished lanes shade and store, idle lanes fetch -- */
						const bool walking = have && shadow_walk_busy(sw);
						if (have && !walking) {
							float cr = 0.f, cg = 0.f, cb = 0.f;
							if (!blocked) {
								const float Nx = ws.entry[E_NX][e_lane], Ny = ws.entry[E_NY][e_lane], Nz = ws.entry[E_NZ][e_lane];
								const float a = dot3(sr.dx, sr.dy, sr.dz, Nx, Ny, Nz);
								/* light-segment attenuation (render.c:191-200) */
								if (o.attenuation == CRT_ATTEN_LINEAR) {
									const float s = 1.f / (o.attenuation_offset + dist);
									ir *= s; ig *= s; ib *= s;
								} else if (o.attenuation == CRT_ATTEN_SQUARE) {
									const float s = 1.f / (o.attenuation_offset + d2);
									ir *= s; ig *= s; ib *= s;
								}
								const float Dx = ws.entry[E_DX][e_lane], Dy = ws.entry[E_DY][e_lane], Dz = ws.entry[E_DZ][e_lane];
								float smul;
								if (o.reflection_model == CRT_REFLECTION_BLINN) {
									/* H = normalize(D - L), s = -N.H (render.c:215-220) */
									float bx = Dx - sr.dx, by = Dy - sr.dy, bz = Dz - sr.dz;
									const float binv = 1.f / sqrtf(bx * bx + by * by + bz * bz);
									bx *= binv; by *= binv; bz *= binv;
									smul = -dot3(Nx, Ny, Nz, bx, by, bz);
								} else {
									/* R = 2 a N - L, s = -R.D (render.c:210-213) */
									const float a2 = 2.f * a;
									smul = -dot3(Nx * a2 - sr.dx, Ny * a2 - sr.dy, Nz * a2 - sr.dz, Dx, Dy, Dz);
								}
								/* raw base: negative bases follow powf's rules; fmaxf drops NaN */
								const float spec = fmaxf(0.f, powf(smul, ws.entry[E_SHIN][e_lane]));
								const float diff = fmaxf(0.f, a);
								cr = ws.entry[E_TR][e_lane] * ir * diff + ws.entry[E_KSR][e_lane] * ir * spec;
								cg = ws.entry[E_TG][e_lane] * ig * diff + ws.entry[E_KSG][e_lane] * ig * spec;
								cb = ws.entry[E_TB][e_lane] * ib * diff + ws.entry[E_KSB][e_lane] * ib * spec;
							}
							slot_r[w] = cr; slot_g[w] = cg; slot_b[w] = cb;
							have = false;
						}
						/* lanes still walking keep their ray; the others take the next indices */
						const unsigned free_mask = __ballot_sync(kFull, !walking);
						if (!walking) {
							w = next + (uint32_t)__popc(free_mask & ((1u << lane) - 1u));
							if (w < nwork) {
								have = true;
								blocked = false;
								shadow_walk_reset(sw);
								const uint32_t q = w / cL;
								const uint32_t li = s0 + (w - q * cL);
								e_lane = nth_set_bit(lit_mask, p0 + q);
								/* which emitter, which sample */
								uint32_t ei, l;
								const bool sampled = light_of_index(sc, li, __float_as_int(ws.entry[E_SKIP][e_lane]), ei, l);
								const DEmitter &em = sc.emitters[ei];
								if (!sampled) {
									blocked = true; /* the hit object is not its own light (render.c:172): no ray */
								} else {
									const float Px = ws.entry[E_PX][e_lane], Py = ws.entry[E_PY][e_lane], Pz = ws.entry[E_PZ][e_lane];
									float lx, ly, lz;
									light_point(em, Px, Py, Pz, __float_as_uint(ws.entry[E_NODE][e_lane]), 2u * l, lx, ly, lz);
									const float vx = lx - Px, vy = ly - Py, vz = lz - Pz;
									d2 = vx * vx + vy * vy + vz * vz;
									dist = sqrtf(d2);
									const float inv = 1.f / dist;
									sr.ox = Px; sr.oy = Py; sr.oz = Pz;
									sr.dx = vx * inv; sr.dy = vy * inv; sr.dz = vz * inv;
									ir = em.ke_n[0]; ig = em.ke_n[1]; ib = em.ke_n[2];
									em_object = em.object;
									Cnt<COUNT>::add(cn.shadow);
									blocked = shadow_planes<COUNT>(sc, sr, dist, ir, ig, ib, cn);
									if (!blocked && sc.num_nodes != 0u) {
										rw = make_wide_ray(sr);
#if CRT_REST_BOX
										/* most rays of an open scene pass the other objects' box by: decided here, with
										 * all fetching lanes in step, instead of by a NODE and a PRIM step of the walk
										 * (the emitter's own primitive) */
										Cnt<COUNT>::add(cn.box);
										if (!ray_misses_box(em.rest_lo, em.rest_hi, sr, rw, dist))
#endif
										sw.inner = 1u; /* the root of the wide tree; both stacks are empty */
									}
								}
							}
						}
						next += (uint32_t)__popc(free_mask);
						if (!__any_sync(kFull, have))
							break;
						/* -- walk until fewer than `keep` lanes are still walking: each
						 * iteration the warp takes the kind of step more lanes are ready for */
						const uint32_t keep = next < nwork ? p.refill : 1u;
						for (;;) {
							const bool ci = have && shadow_can_inner(sw), cp = have && shadow_can_prim(sw);
							const unsigned bi = __ballot_sync(kFull, ci), bp = __ballot_sync(kFull, cp);
							if ((uint32_t)__popc(bi | bp) < keep)
								break; /* a lane is busy iff it can take one of the two steps */
							if (__popc(bi) * kInnerWeight >= __popc(bp) * kPrimWeight) {
								/* several steps per vote; rolled loops keep the walk inside the instruction cache */
#pragma unroll 1
								for (int k = 0; k < kInnerPerVote; k++) {
									const bool go = have && shadow_can_inner(sw);
									if (COUNT) {
										const unsigned g = __ballot_sync(kFull, go), h = __ballot_sync(kFull, have);
										if (lane == 0 && g) {
											cn.node_steps++;
											cn.node_lanes += (uint32_t)__popc(g);
											cn.node_noray += 32u - (uint32_t)__popc(h);
										}
									}
									if (go)
										shadow_inner_step<COUNT>(sc, sr, rw, dist, sw, stacks, cn);
								}
							} else {
#pragma unroll 1
								for (int k = 0; k < kPrimPerVote; k++) {
									const bool go = have && shadow_can_prim(sw);
									if (COUNT) {
										const unsigned g = __ballot_sync(kFull, go), h = __ballot_sync(kFull, have);
										if (lane == 0 && g) {
											cn.prim_steps++;
											cn.prim_lanes += (uint32_t)__popc(g);
											cn.prim_noray += 32u - (uint32_t)__popc(h);
										}
									}
									if (go)
										shadow_prim_step<COUNT>(sc, sr, dist, em_object, sw, stacks, blocked, ir, ig, ib, cn);
								}
							}
						}
						/* lanes whose walk ended are retired at the top of the loop */
					}
					__syncwarp();
					/* fixed-order sums of the chunk's points */
					for (uint32_t q = 0; q < cpts; q++) {
						float a = 0.f, b = 0.f, c = 0.f;
						for (uint32_t i = lane; i < cL; i += 32u) {
							a += slot_r[q * cL + i];
							b += slot_g[q * cL + i];
							c += slot_b[q * cL + i];
						}
#pragma unroll
						for (int off = 16; off; off >>= 1) {
							a += __shfl_xor_sync(kFull, a, off);
							b += __shfl_xor_sync(kFull, b, off);
							c += __shfl_xor_sync(kFull, c, off);
						}
						if (lane == nth_set_bit(lit_mask, p0 + q)) {
							sum_r += a; sum_g += b; sum_b += c;
						}
					}
					__syncwarp();
					if (pts_per_chunk) {
						p0 += cpts;
					} else {
						s0 += cL;
						if (s0 >= L) {
							s0 = 0;
							p0++;
						}
					}
				}
			}

			/* ---- accumulate, then spawn children (render.c:239-340) --------- */
			if (is_hit) {
				if (lit) {
					acc_r += sum_r * wr;
					acc_g += sum_g * wg;
					acc_b += sum_b * wb;
				}
				if (r.remaining > 0u) {
					const bool ghost = primary && lane_group() != 0u;
					const crt_material *m = sc.materials + mat_index;
					/* fog: what the children bring back crosses this segment as well (recomputed, not kept in a register) */
					const float fogT = EXT && o.fog_density > 0.f ? __expf(-o.fog_density * hit.t) : 1.f;
					const uint32_t flags = __ldg(&m->flags);
					uint32_t n_gi = 0;
					if (o.gi_model == CRT_GI_PATH && outside)
						n_gi = (r.remaining == o.max_bounces) ? o.samples_per_pixel : 1u;
					uint32_t ordinal = n_gi;
					bool spawn_r = false, spawn_t = false;
					Pending rr, rt;
					if (!ghost && r.inside != hit.object && (flags & CRT_MAT_REFLECTIVE)) {
						/* render.c:308-317 */
						rr.kr = r.kr * __ldg(&m->kr[0]); rr.kg = r.kg * __ldg(&m->kr[1]); rr.kb = r.kb * __ldg(&m->kr[2]);
						if (EXT && o.fog_density > 0.f) {
							rr.kr *= fogT; rr.kg *= fogT; rr.kb *= fogT;
						}
						if (o.min_intensity_sqr < xdot(rr.kr, rr.kg, rr.kb, rr.kr, rr.kg, rr.kb)) {
							spawn_r = true;
							const float b2 = xmul(2.f, bdot);
							rr.ox = hx; rr.oy = hy; rr.oz = hz;
							rr.dx = xsub(ray.dx, xmul(nx, b2)); rr.dy = xsub(ray.dy, xmul(ny, b2)); rr.dz = xsub(ray.dz, xmul(nz, b2));
							rr.remaining = r.remaining - 1u;
							rr.inside = -1;
							rr.node = crt_node_child(r.node, ordinal++);
							rr.aux = 0;
						}
					}
					if (!ghost && (flags & CRT_MAT_TRANSPARENT)) {
						/* refraction by explicit angles (render.c:320-340) */
						rt.kr = r.kr * __ldg(&m->kt[0]); rt.kg = r.kg * __ldg(&m->kt[1]); rt.kb = r.kb * __ldg(&m->kt[2]);
						if (EXT && o.fog_density > 0.f) {
							rt.kr *= fogT; rt.kg *= fogT; rt.kb *= fogT;
						}
						if (o.min_intensity_sqr < xdot(rt.kr, rt.kg, rt.kb, rt.kr, rt.kg, rt.kb)) {
							spawn_t = true;
							const float ior = __ldg(&m->refractive_index);
							const float incident = acosf(fabsf(bdot));
							const float ratio = outside ? 1.f / ior : ior;
							const float refracted = asinf(sinf(incident) * ratio);
							const float turn = refracted - incident;
							float cx = ray.dy * nz - ray.dz * ny;
							float cy = ray.dz * nx - ray.dx * nz;
							float cz = ray.dx * ny - ray.dy * nx;
							const float cinv = (outside ? 1.f : -1.f) / sqrtf(cx * cx + cy * cy + cz * cz);
							cx *= cinv; cy *= cinv; cz *= cinv;
							const float fx = cy * ray.dz - cz * ray.dy;
							const float fy = cz * ray.dx - cx * ray.dz;
							const float fz = cx * ray.dy - cy * ray.dx;
							float st, ct;
							sincosf(turn, &st, &ct);
							float ox = ray.dx * ct + fx * st, oy = ray.dy * ct + fy * st, oz = ray.dz * ct + fz * st;
							const float oinv = 1.f / sqrtf(ox * ox + oy * oy + oz * oz);
							rt.ox = hx; rt.oy = hy; rt.oz = hz;
							rt.dx = ox * oinv; rt.dy = oy * oinv; rt.dz = oz * oinv;
							rt.remaining = r.remaining - 1u;
							rt.inside = hit.object;
							rt.node = crt_node_child(r.node, ordinal++);
							rt.aux = 0;
						}
					}
					/* push in reverse processing order: refraction, reflection, GI on top */
					if (spawn_t)
						pending_store(pending_slot(p, warp_global, sp++, lane), rt);
					if (spawn_r)
						pending_store(pending_slot(p, warp_global, sp++, lane), rr);
					/* first GI sample of this lane: the primary hit's samples are dealt over the pixel's lanes */
					const uint32_t k0 = primary ? lane_group() : 0u;
					if (n_gi > k0) {
						Pending g;
						g.ox = hx; g.oy = hy; g.oz = hz;
						g.dx = nx; g.dy = ny; g.dz = nz;
						g.kr = wr; g.kg = wg; g.kb = wb;
						g.remaining = REC_GI | k0;
						g.inside = (int32_t)n_gi;
						g.node = r.node;
						uint32_t lights = sc.total_lights;
						if (skip >= 0)
							lights -= sc.emitters[skip].num_lights;
						const bool flip = ny - eps < -1.f; /* render.c:241 */
						g.aux = (2u * lights) | (flip ? 0x80000000u : 0u);
						pending_store(pending_slot(p, warp_global, sp++, lane), g);
					}
				}
			}
			primary = false;
		}
		} /* lens samples */

		if (EXT && lens_n > 1u) { /* the mean over the lens; the rays keep unit weight so that -a cuts them where it would cut the pinhole ray */
			const float inv = 1.f / (float)lens_n;
			acc_r *= inv; acc_g *= inv; acc_b *= inv;
		}
		/* a pixel = the partial sums of its lanes, added in lane order */
		for (uint32_t g = 1; g < (1u << p.gshift); g++) {
			const uint32_t from = lane - lane_group() + g;
			const float tr = __shfl_sync(kFull, acc_r, from);
			const float tg = __shfl_sync(kFull, acc_g, from);
			const float tb = __shfl_sync(kFull, acc_b, from);
			if (lane_group() == 0u) {
				acc_r += tr; acc_g += tg; acc_b += tb;
			}
		}
		if (in_image && lane_group() == 0u) {
			p.raster[3 * (size_t)pixel + 0] = acc_r;
			p.raster[3 * (size_t)pixel + 1] = acc_g;
			p.raster[3 * (size_t)pixel + 2] = acc_b;
			p.z_buffer[pixel] = z_value;
		}
		flush_counters<COUNT>(p, cn);
	}
}

/* Pixel positions on the image plane exactly as the reference accumulates
 * them: position(row, col) = corner + row * vy, then (col + 1) successive
 * FP32 additions of vx (reference render.c:353-361).  One thread
 * per row; 2 x 10^6 dependent adds per 1080p frame are a few microseconds. */
__global__ void raygen_rows_kernel(crt_render_opts o, float *row_pos)
{
	const uint32_t row = blockIdx.x * blockDim.x + threadIdx.x;
	if (row >= o.height)
		return;
	float x = __fadd_rn(__fmul_rn(o.vy[0], (float)row), o.corner[0]);
	float y = __fadd_rn(__fmul_rn(o.vy[1], (float)row), o.corner[1]);
	float z = __fadd_rn(__fmul_rn(o.vy[2], (float)row), o.corner[2]);
	float *out = row_pos + 3 * (size_t)row * o.width;
	for (uint32_t col = 0; col < o.width; col++) {
		x = __fadd_rn(x, o.vx[0]);
		y = __fadd_rn(y, o.vx[1]);
		z = __fadd_rn(z, o.vx[2]);
		out[3 * col + 0] = x;
		out[3 * col + 1] = y;
		out[3 * col + 2] = z;
	}
}

/* ---- ray-batch seams (crt_trace_closest / crt_trace_shadow / crt_texture_eval) */

__global__ void trace_closest_kernel(DScene sc, const float *rays, size_t n, int32_t *object, float *t, float *normal)
{
	const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	Counters cn = {};
	const Ray r = { rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5], rays[6 * i], rays[6 * i + 1], rays[6 * i + 2] };
	uint32_t inner[kDeepStack];
	const PrivateStacks st = { inner, nullptr };
	const Hit h = trace_closest<false>(sc, r, st, cn);
	object[i] = h.object;
	t[i] = h.t;
	if (normal) {
		float nx = 0.f, ny = 0.f, nz = 0.f, eps;
		uint32_t mat;
		if (h.object >= 0)
			hit_normal(sc, h, r, nx, ny, nz, mat, eps);
		normal[3 * i] = nx;
		normal[3 * i + 1] = ny;
		normal[3 * i + 2] = nz;
	}
}

__global__ void trace_shadow_kernel(DScene sc, const float *rays, const float *dist, const int32_t *emitter, size_t n,
	float *light, uint8_t *blocked)
{
	const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	Counters cn = {};
	const Ray r = { rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5], rays[6 * i], rays[6 * i + 1], rays[6 * i + 2] };
	float lr = light[3 * i], lg = light[3 * i + 1], lb = light[3 * i + 2];
	const bool b = trace_shadow_wide<false>(sc, r, dist[i], emitter ? emitter[i] : -1, lr, lg, lb, cn);
	light[3 * i] = lr;
	light[3 * i + 1] = lg;
	light[3 * i + 2] = lb;
	blocked[i] = b ? 1 : 0;
}

__global__ void texture_eval_kernel(DScene sc, int32_t material, const float *points, size_t n, float *colors)
{
	const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
	if (i >= n)
		return;
	float r, g, b;
	texture_color(sc.materials + material, points[3 * i], points[3 * i + 1], points[3 * i + 2], r, g, b);
	colors[3 * i] = r;
	colors[3 * i + 1] = g;
	colors[3 * i + 2] = b;
}

/* FP32 FMA throughput probe: 8 independent chains per thread */
__global__ void fma_peak_kernel(int iters, float *sink)
{
	float a0 = threadIdx.x * 1e-3f, a1 = a0 + 1.f, a2 = a0 + 2.f, a3 = a0 + 3.f;
	float a4 = a0 + 4.f, a5 = a0 + 5.f, a6 = a0 + 6.f, a7 = a0 + 7.f;
	const float m = 0.999f, c = 1e-3f;
	for (int i = 0; i < iters; i++) {
#pragma unroll
		for (int u = 0; u < 8; u++) {
			a0 = fmaf(a0, m, c); a1 = fmaf(a1, m, c); a2 = fmaf(a2, m, c); a3 = fmaf(a3, m, c);
			a4 = fmaf(a4, m, c); a5 = fmaf(a5, m, c); a6 = fmaf(a6, m, c); a7 = fmaf(a7, m, c);
		}
	}
	const float s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
	if (s == 123.456f)
		sink[0] = s;
}

} // namespace

void launch_raygen_rows(const crt_render_opts &o, float *row_pos, cudaStream_t stream)
{
	const int threads = 64;
	raygen_rows_kernel<<<(o.height + threads - 1) / threads, threads, 0, stream>>>(o, row_pos);
}

constexpr size_t kRenderSmem = (sizeof(WarpShared) + sizeof(uint32_t) * kLaneStackWords * 32) * kWarpsPerBlock;

template <bool COUNT, bool EXT>
static int blocks_per_sm_of()
{
	int n = 0;
	cudaFuncSetAttribute(render_kernel<COUNT, EXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRenderSmem);
	cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_kernel<COUNT, EXT>, kBlockThreads, kRenderSmem);
	return n;
}

int render_blocks_per_sm(bool count, bool ext)
{
	if (ext)
		return count ? blocks_per_sm_of<true, true>() : blocks_per_sm_of<false, true>();
	return count ? blocks_per_sm_of<true, false>() : blocks_per_sm_of<false, false>();
}

void launch_render(const RenderParams &p, int grid, bool count, bool ext, cudaStream_t stream)
{
	if (ext) {
		if (count)
			render_kernel<true, true><<<grid, kBlockThreads, kRenderSmem, stream>>>(p);
		else
			render_kernel<false, true><<<grid, kBlockThreads, kRenderSmem, stream>>>(p);
	} else if (count) {
		render_kernel<true, false><<<grid, kBlockThreads, kRenderSmem, stream>>>(p);
	} else {
		render_kernel<false, false><<<grid, kBlockThreads, kRenderSmem, stream>>>(p);
	}
}

void launch_trace_closest(const DScene &sc, const float *rays, size_t n, int32_t *object, float *t, float *normal, cudaStream_t stream)
{
	if (n)
		trace_closest_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(sc, rays, n, object, t, normal);
}

void launch_trace_shadow(const DScene &sc, const float *rays, const float *dist, const int32_t *emitter, size_t n,
	float *light, uint8_t *blocked, cudaStream_t stream)
{
	if (n)
		trace_shadow_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(sc, rays, dist, emitter, n, light, blocked);
}

void launch_texture_eval(const DScene &sc, int32_t material, const float *points, size_t n, float *colors, cudaStream_t stream)
{
	if (n)
		texture_eval_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream>>>(sc, material, points, n, colors);
}

double launch_fma_peak(int blocks, int threads, int iters, float *sink, cudaStream_t stream)
{
	fma_peak_kernel<<<blocks, threads, 0, stream>>>(iters, sink);
	return 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
}

} // namespace crt
