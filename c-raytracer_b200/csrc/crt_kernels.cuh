/*
 * crt_kernels.cuh -- launch interface between the C ABI (crt_api.cu) and the
 * kernels (crt_kernels.cu).
 */
#ifndef CRT_KERNELS_CUH
#define CRT_KERNELS_CUH

#include "crt_device.cuh"

namespace crt {

constexpr int kTileW = CRT_TILE_W, kTileH = CRT_TILE_H;      /* one warp owns an 8x4 pixel tile: lane = pixel */
constexpr int kChunkTiles = CRT_SHARD_CHUNK_TILES;            /* sharding granularity (crt_render_opts.shard_*) */
#ifndef CRT_WARPS_PER_BLOCK
#define CRT_WARPS_PER_BLOCK 8
#endif
constexpr int kWarpsPerBlock = CRT_WARPS_PER_BLOCK;
constexpr int kBlockThreads = kWarpsPerBlock * 32;
#ifndef CRT_INNER_PER_VOTE
#define CRT_INNER_PER_VOTE 3
#endif
#ifndef CRT_PRIM_PER_VOTE
#define CRT_PRIM_PER_VOTE 3
#endif
#ifndef CRT_INNER_WEIGHT
#define CRT_INNER_WEIGHT 1
#endif
#ifndef CRT_PRIM_WEIGHT
#define CRT_PRIM_WEIGHT 1
#endif
constexpr int kInnerPerVote = CRT_INNER_PER_VOTE, kPrimPerVote = CRT_PRIM_PER_VOTE; /* unit steps per warp vote */
constexpr int kInnerWeight = CRT_INNER_WEIGHT, kPrimWeight = CRT_PRIM_WEIGHT;       /* node step iff ready_inner * Wi >= ready_prim * Wp */
constexpr int kNumCounters = 17;
/* instrumented instance: device clock (globaltimer, ns) marks after the event counters --
 * ~min warp start, max warp end, sum of warp lifetimes, ~min time a warp found the queue empty */
constexpr int kTimeStart = kNumCounters, kTimeEnd = kNumCounters + 1, kTimeBusy = kNumCounters + 2, kTimeDry = kNumCounters + 3;
constexpr int kTotalCounters = kNumCounters + 4;
#ifndef CRT_GI_LANES
#define CRT_GI_LANES 4                     /* lanes that share the GI samples of one pixel (power of two; part of a frame's arithmetic) */
#endif
#ifndef CRT_TAPER_PIXELS
#define CRT_TAPER_PIXELS 24                /* units of s pixels are handed out while more than s * this many pixels per resident warp remain */
#endif
constexpr int kPhases = 4;                 /* unit sizes 8, 4, 2, 1 pixels */
#ifndef CRT_MIN_BLOCKS
#define CRT_MIN_BLOCKS 2                   /* resident render blocks per SM the register budget is set for */
#endif
constexpr int kPendingWords = 13;          /* words per pending-ray record */
constexpr int kSlots = 1024;               /* shadow-ray contribution slots per warp (one chunk) */
constexpr int kLaneStackWords = kInnerStack + kLeafStack; /* shared-memory stack words per lane (crt_device.cuh SharedStacks) */
constexpr int kRefill = 8;                 /* lanes fetch new shadow rays when fewer than this many are walking */

struct RenderParams {
	DScene sc;
	crt_render_opts o;
	float *raster;
	float *z_buffer;
	const float *row_pos;           /* W*H*3: running-sum pixel positions (raygen_rows_kernel) */
	uint32_t tiles_x, tiles_y;
	uint32_t num_local_tiles;       /* tiles of this shard */
	uint32_t refill;                /* lanes fetch new shadow rays when fewer than this many are walking */
	uint32_t gshift;                /* log2 of the lanes that trace one pixel (path-traced GI with several samples), else 0 */
	uint32_t fine_pixels;           /* pixels per work unit: min(8, 32 >> gshift) neighbours in an 8x4 tile */
	uint32_t phase_unit[kPhases];   /* phase i = units of fine_pixels >> i pixels: its first unit (0xffffffff: no such phase) ... */
	uint32_t phase_tile[kPhases];   /* ... and its first local tile */
	uint32_t num_units;
	unsigned int *tile_counter;
	uint32_t system_counter;        /* the counter is shared with other devices (peer / IPC mapping): system-scope atomics */
	unsigned long long *counters;   /* kTotalCounters, instrumented instance only */
	float *pending;                 /* per-lane pending-ray stacks */
	float *slots;                   /* per-warp shadow-ray contribution slots: 3 * kSlots floats each */
	uint32_t pending_depth;
	uint32_t *deep;                 /* spill strips of the inner stacks (trees deeper than kInnerStack), or null */
	uint32_t deep_depth;            /* words per lane in a strip */
};

/* grid = number of persistent blocks */
void launch_raygen_rows(const crt_render_opts &o, float *row_pos, cudaStream_t stream);
/* ext: the instance with the lens / fog extensions of crt_render_opts */
void launch_render(const RenderParams &p, int grid, bool count, bool ext, cudaStream_t stream);
int render_blocks_per_sm(bool count, bool ext);

void launch_trace_closest(const DScene &sc, const float *rays, size_t n, int32_t *object, float *t, float *normal, cudaStream_t stream);
void launch_trace_shadow(const DScene &sc, const float *rays, const float *dist, const int32_t *emitter, size_t n,
	float *light, uint8_t *blocked, cudaStream_t stream);
void launch_texture_eval(const DScene &sc, int32_t material, const float *points, size_t n, float *colors, cudaStream_t stream);
/* returns FLOPs executed per launch */
double launch_fma_peak(int blocks, int threads, int iters, float *sink, cudaStream_t stream);

/* shared by the API translation units (crt_api.cu, crt_post.cu): sets
 * crt_last_error() and returns `code`; selects the device or fails with
 * CRT_ERR_NO_DEVICE / CRT_ERR_INVALID */
int api_fail(int code, const char *fmt, ...);
int api_ensure_device(int device);

} // namespace crt

#endif
