/*
 * ref_shim.c -- TEST INFRASTRUCTURE (oracle).  Not part of the product.
 *
 * Link-time replacement for the reference's src/core/system.c when the
 * UNMODIFIED reference sources are compiled into oracle/_ref/ (see
 * oracle/Makefile).  It provides the three functions system.c exports
 * (system_init, system_time, rand_flt: reference src/core/system.h:22-25)
 * and gcc's -finstrument-functions hooks, which is how the structural RNG key
 * of include/crt_rng.h is threaded through the reference's recursive
 * cast_ray() (render.c:136) without touching its source:
 *
 *   - render.c is compiled with -finstrument-functions, so every entry/exit
 *     of cast_ray / get_closest_intersection / is_light_blocked calls the
 *     hooks below (also when gcc inlines or clones them);
 *   - on cast_ray entry a frame {node id, draw counter, child counter} is
 *     pushed; node = crt_node_child(parent.node, parent.children++), or the
 *     pixel's root id when the stack is empty;
 *   - rand_flt() returns crt_rand(top.node, top.draws++).
 *
 * When no pixel context is active (the reference's own render() loop, used
 * for the TIMED cpu baseline) rand_flt() falls back to a thread-local
 * xorshift32 -- the "thread-local RNG patch" of SURVEY F4 -- or, with
 * REF_STOCK_RAND, to glibc rand() exactly as shipped.
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "crt_rng.h"
#include "ref_shim.h"

#ifdef _OPENMP
#include <omp.h>
#endif

/* reference symbols we need (src/core/argv.h) */
extern int myargc;
extern char **myargv;

#define REF_MAX_DEPTH 4096

struct ref_frame {
	uint32_t node;
	uint32_t draws;
	uint32_t children;
};

static _Thread_local struct ref_frame tl_stack[REF_MAX_DEPTH];
static _Thread_local int tl_depth = 0;
static _Thread_local int tl_active = 0;
static _Thread_local uint32_t tl_root = 0;
static _Thread_local uint32_t tl_xs = 0;
static _Thread_local struct ref_counters tl_cnt;

/* addresses of the instrumented functions, filled by ref_shim_bind() */
static void *fn_cast_ray, *fn_closest, *fn_blocked, *fn_box;

static struct timespec start_t;

void ref_shim_bind(void *cast_ray_fn, void *closest_fn, void *blocked_fn)
{
	fn_cast_ray = cast_ray_fn;
	fn_closest = closest_fn;
	fn_blocked = blocked_fn;
}

void ref_shim_bind_box(void *box_fn)
{
	fn_box = box_fn;
}

void ref_shim_begin_pixel(uint32_t seed, uint32_t pixel)
{
	tl_root = crt_node_root(seed, pixel);
	tl_depth = 0;
	tl_active = 1;
}

void ref_shim_end_pixel(void)
{
	tl_active = 0;
	tl_depth = 0;
}

/* A bare node context for ray-batch tests that call helpers directly. */
void ref_shim_begin_node(uint32_t node)
{
	tl_stack[0].node = node;
	tl_stack[0].draws = 0;
	tl_stack[0].children = 0;
	tl_depth = 1;
	tl_active = 1;
}

void ref_shim_get_counters(struct ref_counters *out)
{
	*out = tl_cnt;
}

void ref_shim_reset_counters(void)
{
	memset(&tl_cnt, 0, sizeof tl_cnt);
}

__attribute__((no_instrument_function)) void __cyg_profile_func_enter(void *fn, void *site)
{
	(void)site;
	if (fn == fn_box) { /* by far the most frequent instrumented call */
		tl_cnt.box_tests++;
	} else if (fn == fn_cast_ray) {
		tl_cnt.cast_ray++;
		if (!tl_active)
			return;
		if (tl_depth >= REF_MAX_DEPTH)
			abort();
		struct ref_frame *f = &tl_stack[tl_depth];
		if (tl_depth == 0)
			f->node = tl_root;
		else
			f->node = crt_node_child(tl_stack[tl_depth - 1].node, tl_stack[tl_depth - 1].children++);
		f->draws = 0;
		f->children = 0;
		tl_depth++;
	} else if (fn == fn_blocked) {
		tl_cnt.shadow++;
	} else if (fn == fn_closest) {
		tl_cnt.closest_full++;
	}
}

__attribute__((no_instrument_function)) void __cyg_profile_func_exit(void *fn, void *site)
{
	(void)site;
	if (fn == fn_cast_ray && tl_active && tl_depth > 0)
		tl_depth--;
}

/* ---- replacements for src/core/system.c ------------------------------- */

void system_init(void)
{
	timespec_get(&start_t, TIME_UTC);
#ifdef REF_STOCK_RAND
	srand((unsigned)start_t.tv_sec);
#endif
	/* -m <n>|max, as reference system.c:54-74 */
#ifdef _OPENMP
	int threads = 1;
	for (int i = 1; i + 1 < myargc; i++)
		if (!strcmp(myargv[i], "-m"))
			threads = !strcmp(myargv[i + 1], "max") ? omp_get_max_threads() : abs(atoi(myargv[i + 1]));
	if (threads < 1)
		threads = 1;
	omp_set_num_threads(threads);
	printf("[%08.3f] ref_shim: Using %d threads.\n", 0.0, threads);
#endif
}

double system_time(void)
{
	struct timespec cur_t;
	timespec_get(&cur_t, TIME_UTC);
	return (double)(cur_t.tv_sec - start_t.tv_sec) + (double)(cur_t.tv_nsec - start_t.tv_nsec) * 1e-9;
}

float rand_flt(void)
{
	if (tl_active && tl_depth > 0) {
		struct ref_frame *f = &tl_stack[tl_depth - 1];
		return crt_rand(f->node, f->draws++);
	}
#ifdef REF_STOCK_RAND
	return (float)rand() / (float)RAND_MAX; /* as shipped, system.c:95 */
#else
	uint32_t x = tl_xs;
	if (!x) {
		x = 0x9e3779b9U;
#ifdef _OPENMP
		x ^= 0x85ebca6bU * (uint32_t)(omp_get_thread_num() + 1);
#endif
	}
	x ^= x << 13;
	x ^= x >> 17;
	x ^= x << 5;
	tl_xs = x;
	return (float)(x >> 8) * (1.0f / 16777216.0f);
#endif
}
