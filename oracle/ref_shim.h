/* ref_shim.h -- TEST INFRASTRUCTURE (oracle).  See ref_shim.c. */
#ifndef REF_SHIM_H
#define REF_SHIM_H
#include <stdint.h>

struct ref_counters {
	uint64_t cast_ray;     /* cast_ray() invocations (render.c:136) */
	uint64_t closest_full; /* get_closest_intersection() calls (render.c:118) */
	uint64_t shadow;       /* is_light_blocked() calls (render.c:126) */
	uint64_t box_tests;    /* bounding_cuboid_intersects() calls (accel.c:112), both walks */
};

void ref_shim_bind(void *cast_ray_fn, void *closest_fn, void *blocked_fn);
void ref_shim_bind_box(void *box_fn); /* accel.c compiled with -finstrument-functions as well */
void ref_shim_begin_pixel(uint32_t seed, uint32_t pixel);
void ref_shim_end_pixel(void);
void ref_shim_begin_node(uint32_t node);
void ref_shim_get_counters(struct ref_counters *out);
void ref_shim_reset_counters(void);

#endif
