/*
 * render.h -- render options and the render() seam.
 *
 * render() is the one function of the reference whose body is replaced
 * (reference src/raytracer/render.c:345-368): here it hands the flattened
 * scene to the sm_100a device library through the C ABI in include/crt.h.
 * There is no CPU implementation behind it.
 */
#ifndef RT_RENDER_H
#define RT_RENDER_H

#include "crt.h"

struct rt_args;
struct rt_scene;
struct rt_image;

/* defaults of reference render.c:52-59 */
void rt_render_defaults(crt_render_opts *opts);
/* -b -a -s -g -n -l -o (render.c:61-116) plus the additive --seed <n> and
 * --stats flags; unknown option values are ignored like the reference's
 * switch statements without default */
void rt_render_parse(crt_render_opts *opts, const struct rt_args *args);
/* copy camera position and image plane into opts */
void rt_render_set_view(crt_render_opts *opts, const struct rt_scene *scene, const struct rt_image *img);

/* upload + trace + download; fills img->raster / img->z_buffer */
int rt_render(const struct rt_scene *scene, struct rt_image *img, const crt_render_opts *opts,
	int device, crt_stats *stats, crt_scene_info *info, char *err);

/* the same on several CUDA devices at once (the reference's -m: render.c:349-352 runs the rows on
 * OpenMP threads): scene replicated, one dynamic work queue over all devices (crt_multi_render);
 * device_ms: num_devices values or NULL; dynamic: 1 shared queue over peer access, 0 static shards */
int rt_render_multi(const struct rt_scene *scene, struct rt_image *img, const crt_render_opts *opts,
	const int *devices, int num_devices, crt_stats *stats, crt_scene_info *info, double *device_ms, int *dynamic, char *err);

#endif
