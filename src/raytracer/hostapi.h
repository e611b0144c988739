/*
 * hostapi.h -- handle-style entry points of libcrt_host.so for bindings
 * (Python ctypes in tests/ and bench.py).  Thin wrappers over scene.h,
 * image.h and render.h; nothing here touches the GPU.
 */
#ifndef RT_HOSTAPI_H
#define RT_HOSTAPI_H

#include "crt.h"

struct rt_scene;

/* returns NULL on failure; message via rt_host_last_error() */
struct rt_scene *rt_host_scene_open(const char *filename, const char *rescale);
void rt_host_scene_close(struct rt_scene *scene);
const char *rt_host_last_error(void);
void rt_host_scene_desc(const struct rt_scene *scene, crt_scene_desc *desc);
/* defaults + view for a W x H frame of this scene's camera */
void rt_host_opts(const struct rt_scene *scene, unsigned width, unsigned height, crt_render_opts *opts);
/* apply engine command-line flags (argv[0] is ignored) */
void rt_host_opts_parse(crt_render_opts *opts, int argc, char **argv);
/* the post-processor's flags (src/postprocess/options.h); -1 + rt_host_last_error() on a bad value */
int rt_host_post_opts_parse(crt_post_opts *opts, int argc, char **argv);
/* reads a raw frame written by rt_host_save(..., raw = 1) / the reference engine's -f: returns 0 and
 * malloc'ed raster (W*H*3) and z_buffer (W*H), to be released with rt_host_free() */
int rt_host_load_raw(const char *filename, unsigned *width, unsigned *height, float **raster, float **z_buffer);
void rt_host_free(void *p);
int rt_host_save(const char *filename, unsigned width, unsigned height, const float *raster, const float *z_buffer, int raw);
void rt_host_quantize(const float *raster, size_t samples, unsigned char *out);

#endif
