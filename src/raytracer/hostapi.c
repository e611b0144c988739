#include "hostapi.h"

#include <stdlib.h>
#include <string.h>

#include "argv.h"
#include "error.h"
#include "image.h"
#include "render.h"
#include "scene.h"
#include "tiff.h"
#include "../postprocess/options.h"

static _Thread_local char last_error[RT_ERRLEN];

const char *rt_host_last_error(void)
{
	return last_error;
}

struct rt_scene *rt_host_scene_open(const char *filename, const char *rescale)
{
	struct rt_scene *sc = malloc(sizeof *sc);
	if (!sc)
		return NULL;
	if (rt_scene_load(filename, rescale, sc, last_error)) {
		free(sc);
		return NULL;
	}
	return sc;
}

void rt_host_scene_close(struct rt_scene *scene)
{
	if (scene) {
		rt_scene_free(scene);
		free(scene);
	}
}

void rt_host_scene_desc(const struct rt_scene *scene, crt_scene_desc *desc)
{
	rt_scene_desc(scene, desc);
}

void rt_host_opts(const struct rt_scene *scene, unsigned width, unsigned height, crt_render_opts *opts)
{
	struct rt_image img;
	memset(&img, 0, sizeof img);
	rt_image_plane(&img, &scene->camera, width, height);
	rt_render_defaults(opts);
	rt_render_set_view(opts, scene, &img);
}

void rt_host_opts_parse(crt_render_opts *opts, int argc, char **argv)
{
	struct rt_args args;
	if (rt_args_init(&args, argc, argv))
		return;
	rt_render_parse(opts, &args);
	rt_args_free(&args);
}

int rt_host_post_opts_parse(crt_post_opts *opts, int argc, char **argv)
{
	struct rt_args args;
	if (rt_args_init(&args, argc, argv)) {
		snprintf(last_error, sizeof last_error, "out of memory");
		return -1;
	}
	const int rc = rt_post_options(&args, opts, last_error);
	rt_args_free(&args);
	return rc;
}

int rt_host_load_raw(const char *filename, unsigned *width, unsigned *height, float **raster, float **z_buffer)
{
	struct rt_tiff_image img;
	if (rt_tiff_read(filename, &img, last_error))
		return -1;
	if (img.bits_per_sample != 32 || !img.raster || !img.z_buffer || (uint64_t)img.z_count != (uint64_t)img.width * img.height) {
		snprintf(last_error, sizeof last_error, "[%s] is not a raw frame (32 bits per sample with a z-buffer)", filename);
		rt_tiff_free(&img);
		return -1;
	}
	*width = img.width;
	*height = img.height;
	*raster = img.raster;
	*z_buffer = img.z_buffer;
	free(img.rgb8);
	return 0;
}

void rt_host_free(void *p)
{
	free(p);
}

int rt_host_save(const char *filename, unsigned width, unsigned height, const float *raster, const float *z_buffer, int raw)
{
	struct rt_image img;
	memset(&img, 0, sizeof img);
	img.width = width;
	img.height = height;
	img.pixels = (size_t)width * height;
	img.raster = (float *)raster;
	img.z_buffer = (float *)z_buffer;
	return rt_image_save(&img, filename, raw, last_error);
}

void rt_host_quantize(const float *raster, size_t samples, unsigned char *out)
{
	rt_image_quantize(raster, samples, out);
}
