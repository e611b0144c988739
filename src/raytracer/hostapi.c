#include "hostapi.h"

#include <stdlib.h>
#include <string.h>

#include "argv.h"
#include "error.h"
#include "image.h"
#include "render.h"
#include "scene.h"

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
