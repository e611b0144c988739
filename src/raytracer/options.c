/*
 * options.c -- render option defaults and command-line parsing; kept apart
 * from render.c so that tools which never touch the GPU can link it.
 */
#include "render.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "argv.h"
#include "calc.h"
#include "image.h"
#include "scene.h"

void rt_render_defaults(crt_render_opts *o)
{
	memset(o, 0, sizeof *o);
	o->max_bounces = 10;
	o->min_intensity_sqr = .01f * .01f;
	o->reflection_model = CRT_REFLECTION_PHONG;
	o->gi_model = CRT_GI_AMBIENT;
	o->samples_per_pixel = 1;
	o->attenuation = CRT_ATTEN_SQUARE;
	o->attenuation_offset = 1.f;
	o->seed = 1;
	o->shard_index = 0;
	o->shard_count = 1;
}

static const char *value_of(const struct rt_args *args, const char *flag)
{
	const int i = rt_args_find_with_values(args, flag, 1);
	return i ? args->argv[i + 1] : NULL;
}

void rt_render_parse(crt_render_opts *o, const struct rt_args *args)
{
	const char *v;
	if ((v = value_of(args, "-b")))
		o->max_bounces = (uint32_t)abs(atoi(v));
	if ((v = value_of(args, "-a"))) {
		const float a = (float)atof(v);
		o->min_intensity_sqr = a * a;
	}
	if ((v = value_of(args, "-s"))) {
		if (!strcmp(v, "phong"))
			o->reflection_model = CRT_REFLECTION_PHONG;
		else if (!strcmp(v, "blinn"))
			o->reflection_model = CRT_REFLECTION_BLINN;
	}
	if ((v = value_of(args, "-g"))) {
		if (!strcmp(v, "ambient"))
			o->gi_model = CRT_GI_AMBIENT;
		else if (!strcmp(v, "path"))
			o->gi_model = CRT_GI_PATH;
	}
	if ((v = value_of(args, "-n")))
		o->samples_per_pixel = (uint32_t)abs(atoi(v));
	if ((v = value_of(args, "-l"))) {
		if (!strcmp(v, "none"))
			o->attenuation = CRT_ATTEN_NONE;
		else if (!strcmp(v, "lin"))
			o->attenuation = CRT_ATTEN_LINEAR;
		else if (!strcmp(v, "sqr"))
			o->attenuation = CRT_ATTEN_SQUARE;
	}
	if ((v = value_of(args, "-o")))
		o->attenuation_offset = (float)atof(v);
	if ((v = value_of(args, "--seed")))
		o->seed = (uint32_t)strtoul(v, NULL, 10);
	if (rt_args_find(args, "--stats") || rt_args_find_with_values(args, "--stats-json", 1))
		o->flags |= CRT_RENDER_COUNT_STATS;
	/* extensions without a counterpart in the reference (include/crt.h): off unless asked for */
	if ((v = value_of(args, "--lens-radius")))
		o->lens_radius = (float)atof(v);
	if ((v = value_of(args, "--focus")))
		o->focus_distance = (float)atof(v);
	if ((v = value_of(args, "--lens-samples")))
		o->lens_samples = (uint32_t)abs(atoi(v));
	if ((v = value_of(args, "--fog")))
		o->fog_density = (float)atof(v);
	if ((v = value_of(args, "--fog-color"))) {
		float c[3] = { 0.f, 0.f, 0.f };
		if (sscanf(v, "%f,%f,%f", &c[0], &c[1], &c[2]) == 3)
			v3_copy(o->fog_color, c);
	}
}

void rt_render_set_view(crt_render_opts *o, const struct rt_scene *scene, const struct rt_image *img)
{
	o->width = img->width;
	o->height = img->height;
	v3_copy(o->cam_pos, scene->camera.position);
	v3_copy(o->corner, img->corner);
	v3_copy(o->vx, img->step_x);
	v3_copy(o->vy, img->step_y);
}
