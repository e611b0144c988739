/* loads what it is given through the host library's public entry points; must never crash, only fail */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "crt.h"
struct rt_scene;
const char *rt_host_last_error(void);
struct rt_scene *rt_host_scene_open(const char *filename, const char *rescale);
void rt_host_scene_close(struct rt_scene *scene);
void rt_host_scene_desc(const struct rt_scene *scene, crt_scene_desc *desc);
void rt_host_opts(const struct rt_scene *scene, unsigned width, unsigned height, crt_render_opts *opts);
int rt_host_load_raw(const char *filename, unsigned *width, unsigned *height, float **raster, float **z_buffer);
void rt_host_free(void *p);
void rt_host_opts_parse(crt_render_opts *opts, int argc, char **argv);
int rt_host_post_opts_parse(crt_post_opts *opts, int argc, char **argv);
int main(int argc, char **argv)
{
	if (argc < 3) return 2;
	if (!strcmp(argv[1], "opts")) { crt_render_opts o; memset(&o, 0, sizeof o); rt_host_opts_parse(&o, argc - 2, argv + 2); crt_post_opts p; memset(&p, 0, sizeof p); int rc = rt_host_post_opts_parse(&p, argc - 2, argv + 2); printf("ok %u %u %d\n", o.max_bounces, o.samples_per_pixel, rc); return 0; }
	if (!strcmp(argv[1], "scene")) {
		struct rt_scene *s = rt_host_scene_open(argv[2], argc > 3 ? argv[3] : NULL);
		if (!s) { printf("fail: %s\n", rt_host_last_error()); return 1; }
		crt_scene_desc d; rt_host_scene_desc(s, &d);
		crt_render_opts o; rt_host_opts(s, 64, 36, &o);
		double sum = 0; for (unsigned i = 0; i < d.num_objects; i++) sum += d.objects[i].a[0] + d.objects[i].epsilon + d.objects[i].material;
		for (unsigned i = 0; i < d.num_materials; i++) sum += d.materials[i].shininess;
		printf("ok %u objects %u materials %g\n", d.num_objects, d.num_materials, sum);
		rt_host_scene_close(s);
		return 0;
	}
	if (!strcmp(argv[1], "tiff")) {
		unsigned w, h; float *r = NULL, *z = NULL;
		if (rt_host_load_raw(argv[2], &w, &h, &r, &z)) { printf("fail: %s\n", rt_host_last_error()); return 1; }
		double sum = 0; for (size_t i = 0; i < (size_t)w * h; i++) sum += r[3 * i] + z[i];
		printf("ok %ux%u %g\n", w, h, sum);
		rt_host_free(r); rt_host_free(z);
		return 0;
	}
	return 2;
}
