/*
 * ref_engine.c -- TEST INFRASTRUCTURE (oracle).  Not part of the product.
 *
 * Driver around the UNMODIFIED reference sources compiled in place from
 * /root/reference (oracle/Makefile).  It calls the reference's own
 * scene_load / image_init / accel_init / render_init (main.c:71-77) and then,
 * instead of render() (render.c:345-368), runs the same pixel loop while
 * telling ref_shim.c which pixel is being traced, so the reference's
 * cast_ray() draws from the structural stream of include/crt_rng.h.
 *
 * Modes (argv[1]):
 *   render <scene.json> <out.bin> <W> <H> [reference flags] [--seed S]
 *          [--stock-loop]      dump raster + z-buffer + ray counters
 *   rays   <scene.json> <rays.bin> <out.bin>
 *          closest-hit and shadow queries through the reference's
 *          get_closest_intersection / is_light_blocked
 *   scene  <scene.json> <out.bin> [-r ...]   dump the loaded scene as crt.h records
 *   tex    <scene.json> <material_index> <points.bin> <out.bin>
 *   noise  <points.bin> <out.bin>            simplex_noise()
 *   kat    <in.bin> <out.bin>                primitive known-answer tests
 */
#define _GNU_SOURCE
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* reference headers (from /root/reference, via -I) */
#include "accel.h"
#include "argv.h"
#include "calc.h"
#include "camera.h"
#include "image.h"
#include "material.h"
#include "object.h"
#include "render.h"
#include "scene.h"
#include "system.h"

#include "SimplexNoise.h"

#include "crt.h"
#include "crt_rng.h"
#include "ref_shim.h"

/* non-static reference functions without a header declaration */
float cast_ray(const struct Ray *ray, const v3 kr, v3 color, uint32_t bounce_count, struct Object *inside_object);
void get_closest_intersection(const struct Ray *ray, struct Object **closest_object, v3 closest_normal, float *closest_distance);
bool is_light_blocked(const struct Ray *ray, float distance, v3 light_intensity, const struct Object *emittant_object);
bool moller_trumbore(const v3 vertex, v3 edges[2], const v3 line_position, const v3 line_vector, const float epsilon, float *distance);
bool line_intersects_sphere(const v3 sphere_position, const float sphere_radius, const v3 line_position, const v3 line_vector, const float epsilon, float *distance);

/* private reference layouts re-declared for inspection only
 * (object.c:24-43, material.c:27-53, accel.c:25-28) */
struct RSphere { struct Object object; v3 position; float radius; };
struct RTriangle { struct Object object; v3 vertices[3]; v3 edges[2]; v3 normal; };
struct RPlane { struct Object object; v3 normal; float d; };
struct RTexUniform { struct Texture texture; v3 color; };
struct RTexChecker { struct Texture texture; v3 colors[2]; float scale; };
struct RTexBrick { struct Texture texture; v3 colors[2]; float scale; float mortar_width; };
struct RTexNoisy { struct Texture texture; v3 color; v3 color_gradient; float noise_feature_scale; float noise_scale; float frequency_scale; enum PeriodicFunction func; };
struct RBoundingCuboid { float epsilon; v3 corners[2]; };
bool bounding_cuboid_intersects(const struct RBoundingCuboid *cuboid, const struct Ray *ray, float *tmax, float *tmin);

void texture_get_color_uniform(const struct Texture *tex, const v3 point, v3 color);
void texture_get_color_checkerboard(const struct Texture *tex, const v3 point, v3 color);
void texture_get_color_brick(const struct Texture *tex, const v3 point, v3 color);
void texture_get_color_noisy_periodic(const struct Texture *tex, const v3 point, v3 color);

static void die(const char *msg)
{
	fprintf(stderr, "ref_engine: %s\n", msg);
	exit(2);
}

static void *read_file(const char *path, size_t *len)
{
	FILE *f = fopen(path, "rb");
	if (!f)
		die("cannot open input file");
	fseek(f, 0, SEEK_END);
	long n = ftell(f);
	fseek(f, 0, SEEK_SET);
	void *buf = malloc(n ? (size_t)n : 1);
	if (fread(buf, 1, (size_t)n, f) != (size_t)n)
		die("short read");
	fclose(f);
	*len = (size_t)n;
	return buf;
}

/* Build the reference's global argv from ours: prog scene out W H flags... */
static void set_ref_argv(int argc, char **argv)
{
	myargc = argc;
	myargv = argv;
	argv_init();
}

static int cmp_ptr(const void *a, const void *b)
{
	const void *pa = *(void *const *)a, *pb = *(void *const *)b;
	return (pa > pb) - (pa < pb);
}

struct ptr_index { const void *ptr; int32_t index; };

static struct ptr_index *object_index_table(void)
{
	struct ptr_index *t = malloc(sizeof *t * num_objects);
	for (size_t i = 0; i < num_objects; i++) {
		t[i].ptr = objects[i];
		t[i].index = (int32_t)i;
	}
	qsort(t, num_objects, sizeof *t, cmp_ptr);
	return t;
}

static int32_t object_index(const struct ptr_index *t, const void *obj)
{
	if (!obj)
		return -1;
	size_t lo = 0, hi = num_objects;
	while (lo < hi) {
		size_t mid = (lo + hi) / 2;
		if (t[mid].ptr < obj)
			lo = mid + 1;
		else
			hi = mid;
	}
	return (lo < num_objects && t[lo].ptr == obj) ? t[lo].index : -1;
}

/* ---- mode: render ------------------------------------------------------ */

static int mode_render(int argc, char **argv)
{
	/* argv: render scene out W H flags... ; shift so reference sees argv[1..4] */
	if (argc < 6)
		die("render: need <scene> <out.bin> <W> <H>");
	uint32_t seed = 1;
	int stock_loop = 0;
	for (int i = 6; i < argc; i++) {
		if (!strcmp(argv[i], "--seed") && i + 1 < argc)
			seed = (uint32_t)strtoul(argv[i + 1], NULL, 10);
		if (!strcmp(argv[i], "--stock-loop"))
			stock_loop = 1;
	}
	set_ref_argv(argc - 1, argv + 1);
	system_init();
	scene_load();
	image_init();
	accel_init();
	render_init();
	ref_shim_bind((void *)cast_ray, (void *)get_closest_intersection, (void *)is_light_blocked);
	ref_shim_bind_box((void *)bounding_cuboid_intersects);
	memset(image.raster, 0, image.pixels * sizeof(v3)); /* reference relies on zero pages */
	memset(image.z_buffer, 0, image.pixels * sizeof(float));

	uint32_t max_b = 10;
	int idx = argv_check_with_args("-b", 1);
	if (idx)
		max_b = (uint32_t)abs(atoi(myargv[idx + 1]));

	uint64_t n_cast = 0, n_closest = 0, n_shadow = 0, n_box = 0;
	double t0 = system_time();
	if (stock_loop) {
		render(); /* reference loop, thread-local xorshift stream */
	} else {
		const v3 kr = { 1.f, 1.f, 1.f };
		/* same arithmetic as reference render.c:352-366, incl. the running
		 * sum along the row */
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : n_cast, n_closest, n_shadow, n_box)
		for (uint32_t row = 0; row < image.resolution[Y]; row++) {
			v3 pixel_position;
			mul3s(image.vectors[Y], row, pixel_position);
			add3v(pixel_position, image.corner, pixel_position);
			struct Ray ray;
			assign3(ray.point, camera.position);
			uint32_t pixel_index = image.resolution[X] * row;
			ref_shim_reset_counters();
			for (uint32_t col = 0; col < image.resolution[X]; col++) {
				add3v(pixel_position, image.vectors[X], pixel_position);
				sub3v(pixel_position, camera.position, ray.direction);
				norm3(ray.direction);
				ref_shim_begin_pixel(seed, pixel_index);
				image.z_buffer[pixel_index] = cast_ray(&ray, kr, image.raster[pixel_index], max_b, NULL);
				ref_shim_end_pixel();
				pixel_index++;
			}
			struct ref_counters c;
			ref_shim_get_counters(&c);
			n_cast += c.cast_ray;
			n_closest += c.closest_full;
			n_shadow += c.shadow;
			n_box += c.box_tests;
		}
	}
	double t1 = system_time();

	FILE *f = fopen(argv[3], "wb");
	if (!f)
		die("cannot open output");
	uint32_t hdr[4] = { 0x44545243u /* CRTD */, image.resolution[X], image.resolution[Y], seed };
	fwrite(hdr, sizeof hdr, 1, f);
	fwrite(image.raster, sizeof(v3), image.pixels, f);
	fwrite(image.z_buffer, sizeof(float), image.pixels, f);
	uint64_t cnt[4] = { n_cast, n_closest, n_shadow, n_box };
	fwrite(cnt, sizeof cnt, 1, f);
	double tm[2] = { t1 - t0, 0 };
	fwrite(tm, sizeof tm, 1, f);
	fclose(f);
	printf("ref_engine: render %.3f s, cast_ray %llu, shadow %llu, box tests %llu\n", t1 - t0,
		(unsigned long long)n_cast, (unsigned long long)n_shadow, (unsigned long long)n_box);
	return 0;
}

/* ---- mode: rays -------------------------------------------------------- */

/* rays.bin: u32 n, then n records of 11 floats:
 *   dir[3] point[3] dist emitter_index(as float, -1 = none) light[3]
 * out.bin: n records of: i32 object, f32 t, f32 normal[3], i32 blocked, f32 light[3] */
static int mode_rays(int argc, char **argv)
{
	if (argc < 5)
		die("rays: need <scene> <rays.bin> <out.bin>");
	char *fake[] = { argv[0], argv[2], (char *)"none.tif", (char *)"8", (char *)"8" };
	set_ref_argv(5, fake);
	system_init();
	scene_load();
	image_init();
	accel_init();
	size_t len;
	uint8_t *buf = read_file(argv[3], &len);
	uint32_t n = *(uint32_t *)buf;
	const float *rec = (const float *)(buf + 4);
	struct ptr_index *tab = object_index_table();
	FILE *f = fopen(argv[4], "wb");
	if (!f)
		die("cannot open output");
	for (uint32_t i = 0; i < n; i++) {
		const float *r = rec + (size_t)i * 11;
		struct Ray ray;
		memcpy(ray.direction, r, sizeof(v3));
		memcpy(ray.point, r + 3, sizeof(v3));
		struct Object *obj = NULL;
		v3 normal = { 0, 0, 0 };
		float t = FLT_MAX;
		get_closest_intersection(&ray, &obj, normal, &t);
		int32_t oi = object_index(tab, obj);
		v3 light = { r[8], r[9], r[10] };
		int emi = (int)r[7];
		const struct Object *em = (emi >= 0 && (size_t)emi < num_objects) ? objects[emi] : NULL;
		int32_t blocked = is_light_blocked(&ray, r[6], light, em);
		fwrite(&oi, 4, 1, f);
		fwrite(&t, 4, 1, f);
		fwrite(normal, 4, 3, f);
		fwrite(&blocked, 4, 1, f);
		fwrite(light, 4, 3, f);
	}
	fclose(f);
	return 0;
}

/* ---- mode: scene ------------------------------------------------------- */

static void fill_material(crt_material *m, const struct Material *src)
{
	memset(m, 0, sizeof *m);
	memcpy(m->ks, src->ks, 12);
	memcpy(m->ka, src->ka, 12);
	memcpy(m->kr, src->kr, 12);
	memcpy(m->kt, src->kt, 12);
	memcpy(m->ke, src->ke, 12);
	m->shininess = src->shininess;
	m->refractive_index = src->refractive_index;
	m->flags = (src->reflective ? CRT_MAT_REFLECTIVE : 0) | (src->transparent ? CRT_MAT_TRANSPARENT : 0) | (src->emittant ? CRT_MAT_EMITTANT : 0);
	m->id = src->id;
	void (*gc)(const struct Texture *, const v3, v3) = src->texture->get_color;
	if (gc == texture_get_color_uniform) {
		const struct RTexUniform *t = (const void *)src->texture;
		m->tex_type = CRT_TEX_UNIFORM;
		memcpy(m->color[0], t->color, 12);
	} else if (gc == texture_get_color_checkerboard) {
		const struct RTexChecker *t = (const void *)src->texture;
		m->tex_type = CRT_TEX_CHECKERBOARD;
		memcpy(m->color, t->colors, 24);
		m->scale = t->scale;
	} else if (gc == texture_get_color_brick) {
		const struct RTexBrick *t = (const void *)src->texture;
		m->tex_type = CRT_TEX_BRICK;
		memcpy(m->color, t->colors, 24);
		m->scale = t->scale;
		m->param1 = t->mortar_width;
	} else {
		const struct RTexNoisy *t = (const void *)src->texture;
		m->tex_type = CRT_TEX_NOISY_PERIODIC;
		memcpy(m->color[0], t->color, 12);
		memcpy(m->color[1], t->color_gradient, 12);
		m->scale = t->noise_feature_scale;
		m->param1 = t->noise_scale;
		m->param2 = t->frequency_scale;
		m->tex_func = (int32_t)t->func;
	}
}

/* out.bin: u32 num_objects, u32 num_materials, f32 ambient[3], f32 camera[3+9+2],
 * crt_material[num_materials], crt_object[num_objects] */
static int mode_scene(int argc, char **argv)
{
	if (argc < 4)
		die("scene: need <scene> <out.bin>");
	/* forward optional flags (e.g. -r) after a fake out/W/H */
	char **fake = calloc((size_t)argc + 4, sizeof *fake);
	int n = 0;
	fake[n++] = argv[0];
	fake[n++] = argv[2];
	fake[n++] = (char *)"none.tif";
	fake[n++] = (char *)"8";
	fake[n++] = (char *)"8";
	for (int i = 4; i < argc; i++)
		fake[n++] = argv[i];
	set_ref_argv(n, fake);
	system_init();
	scene_load();
	FILE *f = fopen(argv[3], "wb");
	if (!f)
		die("cannot open output");
	uint32_t hdr[2] = { (uint32_t)num_objects, (uint32_t)num_materials };
	fwrite(hdr, sizeof hdr, 1, f);
	fwrite(global_ambient_light_intensity, 4, 3, f);
	fwrite(camera.position, 4, 3, f);
	fwrite(camera.vectors, 4, 9, f);
	fwrite(&camera.fov, 4, 1, f);
	fwrite(&camera.focal_length, 4, 1, f);
	for (size_t i = 0; i < num_materials; i++) {
		crt_material m;
		fill_material(&m, &materials[i]);
		fwrite(&m, sizeof m, 1, f);
	}
	for (size_t i = 0; i < num_objects; i++) {
		const struct Object *o = objects[i];
		crt_object c;
		memset(&c, 0, sizeof c);
		c.material = (int32_t)(o->material - materials);
		c.num_lights = o->num_lights;
		c.epsilon = o->epsilon;
		switch (o->object_data->type) {
		case OBJECT_SPHERE: {
			const struct RSphere *s = (const void *)o;
			c.type = CRT_OBJ_SPHERE;
			memcpy(c.a, s->position, 12);
			c.b[0] = s->radius;
		} break;
		case OBJECT_TRIANGLE: {
			const struct RTriangle *t = (const void *)o;
			c.type = CRT_OBJ_TRIANGLE;
			memcpy(c.a, t->vertices[0], 12);
			memcpy(c.b, t->edges[0], 12);
			memcpy(c.c, t->edges[1], 12);
			memcpy(c.n, t->normal, 12);
		} break;
		case OBJECT_PLANE: {
			const struct RPlane *p = (const void *)o;
			c.type = CRT_OBJ_PLANE;
			memcpy(c.a, p->normal, 12);
			c.b[0] = p->d;
		} break;
		}
		fwrite(&c, sizeof c, 1, f);
	}
	fclose(f);
	return 0;
}

/* ---- mode: tex / noise / kat ------------------------------------------- */

static int mode_tex(int argc, char **argv)
{
	if (argc < 6)
		die("tex: need <scene> <material_index> <points.bin> <out.bin>");
	char *fake[] = { argv[0], argv[2], (char *)"none.tif", (char *)"8", (char *)"8" };
	set_ref_argv(5, fake);
	system_init();
	scene_load();
	size_t mi = (size_t)atoi(argv[3]);
	if (mi >= num_materials)
		die("material index out of range");
	size_t len;
	float *pts = read_file(argv[4], &len);
	size_t n = len / 12;
	FILE *f = fopen(argv[5], "wb");
	for (size_t i = 0; i < n; i++) {
		v3 c;
		materials[mi].texture->get_color(materials[mi].texture, pts + 3 * i, c);
		fwrite(c, 4, 3, f);
	}
	fclose(f);
	return 0;
}

static int mode_noise(int argc, char **argv)
{
	if (argc < 4)
		die("noise: need <points.bin> <out.bin>");
	size_t len;
	float *pts = read_file(argv[2], &len);
	size_t n = len / 12;
	FILE *f = fopen(argv[3], "wb");
	for (size_t i = 0; i < n; i++) {
		float v = simplex_noise(pts[3 * i], pts[3 * i + 1], pts[3 * i + 2]);
		fwrite(&v, 4, 1, f);
	}
	fclose(f);
	return 0;
}

/* in.bin: u32 n, then n records of 20 floats:
 *   kind (0 tri, 1 sphere, 2 box), eps, p[12], dir[3], point[3]
 *     tri    : p = v0[3] e0[3] e1[3]
 *     sphere : p = centre[3] radius
 *     box    : p = lo[3] hi[3]
 * out.bin: n records of: i32 hit, f32 t, f32 t2   (box: t = tmin, t2 = tmax) */
static int mode_kat(int argc, char **argv)
{
	if (argc < 4)
		die("kat: need <in.bin> <out.bin>");
	size_t len;
	uint8_t *buf = read_file(argv[2], &len);
	uint32_t n = *(uint32_t *)buf;
	const float *rec = (const float *)(buf + 4);
	FILE *f = fopen(argv[3], "wb");
	for (uint32_t i = 0; i < n; i++) {
		const float *r = rec + (size_t)i * 20;
		const float *p = r + 2;
		struct Ray ray;
		memcpy(ray.direction, r + 14, 12);
		memcpy(ray.point, r + 17, 12);
		int32_t hit = 0;
		float t = 0.f, t2 = 0.f;
		switch ((int)r[0]) {
		case 0: {
			v3 edges[2];
			memcpy(edges, p + 3, 24);
			hit = moller_trumbore(p, edges, ray.point, ray.direction, r[1], &t);
		} break;
		case 1:
			hit = line_intersects_sphere(p, p[3], ray.point, ray.direction, r[1], &t);
			break;
		case 2: {
			struct RBoundingCuboid c;
			c.epsilon = r[1];
			memcpy(c.corners, p, 24);
			hit = bounding_cuboid_intersects(&c, &ray, &t2, &t);
		} break;
		}
		if (!hit && (int)r[0] != 2)
			t = 0.f;
		fwrite(&hit, 4, 1, f);
		fwrite(&t, 4, 1, f);
		fwrite(&t2, 4, 1, f);
	}
	fclose(f);
	return 0;
}

int main(int argc, char **argv)
{
	if (argc < 2)
		die("usage: ref_engine <render|rays|scene|tex|noise|kat> ...");
	if (!strcmp(argv[1], "render"))
		return mode_render(argc, argv);
	if (!strcmp(argv[1], "rays"))
		return mode_rays(argc, argv);
	if (!strcmp(argv[1], "scene"))
		return mode_scene(argc, argv);
	if (!strcmp(argv[1], "tex"))
		return mode_tex(argc, argv);
	if (!strcmp(argv[1], "noise"))
		return mode_noise(argc, argv);
	if (!strcmp(argv[1], "kat"))
		return mode_kat(argc, argv);
	die("unknown mode");
	return 2;
}
