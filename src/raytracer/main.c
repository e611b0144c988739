/*
 * main.c -- the `engine` executable: same command line, log lines, scene
 * files and output files as the reference engine (reference
 * src/raytracer/main.c:23-89), with the trace loop running on a B200 through
 * include/crt.h.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "argv.h"
#include "error.h"
#include "image.h"
#include "render.h"
#include "scene.h"
#include "system.h"

static const char *const HELP_TEXT =
	"Render a scene using raytracing.\n"
	"Copyright: (c) 2021-2022 Wojciech Graj\n"
	"License: https://opensource.org/licenses/MIT\n"
	"Features: Multithreading Planes \n"
	"Usage: ./engine <input> <output> <resolution> [OPTIONAL_PARAMETERS]\n"
	"\n"
	"REQUIRED PARAMETERS:\n"
	"<input>      (string)            : .json scene file which will be used to generate the image. Example files can be found in ./scenes.\n"
	"<output>     (string)            : .tif file to which the image will be saved.\n"
	"<resolution> (integer) (integer) : resolution of the output image.\n"
	"OPTIONAL PARAMETERS:\n"
	"[-m] (integer | \"max\")           : DEFAULT = 1       : number of CPU cores\n"
	"[-b] (integer)                   : DEFAULT = 10      : maximum number of times that a light ray can bounce.\n"
	"[-a] (float)                     : DEFAULT = 0.01    : minimum light intensity for which a ray is cast.\n"
	"[-s] (\"phong\" | \"blinn\")         : DEFAULT = phong   : reflection model.\n"
	"[-n] (integer)                   : DEFAULT = 1       : number of samples which are rendered per pixel.\n"
	"[-r] (\"norm\" | float)            : DEFAULT = 1.0     : scene scaling factor.\n"
	"[-l] (\"none\" | \"lin\" | \"sqr\")    : DEFAULT = sqr     : light attenuation.\n"
	"[-p] (\"real\" | \"cpu\")            : DEFAULT = real    : time to print with status messages.\n"
	"[-g] (string)                    : DEFAULT = ambient : global illumination model.\n"
	"    ambient    : ambient lighting\n"
	"    path       : path-tracing\n"
	"[-f]                             : DEFAULT = OFF     : save raw output for post-processing.\n";

static const char *const HELP_EXTRA =
	"B200 BUILD, ADDITIONAL PARAMETERS:\n"
	"[--device] (integer)             : DEFAULT = 0       : CUDA device the trace loop runs on ([-m] is accepted and ignored).\n"
	"[--gpus] (integer | \"max\")       : DEFAULT = 1       : number of CUDA devices that share the frame (devices 0 .. N-1).\n"
	"[--bvh] (host | device)          : DEFAULT = host    : where the acceleration structure is built: binned SAH on a host core, or Morton codes + radix tree on the device.\n"
	"[--devices] (list)               : DEFAULT = none    : comma-separated CUDA devices that share the frame, e.g. 0,2,3.\n"
	"[--lens-radius] (decimal)        : DEFAULT = 0       : thin-lens camera: radius of the lens (0 = the reference's pinhole).\n"
	"[--focus] (decimal)              : DEFAULT = 0       : thin-lens camera: distance of the plane in focus.\n"
	"[--lens-samples] (integer)       : DEFAULT = 1       : thin-lens camera: primary rays per pixel.\n"
	"[--fog] (decimal)                : DEFAULT = 0       : in-tracer fog: extinction per unit length along every eye-path segment.\n"
	"[--fog-color] (r,g,b)            : DEFAULT = 0,0,0   : in-tracer fog: colour gained where light is lost.\n"
	"[--seed] (integer)               : DEFAULT = 1       : seed of the counter-based sampling stream.\n"
	"[--stats]                        : DEFAULT = OFF     : count rays and intersection tests, print Mrays/s.\n"
	"[--stats-json] (string)          : DEFAULT = none    : file the same counters are written to as JSON (rays by class, tests, ms, Mrays/s, FP32 roofline fraction).\n";

/* --stats-json: the counters of the frame, machine readable.  Algorithmic FLOPs by the constants of the scope contract
 * (3 per ray, 12 per box test, triangle 14 / 24 / 40 / 46 by exit stage, sphere 17 / 20, plane 12); the peak is the FFMA
 * rate measured on the same device (crt_measure_fp32_peak). */
static int write_stats_json(const char *path, const char *scene_file, const crt_render_opts *o, const crt_stats *st,
	const crt_scene_info *info, const int *devices, const double *device_ms, int num_devices, int dynamic)
{
	FILE *f = fopen(path, "w");
	if (!f)
		return 1;
	const double rays = (double)(st->rays_closest + st->rays_shadow);
	const double flops = 3.0 * rays + 12.0 * (double)st->box_tests + 14.0 * (double)st->tri_tests[0] + 24.0 * (double)st->tri_tests[1] +
		40.0 * (double)st->tri_tests[2] + 46.0 * (double)st->tri_tests[3] + 17.0 * (double)st->sphere_tests[0] +
		20.0 * (double)st->sphere_tests[1] + 12.0 * (double)st->plane_tests;
	double peak = 0.0;
	if (crt_measure_fp32_peak((int)info->device, &peak) != CRT_OK)
		peak = 0.0;
	const double tflops = st->kernel_ms > 0.0 ? flops / (st->kernel_ms * 1e9) : 0.0;
	fprintf(f, "{\"scene\": \"");
	for (const char *c = scene_file; *c; c++) { /* a path: escape the two characters JSON cannot take raw */
		if (*c == '"' || *c == '\\')
			fputc('\\', f);
		fputc(*c, f);
	}
	fprintf(f, "\", \"width\": %u, \"height\": %u, \"samples_per_pixel\": %u, \"seed\": %u,\n", o->width, o->height, o->samples_per_pixel, o->seed);
	fprintf(f, " \"rays_closest\": %llu, \"rays_shadow\": %llu, \"nodes_shaded\": %llu,\n", (unsigned long long)st->rays_closest,
		(unsigned long long)st->rays_shadow, (unsigned long long)st->nodes_shaded);
	fprintf(f, " \"box_tests\": %llu, \"tri_tests\": [%llu, %llu, %llu, %llu], \"sphere_tests\": [%llu, %llu], \"plane_tests\": %llu,\n",
		(unsigned long long)st->box_tests, (unsigned long long)st->tri_tests[0], (unsigned long long)st->tri_tests[1],
		(unsigned long long)st->tri_tests[2], (unsigned long long)st->tri_tests[3], (unsigned long long)st->sphere_tests[0],
		(unsigned long long)st->sphere_tests[1], (unsigned long long)st->plane_tests);
	fprintf(f, " \"node_steps\": %llu, \"node_step_lanes\": %llu, \"node_step_idle\": %llu, \"prim_steps\": %llu, \"prim_step_lanes\": %llu, \"prim_step_idle\": %llu,\n",
		(unsigned long long)st->node_steps, (unsigned long long)st->node_step_lanes, (unsigned long long)st->node_step_idle,
		(unsigned long long)st->prim_steps, (unsigned long long)st->prim_step_lanes, (unsigned long long)st->prim_step_idle);
	fprintf(f, " \"kernel_ms\": %.6f, \"total_ms\": %.6f, \"queue_ms\": %.6f, \"last_warp_ms\": %.6f, \"warp_busy_ms\": %.6f,\n", st->kernel_ms,
		st->total_ms, st->queue_ms, st->last_warp_ms, st->warp_busy_ms);
	fprintf(f, " \"mrays_per_s\": %.6f, \"box_tests_per_ray\": %.6f, \"flops\": %.0f, \"flops_per_ray\": %.6f,\n",
		st->kernel_ms > 0.0 ? rays / (st->kernel_ms * 1e3) : 0.0, rays > 0.0 ? (double)st->box_tests / rays : 0.0, flops,
		rays > 0.0 ? flops / rays : 0.0);
	fprintf(f, " \"roofline\": {\"bound\": \"fp32\", \"achieved\": %.6f, \"peak\": %.6f, \"unit\": \"TFLOP/s\", \"frac\": %.6f, \"note\": \"kernel_ms is the instrumented instance's (slower than the plain one)\"},\n",
		tflops, peak, peak > 0.0 ? tflops / peak : 0.0);
	fprintf(f, " \"bvh\": {\"nodes\": %u, \"primitives\": %u, \"planes\": %u, \"depth\": %u, \"built_on\": \"%s\", \"build_ms\": %.6f, \"upload_ms\": %.6f, \"device_bytes\": %llu},\n",
		info->num_nodes, info->num_bounded, info->num_planes, info->max_depth, info->built_on_device ? "device" : "host", info->build_ms,
		info->upload_ms, (unsigned long long)info->device_bytes);
	fprintf(f, " \"emitters\": %u, \"light_samples\": %u, \"sm_count\": %u,\n", info->num_emitters, info->total_lights, info->sm_count);
	fprintf(f, " \"devices\": [");
	for (int k = 0; k < num_devices; k++)
		fprintf(f, "%s{\"device\": %d, \"kernel_ms\": %.6f}", k ? ", " : "", devices[k], device_ms[k]);
	fprintf(f, "], \"work_queue\": \"%s\"}\n", num_devices > 1 ? (dynamic ? "shared" : "static shards") : "single device");
	return fclose(f) != 0;
}

int main(int argc, char *argv[])
{
	struct rt_args args;
	if (rt_args_init(&args, argc, argv))
		return 1;

	if (rt_args_find(&args, "--help") || rt_args_find(&args, "-h")) {
		puts(HELP_TEXT);
		puts(HELP_EXTRA);
		return 0;
	}
	if (argc < 5) {
		puts("Too few arguments. Use --help to find out which arguments are required to call this program.");
		return 1;
	}

	rt_system_init(&args);
	char err[RT_ERRLEN] = "";

	rt_log("Loading scene.");
	struct rt_scene scene;
	const int r = rt_args_find_with_values(&args, "-r", 1);
	if (rt_scene_load(argv[RT_ARG_SCENE], r ? argv[r + 1] : NULL, &scene, err))
		rt_fatal("%s", err);
	rt_log("Loaded %zu materials, %zu objects.", scene.num_materials, scene.num_objects);
	const int bv = rt_args_find_with_values(&args, "--bvh", 1);
	if (bv && !strcmp(argv[bv + 1], "device"))
		scene.build = CRT_BUILD_DEVICE;

	rt_log("Initializing image.");
	struct rt_image image;
	rt_image_plane(&image, &scene.camera, (uint32_t)abs(atoi(argv[RT_ARG_WIDTH])), (uint32_t)abs(atoi(argv[RT_ARG_HEIGHT])));
	if (rt_image_alloc(&image, err))
		rt_fatal("%s", err);

	crt_render_opts opts;
	rt_render_defaults(&opts);
	rt_render_parse(&opts, &args);
	rt_render_set_view(&opts, &scene, &image);
	const int d = rt_args_find_with_values(&args, "--device", 1);
	const int device = d ? atoi(argv[d + 1]) : 0;

	/* --gpus N / --devices a,b,c: several devices drain one work queue (crt_multi_render) */
	int devices[64], num_devices = 0;
	const int dl = rt_args_find_with_values(&args, "--devices", 1);
	const int gn = rt_args_find_with_values(&args, "--gpus", 1);
	if (dl) {
		const char *c = argv[dl + 1];
		while (*c && num_devices < 64) {
			devices[num_devices++] = atoi(c);
			while (*c && *c != ',')
				c++;
			if (*c == ',')
				c++;
		}
	} else if (gn) {
		int n = !strcmp(argv[gn + 1], "max") ? crt_device_count() : atoi(argv[gn + 1]);
		if (n > 64)
			n = 64;
		for (int k = 0; k < n; k++)
			devices[num_devices++] = k;
	}

	rt_log("Commencing raytracing.");
	crt_stats stats;
	crt_scene_info info;
	double device_ms[64];
	int dynamic = 0;
	if (num_devices > 1) {
		rt_log("Using %d CUDA devices.", num_devices);
		if (rt_render_multi(&scene, &image, &opts, devices, num_devices, &stats, &info, device_ms, &dynamic, err))
			rt_fatal("%s", err);
		for (int k = 0; k < num_devices; k++)
			rt_log("Device %d: trace kernel %.3f ms (%s).", devices[k], device_ms[k], dynamic ? "shared work queue" : "static shard");
	} else if (rt_render(&scene, &image, &opts, num_devices == 1 ? devices[0] : device, &stats, &info, err)) {
		rt_fatal("%s", err);
	}
	rt_log("BVH: %u nodes over %u primitives, depth %u, built on the %s in %.1f ms; trace kernel %.3f ms, call %.3f ms.",
		info.num_nodes, info.num_bounded, info.max_depth, info.built_on_device ? "device" : "host", info.build_ms, stats.kernel_ms, stats.total_ms);
	if (opts.flags & CRT_RENDER_COUNT_STATS) {
		const double rays = (double)(stats.rays_closest + stats.rays_shadow);
		rt_log("Rays: %llu closest-hit + %llu shadow = %.1f Mrays/s; %.1f box tests per ray.",
			(unsigned long long)stats.rays_closest, (unsigned long long)stats.rays_shadow,
			rays / (stats.kernel_ms * 1e3), rays > 0 ? (double)stats.box_tests / rays : 0.0);
	}
	const int sj = rt_args_find_with_values(&args, "--stats-json", 1);
	if (sj) {
		if (num_devices <= 1) {
			devices[0] = num_devices == 1 ? devices[0] : device;
			device_ms[0] = stats.kernel_ms;
		}
		if (write_stats_json(argv[sj + 1], argv[RT_ARG_SCENE], &opts, &stats, &info, devices, device_ms, num_devices > 1 ? num_devices : 1, dynamic))
			rt_fatal("Failed to write statistics to [%s].", argv[sj + 1]);
		rt_log("Statistics written to [%s].", argv[sj + 1]);
	}

	rt_log("Saving image.");
	const char *out = argv[RT_ARG_OUTPUT];
	if (!strstr(out, ".tif"))
		rt_log("Expected output file [%s] with extension .tif.", out);
	if (rt_image_save(&image, out, rt_args_find(&args, "-f") != 0, err))
		rt_fatal("%s", err);

	rt_log("Terminating.");
	rt_image_free(&image);
	rt_scene_free(&scene);
	rt_args_free(&args);
	return 0;
}
