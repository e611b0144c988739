/*
 * main.c -- the `postprocess` executable: same command line, log lines, input
 * and output files as the reference post-processor (reference
 * src/postprocess/main.c:46-71, image.c:30-124), with the operations running
 * on a B200 through crt_postprocess() of include/crt.h.
 *
 *   postprocess <input raw .tif> <output .tif> [-b f] [--dof s b | --dof-camera a f p]
 *               [--mist start depth quad|lin|inv-quad r g b] [--device n] [-f]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "argv.h"
#include "crt.h"
#include "error.h"
#include "image.h"
#include "options.h"
#include "system.h"
#include "tiff.h"

static const char *const HELP_TEXT =
	"Apply post-processing effects to raytraced image.\n"
	"\n"
	"Usage: ./postprocess <input> <output> [OPTIONAL_PARAMETERS]\n"
	"\n"
	"REQUIRED PARAMETERS:\n"
	"<input>      (string)            : .tif raw file created by raytracer.\n"
	"<output>     (string)            : .tif file to which the image will be saved.\n"
	"OPTIONAL PARAMETERS:\n"
	"[-b] (float)                     : DEFAULT = 1.0     : brighten.\n"
	"[--dof] <scale> <bias>           : Apply depth of field effect (incompatible with --dof-camera).\n"
	"[--dof-camera] <aperture> <focal length> <plane in focus> : Apply depth of field effect (incompatible with --dof).\n"
	"[--mist] <start> <depth> <falloff> <color> : Apply colored mist based on distance.\n"
	"  <falloff> (\"quad\"|\"lin\"|\"inv-quad\") : Decay rate.\n"
	"  <color> (float) (float) (float) : \n"
	"B200 BUILD, ADDITIONAL PARAMETERS:\n"
	"[--device] (integer)             : DEFAULT = 0       : CUDA device the operations run on.\n"
	"[-f]                             : DEFAULT = OFF     : save the processed frame raw (FP32 + z-buffer) instead of 8-bit.\n";

int main(int argc, char *argv[])
{
	struct rt_args args;
	if (rt_args_init(&args, argc, argv))
		return 1;
	if (rt_args_find(&args, "--help") || rt_args_find(&args, "-h")) {
		puts(HELP_TEXT);
		return 0;
	}
	if (argc < 3) {
		puts("Too few arguments. Use --help to find out which arguments are required to call this program.");
		return 1;
	}
	rt_system_init(&args);
	char err[RT_ERRLEN] = "";

	rt_log("Loading image.");
	struct rt_tiff_image in;
	if (rt_tiff_read(argv[1], &in, err))
		rt_fatal("%s", err);
	/* the checks of postprocess/image.c:48-50, :68 */
	if (in.bits_per_sample != 32 || !in.raster)
		rt_fatal("Expected 32 bits per sample but got [%u].", in.bits_per_sample);
	if (!in.z_buffer || (uint64_t)in.z_count != (uint64_t)in.width * in.height)
		rt_fatal("Corrupted Z-Buffer.");

	crt_post_opts opts;
	if (rt_post_options(&args, &opts, err))
		rt_fatal("%s", err);
	int device = 0;
	const int d = rt_args_find_with_values(&args, "--device", 1);
	if (d)
		device = atoi(argv[d + 1]);

	rt_log("Commencing Postprocessing");
	if (opts.ops & CRT_POST_BRIGHTEN)
		rt_log("Brightening by factor %f.", (double)opts.brighten);
	crt_post_stats stats;
	if (crt_postprocess(device, &opts, in.width, in.height, in.raster, in.z_buffer, &stats) != CRT_OK)
		rt_fatal("%s", crt_last_error());
	if (opts.ops & (CRT_POST_DOF | CRT_POST_DOF_CAMERA))
		rt_log("Applying depth of field with scale [%f] and bias [%f].", (double)stats.dof_scale, (double)stats.dof_bias);
	rt_log("Postprocessed on the GPU in %.3f ms (%.3f ms with copies).", stats.kernel_ms, stats.total_ms);

	rt_log("Saving image.");
	const char *filename = argv[2];
	if (!strstr(filename, ".tif"))
		rt_log("Expected output file [%s] with extension .tif.", filename);
	int rc;
	if (rt_args_find(&args, "-f")) {
		rc = rt_tiff_write_raw(filename, in.width, in.height, in.raster, in.z_buffer, err);
	} else {
		const size_t samples = (size_t)in.width * in.height * 3;
		uint8_t *rgb = malloc(samples ? samples : 1);
		if (!rgb)
			rt_fatal("Out of memory.");
		rt_image_quantize(in.raster, samples, rgb); /* postprocess/image.c:104-110 */
		rc = rt_tiff_write_rgb8(filename, in.width, in.height, rgb, err);
		free(rgb);
	}
	if (rc)
		rt_fatal("%s", err);

	rt_log("Terminating.");
	rt_tiff_free(&in);
	rt_args_free(&args);
	return 0;
}
