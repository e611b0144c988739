/*
 * ref_post.c -- TEST INFRASTRUCTURE (oracle).  Drives the UNMODIFIED reference
 * post-processor (src/postprocess/image.c image_load(), postproc.c
 * postprocess()) and dumps the FP32 raster it leaves in `image.raster`: the
 * reference executable itself only saves the 8-bit quantisation
 * (postprocess/image.c:84-124).
 *
 *   ref_post <input raw .tif> <output .f32> [-b f] [--dof s b | --dof-camera a f p] [--mist s d falloff r g b]
 *
 * Output: width (u32), height (u32), then width*height*3 floats.
 */
#include <stdint.h>
#include <stdio.h>

#include "argv.h"
#include "image.h"
#include "postproc.h"
#include "system.h"

int main(int argc, char *argv[])
{
	if (argc < 3)
		return 2;
	myargc = argc;
	myargv = argv;
	argv_init();
	system_init();
	image_load();
	postprocess();
	FILE *f = fopen(argv[2], "wb");
	if (!f)
		return 1;
	const uint32_t dims[2] = { image.resolution[0], image.resolution[1] };
	fwrite(dims, sizeof dims, 1, f);
	fwrite(image.raster, sizeof(float) * 3, image.pixels, f);
	fclose(f);
	argv_deinit();
	image_deinit();
	return 0;
}
