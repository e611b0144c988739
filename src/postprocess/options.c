/*
 * options.c -- see options.h.  Values go through atof() like the reference's
 * (postproc.c:43-71), so "1e-2" or trailing junk parse the same way.
 */
#include "options.h"

#include <stdlib.h>
#include <string.h>

#include "error.h"

int rt_post_options(const struct rt_args *args, crt_post_opts *opts, char *err)
{
	memset(opts, 0, sizeof *opts);
	char **argv = args->argv;
	int i = rt_args_find_with_values(args, "-b", 1);
	if (i) {
		opts->ops |= CRT_POST_BRIGHTEN;
		opts->brighten = (float)atof(argv[i + 1]);
	}
	if ((i = rt_args_find_with_values(args, "--dof", 2))) {
		opts->ops |= CRT_POST_DOF;
		opts->dof_scale = (float)atof(argv[i + 1]);
		opts->dof_bias = (float)atof(argv[i + 2]);
	} else if ((i = rt_args_find_with_values(args, "--dof-camera", 3))) {
		opts->ops |= CRT_POST_DOF_CAMERA;
		opts->dof_aperture = (float)atof(argv[i + 1]);
		opts->dof_focal_length = (float)atof(argv[i + 2]);
		opts->dof_plane_in_focus = (float)atof(argv[i + 3]);
	}
	if ((i = rt_args_find_with_values(args, "--mist", 6))) {
		opts->ops |= CRT_POST_MIST;
		opts->mist_start = (float)atof(argv[i + 1]);
		opts->mist_depth = (float)atof(argv[i + 2]);
		const char *falloff = argv[i + 3];
		if (!strcmp(falloff, "quad"))
			opts->mist_falloff = CRT_FALLOFF_QUADRATIC;
		else if (!strcmp(falloff, "lin"))
			opts->mist_falloff = CRT_FALLOFF_LINEAR;
		else if (!strcmp(falloff, "inv-quad"))
			opts->mist_falloff = CRT_FALLOFF_INV_QUADRATIC;
		else
			rt_fail(err, "Unrecognized falloff type [%s].", falloff);
		opts->mist_color[0] = (float)atof(argv[i + 4]);
		opts->mist_color[1] = (float)atof(argv[i + 5]);
		opts->mist_color[2] = (float)atof(argv[i + 6]);
	}
	return 0;
}
