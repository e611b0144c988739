/*
 * options.h -- the post-processor's command line as a crt_post_opts
 * (reference src/postprocess/postproc.c:36-92 parses the same flags inline).
 */
#ifndef RT_POST_OPTIONS_H
#define RT_POST_OPTIONS_H

#include "argv.h"
#include "crt.h"

/* -b <factor>, --dof <scale> <bias>, --dof-camera <aperture> <focal length> <plane in focus>,
 * --mist <start> <depth> quad|lin|inv-quad <r> <g> <b>.  --dof wins over --dof-camera
 * (postproc.c:47-52).  0 on success; -1 with a message for an unknown falloff. */
int rt_post_options(const struct rt_args *args, crt_post_opts *opts, char *err);

#endif
