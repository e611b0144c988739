/*
 * type.h -- common scalar/vector typedefs of the host code.
 * (same roles as reference src/core/type.h:19-38; PI keeps the reference's
 * single-precision literal because image-plane set-up and the samplers use it)
 */
#ifndef RT_TYPE_H
#define RT_TYPE_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

typedef float vec3f[3];

#define RT_PI 3.1415927f
#define RT_ARRAY_LEN(a) (sizeof(a) / sizeof((a)[0]))

#endif
