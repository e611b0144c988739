/*
 * calc.h -- float 3-vector helpers for the host code (header-only inlines).
 * Evaluation order follows reference src/core/calc.c so that set-up values
 * (camera basis, image plane, triangle normals, epsilons) round identically:
 * dot = x*x' + y*y' + z*z' left to right (calc.c:36), normalise multiplies by
 * 1/sqrtf (calc.c:139-142).  Build the host with -ffp-contract=off.
 */
#ifndef RT_CALC_H
#define RT_CALC_H

#include <math.h>

#include "type.h"

static inline float v3_dot(const float *a, const float *b)
{
	return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

static inline float v3_len(const float *a)
{
	return sqrtf(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]);
}

static inline void v3_set(float *d, float x, float y, float z)
{
	d[0] = x;
	d[1] = y;
	d[2] = z;
}

static inline void v3_copy(float *d, const float *s)
{
	d[0] = s[0];
	d[1] = s[1];
	d[2] = s[2];
}

static inline void v3_add(float *d, const float *a, const float *b)
{
	d[0] = a[0] + b[0];
	d[1] = a[1] + b[1];
	d[2] = a[2] + b[2];
}

static inline void v3_sub(float *d, const float *a, const float *b)
{
	d[0] = a[0] - b[0];
	d[1] = a[1] - b[1];
	d[2] = a[2] - b[2];
}

static inline void v3_scale(float *d, const float *a, float s)
{
	d[0] = a[0] * s;
	d[1] = a[1] * s;
	d[2] = a[2] * s;
}

static inline void v3_cross(float *d, const float *a, const float *b)
{
	const float x = a[1] * b[2] - a[2] * b[1];
	const float y = a[2] * b[0] - a[0] * b[2];
	const float z = a[0] * b[1] - a[1] * b[0];
	d[0] = x;
	d[1] = y;
	d[2] = z;
}

static inline void v3_normalize(float *a)
{
	const float inv = 1.f / v3_len(a);
	v3_scale(a, a, inv);
}

/* d = M a, M row-major 3x3 (calc.c:182-187) */
static inline void m3_apply(float *d, const float m[3][3], const float *a)
{
	const float x = v3_dot(m[0], a);
	const float y = v3_dot(m[1], a);
	const float z = v3_dot(m[2], a);
	d[0] = x;
	d[1] = y;
	d[2] = z;
}

#endif
