#include "object.h"

#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "calc.h"
#include "error.h"

static void object_base(crt_object *o, int type, int material, float epsilon, unsigned lights)
{
	memset(o, 0, sizeof *o);
	o->type = type;
	o->material = material;
	o->epsilon = epsilon;
	o->num_lights = lights;
}

void rt_sphere_make(crt_object *o, const float centre[3], float radius, int material, float epsilon, unsigned lights)
{
	object_base(o, CRT_OBJ_SPHERE, material, epsilon, lights);
	v3_copy(o->a, centre);
	o->b[0] = radius;
	if (epsilon == RT_EPSILON_AUTO)
		o->epsilon = radius * 0.0003f;
}

void rt_triangle_make(crt_object *o, const float v[3][3], int material, float epsilon, unsigned lights)
{
	object_base(o, CRT_OBJ_TRIANGLE, material, epsilon, lights);
	v3_copy(o->a, v[0]);
	v3_sub(o->b, v[1], v[0]);
	v3_sub(o->c, v[2], v[0]);
	v3_cross(o->n, o->b, o->c);
	v3_normalize(o->n);
	if (epsilon == RT_EPSILON_AUTO) {
		/* 0.003 * area^(3/4), the area written as |b||c| sin(angle)/2 */
		const float len_prod = v3_len(o->b) * v3_len(o->c);
		const float angle = acosf(v3_dot(o->b, o->c) / len_prod);
		o->epsilon = 0.003f * powf(0.5f * len_prod * sinf(angle), 0.75f);
	}
}

void rt_plane_make(crt_object *o, const float point[3], const float normal[3], int material, float epsilon, unsigned lights)
{
	object_base(o, CRT_OBJ_PLANE, material, epsilon, lights);
	v3_copy(o->a, normal);
	v3_normalize(o->a);
	o->b[0] = v3_dot(o->a, point);
	if (epsilon == RT_EPSILON_AUTO)
		o->epsilon = 1.e-6f;
}

void rt_object_corners(const crt_object *o, float lo[3], float hi[3])
{
	if (o->type == CRT_OBJ_SPHERE) {
		for (int k = 0; k < 3; k++) {
			lo[k] = o->a[k] - o->b[0];
			hi[k] = o->a[k] + o->b[0];
		}
		return;
	}
	/* triangle: the three vertices are a, a+b, a+c */
	for (int k = 0; k < 3; k++) {
		const float p0 = o->a[k], p1 = o->a[k] + o->b[k], p2 = o->a[k] + o->c[k];
		lo[k] = fminf(p0, fminf(p1, p2));
		hi[k] = fmaxf(p0, fmaxf(p1, p2));
	}
}

void rt_objects_extents(const crt_object *objects, size_t count, float lo[3], float hi[3])
{
	v3_set(lo, FLT_MAX, FLT_MAX, FLT_MAX);
	v3_set(hi, FLT_MIN, FLT_MIN, FLT_MIN); /* sic: smallest positive, as the reference */
	for (size_t i = 0; i < count; i++) {
		if (objects[i].type == CRT_OBJ_PLANE)
			continue;
		float a[3], b[3];
		rt_object_corners(&objects[i], a, b);
		for (int k = 0; k < 3; k++) {
			if (a[k] < lo[k])
				lo[k] = a[k];
			if (b[k] > hi[k])
				hi[k] = b[k];
		}
	}
}

void rt_object_scale(crt_object *o, float scale)
{
	switch (o->type) {
	case CRT_OBJ_SPHERE:
		o->epsilon *= scale;
		o->b[0] *= scale;
		v3_scale(o->a, o->a, scale);
		break;
	case CRT_OBJ_TRIANGLE:
		o->epsilon *= scale;
		v3_scale(o->a, o->a, scale);
		v3_scale(o->b, o->b, scale);
		v3_scale(o->c, o->c, scale);
		break;
	case CRT_OBJ_PLANE: {
		/* move one point of the plane and recompute d; the axis search uses
		 * the unscaled epsilon, which is scaled last */
		const float eps = o->epsilon;
		float p[3] = { 1.f, 1.f, 1.f };
		int k = 0;
		while (k < 3 && !(fabsf(o->a[k]) > eps))
			k++;
		if (k == 3)
			k = 2;
		p[k] = 0.f;
		p[k] = (o->b[0] - v3_dot(p, o->a)) / o->a[k];
		v3_scale(p, p, scale);
		o->b[0] = v3_dot(o->a, p);
		o->epsilon *= scale;
	} break;
	}
}

/* ---- binary STL -------------------------------------------------------- */

#define STL_HEADER_BYTES 80
#define STL_RECORD_BYTES 50

static void euler_zyx(float m[3][3], const float r[3])
{
	const float sx = sinf(r[0]), cx = cosf(r[0]);
	const float sy = sinf(r[1]), cy = cosf(r[1]);
	const float sz = sinf(r[2]), cz = cosf(r[2]);
	const float czsy = cz * sy, szsy = sz * sy;
	m[0][0] = cz * cy;
	m[0][1] = czsy * sx - sz * cx;
	m[0][2] = czsy * cx + sz * sx;
	m[1][0] = sz * cy;
	m[1][1] = szsy * sx + cz * cx;
	m[1][2] = szsy * cx - cz * sx;
	m[2][0] = -sy;
	m[2][1] = cy * sx;
	m[2][2] = cy * cx;
}

int rt_stl_load(const char *filename, const float position[3], const float rotation[3], float scale,
	int material, float epsilon, unsigned lights, crt_object **out, size_t *count_out, char *err)
{
	*out = NULL;
	*count_out = 0;
	FILE *f = fopen(filename, "rb");
	if (!f)
		rt_fail(err, "Failed to open mesh file %s.", filename);

	unsigned char header[STL_HEADER_BYTES + 4];
	if (fread(header, 1, sizeof header, f) != sizeof header) {
		fclose(f);
		rt_fail(err, "Failed to read header of mesh file [%s].", filename);
	}
	if (!memcmp(header, "solid", 5)) {
		fclose(f);
		rt_fail(err, "Mesh file [%s] does not use binary encoding.", filename);
	}
	uint32_t n;
	memcpy(&n, header + STL_HEADER_BYTES, 4);
	/* the triangle count is the file's own claim: hold it against the file's size before anything is allocated for it */
	{
		long size = -1;
		if (!fseek(f, 0, SEEK_END))
			size = ftell(f);
		if (size < 0 || fseek(f, (long)sizeof header, SEEK_SET)) {
			fclose(f);
			rt_fail(err, "Failed to read triangle in mesh file [%s].", filename);
		}
		if ((uint64_t)n * STL_RECORD_BYTES > (uint64_t)size - sizeof header) {
			fclose(f);
			rt_fail(err, "Failed to read triangle in mesh file [%s].", filename);
		}
	}

	float rot[3][3];
	euler_zyx(rot, rotation);

	crt_object *tris = malloc(sizeof *tris * (n ? n : 1));
	unsigned char *raw = malloc((size_t)STL_RECORD_BYTES * (n ? n : 1));
	if (!tris || !raw) {
		free(tris);
		free(raw);
		fclose(f);
		rt_fail(err, "Unable to allocate memory for mesh file [%s].", filename);
	}
	if (fread(raw, STL_RECORD_BYTES, n, f) != n) {
		free(tris);
		free(raw);
		fclose(f);
		rt_fail(err, "Failed to read triangle in mesh file [%s].", filename);
	}
	fclose(f);

	for (uint32_t i = 0; i < n; i++) {
		float file_v[3][3], v[3][3];
		memcpy(file_v, raw + (size_t)STL_RECORD_BYTES * i + 12, sizeof file_v); /* skip the stored normal */
		for (int k = 0; k < 3; k++) {
			float t[3];
			m3_apply(t, rot, file_v[k]);
			v3_scale(t, t, scale);
			v3_add(v[k], t, position);
		}
		rt_triangle_make(&tris[i], v, material, epsilon, lights);
	}
	free(raw);
	*out = tris;
	*count_out = n;
	return 0;
}
