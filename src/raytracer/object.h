/*
 * object.h -- construction of scene objects as crt_object records
 * (include/crt.h) and the binary-STL mesh loader.
 */
#ifndef RT_OBJECT_H
#define RT_OBJECT_H

#include <stddef.h>

#include "crt.h"

#define RT_EPSILON_AUTO (-1.f)

/* Constructors resolve epsilon == RT_EPSILON_AUTO with the reference's rules:
 * sphere 0.0003 r (object.c:235-236), triangle 0.003 area^0.75
 * (object.c:336-339), plane 1e-6 (object.c:453-454). */
void rt_sphere_make(crt_object *o, const float centre[3], float radius, int material, float epsilon, unsigned lights);
void rt_triangle_make(crt_object *o, const float v[3][3], int material, float epsilon, unsigned lights);
void rt_plane_make(crt_object *o, const float point[3], const float normal[3], int material, float epsilon, unsigned lights);

/* Binary STL: 80-byte header, u32 count, 50-byte records; normals and
 * attribute bytes ignored; rejected if the file starts with "solid"
 * (object.c:45-49, :542-587).  v' = (Rz Ry Rx v) * scale + position.
 * Appends `*count_out` triangles to a malloc'ed array returned in *out. */
int rt_stl_load(const char *filename, const float position[3], const float rotation[3], float scale,
	int material, float epsilon, unsigned lights, crt_object **out, size_t *count_out, char *err);

/* axis-aligned corners of a bounded object (object.c:267-272, :375-388) */
void rt_object_corners(const crt_object *o, float lo[3], float hi[3]);
/* extents over bounded objects, with the reference's FLT_MIN initial maximum
 * (object.c:200-225) */
void rt_objects_extents(const crt_object *objects, size_t count, float lo[3], float hi[3]);
/* uniform rescale about the origin (object.c:274-281, :390-401, :500-514) */
void rt_object_scale(crt_object *o, float scale);

#endif
