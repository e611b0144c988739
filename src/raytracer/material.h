/*
 * material.h -- material records are stored directly as crt_material
 * (include/crt.h), the layout handed to the device library.
 */
#ifndef RT_MATERIAL_H
#define RT_MATERIAL_H

#include "crt.h"

/* derive the reflective / transparent / emittant flags: |k| > 1e-6
 * (reference material.c:25, :81-83) */
void rt_material_finalize(crt_material *m);

/* index of the material with scene-file id `id`, first match, -1 if none
 * (reference get_material, material.c:94-102) */
int rt_material_find(const crt_material *materials, size_t count, int id);

#endif
