#include "material.h"

#include "calc.h"

#define RT_MATERIAL_THRESHOLD 1e-6f

void rt_material_finalize(crt_material *m)
{
	m->flags = 0;
	if (v3_len(m->kr) > RT_MATERIAL_THRESHOLD)
		m->flags |= CRT_MAT_REFLECTIVE;
	if (v3_len(m->kt) > RT_MATERIAL_THRESHOLD)
		m->flags |= CRT_MAT_TRANSPARENT;
	if (v3_len(m->ke) > RT_MATERIAL_THRESHOLD)
		m->flags |= CRT_MAT_EMITTANT;
}

int rt_material_find(const crt_material *materials, size_t count, int id)
{
	for (size_t i = 0; i < count; i++)
		if (materials[i].id == id)
			return (int)i;
	return -1;
}
