/*
 * scene.h -- JSON scene file -> flat scene (schema: SURVEY A.1, enforced by
 * reference src/raytracer/scene.c:70-470).
 */
#ifndef RT_SCENE_H
#define RT_SCENE_H

#include <stddef.h>

#include "camera.h"
#include "crt.h"

struct rt_scene {
	struct rt_camera camera;
	crt_material *materials;
	size_t num_materials;
	crt_object *objects; /* scene order, meshes expanded in place */
	size_t num_objects;
	size_t num_emitters;
	size_t num_planes;
	float ambient[3];
	uint32_t build; /* enum crt_build (--bvh host | device); 0 after rt_scene_load() */
};

/* rescale: NULL, "norm" (1 / largest extent) or a decimal factor: the -r flag
 * (scene.c:109-121).  Mesh file names are resolved relative to the current
 * working directory.  Returns 0 or -1 with a message in err
 * (RT_ERRLEN bytes). */
int rt_scene_load(const char *filename, const char *rescale, struct rt_scene *scene, char *err);
void rt_scene_free(struct rt_scene *scene);
void rt_scene_scale(struct rt_scene *scene, float factor);
/* view of the scene as the device library's input */
void rt_scene_desc(const struct rt_scene *scene, crt_scene_desc *desc);

#endif
