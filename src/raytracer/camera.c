#include "camera.h"

#include "calc.h"
#include "error.h"

int rt_camera_init(struct rt_camera *cam, const float position[3], const float x[3], const float y[3],
	float fov, float focal_length, char *err)
{
	if (!(fov > 0.f && fov < 180.f))
		rt_fail(err, "Expected camera fov [%.2f] between [0.] and [180.].", (double)fov);
	cam->fov = fov;
	cam->focal_length = focal_length;
	v3_copy(cam->position, position);
	v3_copy(cam->axis[0], x);
	v3_copy(cam->axis[1], y);
	v3_normalize(cam->axis[0]);
	v3_normalize(cam->axis[1]);
	v3_cross(cam->axis[2], cam->axis[0], cam->axis[1]);
	return 0;
}

/* uniform rescale about the origin (reference camera.c:35-40 with a zero shift) */
void rt_camera_scale(struct rt_camera *cam, float scale)
{
	v3_scale(cam->position, cam->position, scale);
	cam->focal_length *= scale;
}
