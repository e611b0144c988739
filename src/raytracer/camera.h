/*
 * camera.h -- pinhole camera basis (reference src/raytracer/camera.h:17-22).
 */
#ifndef RT_CAMERA_H
#define RT_CAMERA_H

struct rt_camera {
	float position[3];
	float axis[3][3]; /* unit x, unit y, forward = x cross y (not re-orthogonalised) */
	float fov;        /* horizontal, degrees, in (0,180) */
	float focal_length;
};

/* returns -1 and fills err when fov is out of range (camera.c:23) */
int rt_camera_init(struct rt_camera *cam, const float position[3], const float x[3], const float y[3],
	float fov, float focal_length, char *err);
void rt_camera_scale(struct rt_camera *cam, float scale);

#endif
