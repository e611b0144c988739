#include "image.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "calc.h"
#include "camera.h"
#include "error.h"
#include "tiff.h"

void rt_image_plane(struct rt_image *img, const struct rt_camera *cam, uint32_t width, uint32_t height)
{
	img->width = width;
	img->height = height;
	img->pixels = (size_t)width * height;
	img->size[0] = 2 * cam->focal_length * tanf(cam->fov * RT_PI / 360.f);
	img->size[1] = img->size[0] * height / width;

	float centre[3], to_plane[3], off_x[3], off_y[3];
	v3_scale(to_plane, cam->axis[2], cam->focal_length);
	v3_add(centre, to_plane, cam->position);
	v3_scale(img->step_x, cam->axis[0], img->size[0] / width);
	v3_scale(img->step_y, cam->axis[1], img->size[1] / height);
	v3_scale(off_x, img->step_x, .5f - width / 2.f);
	v3_scale(off_y, img->step_y, .5f - height / 2.f);
	for (int k = 0; k < 3; k++)
		img->corner[k] = centre[k] + off_x[k] + off_y[k];
}

int rt_image_alloc(struct rt_image *img, char *err)
{
	img->raster = calloc(img->pixels ? img->pixels : 1, 3 * sizeof(float));
	img->z_buffer = calloc(img->pixels ? img->pixels : 1, sizeof(float));
	if (!img->raster || !img->z_buffer) {
		rt_image_free(img);
		rt_fail(err, "Unable to allocate [%zu] bytes on heap.", img->pixels * 16);
	}
	return 0;
}

void rt_image_free(struct rt_image *img)
{
	free(img->raster);
	free(img->z_buffer);
	img->raster = NULL;
	img->z_buffer = NULL;
}

void rt_image_quantize(const float *raster, size_t samples, uint8_t *out)
{
	for (size_t i = 0; i < samples; i++)
		out[i] = (uint8_t)fmaxf(fminf(raster[i] * 255.f, 255.f), 0.f);
}

int rt_image_save(const struct rt_image *img, const char *filename, int raw, char *err)
{
	if (raw)
		return rt_tiff_write_raw(filename, img->width, img->height, img->raster, img->z_buffer, err);
	uint8_t *rgb = malloc(img->pixels ? img->pixels * 3 : 1);
	if (!rgb)
		rt_fail(err, "Unable to allocate [%zu] bytes on heap.", img->pixels * 3);
	rt_image_quantize(img->raster, img->pixels * 3, rgb);
	const int rc = rt_tiff_write_rgb8(filename, img->width, img->height, rgb, err);
	free(rgb);
	return rc;
}
