/*
 * image.h -- image plane derived from the camera, the frame buffer the trace
 * loop fills, and the two output encodings (reference
 * src/raytracer/image.h:17-26, image.c:34-139).
 */
#ifndef RT_IMAGE_H
#define RT_IMAGE_H

#include <stddef.h>
#include <stdint.h>

struct rt_camera;

struct rt_image {
	uint32_t width, height;
	size_t pixels;
	float size[2];   /* extent of the image plane in scene units */
	float corner[3]; /* centre of the top-left pixel, minus one x step: see render */
	float step_x[3]; /* one pixel to the right */
	float step_y[3]; /* one pixel down */
	float *raster;   /* pixels * 3, zero-initialised */
	float *z_buffer; /* pixels, zero-initialised */
};

/* Image-plane geometry only (no allocation): size_x = 2 f tan(fov pi / 360),
 * size_y = size_x H / W, corner = centre + (1/2 - W/2) step_x + (1/2 - H/2) step_y. */
void rt_image_plane(struct rt_image *img, const struct rt_camera *cam, uint32_t width, uint32_t height);
int rt_image_alloc(struct rt_image *img, char *err);
void rt_image_free(struct rt_image *img);

/* 8-bit quantiser of the reference: (uint8_t)max(min(255 v, 255), 0), i.e.
 * truncation, not rounding (image.c:96-98). */
void rt_image_quantize(const float *raster, size_t samples, uint8_t *out);

/* raw != 0: FP32 raster + z-buffer in tag 65000 (the -f flag); else 8-bit RGB */
int rt_image_save(const struct rt_image *img, const char *filename, int raw, char *err);

#endif
