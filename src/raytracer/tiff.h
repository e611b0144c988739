/*
 * tiff.h -- minimal baseline-TIFF writer/reader (libtiff is not available,
 * SURVEY F2).  Writes exactly the layout the reference produces through
 * libtiff (reference src/raytracer/image.c:114-139):
 * little-endian, uncompressed, RGB, 3 samples per pixel, contiguous,
 * orientation top-left, RowsPerStrip = 1, and either 8 bits per sample or --
 * raw mode -- 32 bits per sample (IEEE floats, no SampleFormat tag) plus the
 * private tag 65000 (type FLOAT, count W*H) holding the z-buffer, which is
 * what the reference post-processor reads (src/postprocess/image.c:30-73).
 */
#ifndef RT_TIFF_H
#define RT_TIFF_H

#include <stddef.h>
#include <stdint.h>

#define RT_TIFFTAG_Z_BUFFER 65000

/* rgb: width*height*3 bytes */
int rt_tiff_write_rgb8(const char *filename, uint32_t width, uint32_t height, const uint8_t *rgb, char *err);
/* raster: width*height*3 floats, z_buffer: width*height floats */
int rt_tiff_write_raw(const char *filename, uint32_t width, uint32_t height, const float *raster,
	const float *z_buffer, char *err);

struct rt_tiff_image {
	uint32_t width, height;
	uint32_t bits_per_sample; /* 8 or 32 */
	uint8_t *rgb8;   /* bits_per_sample == 8 */
	float *raster;   /* bits_per_sample == 32 */
	float *z_buffer; /* tag 65000 if present, else NULL */
	uint32_t z_count;
};

/* Reads files with the layout above (any strip layout, uncompressed, II or MM). */
int rt_tiff_read(const char *filename, struct rt_tiff_image *img, char *err);
void rt_tiff_free(struct rt_tiff_image *img);

#endif
