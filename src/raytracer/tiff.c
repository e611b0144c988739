#include "tiff.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "error.h"

enum { T_SHORT = 3, T_LONG = 4, T_FLOAT = 11 };

struct ifd_entry {
	uint16_t tag, type;
	uint32_t count, value;
};

static void store_entry(uint8_t *dst, const struct ifd_entry *e)
{
	memcpy(dst, &e->tag, 2);
	memcpy(dst + 2, &e->type, 2);
	memcpy(dst + 4, &e->count, 4);
	if (e->type == T_SHORT && e->count == 1) {
		uint16_t v = (uint16_t)e->value;
		memset(dst + 8, 0, 4);
		memcpy(dst + 8, &v, 2);
	} else {
		memcpy(dst + 8, &e->value, 4);
	}
}

static int write_tiff(const char *filename, uint32_t width, uint32_t height, uint32_t bits,
	const void *pixels, const float *z_buffer, char *err)
{
	const uint64_t row_bytes = (uint64_t)width * 3u * (bits / 8u);
	const uint64_t pixel_count = (uint64_t)width * height;
	const int have_z = z_buffer != NULL;
	const uint32_t n_entries = have_z ? 12u : 11u;

	/* layout: header | IFD | bits array | strip offsets | strip counts | z | pixel rows */
	const uint64_t off_ifd = 8;
	const uint64_t off_bits = off_ifd + 2 + 12ull * n_entries + 4;
	const uint64_t off_strips = off_bits + 6 + 2; /* keep 4-byte alignment */
	const uint64_t off_counts = off_strips + 4ull * height;
	const uint64_t off_z = off_counts + 4ull * height;
	const uint64_t off_pixels = off_z + (have_z ? 4ull * pixel_count : 0);
	if (off_pixels + row_bytes * height > 0xffffffffull)
		rt_fail(err, "Image too large for a classic TIFF file [%s].", filename);

	FILE *f = fopen(filename, "wb");
	if (!f)
		rt_fail(err, "Failed to open output file [%s].", filename);

	uint8_t *head = calloc(1, (size_t)off_z);
	if (!head) {
		fclose(f);
		rt_fail(err, "Unable to allocate TIFF header for [%s].", filename);
	}
	const uint32_t ifd32 = (uint32_t)off_ifd;
	memcpy(head, "II*\0", 4);
	memcpy(head + 4, &ifd32, 4);

	const struct ifd_entry entries[12] = {
		{ 256, T_LONG, 1, width },
		{ 257, T_LONG, 1, height },
		{ 258, T_SHORT, 3, (uint32_t)off_bits },
		{ 259, T_SHORT, 1, 1 },  /* no compression */
		{ 262, T_SHORT, 1, 2 },  /* RGB */
		{ 273, T_LONG, height, (uint32_t)off_strips },
		{ 274, T_SHORT, 1, 1 },  /* top-left */
		{ 277, T_SHORT, 1, 3 },
		{ 278, T_LONG, 1, 1 },   /* one row per strip */
		{ 279, T_LONG, height, (uint32_t)off_counts },
		{ 284, T_SHORT, 1, 1 },  /* contiguous */
		{ RT_TIFFTAG_Z_BUFFER, T_FLOAT, (uint32_t)pixel_count, (uint32_t)off_z },
	};
	uint8_t *p = head + off_ifd;
	const uint16_t ne16 = (uint16_t)n_entries;
	memcpy(p, &ne16, 2);
	p += 2;
	for (uint32_t i = 0; i < n_entries; i++, p += 12)
		store_entry(p, &entries[i]);
	/* next-IFD offset stays 0 */
	if (height == 1) {
		/* a single strip keeps its offset/count inline */
		const uint32_t o = (uint32_t)off_pixels, c = (uint32_t)row_bytes;
		memcpy(head + off_ifd + 2 + 12 * 5 + 8, &o, 4);
		memcpy(head + off_ifd + 2 + 12 * 9 + 8, &c, 4);
	}
	const uint16_t bits16 = (uint16_t)bits;
	for (int i = 0; i < 3; i++)
		memcpy(head + off_bits + 2 * i, &bits16, 2);
	for (uint32_t r = 0; r < height; r++) {
		const uint32_t o = (uint32_t)(off_pixels + row_bytes * r), c = (uint32_t)row_bytes;
		memcpy(head + off_strips + 4ull * r, &o, 4);
		memcpy(head + off_counts + 4ull * r, &c, 4);
	}

	int ok = fwrite(head, 1, (size_t)off_z, f) == (size_t)off_z;
	if (ok && have_z)
		ok = fwrite(z_buffer, 4, (size_t)pixel_count, f) == (size_t)pixel_count;
	if (ok)
		ok = fwrite(pixels, 1, (size_t)(row_bytes * height), f) == (size_t)(row_bytes * height);
	free(head);
	if (fclose(f) != 0)
		ok = 0;
	if (!ok)
		rt_fail(err, "Failed to write output file [%s].", filename);
	return 0;
}

int rt_tiff_write_rgb8(const char *filename, uint32_t width, uint32_t height, const uint8_t *rgb, char *err)
{
	return write_tiff(filename, width, height, 8, rgb, NULL, err);
}

int rt_tiff_write_raw(const char *filename, uint32_t width, uint32_t height, const float *raster,
	const float *z_buffer, char *err)
{
	return write_tiff(filename, width, height, 32, raster, z_buffer, err);
}

/* ---- reader ------------------------------------------------------------ */

struct reader {
	const uint8_t *data;
	size_t size;
	int big_endian;
};

static uint32_t rd(const struct reader *r, size_t off, int bytes)
{
	uint32_t v = 0;
	if (off + (size_t)bytes > r->size)
		return 0;
	for (int i = 0; i < bytes; i++) {
		const uint32_t b = r->data[off + (size_t)i];
		v |= r->big_endian ? b << (8 * (bytes - 1 - i)) : b << (8 * i);
	}
	return v;
}

static uint32_t entry_value(const struct reader *r, size_t entry, uint32_t index)
{
	const uint32_t type = rd(r, entry + 2, 2), count = rd(r, entry + 4, 4);
	const int width = type == T_SHORT ? 2 : 4;
	const size_t base = ((uint64_t)count * (uint32_t)width <= 4) ? entry + 8 : rd(r, entry + 8, 4);
	return rd(r, base + (size_t)index * (size_t)width, width);
}

void rt_tiff_free(struct rt_tiff_image *img)
{
	free(img->rgb8);
	free(img->raster);
	free(img->z_buffer);
	memset(img, 0, sizeof *img);
}

int rt_tiff_read(const char *filename, struct rt_tiff_image *img, char *err)
{
	memset(img, 0, sizeof *img);
	FILE *f = fopen(filename, "rb");
	if (!f)
		rt_fail(err, "Failed to open input file [%s].", filename);
	fseek(f, 0, SEEK_END);
	const long len = ftell(f);
	fseek(f, 0, SEEK_SET);
	uint8_t *data = malloc(len > 0 ? (size_t)len : 1);
	if (!data || len < 8 || fread(data, 1, (size_t)len, f) != (size_t)len) {
		free(data);
		fclose(f);
		rt_fail(err, "Failed to read input file [%s].", filename);
	}
	fclose(f);

	struct reader r = { data, (size_t)len, 0 };
	if (!memcmp(data, "MM", 2))
		r.big_endian = 1;
	else if (memcmp(data, "II", 2)) {
		free(data);
		rt_fail(err, "Input file [%s] is not a TIFF file.", filename);
	}
	if (rd(&r, 2, 2) != 42) {
		free(data);
		rt_fail(err, "Input file [%s] is not a classic TIFF file.", filename);
	}
	const size_t ifd = rd(&r, 4, 4);
	const uint32_t n = rd(&r, ifd, 2);
	uint32_t spp = 1, bits = 1, planar = 1, compression = 1, rows_per_strip = 0xffffffffu;
	size_t e_offsets = 0, e_counts = 0, e_z = 0;
	for (uint32_t i = 0; i < n; i++) {
		const size_t e = ifd + 2 + 12ull * i;
		switch (rd(&r, e, 2)) {
		case 256: img->width = entry_value(&r, e, 0); break;
		case 257: img->height = entry_value(&r, e, 0); break;
		case 258: bits = entry_value(&r, e, 0); break;
		case 259: compression = entry_value(&r, e, 0); break;
		case 273: e_offsets = e; break;
		case 277: spp = entry_value(&r, e, 0); break;
		case 278: rows_per_strip = entry_value(&r, e, 0); break;
		case 279: e_counts = e; break;
		case 284: planar = entry_value(&r, e, 0); break;
		case RT_TIFFTAG_Z_BUFFER: e_z = e; break;
		default: break;
		}
	}
	int status = -1;
	if (spp != 3)
		snprintf(err, RT_ERRLEN, "Expected 3 samples per pixel but got [%u].", spp);
	else if (bits != 8 && bits != 32)
		snprintf(err, RT_ERRLEN, "Expected 8 or 32 bits per sample but got [%u].", bits);
	else if (planar != 1)
		snprintf(err, RT_ERRLEN, "Expected contiguous planar configuration");
	else if (compression != 1)
		snprintf(err, RT_ERRLEN, "Expected an uncompressed TIFF file.");
	else if (!e_offsets || !img->width || !img->height)
		snprintf(err, RT_ERRLEN, "Input file [%s] lacks image data.", filename);
	else
		status = 0;
	if (status) {
		free(data);
		return -1;
	}

	img->bits_per_sample = bits;
	/* Width and height come from the file: the pixel data of an uncompressed image cannot be
	 * larger than the file itself, which also keeps row_bytes * height far from wrapping. */
	const uint64_t row_bytes64 = (uint64_t)img->width * 3u * (bits / 8u);
	if (row_bytes64 > (uint64_t)r.size || (uint64_t)img->height > (uint64_t)r.size / row_bytes64) {
		free(data);
		rt_fail(err, "Image dimensions %ux%u of [%s] exceed the file size.", img->width, img->height, filename);
	}
	const size_t row_bytes = (size_t)row_bytes64;
	uint8_t *pixels = calloc(row_bytes * img->height, 1);
	if (!pixels) {
		free(data);
		rt_fail(err, "Unable to allocate image [%s].", filename);
	}
	if (rows_per_strip > img->height)
		rows_per_strip = img->height;
	const uint32_t n_strips = rd(&r, e_offsets + 4, 4);
	uint32_t row = 0;
	for (uint32_t s = 0; s < n_strips && row < img->height; s++) {
		const size_t off = entry_value(&r, e_offsets, s);
		uint32_t rows = img->height - row < rows_per_strip ? img->height - row : rows_per_strip;
		size_t bytes = row_bytes * rows;
		if (e_counts) {
			const size_t c = entry_value(&r, e_counts, s);
			if (c < bytes)
				bytes = c;
		}
		if (off > r.size || bytes > r.size - off) {
			free(pixels);
			free(data);
			rt_fail(err, "Truncated strip in [%s].", filename);
		}
		memcpy(pixels + row_bytes * row, data + off, bytes);
		row += rows;
	}
	if (row < img->height) {
		free(pixels);
		free(data);
		rt_fail(err, "Strips of [%s] cover %u of %u rows.", filename, row, img->height);
	}
	if (bits == 32 && r.big_endian) {
		for (size_t i = 0; i + 3 < row_bytes * img->height; i += 4) {
			uint8_t *q = pixels + i, t0 = q[0], t1 = q[1];
			q[0] = q[3]; q[1] = q[2]; q[2] = t1; q[3] = t0;
		}
	}
	if (bits == 8)
		img->rgb8 = pixels;
	else
		img->raster = (float *)pixels;

	if (e_z) {
		img->z_count = rd(&r, e_z + 4, 4);
		const size_t off = img->z_count == 1 ? e_z + 8 : rd(&r, e_z + 8, 4);
		/* one depth per pixel, compared in 64 bits (a wrapped product must not pass) */
		if ((uint64_t)img->z_count != (uint64_t)img->width * img->height || off > r.size ||
		    4ull * img->z_count > r.size - off) {
			free(data);
			rt_tiff_free(img);
			rt_fail(err, "Corrupted Z-Buffer.");
		}
		img->z_buffer = malloc(4ull * (img->z_count ? img->z_count : 1));
		if (!img->z_buffer) {
			free(data);
			rt_tiff_free(img);
			rt_fail(err, "Unable to allocate the Z-Buffer of [%s].", filename);
		}
		for (uint32_t i = 0; i < img->z_count; i++) {
			const uint32_t bitsv = rd(&r, off + 4ull * i, 4);
			memcpy(&img->z_buffer[i], &bitsv, 4);
		}
	}
	free(data);
	return 0;
}
