/*
 * tiff_stub.c -- TEST INFRASTRUCTURE (oracle).  Implements the handful of
 * libtiff calls the reference's raytracer/image.c:64-139 makes, writing an
 * uncompressed little-endian baseline TIFF with RowsPerStrip=1 and, when set,
 * the private tag 65000 (FLOAT array) carrying the z-buffer; and the calls of
 * postprocess/image.c:30-73, reading such a file back (little-endian,
 * uncompressed, any strip layout).
 */
#include "tiffio.h"

#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

struct tiff_stub {
	FILE *f;
	uint32_t width, height, bps, spp, photometric, orientation, planar, rps;
	uint32_t zcount;
	const float *z;
	uint8_t *rows;
	/* read side */
	int reading;
	uint8_t *file;
	size_t file_size;
	uint32_t *strip_off, strips;
	float *zread;
};

static uint32_t rd(const TIFF *t, size_t off, int bytes)
{
	uint32_t v = 0;
	if (off + (size_t)bytes <= t->file_size)
		memcpy(&v, t->file + off, (size_t)bytes);
	return v;
}

static TIFF *open_read(const char *filename)
{
	FILE *f = fopen(filename, "rb");
	if (!f)
		return NULL;
	TIFF *t = calloc(1, sizeof *t);
	t->reading = 1;
	fseek(f, 0, SEEK_END);
	t->file_size = (size_t)ftell(f);
	fseek(f, 0, SEEK_SET);
	t->file = malloc(t->file_size ? t->file_size : 1);
	if (fread(t->file, 1, t->file_size, f) != t->file_size || t->file_size < 8 || memcmp(t->file, "II*\0", 4)) {
		fclose(f);
		free(t->file);
		free(t);
		return NULL;
	}
	fclose(f);
	t->rps = 0xffffffffu;
	t->planar = 1;
	t->spp = 1;
	t->bps = 1;
	const size_t ifd = rd(t, 4, 4);
	const uint32_t n = rd(t, ifd, 2);
	for (uint32_t k = 0; k < n; k++) {
		const size_t e = ifd + 2 + 12 * (size_t)k;
		const uint32_t tag = rd(t, e, 2), type = rd(t, e + 2, 2), count = rd(t, e + 4, 4);
		const int size = type == 3 ? 2 : 4;
		const size_t data = (size_t)count * size <= 4 ? e + 8 : rd(t, e + 8, 4);
		const uint32_t first = rd(t, data, size);
		switch (tag) {
		case TIFFTAG_IMAGEWIDTH: t->width = first; break;
		case TIFFTAG_IMAGELENGTH: t->height = first; break;
		case TIFFTAG_BITSPERSAMPLE: t->bps = first; break;
		case TIFFTAG_SAMPLESPERPIXEL: t->spp = first; break;
		case TIFFTAG_PHOTOMETRIC: t->photometric = first; break;
		case TIFFTAG_ORIENTATION: t->orientation = first; break;
		case TIFFTAG_PLANARCONFIG: t->planar = first; break;
		case TIFFTAG_ROWSPERSTRIP: t->rps = first; break;
		case 273:
			t->strips = count;
			t->strip_off = malloc(sizeof(uint32_t) * (count ? count : 1));
			for (uint32_t i = 0; i < count; i++)
				t->strip_off[i] = rd(t, data + (size_t)i * size, size);
			break;
		case 65000:
			t->zcount = count;
			t->zread = malloc(sizeof(float) * (count ? count : 1));
			if (data + 4 * (size_t)count <= t->file_size)
				memcpy(t->zread, t->file + data, 4 * (size_t)count);
			break;
		default: break;
		}
	}
	if (t->rps > t->height)
		t->rps = t->height;
	return t;
}

TIFFErrorHandler TIFFSetWarningHandler(TIFFErrorHandler handler)
{
	(void)handler;
	return NULL;
}

/* the reference reads every scalar tag into a uint16_t (postprocess/image.c:41-46) */
int TIFFGetField(TIFF *t, ttag_t tag, ...)
{
	va_list ap;
	va_start(ap, tag);
	int ok = 1;
	switch (tag) {
	case TIFFTAG_IMAGEWIDTH: *va_arg(ap, uint16_t *) = (uint16_t)t->width; break;
	case TIFFTAG_IMAGELENGTH: *va_arg(ap, uint16_t *) = (uint16_t)t->height; break;
	case TIFFTAG_BITSPERSAMPLE: *va_arg(ap, uint16_t *) = (uint16_t)t->bps; break;
	case TIFFTAG_SAMPLESPERPIXEL: *va_arg(ap, uint16_t *) = (uint16_t)t->spp; break;
	case TIFFTAG_PLANARCONFIG: *va_arg(ap, uint16_t *) = (uint16_t)t->planar; break;
	case 65000: {
		uint32_t *count = va_arg(ap, uint32_t *);
		float **data = va_arg(ap, float **);
		*count = t->zcount;
		*data = t->zread;
		ok = t->zread != NULL;
		break;
	}
	default: ok = 0; break;
	}
	va_end(ap);
	return ok;
}

tsize_t TIFFScanlineSize(TIFF *t) { return (tsize_t)t->width * t->spp * (t->bps / 8); }

int TIFFReadScanline(TIFF *t, tdata_t buf, uint32_t row, uint16_t sample)
{
	(void)sample;
	const size_t rb = (size_t)TIFFScanlineSize(t);
	const uint32_t strip = row / t->rps;
	if (strip >= t->strips)
		return -1;
	const size_t off = t->strip_off[strip] + rb * (row % t->rps);
	if (off + rb > t->file_size)
		return -1;
	memcpy(buf, t->file + off, rb);
	return 1;
}

TIFF *TIFFOpen(const char *filename, const char *mode)
{
	if (mode && mode[0] == 'r')
		return open_read(filename);
	if (!mode || mode[0] != 'w')
		return NULL;
	TIFF *t = calloc(1, sizeof *t);
	t->f = fopen(filename, "wb");
	if (!t->f) {
		free(t);
		return NULL;
	}
	return t;
}

int TIFFSetField(TIFF *t, ttag_t tag, ...)
{
	va_list ap;
	va_start(ap, tag);
	switch (tag) {
	case TIFFTAG_IMAGEWIDTH: t->width = va_arg(ap, uint32_t); break;
	case TIFFTAG_IMAGELENGTH: t->height = va_arg(ap, uint32_t); break;
	case TIFFTAG_BITSPERSAMPLE: t->bps = (uint32_t)va_arg(ap, int); break;
	case TIFFTAG_SAMPLESPERPIXEL: t->spp = (uint32_t)va_arg(ap, int); break;
	case TIFFTAG_PHOTOMETRIC: t->photometric = (uint32_t)va_arg(ap, int); break;
	case TIFFTAG_ORIENTATION: t->orientation = (uint32_t)va_arg(ap, int); break;
	case TIFFTAG_PLANARCONFIG: t->planar = (uint32_t)va_arg(ap, int); break;
	case TIFFTAG_ROWSPERSTRIP: t->rps = (uint32_t)va_arg(ap, int); break;
	case 65000:
		t->zcount = (uint32_t)va_arg(ap, size_t); /* image.pixels is size_t (image.c:73) */
		t->z = va_arg(ap, const float *);
		break;
	default: break;
	}
	va_end(ap);
	return 1;
}

int TIFFMergeFieldInfo(TIFF *t, const TIFFFieldInfo info[], uint32_t n)
{
	(void)t; (void)info; (void)n;
	return 0;
}

tsize_t TIFFStripSize(TIFF *t) { return (tsize_t)t->width * t->spp * (t->bps / 8) * (t->rps ? t->rps : 1); }
tstrip_t TIFFNumberOfStrips(TIFF *t) { return t->height; }
void *_TIFFmalloc(tsize_t size) { return malloc((size_t)size); }
void _TIFFfree(void *p) { free(p); }

int TIFFWriteScanline(TIFF *t, tdata_t buf, uint32_t row, uint16_t sample)
{
	(void)sample;
	size_t rb = (size_t)t->width * t->spp * (t->bps / 8);
	if (!t->rows)
		t->rows = calloc(t->height, rb);
	memcpy(t->rows + rb * row, buf, rb);
	return 1;
}

static void put_entry(uint8_t **p, uint16_t tag, uint16_t type, uint32_t count, uint32_t value)
{
	memcpy(*p, &tag, 2); memcpy(*p + 2, &type, 2); memcpy(*p + 4, &count, 4); memcpy(*p + 8, &value, 4);
	*p += 12;
}

void TIFFClose(TIFF *t)
{
	if (t->reading) {
		free(t->file);
		free(t->strip_off);
		free(t->zread);
		free(t);
		return;
	}
	size_t rb = (size_t)t->width * t->spp * (t->bps / 8);
	uint32_t n_entries = 11 + (t->z ? 1 : 0);
	uint32_t off_ifd = 8;
	uint32_t off_bps = off_ifd + 2 + 12 * n_entries + 4;
	uint32_t off_stripoff = off_bps + 8;
	uint32_t off_stripcnt = off_stripoff + 4 * t->height;
	uint32_t off_z = off_stripcnt + 4 * t->height;
	uint32_t off_data = off_z + (t->z ? 4 * t->zcount : 0);
	uint8_t *hdr = calloc(1, off_data), *p = hdr;
	memcpy(p, "II*\0", 4); memcpy(p + 4, &off_ifd, 4);
	p = hdr + off_ifd;
	uint16_t ne = (uint16_t)n_entries;
	memcpy(p, &ne, 2); p += 2;
	put_entry(&p, 256, 4, 1, t->width);
	put_entry(&p, 257, 4, 1, t->height);
	put_entry(&p, 258, 3, 3, off_bps);
	put_entry(&p, 259, 3, 1, 1);
	put_entry(&p, 262, 3, 1, t->photometric);
	put_entry(&p, 273, 4, t->height, off_stripoff);
	put_entry(&p, 274, 3, 1, t->orientation);
	put_entry(&p, 277, 3, 1, t->spp);
	put_entry(&p, 278, 4, 1, t->rps);
	put_entry(&p, 279, 4, t->height, off_stripcnt);
	put_entry(&p, 284, 3, 1, t->planar);
	if (t->z)
		put_entry(&p, 65000, 11, t->zcount, off_z);
	uint16_t b = (uint16_t)t->bps;
	for (int i = 0; i < 3; i++) memcpy(hdr + off_bps + 2 * i, &b, 2);
	for (uint32_t r = 0; r < t->height; r++) {
		uint32_t o = off_data + (uint32_t)(rb * r), c = (uint32_t)rb;
		memcpy(hdr + off_stripoff + 4 * r, &o, 4);
		memcpy(hdr + off_stripcnt + 4 * r, &c, 4);
	}
	if (t->z) memcpy(hdr + off_z, t->z, 4 * (size_t)t->zcount);
	fwrite(hdr, 1, off_data, t->f);
	if (t->rows) fwrite(t->rows, rb, t->height, t->f);
	fclose(t->f);
	free(hdr); free(t->rows); free(t);
}
