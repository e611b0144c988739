/*
 * tiffio.h -- TEST INFRASTRUCTURE (oracle).  Minimal stand-in for libtiff's
 * header so that the unmodified reference src/raytracer/image.c (which
 * includes <tiffio.h>, image.c:25) compiles in a container without libtiff
 * (SURVEY F2).  Only the calls raytracer/image.c:64-139 (write side) and
 * postprocess/image.c:30-73 (read side) make are provided; they are
 * implemented in tiff_stub.c and write / read a real baseline TIFF.
 */
#ifndef ORACLE_TIFFIO_STUB_H
#define ORACLE_TIFFIO_STUB_H

#include <stddef.h>
#include <stdint.h>

typedef struct tiff_stub TIFF;
typedef void *tdata_t;
typedef uint32_t tstrip_t;
typedef int64_t tsize_t;
typedef uint32_t ttag_t;

enum {
	TIFFTAG_IMAGEWIDTH = 256,
	TIFFTAG_IMAGELENGTH = 257,
	TIFFTAG_BITSPERSAMPLE = 258,
	TIFFTAG_PHOTOMETRIC = 262,
	TIFFTAG_ORIENTATION = 274,
	TIFFTAG_SAMPLESPERPIXEL = 277,
	TIFFTAG_ROWSPERSTRIP = 278,
	TIFFTAG_PLANARCONFIG = 284,
};
enum { ORIENTATION_TOPLEFT = 1, PLANARCONFIG_CONTIG = 1, PHOTOMETRIC_RGB = 2 };

typedef enum { TIFF_FLOAT = 11 } TIFFDataType;
#define TIFF_VARIABLE (-1)
#define FIELD_CUSTOM 65

typedef struct {
	ttag_t field_tag;
	short field_readcount;
	short field_writecount;
	TIFFDataType field_type;
	unsigned short field_bit;
	unsigned char field_oktochange;
	unsigned char field_passcount;
	const char *field_name;
} TIFFFieldInfo;

TIFF *TIFFOpen(const char *filename, const char *mode);
void TIFFClose(TIFF *tif);
int TIFFSetField(TIFF *tif, ttag_t tag, ...);
int TIFFMergeFieldInfo(TIFF *tif, const TIFFFieldInfo info[], uint32_t n);
tsize_t TIFFStripSize(TIFF *tif);
tstrip_t TIFFNumberOfStrips(TIFF *tif);
int TIFFWriteScanline(TIFF *tif, tdata_t buf, uint32_t row, uint16_t sample);
/* read side */
typedef void (*TIFFErrorHandler)(const char *, const char *, void *);
TIFFErrorHandler TIFFSetWarningHandler(TIFFErrorHandler handler);
int TIFFGetField(TIFF *tif, ttag_t tag, ...);
tsize_t TIFFScanlineSize(TIFF *tif);
int TIFFReadScanline(TIFF *tif, tdata_t buf, uint32_t row, uint16_t sample);
void *_TIFFmalloc(tsize_t size);
void _TIFFfree(void *ptr);

#endif
