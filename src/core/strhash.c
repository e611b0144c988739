#include "strhash.h"

uint32_t rt_hash_djb(const char *s)
{
	uint32_t h = 5381u;
	for (const unsigned char *p = (const unsigned char *)s; *p; p++)
		h = (h * 33u) ^ *p;
	return h;
}
