/* strhash.h -- DJB2-xor string hash used for flag and keyword matching
 * (reference src/core/strhash.c:14-20: hash = 33*hash ^ byte, seed 5381). */
#ifndef RT_STRHASH_H
#define RT_STRHASH_H
#include <stdint.h>
uint32_t rt_hash_djb(const char *s);
#endif
