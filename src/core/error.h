/*
 * error.h -- error convention of the host code.
 * Library functions (scene loader, image writer ...) return 0 / -1 and leave a
 * message in a caller-supplied buffer, so they can be used from tests and
 * bindings; the `engine` executable turns a failure into the reference's
 * behaviour: one timestamped log line, then exit(1)
 * (reference src/core/error.h:21-33).
 */
#ifndef RT_ERROR_H
#define RT_ERROR_H

#include <stdio.h>
#include <stdlib.h>

#include "system.h"

#define RT_ERRLEN 512

#define rt_fatal(...)            \
	do {                     \
		rt_log(__VA_ARGS__); \
		exit(1);         \
	} while (0)

/* format into an error buffer and return -1 from the calling function */
#define rt_fail(err, ...)                                   \
	do {                                                \
		if (err)                                    \
			snprintf((err), RT_ERRLEN, __VA_ARGS__); \
		return -1;                                  \
	} while (0)

#endif
