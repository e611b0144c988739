/*
 * system.h -- process clock and the timestamped log line.
 * Log format and the -p real|cpu switch follow reference
 * src/core/system.h:28-32 and system.c:42-52, :77-91.
 */
#ifndef RT_SYSTEM_H
#define RT_SYSTEM_H

#include <stdio.h>
#include <string.h>

struct rt_args;

void rt_system_init(const struct rt_args *args);
double rt_system_time(void);

#define RT_BASENAME(path) (strrchr((path), '/') ? strrchr((path), '/') + 1 : (path))

#define rt_log(...)                                                                 \
	do {                                                                        \
		printf("[%08.3f] %10s:%18s:%3u: ", rt_system_time(),                 \
			RT_BASENAME(__FILE__), __func__, (unsigned)__LINE__);        \
		printf(__VA_ARGS__);                                                 \
		putchar('\n');                                                       \
	} while (0)

#endif
