/*
 * argv.h -- command-line lookup.  Flags may appear anywhere on the command
 * line; they are found by comparing DJB hashes of whole argv strings, and a
 * flag whose values would run past the end of argv counts as absent
 * (reference src/core/argv.c:38-55).  Positional arguments 1-4 are
 * <scene> <output> <width> <height> (argv.h:17-22).
 */
#ifndef RT_ARGV_H
#define RT_ARGV_H

#include <stdint.h>

enum rt_positional {
	RT_ARG_SCENE = 1,
	RT_ARG_OUTPUT = 2,
	RT_ARG_WIDTH = 3,
	RT_ARG_HEIGHT = 4,
};

struct rt_args {
	int argc;
	char **argv;
	uint32_t *hashes;
};

int rt_args_init(struct rt_args *args, int argc, char **argv);
void rt_args_free(struct rt_args *args);
/* index of the flag in argv, 0 if absent */
int rt_args_find(const struct rt_args *args, const char *flag);
/* as above, but 0 unless `count` values follow the flag */
int rt_args_find_with_values(const struct rt_args *args, const char *flag, int count);

#endif
