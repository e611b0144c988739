#include "argv.h"

#include <stdlib.h>

#include "strhash.h"

int rt_args_init(struct rt_args *args, int argc, char **argv)
{
	args->argc = argc;
	args->argv = argv;
	args->hashes = malloc(sizeof(uint32_t) * (size_t)(argc > 0 ? argc : 1));
	if (!args->hashes)
		return -1;
	for (int i = 0; i < argc; i++)
		args->hashes[i] = rt_hash_djb(argv[i]);
	return 0;
}

void rt_args_free(struct rt_args *args)
{
	free(args->hashes);
	args->hashes = NULL;
}

int rt_args_find(const struct rt_args *args, const char *flag)
{
	const uint32_t h = rt_hash_djb(flag);
	for (int i = 1; i < args->argc; i++)
		if (args->hashes[i] == h)
			return i;
	return 0;
}

int rt_args_find_with_values(const struct rt_args *args, const char *flag, int count)
{
	const int i = rt_args_find(args, flag);
	return (i + count < args->argc) ? i : 0;
}
