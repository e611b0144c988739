#include "system.h"

#include <time.h>

#include "argv.h"
#include "strhash.h"

static struct timespec t_start;
static clock_t c_start;
static int use_cpu_clock;

void rt_system_init(const struct rt_args *args)
{
	timespec_get(&t_start, TIME_UTC);
	c_start = clock();
	use_cpu_clock = 0;
	int i = rt_args_find_with_values(args, "-p", 1);
	if (i) {
		/* unknown values are ignored, like the reference's switch without default */
		if (!strcmp(args->argv[i + 1], "cpu"))
			use_cpu_clock = 1;
		else if (!strcmp(args->argv[i + 1], "real"))
			use_cpu_clock = 0;
	}
}

double rt_system_time(void)
{
	if (use_cpu_clock)
		return (double)(clock() - c_start) / CLOCKS_PER_SEC;
	struct timespec now;
	timespec_get(&now, TIME_UTC);
	return (double)(now.tv_sec - t_start.tv_sec) + (double)(now.tv_nsec - t_start.tv_nsec) * 1e-9;
}
