#include "render.h"

#include <stdio.h>

#include "error.h"
#include "image.h"
#include "scene.h"

int rt_render(const struct rt_scene *scene, struct rt_image *img, const crt_render_opts *opts,
	int device, crt_stats *stats, crt_scene_info *info, char *err)
{
	crt_scene_desc desc;
	rt_scene_desc(scene, &desc);

	crt_scene *dev = NULL;
	if (crt_scene_create(&desc, device, &dev) != CRT_OK)
		rt_fail(err, "Device scene set-up failed: %s", crt_last_error());
	if (info)
		crt_scene_get_info(dev, info);
	const int rc = crt_render(dev, opts, img->raster, img->z_buffer, stats);
	if (rc != CRT_OK)
		snprintf(err, RT_ERRLEN, "Device render failed: %s", crt_last_error());
	crt_scene_destroy(dev);
	return rc == CRT_OK ? 0 : -1;
}

int rt_render_multi(const struct rt_scene *scene, struct rt_image *img, const crt_render_opts *opts,
	const int *devices, int num_devices, crt_stats *stats, crt_scene_info *info, double *device_ms, int *dynamic, char *err)
{
	crt_scene_desc desc;
	rt_scene_desc(scene, &desc);

	crt_multi *multi = NULL;
	if (crt_multi_create(&desc, devices, num_devices, &multi) != CRT_OK)
		rt_fail(err, "Device scene set-up failed: %s", crt_last_error());
	if (info)
		crt_multi_get_info(multi, 0, info);
	if (dynamic)
		*dynamic = crt_multi_is_dynamic(multi);
	const int rc = crt_multi_render(multi, opts, img->raster, img->z_buffer, stats, device_ms);
	if (rc != CRT_OK)
		snprintf(err, RT_ERRLEN, "Device render failed: %s", crt_last_error());
	crt_multi_destroy(multi);
	return rc == CRT_OK ? 0 : -1;
}
