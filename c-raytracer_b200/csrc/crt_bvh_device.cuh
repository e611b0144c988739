/*
 * crt_bvh_device.cuh -- interface of the device-side tree builder (crt_bvh_device.cu) to crt_api.cu.
 */
#ifndef CRT_BVH_DEVICE_CUH
#define CRT_BVH_DEVICE_CUH

#include "crt_kernels.cuh"

namespace crt {

struct DeviceTree {
	float4 *wide = nullptr;  /* num_nodes WideNode records (bvh_build.h), cudaMalloc'd: the caller owns it */
	uint32_t num_nodes = 0;
	uint32_t max_depth = 0;
};

/* d_prims_in: n primitive records (4 float4 each, crt_device.cuh) in scene order of the bounded objects.
 * Writes the records in tree order to d_prims_out and obj_slot[object] = slot for every bounded object.
 * Synchronous (a handful of 4-byte read-backs, one per level of the wide tree). */
int build_device_tree(const float4 *d_prims_in, uint32_t n, int32_t *d_obj_slot, float4 *d_prims_out, DeviceTree &out);

} // namespace crt

#endif
