// Per-drawable matrix records for one camera's visibility list ("drawable uniform packing").
//
// In the reference a renderer walks the camera's drawable list and, per drawable, writes the matrix uniforms its
// shaders ask for: model and inverse model matrix (Drawable::Update, Source/Rendering/RNRenderer.h:65-76),
// view * model and projection * view * model (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168, same products
// in the D3D12 / Metal renderers).  Here one kernel gathers the list's nodes and writes a 256-byte record per entry
// into a caller-owned device buffer, in list order, ready to be bound as an instance / storage buffer.
//
// (The other per-drawable step of Entity::Render, LOD selection, is compiled out in this reference:
// RN_MODEL_LOD_DISABLED 1, Source/Rendering/RNModel.h:9 - there is nothing live to accelerate.)
//
// HBM-bound gather: per entry 4 B (list id) + 64 B model + 64 B inverse read, 256 B written.  16 lanes per entry,
// one float4 each, so every record is two fully coalesced 128-byte stores per half-warp.
#include "rn_internal.h"
#include <algorithm>
#include <string.h>

namespace rn {

struct InstanceMats
{
	float view[16];
	float proj_view[16];
};

// Matrix::operator* (Source/Math/RNMatrix.h:59-90), column `c` of a * o: separately rounded products summed left to
// right (this file is compiled with -fmad=false).
__device__ __forceinline__ float4 mat_mul_column(const float *a, const float4 &o)
{
	float4 r;
	r.x = a[0] * o.x + a[4] * o.y + a[8] * o.z + a[12] * o.w;
	r.y = a[1] * o.x + a[5] * o.y + a[9] * o.z + a[13] * o.w;
	r.z = a[2] * o.x + a[6] * o.y + a[10] * o.z + a[14] * o.w;
	r.w = a[3] * o.x + a[7] * o.y + a[11] * o.z + a[15] * o.w;
	return r;
}

__global__ void __launch_bounds__(256) pack_instances_kernel(const uint32_t *__restrict__ list,
                                                             const uint32_t *__restrict__ list_count, uint32_t capacity,
                                                             const uint32_t *__restrict__ slot_of_rank, uint32_t rank_mode,
                                                             uint32_t rank_offset, const float4 *__restrict__ wmat,
                                                             const float4 *__restrict__ winv,
                                                             const __grid_constant__ InstanceMats mats,
                                                             float4 *__restrict__ out)
{
	const uint32_t total = min(*list_count, capacity);
	const uint32_t lane16 = threadIdx.x & 15u;
	const uint32_t which = lane16 >> 2, col = lane16 & 3u;
	const uint32_t stride = (gridDim.x * blockDim.x) >> 4;
	for(uint32_t e = (blockIdx.x * blockDim.x + threadIdx.x) >> 4; e < total; e += stride)
	{
		uint32_t id = __ldg(list + e);
		const uint32_t slot = rank_mode ? __ldg(slot_of_rank + (id - rank_offset)) : id;
		const float4 *src = (which == 1) ? winv : wmat;
		float4 v = __ldg(src + (size_t)slot * 4 + col);
		if(which == 2) v = mat_mul_column(mats.view, v);
		else if(which == 3) v = mat_mul_column(mats.proj_view, v);
		__stcs(out + (size_t)e * 16 + lane16, v);
	}
}

} // namespace rn

using namespace rn;

extern "C" int rn_cull_pack_instances(rn_scene_t *s, uint32_t camera, const float *view, const float *proj_view,
                                      void *device_out, uint32_t capacity, void *cuda_stream)
{
	RN_REQUIRE(s && view && proj_view && device_out, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_pack_instances needs the lists of a cull");
	RN_REQUIRE(camera < s->last_ncams, RN_ERR_INVALID, "camera %u of %u", camera, s->last_ncams);
	RN_REQUIRE(s->have_world, RN_ERR_STATE, "no world transforms: run rn_transform_propagate first");
	RN_REQUIRE(s->have_inverse, RN_ERR_STATE, "no inverse world matrices: run rn_transform_inverse first");
	RN_REQUIRE((reinterpret_cast<uintptr_t>(device_out) & 15u) == 0, RN_ERR_INVALID, "device_out must be 16-byte aligned");
	if(capacity == 0 || s->n == 0)
		return RN_OK;
	RN_CUDA(cudaSetDevice(s->device));
	InstanceMats mats;
	memcpy(mats.view, view, sizeof(mats.view));
	memcpy(mats.proj_view, proj_view, sizeof(mats.proj_view));
	// grid-stride over min(count, capacity) entries; the count stays on the device (no host round trip)
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
	const uint64_t want = ((uint64_t)std::min(capacity, s->n) * 16 + 255) / 256;
	const uint32_t grid = (uint32_t)std::min<uint64_t>(want, (uint64_t)sms * 8);
	pack_instances_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
	    s->lists[camera], s->list_counts + camera, capacity, s->d_slot_of_rank, s->list_rank_mode, s->list_rank_offset,
	    s->wmat, s->winv, mats, static_cast<float4 *>(device_out));
	RN_CUDA(cudaGetLastError());
	s->launches++;
	return RN_OK;
}
