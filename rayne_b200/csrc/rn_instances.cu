// Per-drawable matrix records for one camera's visibility list ("drawable uniform packing").
//
// In the reference a renderer walks the camera's drawable list and, per drawable, writes the matrix uniforms its
// shaders ask for: model and inverse model matrix (Drawable::Update, Source/Rendering/RNRenderer.h:65-76),
// view * model and projection * view * model (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168, same products
// in the D3D12 / Metal renderers).  Here one kernel gathers the list's nodes and writes a 256-byte record per entry
// into a caller-owned device buffer, in list order, ready to be bound as an instance / storage buffer.
//
// The other per-drawable step of Entity::Render, LOD selection (Source/Scene/RNEntity.cpp:98-106 ->
// Model::GetLODStageForDistance, Source/Rendering/RNModel.cpp:260-281), is compiled out in this reference
// (RN_MODEL_LOD_DISABLED 1, Source/Rendering/RNModel.h:9); its #else branch is restated here as an extension
// (lod_select_kernel, rn_cull_select_lod) and checked against the same restatement in the oracle.
//
// HBM-bound gather: per entry 4 B (list id) + 64 B model + 64 B inverse read, 256 B written.  16 lanes per entry,
// one float4 each, so every record is two fully coalesced 128-byte stores per half-warp.
#include "rn_internal.h"
#include <algorithm>
#include <string.h>
#include <vector>

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

// Entity::Render's #else branch: lodDistance = |worldPosition - lodCamera.worldPosition| / lodCamera.clipFar, then
// Model::GetLODStageForDistance: stage 0 for <= 1 stage or distance <= epsilon, else the first stage whose distance is
// >= lodDistance, else the last.  One thread per list entry; 4 B + 16 B + 2 B read, 1 B written per entry.
__global__ void __launch_bounds__(256) lod_select_kernel(const uint32_t *__restrict__ list, const uint32_t *__restrict__ list_count,
                                                         uint32_t capacity, const uint32_t *__restrict__ slot_of_rank,
                                                         uint32_t rank_mode, uint32_t rank_offset, const float4 *__restrict__ wpos,
                                                         const uint16_t *__restrict__ node_model, const uint32_t *__restrict__ stage_ofs,
                                                         const float *__restrict__ stage_dist, uint32_t n_models, float cx, float cy,
                                                         float cz, float clip_far, uint8_t *__restrict__ out)
{
	const uint32_t total = min(*list_count, capacity);
	for(uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x)
	{
		const uint32_t id = __ldg(list + e);
		const uint32_t slot = rank_mode ? __ldg(slot_of_rank + (id - rank_offset)) : id;
		const uint32_t model = node_model[slot];
		uint32_t stage = 0;
		if(model < n_models)
		{
			const uint32_t first = stage_ofs[model], count = stage_ofs[model + 1] - first;
			const float4 p = __ldg(wpos + slot);
			// Vector3::GetDistance (Source/Math/RNVector.h): difference.GetLength() = sqrt(x*x + y*y + z*z)
			const float dx = p.x - cx, dy = p.y - cy, dz = p.z - cz;
			float d = sqrtf(dx * dx + dy * dy + dz * dz);
			d /= clip_far;
			if(!(count <= 1u || d <= 1.1920928955078125e-7f))
			{
				stage = count - 1u;
				for(uint32_t i = 0; i < count; i++)
				{
					if(d <= stage_dist[first + i]) { stage = i; break; }
				}
			}
		}
		out[e] = (uint8_t)stage;
	}
}

} // namespace rn

using namespace rn;

extern "C" int rn_lod_set_models(rn_scene_t *s, const uint32_t *stage_offsets, const float *stage_distances, uint32_t n_models)
{
	RN_REQUIRE(s && (n_models == 0 || (stage_offsets && stage_distances)), RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_models < 0xffffu, RN_ERR_CAPACITY, "%u models; rn_lod supports 65534", n_models);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	if(s->lod_stage_ofs) { cudaFree(s->lod_stage_ofs); s->lod_stage_ofs = nullptr; }
	if(s->lod_stage_dist) { cudaFree(s->lod_stage_dist); s->lod_stage_dist = nullptr; }
	s->lod_models = 0;
	if(n_models == 0) return RN_OK;
	RN_REQUIRE(stage_offsets[0] == 0, RN_ERR_INVALID, "stage_offsets[0] must be 0");
	for(uint32_t m = 0; m < n_models; m++)
	{
		RN_REQUIRE(stage_offsets[m + 1] > stage_offsets[m], RN_ERR_INVALID, "model %u has no LOD stage", m);
		RN_REQUIRE(stage_offsets[m + 1] - stage_offsets[m] <= 255, RN_ERR_CAPACITY, "model %u has more than 255 LOD stages", m);
	}
	const uint32_t total = stage_offsets[n_models];
	int rc = dev_alloc(s, &s->lod_stage_ofs, (size_t)n_models + 1);
	if(!rc) rc = dev_alloc(s, &s->lod_stage_dist, (size_t)total);
	if(rc) return rc;
	RN_CUDA(cudaMemcpy(s->lod_stage_ofs, stage_offsets, sizeof(uint32_t) * (n_models + 1), cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->lod_stage_dist, stage_distances, sizeof(float) * total, cudaMemcpyHostToDevice));
	s->lod_models = n_models;
	return RN_OK;
}

extern "C" int rn_lod_set_node_models(rn_scene_t *s, const uint16_t *model_of_rank, uint32_t first, uint32_t count)
{
	RN_REQUIRE(s && model_of_rank, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "range [%u, %u) exceeds the %u uploaded nodes", first, first + count, s->n);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	if(!s->lod_node_model)
	{
		int rc = dev_alloc(s, &s->lod_node_model, (size_t)s->desc.max_nodes);
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->lod_node_model, 0xff, sizeof(uint16_t) * s->desc.max_nodes));   // 0xffff: no model
	}
	// rank order -> slot order through the host mirror (a tile's slots are a permutation of its ranks)
	std::vector<uint16_t> host(s->n);
	RN_CUDA(cudaMemcpy(host.data(), s->lod_node_model, sizeof(uint16_t) * s->n, cudaMemcpyDeviceToHost));
	for(uint32_t i = 0; i < count; i++)
		host[s->slot_of_rank[first + i]] = model_of_rank[i];
	RN_CUDA(cudaMemcpy(s->lod_node_model, host.data(), sizeof(uint16_t) * s->n, cudaMemcpyHostToDevice));
	return RN_OK;
}

extern "C" int rn_cull_select_lod(rn_scene_t *s, uint32_t camera, const float *lod_camera_world_position, float lod_camera_clip_far,
                                  uint8_t *device_out, uint32_t capacity, void *cuda_stream)
{
	RN_REQUIRE(s && lod_camera_world_position && device_out, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_select_lod needs the lists of a cull");
	RN_REQUIRE(camera < s->last_ncams, RN_ERR_INVALID, "camera %u of %u", camera, s->last_ncams);
	RN_REQUIRE(s->have_world, RN_ERR_STATE, "no world transforms: run rn_transform_propagate first");
	RN_REQUIRE(s->lod_models && s->lod_node_model, RN_ERR_STATE, "rn_lod_set_models / rn_lod_set_node_models have not been called");
	if(capacity == 0 || s->n == 0)
		return RN_OK;
	RN_CUDA(cudaSetDevice(s->device));
	int sms = 148;
	cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s->device);
	const uint64_t want = ((uint64_t)std::min(capacity, s->n) + 255) / 256;
	const uint32_t grid = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(want, (uint64_t)sms * 8));
	lod_select_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
	    s->lists[camera], s->list_counts + camera, capacity, s->d_slot_of_rank, s->list_rank_mode, s->list_rank_offset, s->wpos,
	    s->lod_node_model, s->lod_stage_ofs, s->lod_stage_dist, s->lod_models, lod_camera_world_position[0],
	    lod_camera_world_position[1], lod_camera_world_position[2], lod_camera_clip_far, device_out);
	RN_CUDA(cudaGetLastError());
	s->launches++;
	return RN_OK;
}

extern "C" int rn_cull_pack_instances(rn_scene_t *s, uint32_t camera, const float *view, const float *proj_view,
                                      void *device_out, uint32_t capacity, void *cuda_stream)
{
	RN_RANGE("rn_cull_pack_instances");
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
