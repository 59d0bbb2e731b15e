// librn_cuda - optional per-node outputs next to the transform + cull path (sm_100a):
//   rn_transform_inverse   SceneNode::UpdateInternalInverseTransformData   Source/Scene/RNSceneNode.cpp:548-568
//   rn_bounds_update       SceneNode::UpdateInternalBoundsData             Source/Scene/RNSceneNode.cpp:570-588
//   RN_CAMERA_TEST_AABB    extension restated from the commented body of Camera::InFrustum(const AABB&),
//                          Source/Scene/RNCamera.cpp:479-496 (the live function returns true)
//
// The inverse pass walks the same tiles as the transform kernel (rn_scene.cu): one warp per tile, the tile's
// sublevels in order, the inverse world matrices of the tile's inner nodes in the warp's shared memory (48 bytes
// each), so a parent never comes back from HBM.  Algorithmic bytes: 48 B read (local TRS + parent reference) +
// 64 B written per node.  The bounds pass is a flat one-thread-per-slot kernel over the stored world TRS:
// 40 B (world TRS) + 16 B (local sphere) [+ 24 B local box] read, 16 B + 24 B written.
//
// Compiled with -fmad=false like the rest of the library (rn_math.cuh).
#include "rn_internal.h"
#include <vector>
#include <string.h>
#include "rn_math.cuh"

#include <cstring>

namespace rn {

struct ExtraView
{
	const float4 *c0, *c1, *c2, *c3;
	const float4 *wpos, *wrot;
	const float2 *wsyz;
	const float4 *box0;
	const float2 *box1;
	float4 *winv;
	float4 *bsphere;
	float2 *baabb;
	uint8_t *vis;
	uint32_t *tile_counts, *super_totals;
	const TileMeta *tiles;
	uint32_t n_tiles, inner_cap, n;
};

static ExtraView make_extra_view(const rn_scene *s)
{
	ExtraView v;
	v.c0 = s->c0; v.c1 = s->c1; v.c2 = s->c2; v.c3 = s->c3;
	v.wpos = s->wpos; v.wrot = s->wrot; v.wsyz = s->wsyz;
	v.box0 = s->box0; v.box1 = s->box1;
	v.winv = s->winv; v.bsphere = s->bsphere; v.baabb = s->baabb;
	v.vis = s->vis; v.tile_counts = s->tile_counts; v.super_totals = s->super_totals;
	v.tiles = s->tiles; v.n_tiles = s->n_tiles; v.inner_cap = s->inner_cap; v.n = s->n;
	return v;
}

__device__ __forceinline__ void store_inverse(float4 *winv, uint32_t slot, const Affine &a)
{
	winv[slot * 4u + 0] = make_float4(a.m[0], a.m[1], a.m[2], 0.0f);
	winv[slot * 4u + 1] = make_float4(a.m[3], a.m[4], a.m[5], 0.0f);
	winv[slot * 4u + 2] = make_float4(a.m[6], a.m[7], a.m[8], 0.0f);
	winv[slot * 4u + 3] = make_float4(a.m[9], a.m[10], a.m[11], 1.0f);
}

__device__ __forceinline__ Affine load_inverse(const float4 *winv, uint32_t slot)
{
	const float4 m0 = winv[slot * 4u + 0], m1 = winv[slot * 4u + 1], m2 = winv[slot * 4u + 2], m3 = winv[slot * 4u + 3];
	Affine a;
	a.m[0] = m0.x; a.m[1] = m0.y; a.m[2] = m0.z;
	a.m[3] = m1.x; a.m[4] = m1.y; a.m[5] = m1.z;
	a.m[6] = m2.x; a.m[7] = m2.y; a.m[8] = m2.z;
	a.m[9] = m3.x; a.m[10] = m3.y; a.m[11] = m3.z;
	return a;
}

__device__ __forceinline__ Affine node_inverse_local(const ExtraView &v, uint32_t slot, uint32_t &parent)
{
	const float4 a = v.c0[slot], b = v.c1[slot], c = v.c2[slot];
	const Vec3 pos = {a.x, a.y, a.z};
	const Vec3 scale = {a.w, c.x, c.y};
	const Quat rot = {b.x, b.y, b.z, b.w};
	parent = __float_as_uint(c.z);
	return inverse_local_matrix(pos, rot, scale);
}

// Spine nodes of one hierarchy depth: the parent (another spine node) is read from global memory.
__global__ void __launch_bounds__(256) inverse_spine_kernel(ExtraView v, const uint32_t *__restrict__ slots, uint32_t count)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= count)
		return;
	const uint32_t slot = slots[i];
	uint32_t parent;
	const Affine il = node_inverse_local(v, slot, parent);
	// _inverseWorldTransform = _inverseLocalTransform * _parent->GetInverseWorldTransform()  (RNSceneNode.cpp:559)
	store_inverse(v.winv, slot, parent != RN_NO_PARENT ? affine_mul(il, load_inverse(v.winv, parent)) : il);
}

// One warp per tile; sublevel 0 (spine nodes stored inside the tile) was done by the spine launches.
__global__ void __launch_bounds__(kTileThreads) inverse_tile_kernel(const __grid_constant__ ExtraView v)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	const uint32_t tile = blockIdx.x * kWarpsPerCta + warp;
	if(tile >= v.n_tiles)
		return;
	float4 *sm = reinterpret_cast<float4 *>(smem_raw) + (size_t)warp * v.inner_cap * 3u;
	const TileMeta &tm = v.tiles[tile];
	const uint32_t start = tm.start, nlev = tm.nlev;
	for(uint32_t lev = 1; lev < nlev; lev++)
	{
		const uint32_t jb = tm.lvl_end[lev - 1], je = tm.lvl_end[lev];
		const uint32_t ib = tm.inner_end[lev - 1], ni = tm.inner_end[lev] - ib;
		for(uint32_t j = jb + lane; j < je; j += 32)
		{
			const uint32_t slot = start + j;
			uint32_t parent;
			Affine w = node_inverse_local(v, slot, parent);
			if(parent != RN_NO_PARENT)
			{
				Affine P;
				if(parent & kParentInTile)
				{
					const uint32_t k = (parent & 0xffffu) * 3u;
					const float4 q0 = sm[k], q1 = sm[k + 1], q2 = sm[k + 2];
					P.m[0] = q0.x; P.m[1] = q0.y; P.m[2] = q0.z; P.m[3] = q0.w;
					P.m[4] = q1.x; P.m[5] = q1.y; P.m[6] = q1.z; P.m[7] = q1.w;
					P.m[8] = q2.x; P.m[9] = q2.y; P.m[10] = q2.z; P.m[11] = q2.w;
				}
				else
					P = load_inverse(v.winv, parent);
				w = affine_mul(w, P);
			}
			store_inverse(v.winv, slot, w);
			if(j - jb < ni)
			{
				const uint32_t k = (ib + (j - jb)) * 3u;
				sm[k] = make_float4(w.m[0], w.m[1], w.m[2], w.m[3]);
				sm[k + 1] = make_float4(w.m[4], w.m[5], w.m[6], w.m[7]);
				sm[k + 2] = make_float4(w.m[8], w.m[9], w.m[10], w.m[11]);
			}
		}
		__syncwarp();
	}
}

__device__ __forceinline__ Box load_local_box(const ExtraView &v, uint32_t slot)
{
	Box b = {{-1.0f, -1.0f, -1.0f}, {1.0f, 1.0f, 1.0f}};   // AABB(Vector3(-1), Vector3(1)), RNSceneNode.cpp:34
	if(v.box0)
	{
		const float4 p = v.box0[slot];
		const float2 q = v.box1[slot];
		b.mn = {p.x, p.y, p.z};
		b.mx = {p.w, q.x, q.y};
	}
	return b;
}

// SceneNode::UpdateInternalBoundsData: both transformed volumes of every node.
__global__ void __launch_bounds__(256) bounds_kernel(const __grid_constant__ ExtraView v)
{
	const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
	if(slot >= v.n)
		return;
	const float4 a = v.wpos[slot], b = v.wrot[slot], sph = v.c3[slot];
	const float2 c = v.wsyz[slot];
	const Vec3 ws = {a.w, c.x, c.y};
	const Quat wr = {b.x, b.y, b.z, b.w};
	// Sphere::operator*=(Vector3), SetRotation (Source/Math/RNSphere.h:62-73); centre = position + offset
	const float s = (((ws.x < ws.y) ? ws.y : ws.x) < ws.z) ? ws.z : ((ws.x < ws.y) ? ws.y : ws.x);
	Vec3 off = {sph.x * s, sph.y * s, sph.z * s};
	off = quat_rotate(wr, off);
	v.bsphere[slot] = make_float4(a.x + off.x, a.y + off.y, a.z + off.z, sph.w * s);
	const Box w = world_aabb(load_local_box(v, slot), ws, wr);
	v.baabb[slot * 3u + 0] = make_float2(w.mn.x, w.mn.y);
	v.baabb[slot * 3u + 1] = make_float2(w.mn.z, w.mx.x);
	v.baabb[slot * 3u + 2] = make_float2(w.mx.y, w.mx.z);
}

// ---------------------------------------------------------------------------------------------------------------
// Extension RN_CAMERA_TEST_AABB.  Restated on the live plane convention (Plane::GetDistance = p.n - d, outside when
// positive, Source/Scene/RNCamera.cpp:456-464): centre = position + (min + max) * 0.5, half = (max - min) * 0.5,
// outside plane k <=> GetDistance_k(centre) > half . |n_k|.  Applied after the reference's sphere predicate to the
// cameras that carry the flag; NoCulling nodes and UseSimpleCulling cameras pass.  One warp per tile: filters the
// mask bytes the cull kernel left and recounts the tile's per-camera totals for the compaction.
// ---------------------------------------------------------------------------------------------------------------
struct AabbCams
{
	rn_camera_frustum cam[RN_MAX_CAMERAS];
	uint32_t n;
	uint32_t test_mask;   // cameras with RN_CAMERA_TEST_AABB and without UseSimpleCulling
};

__global__ void __launch_bounds__(256) aabb_filter_kernel(const __grid_constant__ ExtraView v, const __grid_constant__ AabbCams ac)
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(tile >= v.n_tiles)
		return;
	const uint32_t start = v.tiles[tile].start, count = v.tiles[tile].count;
	uint32_t total[RN_MAX_CAMERAS];
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++) total[c] = 0;
	for(uint32_t j0 = 0; j0 < count; j0 += 32)
	{
		const bool active = j0 + lane < count;
		const uint32_t slot = start + j0 + lane;
		// the mask bytes are stored in rank order: the node's rank offset inside the tile sits in its flags word
		const uint32_t fw = active ? __float_as_uint(v.c2[slot].w) : 0u;
		const uint32_t vidx = start + ((fw >> kRankShift) & kRankMask);
		uint32_t mask = active ? v.vis[vidx] : 0u;
		const uint32_t todo = mask & ac.test_mask;
		if(todo && !(fw & RN_NODE_NOCULLING))
		{
			const float4 a = v.wpos[slot], b = v.wrot[slot];
			const float2 c = v.wsyz[slot];
			const Vec3 ws = {a.w, c.x, c.y};
			const Quat wr = {b.x, b.y, b.z, b.w};
			const Box w = world_aabb(load_local_box(v, slot), ws, wr);
			const float cx = a.x + (w.mn.x + w.mx.x) * 0.5f, cy = a.y + (w.mn.y + w.mx.y) * 0.5f, cz = a.z + (w.mn.z + w.mx.z) * 0.5f;
			const float hx = (w.mx.x - w.mn.x) * 0.5f, hy = (w.mx.y - w.mn.y) * 0.5f, hz = (w.mx.z - w.mn.z) * 0.5f;
#pragma unroll
			for(int cc = 0; cc < RN_MAX_CAMERAS; cc++)
			{
				if((todo >> cc) & 1u)
				{
					bool inside = true;
#pragma unroll
					for(int k = 0; k < 6; k++)
					{
						const float *pl = ac.cam[cc].planes[k];
						const float d = (cx * pl[0] + cy * pl[1] + cz * pl[2]) - pl[3];
						const float r = hx * fabsf(pl[0]) + hy * fabsf(pl[1]) + hz * fabsf(pl[2]);
						if(d > r) inside = false;
					}
					if(!inside) mask &= ~(1u << cc);
				}
			}
			v.vis[vidx] = (uint8_t)mask;
		}
#pragma unroll
		for(int c = 0; c < RN_MAX_CAMERAS; c++)
			total[c] += __popc(__ballot_sync(0xffffffffu, (mask >> c) & 1u));
	}
	uint32_t mine = 0;
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++)
		if(lane == (uint32_t)c) mine = total[c];
	if(lane < RN_MAX_CAMERAS)
	{
		v.tile_counts[tile * RN_MAX_CAMERAS + lane] = mine;
		if(mine)
			atomicAdd(&v.super_totals[(tile / kSuperTiles) * RN_MAX_CAMERAS + lane], mine);
	}
}

int aabb_filter_launch(rn_scene *s, const rn_camera_frustum *cams, uint32_t ncams, cudaStream_t stream)
{
	AabbCams ac;
	memset(&ac, 0, sizeof(ac));
	memcpy(ac.cam, cams, sizeof(rn_camera_frustum) * ncams);
	ac.n = ncams;
	for(uint32_t c = 0; c < ncams; c++)
		if((cams[c].flags & RN_CAMERA_TEST_AABB) && !(cams[c].flags & RN_CAMERA_USE_SIMPLE_CULLING)) ac.test_mask |= 1u << c;
	aabb_filter_kernel<<<(s->n_tiles + 7) / 8, 256, 0, stream>>>(make_extra_view(s), ac);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

// SceneWithVisibilityLists::Render (Source/Scene/RNSceneWithVisibilityLists.cpp:157-175): a camera submits the default
// volume's nodes and then those of the first volume that contains its position.  With the nodes uploaded as
// [default volume][volume 1][volume 2]... the list of the cull over ALL nodes, minus the entries of the volumes the
// camera is not in, is exactly that sequence.  One warp per tile, same recount as the AABB filter.
struct VolumeCams
{
	uint32_t keep[8];      // 256 bits per camera would be overkill: keep[c] = the volume id camera c may see besides 0
	uint32_t ncams;
};

__global__ void __launch_bounds__(256) volume_filter_kernel(const __grid_constant__ ExtraView v, const uint8_t *__restrict__ node_volume,
                                                            const __grid_constant__ VolumeCams vc)
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(tile >= v.n_tiles)
		return;
	const uint32_t start = v.tiles[tile].start, count = v.tiles[tile].count;
	uint32_t total[RN_MAX_CAMERAS];
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++) total[c] = 0;
	for(uint32_t j0 = 0; j0 < count; j0 += 32)
	{
		const bool active = j0 + lane < count;
		const uint32_t slot = start + j0 + lane;
		const uint32_t vidx = start + (active ? (__float_as_uint(v.c2[slot].w) >> kRankShift) & kRankMask : 0u);
		uint32_t mask = active ? v.vis[vidx] : 0u;
		if(mask)
		{
			const uint32_t vol = node_volume[slot];
			if(vol)
			{
				uint32_t allowed = 0;
#pragma unroll
				for(int c = 0; c < RN_MAX_CAMERAS; c++)
					if(vc.keep[c] == vol) allowed |= 1u << c;
				if((mask & allowed) != mask)
				{
					mask &= allowed;
					v.vis[vidx] = (uint8_t)mask;
				}
			}
		}
#pragma unroll
		for(int c = 0; c < RN_MAX_CAMERAS; c++)
			total[c] += __popc(__ballot_sync(0xffffffffu, (mask >> c) & 1u));
	}
	uint32_t mine = 0;
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++)
		if(lane == (uint32_t)c) mine = total[c];
	if(lane < RN_MAX_CAMERAS)
	{
		v.tile_counts[tile * RN_MAX_CAMERAS + lane] = mine;
		if(mine)
			atomicAdd(&v.super_totals[(tile / kSuperTiles) * RN_MAX_CAMERAS + lane], mine);
	}
}

int volume_filter_launch(rn_scene *s, uint32_t ncams, cudaStream_t stream)
{
	VolumeCams vc;
	memset(&vc, 0, sizeof(vc));
	vc.ncams = ncams;
	for(uint32_t c = 0; c < RN_MAX_CAMERAS; c++)
		vc.keep[c] = c < ncams && s->cam_volume[c] ? s->cam_volume[c] : 0x100u;   // 0x100: matches no node
	volume_filter_kernel<<<(s->n_tiles + 7) / 8, 256, 0, stream>>>(make_extra_view(s), s->node_volume, vc);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

void extras_free(rn_scene *s)
{
	if(s->node_volume) { cudaFree(s->node_volume); s->node_volume = nullptr; }
	s->volumes_on = false;
	void *ptrs[] = {s->box0, s->box1, s->winv, s->bsphere, s->baabb, s->prio_group, s->sort_scratch, s->sort_hist, s->sort_ranges,
	                s->rank_masks, s->expand_totals, s->lod_node_model, s->lod_stage_ofs, s->lod_stage_dist};
	for(void *p : ptrs)
		if(p) cudaFree(p);
	s->lod_node_model = nullptr; s->lod_stage_ofs = nullptr; s->lod_stage_dist = nullptr; s->lod_models = 0;
	s->box0 = nullptr; s->box1 = nullptr; s->winv = nullptr; s->bsphere = nullptr; s->baabb = nullptr;
	s->prio_group = nullptr; s->sort_scratch = s->sort_hist = s->sort_ranges = nullptr;
	s->rank_masks = s->expand_totals = nullptr;
	s->expand_blocks_cap = 0;
	s->have_inverse = s->have_bounds = s->have_masks = false;
}

} // namespace rn

using namespace rn;

extern "C" int rn_transform_inverse(rn_scene_t *s, void *cuda_stream)
{
	RN_RANGE("rn_transform_inverse");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "no nodes uploaded");
	RN_CUDA(cudaSetDevice(s->device));
	if(!s->winv)
	{
		int rc = dev_alloc(s, &s->winv, (size_t)s->desc.max_nodes * 4);
		if(rc) return rc;
	}
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const ExtraView v = make_extra_view(s);
	for(size_t d = 0; d + 1 < s->spine_level_ofs.size(); d++)
	{
		const uint32_t cnt = s->spine_level_ofs[d + 1] - s->spine_level_ofs[d];
		if(!cnt) continue;
		inverse_spine_kernel<<<(cnt + 255) / 256, 256, 0, stream>>>(v, s->spine_slots + s->spine_level_ofs[d], cnt);
		s->launches++;
	}
	const size_t smem = (size_t)s->inner_cap * 48u * kWarpsPerCta;
	static size_t configured = 0;
	if(configured < smem)
	{
		RN_CUDA(cudaFuncSetAttribute(inverse_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		configured = smem;
	}
	inverse_tile_kernel<<<(s->n_tiles + kWarpsPerCta - 1) / kWarpsPerCta, kTileThreads, smem, stream>>>(v);
	s->launches++;
	if(!s->late_tiles.empty())
	{
		// tiles added under tile-computed nodes (rn_scene_add_nodes) read their parent's inverse from memory: a second
		// sweep, now that every other tile has written its results, makes them right (rare, so no separate tile list)
		inverse_tile_kernel<<<(s->n_tiles + kWarpsPerCta - 1) / kWarpsPerCta, kTileThreads, smem, stream>>>(v);
		s->launches++;
	}
	RN_CUDA(cudaGetLastError());
	s->have_inverse = true;
	return RN_OK;
}

extern "C" int rn_bounds_update(rn_scene_t *s, void *cuda_stream)
{
	RN_RANGE("rn_bounds_update");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "no nodes uploaded");
	RN_REQUIRE(s->have_world, RN_ERR_STATE, "rn_bounds_update needs world transforms: call rn_transform_propagate / rn_transform_cull first");
	RN_CUDA(cudaSetDevice(s->device));
	int rc = RN_OK;
	if(!s->bsphere) rc = dev_alloc(s, &s->bsphere, (size_t)s->desc.max_nodes);
	if(!rc && !s->baabb) rc = dev_alloc(s, &s->baabb, (size_t)s->desc.max_nodes * 3);
	if(rc) return rc;
	bounds_kernel<<<(s->n + 255) / 256, 256, 0, static_cast<cudaStream_t>(cuda_stream)>>>(make_extra_view(s));
	s->launches++;
	RN_CUDA(cudaGetLastError());
	s->have_bounds = true;
	return RN_OK;
}

extern "C" int rn_scene_device_extras(rn_scene_t *s, const float **inverse_world_matrix, const float **bounding_sphere,
                                      const float **bounding_box)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	if(inverse_world_matrix) *inverse_world_matrix = s->have_inverse ? reinterpret_cast<const float *>(s->winv) : nullptr;
	if(bounding_sphere) *bounding_sphere = s->have_bounds ? reinterpret_cast<const float *>(s->bsphere) : nullptr;
	if(bounding_box) *bounding_box = s->have_bounds ? reinterpret_cast<const float *>(s->baabb) : nullptr;
	return RN_OK;
}

extern "C" int rn_scenevis_set_node_volumes(rn_scene_t *s, const uint8_t *volume_of_rank, uint32_t first, uint32_t count)
{
	RN_REQUIRE(s && volume_of_rank, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "ranks [%u, %u) outside the %u uploaded nodes", first, first + count, s->n);
	RN_CUDA(cudaSetDevice(s->device));
	const bool fresh = !s->node_volume;
	if(fresh)
	{
		int rc = dev_alloc(s, &s->node_volume, (size_t)s->desc.max_nodes);
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->node_volume, 0, s->desc.max_nodes));
	}
	if(count == 0)
		return RN_OK;
	// rank order -> slot order through the host mirror; a tile is a rank range AND a slot range, so the touched slots are
	// the range [min slot, max slot] plus untouched neighbours that are re-written with their old value: copy back first
	std::vector<uint8_t> all(s->n, 0);
	if(!fresh)
		RN_CUDA(cudaMemcpy(all.data(), s->node_volume, s->n, cudaMemcpyDeviceToHost));
	for(uint32_t k = 0; k < count; k++)
		all[s->slot_of_rank[first + k]] = volume_of_rank[k];
	RN_CUDA(cudaMemcpy(s->node_volume, all.data(), s->n, cudaMemcpyHostToDevice));
	return RN_OK;
}

extern "C" int rn_scenevis_set_camera_volumes(rn_scene_t *s, const uint8_t *volume_per_camera, uint32_t n_cameras)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(n_cameras <= RN_MAX_CAMERAS, RN_ERR_INVALID, "%u cameras, at most %u", n_cameras, (unsigned)RN_MAX_CAMERAS);
	memset(s->cam_volume, 0, sizeof(s->cam_volume));
	s->volumes_on = volume_per_camera != nullptr;
	if(volume_per_camera)
	{
		RN_REQUIRE(s->node_volume, RN_ERR_STATE, "rn_scenevis_set_node_volumes has not been called");
		for(uint32_t c = 0; c < n_cameras; c++) s->cam_volume[c] = volume_per_camera[c];
	}
	return RN_OK;
}
