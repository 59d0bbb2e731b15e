// Internal declarations shared by the translation units of librn_cuda (not part of the ABI).
#pragma once
#include "../../include/rn_cuda.h"
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace rn {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t err, const char *what, const char *file, int line);

#define RN_CUDA(call)                                                                  \
	do {                                                                               \
		cudaError_t rn_err__ = (call);                                                 \
		if(rn_err__ != cudaSuccess) return rn::cuda_fail(rn_err__, #call, __FILE__, __LINE__); \
	} while(0)

#define RN_REQUIRE(cond, status, ...)                                                  \
	do {                                                                               \
		if(!(cond)) { rn::set_error(__VA_ARGS__); return (status); }                   \
	} while(0)

// NVTX range around a C-ABI call (header-only NVTX 3: a no-op unless a profiler is attached) - the counterpart of the
// reference's own frame markers
struct NvtxRange
{
	explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
	~NvtxRange() { nvtxRangePop(); }
};
#define RN_RANGE(name) rn::NvtxRange rn_nvtx_range__(name)

constexpr uint32_t kMaxSublevels = 12;   // sublevel 0 = spine nodes of the tile, 1.. = depth inside a unit
constexpr uint32_t kWarpsPerCta = 4;     // tile kernel: one tile per warp
constexpr uint32_t kTileThreads = kWarpsPerCta * 32;
constexpr uint32_t kSuperTiles = 256;    // tiles per superblock of the compaction offsets
constexpr uint32_t kRankShift = 16;      // c2.w bits 16-25: rank - tile start (the tile kernel stores mask bytes in RANK order)
constexpr uint32_t kRankMask = 0x3ffu;
constexpr uint32_t kNodeRemoved = 0x80u;          // c2.w bit 7 (unused by RN::SceneNode::Flags): the slot is a hole - rn_scene_remove_nodes
constexpr uint32_t kParentInTile = 0x80000000u;   // c2.z: bit 31 set (and not RN_NO_PARENT) -> low bits = inner index

// One tile = a contiguous range of ranks [start, start+count) stored at the same slots, permuted so that each
// sublevel is contiguous and, inside a sublevel, the inner nodes (nodes with children) come first.  64 bytes.
struct TileMeta
{
	uint32_t start;
	uint16_t count;
	uint16_t nlev;
	uint16_t lvl_end[kMaxSublevels];   // end offset (exclusive) of sublevel k inside the tile
	uint16_t inner_end[kMaxSublevels]; // number of inner nodes in sublevels 1..k (sublevel 0 holds none)
	uint32_t late;        // 1: holds nodes whose parent is computed by another tile -> second tile launch (rn_scene_add_nodes)
	uint32_t pad;
};
static_assert(sizeof(TileMeta) == 64, "TileMeta must be 64 bytes");

struct ParticleStore;

} // namespace rn

struct rn_scene
{
	rn_scene_desc desc{};
	int device = 0;
	uint32_t tile_cap = 512;
	uint32_t inner_cap = 128;    // inner nodes (parents kept in shared memory) per tile
	uint32_t unit_cap = 128;     // largest subtree kept whole inside a tile
	uint32_t n = 0;              // nodes
	uint32_t n_tiles = 0;
	uint32_t max_depth = 0;
	uint32_t spine_nodes = 0;
	uint64_t device_bytes = 0;
	uint64_t launches = 0;
	int sm_count = 0;
	bool fma_matrices = true;    // world matrices with fused multiply-adds (rn_scene_desc.flags & RN_SCENE_STRICT_MATRICES clears it)
	size_t tile_smem_configured[8] = {};   // dynamic shared memory opted in per tile_kernel instantiation on THIS device
	bool have_world = false;     // world TRS valid (a propagate ran since the last upload)
	uint32_t last_ncams = 0;
	uint32_t list_rank_mode = 0;     // 1: visibility lists hold rank + list_rank_offset instead of slots
	uint32_t list_rank_offset = 0;

	// node columns, slot order (float4 each): c0 = pos.xyz, scale.x ; c1 = rot ; c2 = scale.y, scale.z,
	// parent slot (bits), flags|group<<8 (bits) ; c3 = sphere offset.xyz, radius
	float4 *c0 = nullptr, *c1 = nullptr, *c2 = nullptr, *c3 = nullptr;
	// outputs, slot order
	float4 *wpos = nullptr;      // world position xyz, world scale x
	float4 *wrot = nullptr;      // world rotation
	float2 *wsyz = nullptr;      // world scale y, z
	float4 *wmat = nullptr;      // 4 x float4 per slot (column-major 4x4)
	uint32_t *d_slot_of_rank = nullptr; // rank -> slot
	uint8_t *vis = nullptr;      // RANK -> camera bit mask of the last cull (padded to a multiple of 32 bytes, padding zero)
	rn::TileMeta *tiles = nullptr;
	uint32_t *tile_counts = nullptr;   // [n_tiles][RN_MAX_CAMERAS] visible nodes per tile (last cull)
	uint32_t *super_totals = nullptr;  // [n_tiles / 256][RN_MAX_CAMERAS]
	uint32_t *list_counts = nullptr;   // [RN_MAX_CAMERAS]
	uint32_t *lists[RN_MAX_CAMERAS] = {};
	// second set of cull outputs (lists, counts, rank masks): rn_pipeline's single-range mode flips between the two so that
	// the copy-out of frame i can run beside the kernels of frame i + 1 (scene_flip_outputs, rn_scene.cu)
	uint32_t *lists_alt[RN_MAX_CAMERAS] = {};
	uint32_t *list_counts_alt = nullptr;
	uint32_t *rank_masks_alt = nullptr;
	uint32_t list_capacity = 0;
	// optional columns / outputs (allocated on first use; rn_bounds.cu, rn_sort.cu)
	float4 *box0 = nullptr;      // local AABB: minExtend xyz, maxExtend x   (nullptr = default AABB(-1, 1))
	float2 *box1 = nullptr;      // local AABB: maxExtend y, z
	float4 *winv = nullptr;      // inverse world matrix, 4 x float4 per slot (rn_transform_inverse)
	float4 *bsphere = nullptr;   // transformed bounding sphere: centre xyz, radius (rn_bounds_update)
	float2 *baabb = nullptr;     // transformed AABB: 3 x float2 per slot = min xyz, max xyz (rn_bounds_update)
	bool have_inverse = false, have_bounds = false;
	uint8_t *prio_group = nullptr;     // per slot: dense index of the node's render priority (ascending)
	uint32_t n_prio_groups = 1;
	uint32_t transparent_group = 0xffffffffu;  // dense index of RN_RENDER_TRANSPARENT, or none
	uint32_t first_sky_group = 0xffffffffu;    // first dense index whose priority is >= RN_RENDER_SKY, or none
	uint32_t last_cam_flags[RN_MAX_CAMERAS] = {};
	uint32_t *sort_scratch = nullptr;  // radix sort: 4 arrays of max_nodes + histograms
	uint32_t *sort_hist = nullptr;
	uint32_t *sort_ranges = nullptr;   // [RN_MAX_CAMERAS][2]: transparent range begin, end (device)
	uint32_t sort_blocks_cap = 0;
	uint8_t *node_volume = nullptr;    // per slot: 0 = default volume, 1..254 = user volume, 255 = in no volume (rn_bounds.cu)
	uint8_t cam_volume[RN_MAX_CAMERAS] = {};   // per camera of the next culls: the user volume that contains it, 0 = none
	bool volumes_on = false;
	// LOD extension (rn_instances.cu): per slot the model id (0xffff = none), per model its stages' distances
	uint16_t *lod_node_model = nullptr;
	uint32_t *lod_stage_ofs = nullptr;
	float *lod_stage_dist = nullptr;
	uint32_t lod_models = 0;
	// software occlusion culling (rn_occlusion.cu), allocated on first use
	uint16_t *occ_counter = nullptr;       // per slot: SceneBasicInfo::occludedFrameCounter
	int32_t *occ_mesh_of_slot = nullptr;   // per slot: mesh index of the Entity, -1 = none
	uint32_t *occ_tri_ofs = nullptr;       // [occ_meshes + 1]
	float *occ_tris = nullptr;             // 9 floats per triangle
	uint32_t occ_meshes = 0;
	uint32_t *occ_cand = nullptr;          // candidate occluders of the camera being processed
	uint32_t *occ_sel = nullptr;           // [RN_MAX_CAMERAS] selected occluders + flags
	float *occ_depth = nullptr;            // [RN_MAX_CAMERAS][40 * 40]
	uint32_t *rank_masks = nullptr;    // [RN_MAX_CAMERAS][ceil(max_nodes / 32)] visibility bits in rank order (rn_masks.cu)
	bool have_masks = false;
	uint32_t mask_ncams = 0;
	uint32_t *expand_totals = nullptr; // rn_masks_expand: per-CTA totals / offsets
	uint32_t expand_blocks_cap = 0;
	uint32_t *spine_slots = nullptr;   // all spine slots grouped by depth
	std::vector<uint32_t> spine_level_ofs; // host: offsets into spine_slots per depth (size levels+1)
	float *scratch_cols = nullptr;     // device staging for rn_scene_update_nodes
	float *scratch_host = nullptr;     // pinned host staging
	uint32_t scratch_cap = 0;
	cudaEvent_t scratch_done = nullptr;

	// optional profiling of the tile kernel
	bool profile = false;
	static constexpr int kProfileSlots = 256;
	cudaEvent_t prof_ev[kProfileSlots][2] = {};
	int prof_used = 0;

	// host mirrors
	std::vector<uint32_t> slot_of_rank, rank_of_slot;

	// incremental edits (rn_scene_add_nodes / rn_scene_remove_nodes, rn_scene.cu)
	rn::TileMeta *tiles_base = nullptr;   // the allocation `tiles` points into: [tile_front_room][n_tiles][tile_back_room]
	uint32_t tile_head = 0;               // tiles == tiles_base + tile_head
	uint32_t tile_capacity = 0;
	uint32_t slot_high = 0;               // slots [0, slot_high) are in use (nodes and holes); == n
	uint32_t live_nodes = 0;              // nodes that are not holes
	std::vector<uint8_t> slot_state;      // per slot: bit 0 spine, bit 1 in a late tile, bit 2 removed, bit 3 in use
	std::vector<int32_t> prio_of_slot;    // only when the upload carried render priorities
	std::vector<std::pair<int32_t, uint32_t>> live_by_prio;   // sorted by priority: live nodes per priority
	std::vector<uint32_t> late_tiles;     // absolute indices (into tiles_base) of the late tiles
	uint32_t *d_late_tiles = nullptr;
	uint32_t d_late_cap = 0;
	uint64_t edits_fast = 0, edits_nodes = 0;

	rn::ParticleStore *particles = nullptr;
};

namespace rn {
int particles_destroy(rn_scene *scene);
void extras_free(rn_scene *scene);      // rn_bounds.cu / rn_sort.cu storage
// Extension RN_CAMERA_TEST_AABB: clears the visibility bits of the flagged cameras for nodes whose transformed AABB
// fails the plane test and recounts tile_counts / super_totals (between the cull kernel and the emit kernel).
int aabb_filter_launch(rn_scene *scene, const rn_camera_frustum *cams, uint32_t ncams, cudaStream_t stream);
// SceneWithVisibilityLists volumes: keeps camera c's bit only for nodes of the default volume or of the volume chosen for
// c, recounts like the AABB filter.
int volume_filter_launch(rn_scene *scene, uint32_t ncams, cudaStream_t stream);
// rn_scene.cu: zeroes the per-superblock totals (every mask filter recounts them) / emits the lists of the last cull
int clear_super_totals(rn_scene *scene, cudaStream_t stream);
int launch_compaction(rn_scene *scene, uint32_t ncams, cudaStream_t stream);
void occlusion_free(rn_scene *scene);   // rn_occlusion.cu
int scene_flip_outputs(rn_scene *scene);   // rn_scene.cu: swaps the two sets of cull outputs
// rn_masks.cu: the rank-order masks of the last cull stored into n_dst buffers at once (peer-mapped pointers)
int build_masks_fanout(rn_scene *scene, uint32_t *const *dsts, uint32_t n_dst, uint32_t stride, cudaStream_t stream);

template <class T>
inline int dev_alloc(rn_scene *s, T **p, size_t count)
{
	*p = nullptr;
	if(count == 0) count = 1;
	RN_CUDA(cudaMalloc(reinterpret_cast<void **>(p), count * sizeof(T)));
	s->device_bytes += count * sizeof(T);
	return RN_OK;
}
}
