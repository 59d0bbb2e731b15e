// Internal declarations shared by the translation units of librn_cuda (not part of the ABI).
#pragma once
#include "../../include/rn_cuda.h"
#include <cuda_runtime.h>
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

constexpr uint32_t kMaxSublevels = 12;   // sublevel 0 = spine nodes of the tile, 1.. = depth inside a unit
constexpr uint32_t kWarpsPerCta = 4;     // tile kernel: one tile per warp
constexpr uint32_t kTileThreads = kWarpsPerCta * 32;

// One tile = a contiguous range of ranks [start, start+count) stored at the same slots, permuted so that each
// sublevel is contiguous and, inside a sublevel, the inner nodes (nodes with children) come first.  64 bytes.
struct TileMeta
{
	uint32_t start;
	uint16_t count;
	uint16_t nlev;
	uint16_t lvl_end[kMaxSublevels];   // end offset (exclusive) of sublevel k inside the tile
	uint16_t inner_end[kMaxSublevels]; // number of inner nodes in sublevels 1..k (sublevel 0 holds none)
	uint32_t pad[2];
};
static_assert(sizeof(TileMeta) == 64, "TileMeta must be 64 bytes");

struct ParticleStore;

} // namespace rn

struct rn_scene
{
	rn_scene_desc desc{};
	int device = 0;
	uint32_t tile_cap = 512;
	uint32_t inner_cap = 96;     // inner nodes (parents kept in shared memory) per tile
	uint32_t n = 0;              // nodes
	uint32_t n_tiles = 0;
	uint32_t max_depth = 0;
	uint32_t spine_nodes = 0;
	uint64_t device_bytes = 0;
	uint64_t launches = 0;
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
	uint8_t *vis = nullptr;      // slot -> camera bit mask of the last cull
	rn::TileMeta *tiles = nullptr;
	uint32_t *tile_counts = nullptr;   // [n_tiles][RN_MAX_CAMERAS] visible nodes per tile (last cull)
	uint32_t *super_totals = nullptr;  // [n_tiles / 256][RN_MAX_CAMERAS]
	uint32_t *list_counts = nullptr;   // [RN_MAX_CAMERAS]
	uint32_t *lists[RN_MAX_CAMERAS] = {};
	uint32_t list_capacity = 0;
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

	rn::ParticleStore *particles = nullptr;
};

namespace rn {
int particles_destroy(rn_scene *scene);
}
