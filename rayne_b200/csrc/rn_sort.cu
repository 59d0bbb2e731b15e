// librn_cuda - the post-cull sorts of SceneBasic::Render on the device (sm_100a).
//
// Reference behaviour being replaced (Source/Scene/RNSceneBasic.cpp:576-616):
//   SortFrontToBack            std::sort of the whole list by (render priority, squared distance ascending inside equal
//                              priorities below RenderSky)
//   SortTransparentBackToFront std::sort of [firstTransparentIndex, lastTransparentIndex] by squared distance
//                              descending; firstTransparentIndex == 0 doubles as "unset" (:583-587)
//   squared distance           Vector3::GetSquaredDistance, Source/Math/RNVector.h:543-547 (fp32, left to right)
//
// Implementation: a stable LSD radix sort (8-bit digits) of (key, list entry) pairs whose range [begin, end) lives
// in device memory, so rn_cull_sort never synchronises with the host: every launch is sized for the node count and
// blocks outside the range leave at once.  Keys: the fp32 bits of the squared distance (non-negative, so the
// unsigned order is the float order; complemented for the descending sort); the priority of an entry is a fifth
// digit looked up through the entry (dense index of the node's render priority, built at upload).  Stable means
// equal keys keep render-list order - one of the outcomes std::sort permits.
//
// Per pass: histogram (one 2048-entry block per CTA) -> per-digit exclusive scan over the blocks -> scatter with an
// in-block stable rank (match.any per 32 consecutive entries + per-warp digit counters in shared memory).
// HBM-bound in principle: 16 B per entry and pass (key + entry read and written) + the scattered writes.
#include "rn_internal.h"

#include <cstring>

namespace rn {

constexpr uint32_t kSortThreads = 256;
constexpr uint32_t kSortItems = 8;
constexpr uint32_t kSortTile = kSortThreads * kSortItems;   // 2048 entries per CTA
constexpr uint32_t kSortWarps = kSortThreads / 32;

struct SortView
{
	const float4 *wpos;
	const uint32_t *slot_of_rank;
	const uint8_t *prio_group;      // may be nullptr (single priority)
	uint32_t *list;                 // the camera's visibility list
	const uint32_t *list_count;
	uint32_t *range;                // [2]: begin, end of the current sort
	uint32_t *keys[2], *vals[2];
	uint32_t *hist;                 // [256][nb_max] block histograms, then [256] digit totals
	uint32_t nb_max;
	uint32_t rank_mode, rank_offset;
	uint32_t transparent_group, first_sky_group;
	float cam[3];
};

__device__ __forceinline__ uint32_t entry_slot(const SortView &v, uint32_t entry)
{
	return v.rank_mode ? v.slot_of_rank[entry - v.rank_offset] : entry;
}

__device__ __forceinline__ uint32_t entry_group(const SortView &v, uint32_t entry)
{
	return v.prio_group ? (uint32_t)v.prio_group[entry_slot(v, entry)] : 0u;
}

// range = whole list (front-to-back sort)
__global__ void sort_range_all_kernel(SortView v)
{
	v.range[0] = 0;
	v.range[1] = *v.list_count;
}

// range = [first, last + 1) of the RenderTransparent entries with the reference's sentinel semantics.
// Phase 0 resets, phase 1 scans the list, phase 2 closes the range.
__global__ void sort_range_reset_kernel(SortView v)
{
	v.range[0] = 0xffffffffu;
	v.range[1] = 0;
}

__global__ void __launch_bounds__(256) sort_range_scan_kernel(SortView v)
{
	const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
	if(q >= *v.list_count)
		return;
	if(entry_group(v, v.list[q]) == v.transparent_group)
	{
		if(q != 0)
			atomicMin(&v.range[0], q);   // firstTransparentIndex: an entry at index 0 leaves it "unset"
		atomicMax(&v.range[1], q);       // lastTransparentIndex
	}
}

__global__ void sort_range_close_kernel(SortView v)
{
	const uint32_t first = v.range[0] == 0xffffffffu ? 0u : v.range[0], last = v.range[1];
	if(first < last) { v.range[0] = first; v.range[1] = last + 1; }
	else { v.range[0] = 0; v.range[1] = 0; }
}

// keys + entries of the range into the scratch arrays.  DESC = back to front.
template <bool DESC>
__global__ void __launch_bounds__(256) sort_build_kernel(SortView v)
{
	const uint32_t begin = v.range[0], end = v.range[1];
	const uint32_t e = blockIdx.x * blockDim.x + threadIdx.x;
	if(e >= end - begin)
		return;
	const uint32_t entry = v.list[begin + e];
	const float4 p = v.wpos[entry_slot(v, entry)];
	// (a - b).GetDotProduct(a - b): x*x + y*y + z*z, separately rounded (-fmad=false)
	const float dx = p.x - v.cam[0], dy = p.y - v.cam[1], dz = p.z - v.cam[2];
	uint32_t key = __float_as_uint(dx * dx + dy * dy + dz * dz);
	if(DESC)
		key = ~key;
	else if(entry_group(v, entry) >= v.first_sky_group)
		key = 0;   // priorities >= RenderSky compare equal inside their priority (RNSceneBasic.cpp:601-605)
	v.keys[0][e] = key;
	v.vals[0][e] = entry;
}

struct PassArgs
{
	uint32_t src;      // 0 / 1: which scratch pair is read
	uint32_t shift;    // bit offset of the digit inside the key; 32 = the priority digit
	uint32_t last;     // 1: final pass - the entries go straight back into the list, the keys are dropped
};

__device__ __forceinline__ uint32_t digit_of(const SortView &v, const PassArgs &a, uint32_t key, uint32_t val)
{
	return a.shift < 32 ? (key >> a.shift) & 255u : entry_group(v, val);
}

__global__ void __launch_bounds__(kSortThreads) sort_hist_kernel(SortView v, PassArgs a)
{
	__shared__ uint32_t h[256];
	const uint32_t m = v.range[1] - v.range[0];
	const uint32_t nb = (m + kSortTile - 1) / kSortTile;
	if(blockIdx.x >= nb)
		return;
	h[threadIdx.x] = 0;
	__syncthreads();
	const uint32_t base = blockIdx.x * kSortTile;
	const uint32_t *keys = a.src ? v.keys[1] : v.keys[0], *vals = a.src ? v.vals[1] : v.vals[0];
#pragma unroll
	for(uint32_t i = 0; i < kSortItems; i++)
	{
		const uint32_t e = base + i * kSortThreads + threadIdx.x;
		if(e < m)
			atomicAdd(&h[digit_of(v, a, keys[e], a.shift < 32 ? 0u : vals[e])], 1u);
	}
	__syncthreads();
	v.hist[threadIdx.x * v.nb_max + blockIdx.x] = h[threadIdx.x];
}

// One CTA per digit: exclusive scan of that digit's block counts, total behind the table.
__global__ void __launch_bounds__(256) sort_scan_kernel(SortView v)
{
	__shared__ uint32_t warp_sum[8];
	__shared__ uint32_t carry_s;
	const uint32_t m = v.range[1] - v.range[0];
	const uint32_t nb = (m + kSortTile - 1) / kSortTile;
	uint32_t *row = v.hist + (size_t)blockIdx.x * v.nb_max;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if(threadIdx.x == 0) carry_s = 0;
	__syncthreads();
	for(uint32_t b0 = 0; b0 < nb; b0 += 256)
	{
		const uint32_t b = b0 + threadIdx.x;
		const uint32_t x = b < nb ? row[b] : 0u;
		uint32_t s = x;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
			if(lane >= (uint32_t)o) s += y;
		}
		if(lane == 31) warp_sum[warp] = s;
		__syncthreads();
		uint32_t before = carry_s;
		for(uint32_t w = 0; w < warp; w++) before += warp_sum[w];
		if(b < nb) row[b] = before + s - x;
		__syncthreads();
		if(threadIdx.x == 255) carry_s = before + s;
		__syncthreads();
	}
	if(threadIdx.x == 0)
		v.hist[(size_t)256 * v.nb_max + blockIdx.x] = carry_s;
}

__global__ void __launch_bounds__(kSortThreads) sort_scatter_kernel(SortView v, PassArgs a)
{
	__shared__ uint32_t wcount[kSortWarps][256];
	__shared__ uint32_t dbase[256];
	__shared__ uint32_t lpre[256];
	__shared__ uint32_t wsum[8];
	__shared__ uint32_t s_valid;
	__shared__ uint32_t s_key[kSortTile], s_val[kSortTile];
	__shared__ uint8_t s_dig[kSortTile];
	const uint32_t m = v.range[1] - v.range[0];
	const uint32_t nb = (m + kSortTile - 1) / kSortTile;
	if(blockIdx.x >= nb)
		return;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	for(uint32_t i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads)
		(&wcount[0][0])[i] = 0;
	// exclusive scan of the 256 digit totals -> first output index of every digit
	{
		const uint32_t x = v.hist[(size_t)256 * v.nb_max + threadIdx.x];
		uint32_t s = x;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
			if(lane >= (uint32_t)o) s += y;
		}
		if(lane == 31) wsum[warp] = s;
		__syncthreads();
		uint32_t before = 0;
		for(uint32_t w = 0; w < warp; w++) before += wsum[w];
		dbase[threadIdx.x] = before + s - x + v.hist[(size_t)threadIdx.x * v.nb_max + blockIdx.x];
	}
	__syncthreads();
	// selects, not v.keys[a.src]: a dynamically indexed kernel parameter is copied to local memory by every thread
	const uint32_t *keys = a.src ? v.keys[1] : v.keys[0], *vals = a.src ? v.vals[1] : v.vals[0];
	uint32_t *okeys = a.src ? v.keys[0] : v.keys[1];
	uint32_t *ovals = a.last ? v.list + v.range[0] : (a.src ? v.vals[0] : v.vals[1]);
	// warp w owns entries [w*256, (w+1)*256) of the block, 32 consecutive entries per round: stable inside the warp
	const uint32_t wbase = blockIdx.x * kSortTile + warp * (kSortItems * 32u);
	uint32_t key[kSortItems], val[kSortItems], dig[kSortItems], rank[kSortItems];
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		const uint32_t e = wbase + r * 32u + lane;
		const bool valid = e < m;
		key[r] = valid ? keys[e] : 0u;
		val[r] = valid ? vals[e] : 0u;
		dig[r] = valid ? digit_of(v, a, key[r], val[r]) : 0xffffu;
		const uint32_t peers = __match_any_sync(0xffffffffu, dig[r]);
		const uint32_t leader = __ffs(peers) - 1u;
		uint32_t base = 0;
		if(valid && lane == leader)
		{
			base = wcount[warp][dig[r]];
			wcount[warp][dig[r]] = base + __popc(peers);
		}
		base = __shfl_sync(0xffffffffu, base, leader);
		rank[r] = base + __popc(peers & ((1u << lane) - 1u));
		__syncwarp();
	}
	__syncthreads();
	// per digit: exclusive prefix over the warps (thread = digit); `run` ends as the digit's total in this block
	uint32_t run = 0;
#pragma unroll
	for(uint32_t w = 0; w < kSortWarps; w++)
	{
		const uint32_t t = wcount[w][threadIdx.x];
		wcount[w][threadIdx.x] = run;
		run += t;
	}
	// block-local exclusive prefix over the digits: where a digit's run starts in the reordered tile
	{
		uint32_t sc = run;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, sc, o);
			if(lane >= (uint32_t)o) sc += y;
		}
		if(lane == 31) wsum[warp] = sc;
		__syncthreads();
		uint32_t before = 0;
		for(uint32_t w = 0; w < warp; w++) before += wsum[w];
		lpre[threadIdx.x] = before + sc - run;
		if(threadIdx.x == kSortThreads - 1) s_valid = before + sc;
	}
	__syncthreads();
	// reorder through shared memory, then write digit runs with consecutive threads on consecutive addresses
	// (a thread-per-entry scatter leaves 32 partial sectors per store instruction)
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		if(dig[r] != 0xffffu)
		{
			const uint32_t lp = lpre[dig[r]] + wcount[warp][dig[r]] + rank[r];
			s_key[lp] = key[r];
			s_val[lp] = val[r];
			s_dig[lp] = (uint8_t)dig[r];
		}
	}
	__syncthreads();
	const uint32_t n_valid = s_valid;
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		const uint32_t idx = r * kSortThreads + threadIdx.x;
		if(idx < n_valid)
		{
			const uint32_t d = s_dig[idx];
			const uint32_t dst = dbase[d] + (idx - lpre[d]);
			if(!a.last) okeys[dst] = s_key[idx];
			ovals[dst] = s_val[idx];
		}
	}
}

static int run_passes(rn_scene *s, const SortView &v, bool with_priority_digit, cudaStream_t stream)
{
	const uint32_t grid_tiles = (s->n + kSortTile - 1) / kSortTile;
	uint32_t src = 0;
	const uint32_t n_pass = with_priority_digit ? 5u : 4u;
	for(uint32_t p = 0; p < n_pass; p++)
	{
		PassArgs a;
		a.src = src;
		a.shift = p * 8u;
		a.last = p + 1 == n_pass ? 1u : 0u;
		sort_hist_kernel<<<grid_tiles, kSortThreads, 0, stream>>>(v, a);
		sort_scan_kernel<<<256, 256, 0, stream>>>(v);
		sort_scatter_kernel<<<grid_tiles, kSortThreads, 0, stream>>>(v, a);
		s->launches += 3;
		src ^= 1u;
	}
	// no write-back pass: the last scatter wrote the entries straight into the list
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

} // namespace rn

using namespace rn;

extern "C" int rn_cull_sort(rn_scene_t *s, const float *camera_world_positions, uint32_t n_cameras, void *cuda_stream)
{
	RN_REQUIRE(s && camera_world_positions, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_sort needs the lists of a cull");
	RN_REQUIRE(n_cameras == s->last_ncams, RN_ERR_INVALID, "the last cull had %u cameras, %u positions given", s->last_ncams, n_cameras);
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const uint32_t nb_max = (s->desc.max_nodes + kSortTile - 1) / kSortTile;
	bool any = false;
	for(uint32_t c = 0; c < n_cameras; c++)
	{
		const uint32_t f = s->last_cam_flags[c];
		if((f & RN_CAMERA_SORT_FRONT_TO_BACK) || ((f & RN_CAMERA_SORT_TRANSPARENT_BACK_TO_FRONT) && s->transparent_group != 0xffffffffu))
			any = true;
	}
	if(!any)
		return RN_OK;   // nothing to do: no front-to-back camera and no RenderTransparent node in the scene
	if(!s->sort_scratch)
	{
		int rc = dev_alloc(s, &s->sort_scratch, (size_t)s->desc.max_nodes * 4);
		if(!rc) rc = dev_alloc(s, &s->sort_hist, (size_t)256 * nb_max + 256);
		if(!rc) rc = dev_alloc(s, &s->sort_ranges, (size_t)RN_MAX_CAMERAS * 2);
		if(rc) return rc;
	}
	const uint32_t grid_flat = (s->n + 255) / 256;
	for(uint32_t c = 0; c < n_cameras; c++)
	{
		const uint32_t f = s->last_cam_flags[c];
		SortView v;
		memset(&v, 0, sizeof(v));
		v.wpos = s->wpos;
		v.slot_of_rank = s->d_slot_of_rank;
		v.prio_group = s->prio_group;
		v.list = s->lists[c];
		v.list_count = s->list_counts + c;
		v.range = s->sort_ranges + 2 * c;
		const size_t cap = s->desc.max_nodes;
		v.keys[0] = s->sort_scratch; v.vals[0] = s->sort_scratch + cap;
		v.keys[1] = s->sort_scratch + 2 * cap; v.vals[1] = s->sort_scratch + 3 * cap;
		v.hist = s->sort_hist;
		v.nb_max = nb_max;
		v.rank_mode = s->list_rank_mode;
		v.rank_offset = s->list_rank_offset;
		v.transparent_group = s->transparent_group;
		v.first_sky_group = s->first_sky_group;
		memcpy(v.cam, camera_world_positions + 3 * c, sizeof(float) * 3);
		if(f & RN_CAMERA_SORT_FRONT_TO_BACK)
		{
			sort_range_all_kernel<<<1, 1, 0, stream>>>(v);
			sort_build_kernel<false><<<grid_flat, 256, 0, stream>>>(v);
			s->launches += 2;
			int rc = run_passes(s, v, s->n_prio_groups > 1, stream);
			if(rc) return rc;
		}
		if((f & RN_CAMERA_SORT_TRANSPARENT_BACK_TO_FRONT) && s->transparent_group != 0xffffffffu)
		{
			sort_range_reset_kernel<<<1, 1, 0, stream>>>(v);
			sort_range_scan_kernel<<<grid_flat, 256, 0, stream>>>(v);
			sort_range_close_kernel<<<1, 1, 0, stream>>>(v);
			sort_build_kernel<true><<<grid_flat, 256, 0, stream>>>(v);
			s->launches += 4;
			int rc = run_passes(s, v, false, stream);
			if(rc) return rc;
		}
	}
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}
