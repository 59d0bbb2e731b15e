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
// The digit totals of all passes are counted while the keys are built; every pass is then ONE kernel (one sweep): tile
// ticket, in-block stable rank (match.any per 32 consecutive entries + per-warp digit counters in shared memory),
// decoupled look-back over the earlier tiles' digit counts, reorder through shared memory, coalesced stores.
// HBM-bound in principle: 16 B per entry and pass (key + entry read once and written once).
#include "rn_internal.h"

#include <algorithm>
#include <cstring>

namespace rn {

constexpr uint32_t kSortThreads = 256;
constexpr uint32_t kSortItems = 16;
constexpr uint32_t kSortTile = kSortThreads * kSortItems;   // 4096 entries per CTA
constexpr uint32_t kSortWarps = kSortThreads / 32;
constexpr uint32_t kSortDigits = 5;                         // four key bytes + the render-priority digit

struct SortView
{
	const float4 *wpos;
	const uint32_t *slot_of_rank;
	const uint8_t *prio_group;      // may be nullptr (single priority)
	uint32_t *list;                 // the camera's visibility list
	const uint32_t *list_count;
	uint32_t *range;                // [2]: begin, end of the current sort
	uint2 *kv[2];                   // (key, entry) pairs, ping-pong: one 8-byte store per entry keeps a digit run contiguous
	uint32_t *ghist;                // [kSortDigits][256] digit totals of the range, then [kSortDigits][256] their exclusive scans
	uint32_t *ticket;               // [kSortDigits] tile hand-out counters, one per pass
	uint32_t *status;               // [tiles][256] decoupled look-back words of the running pass
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

// keys + entries of the range into the scratch arrays, and the digit histograms of ALL passes in the same sweep (they do
// not depend on the order of the entries).  DESC = back to front.  Persistent grid, shared-memory histograms.
template <bool DESC>
__global__ void __launch_bounds__(256) sort_build_kernel(SortView v, uint32_t with_priority)
{
	__shared__ uint32_t h[kSortDigits][256];
	for(uint32_t i = threadIdx.x; i < kSortDigits * 256; i += blockDim.x) (&h[0][0])[i] = 0;
	__syncthreads();
	const uint32_t begin = v.range[0], m = v.range[1] - v.range[0];
	for(uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < m; e += gridDim.x * blockDim.x)
	{
		const uint32_t entry = v.list[begin + e];
		const float4 p = v.wpos[entry_slot(v, entry)];
		// (a - b).GetDotProduct(a - b): x*x + y*y + z*z, separately rounded (-fmad=false)
		const float dx = p.x - v.cam[0], dy = p.y - v.cam[1], dz = p.z - v.cam[2];
		uint32_t key = __float_as_uint(dx * dx + dy * dy + dz * dz);
		uint32_t g = 0;
		if(DESC)
			key = ~key;
		else
		{
			g = entry_group(v, entry);
			if(g >= v.first_sky_group)
				key = 0;   // priorities >= RenderSky compare equal inside their priority (RNSceneBasic.cpp:601-605)
		}
		v.kv[0][e] = make_uint2(key, entry);
		atomicAdd(&h[0][key & 255u], 1u);
		atomicAdd(&h[1][(key >> 8) & 255u], 1u);
		// the upper digits take few values (exponent bits, a handful of priorities): count warp-wide, one atomic per value
		const uint32_t act = __activemask();
		{
			const uint32_t d = (key >> 16) & 255u, peers = __match_any_sync(act, d);
			if((uint32_t)(__ffs(peers) - 1) == (threadIdx.x & 31u)) atomicAdd(&h[2][d], __popc(peers));
		}
		{
			const uint32_t d = key >> 24, peers = __match_any_sync(act, d);
			if((uint32_t)(__ffs(peers) - 1) == (threadIdx.x & 31u)) atomicAdd(&h[3][d], __popc(peers));
		}
		if(with_priority)
		{
			const uint32_t peers = __match_any_sync(act, g);
			if((uint32_t)(__ffs(peers) - 1) == (threadIdx.x & 31u)) atomicAdd(&h[4][g], __popc(peers));
		}
	}
	__syncthreads();
	for(uint32_t i = threadIdx.x; i < kSortDigits * 256; i += blockDim.x)
	{
		const uint32_t c = (&h[0][0])[i];
		if(c) atomicAdd(&v.ghist[i], c);
	}
}

// exclusive scan of every pass's 256 digit totals (one warp per pass)
__global__ void __launch_bounds__(32 * kSortDigits) sort_prefix_kernel(SortView v)
{
	const uint32_t lane = threadIdx.x & 31u, pass = threadIdx.x >> 5;
	const uint32_t *src = v.ghist + pass * 256u;
	uint32_t *dst = v.ghist + (kSortDigits + pass) * 256u;
	uint32_t carry = 0;
	for(uint32_t b = 0; b < 256; b += 32)
	{
		const uint32_t x = src[b + lane];
		uint32_t s = x;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
			if(lane >= (uint32_t)o) s += y;
		}
		dst[b + lane] = carry + s - x;
		carry += __shfl_sync(0xffffffffu, s, 31);
	}
}

struct PassArgs
{
	uint32_t src;      // 0 / 1: which scratch pair is read
	uint32_t shift;    // bit offset of the digit inside the key; 32 = the priority digit
	uint32_t last;     // 1: final pass - the entries go straight back into the list, the keys are dropped
	uint32_t pass;     // histogram row / ticket of this pass
};

__device__ __forceinline__ uint32_t digit_of(const SortView &v, const PassArgs &a, uint32_t key, uint32_t val)
{
	return a.shift < 32 ? (key >> a.shift) & 255u : entry_group(v, val);
}

// One pass of the LSD radix sort in ONE kernel ("one sweep"): a CTA takes the next tile of 4 096 entries (ticket), ranks
// them by digit (stable: warp w owns entries [512 w, 512 w + 512) of the tile, 32 consecutive entries per round, match.any
// + per-warp digit counters), publishes its 256 digit counts and finds the counts of all earlier tiles by decoupled
// look-back (thread d follows digit d: status word = 2 flag bits + 30 count bits; 1 = this tile's count, 2 = inclusive
// prefix), then writes the tile reordered by digit so that consecutive threads store to consecutive addresses.
// 16 bytes per entry and pass move through HBM - the keys and entries are read once and written once.
__global__ void __launch_bounds__(kSortThreads) sort_onesweep_kernel(SortView v, PassArgs a)
{
	__shared__ uint32_t wcount[kSortWarps][256];
	__shared__ uint32_t dbase[256];
	__shared__ uint32_t lpre[256];
	__shared__ uint32_t wsum[8];
	__shared__ uint32_t s_valid, s_tile;
	__shared__ uint2 s_kv[kSortTile];
	__shared__ uint8_t s_dig[kSortTile];
	const uint32_t m = v.range[1] - v.range[0];
	const uint32_t nb = (m + kSortTile - 1) / kSortTile;
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	if(threadIdx.x == 0)
		s_tile = atomicAdd(&v.ticket[a.pass], 1u);
	for(uint32_t i = threadIdx.x; i < kSortWarps * 256; i += kSortThreads)
		(&wcount[0][0])[i] = 0;
	__syncthreads();
	const uint32_t tile = s_tile;
	if(tile >= nb)
		return;
	// selects, not v.keys[a.src]: a dynamically indexed kernel parameter is copied to local memory by every thread
	const uint2 *__restrict__ kv = a.src ? v.kv[1] : v.kv[0];
	uint2 *__restrict__ okv = a.src ? v.kv[0] : v.kv[1];
	uint32_t *__restrict__ olist = v.list + v.range[0];
	const uint32_t wbase = tile * kSortTile + warp * (kSortItems * 32u);
	uint32_t key[kSortItems], val[kSortItems], dr[kSortItems];     // dr = digit | rank inside the warp's digit run << 8
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		const uint32_t e = wbase + r * 32u + lane;
		const uint2 x = e < m ? __ldg(kv + e) : make_uint2(0u, 0u);
		key[r] = x.x;
		val[r] = x.y;
	}
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		const bool valid = wbase + r * 32u + lane < m;
		const uint32_t dig = valid ? digit_of(v, a, key[r], val[r]) : 0xffffu;
		const uint32_t peers = __match_any_sync(0xffffffffu, dig);
		const uint32_t leader = __ffs(peers) - 1u;
		uint32_t base = 0;
		if(valid && lane == leader)
		{
			base = wcount[warp][dig];
			wcount[warp][dig] = base + __popc(peers);
		}
		base = __shfl_sync(0xffffffffu, base, leader);
		dr[r] = valid ? (dig | ((base + __popc(peers & ((1u << lane) - 1u))) << 8)) : 0xffffffffu;
		__syncwarp();
	}
	__syncthreads();
	// per digit (thread = digit): exclusive prefix over the warps; `run` ends as the digit's count in this tile
	uint32_t run = 0;
#pragma unroll
	for(uint32_t w = 0; w < kSortWarps; w++)
	{
		const uint32_t t = wcount[w][threadIdx.x];
		wcount[w][threadIdx.x] = run;
		run += t;
	}
	// decoupled look-back: the number of entries with this digit in all earlier tiles
	{
		volatile uint32_t *status = v.status;
		const uint32_t d = threadIdx.x;
		uint32_t excl = 0;
		if(tile > 0)
		{
			status[(size_t)tile * 256u + d] = (1u << 30) | run;
			// walk back eight tiles at a time: the eight loads are in flight together, so a window of several hundred
			// concurrently running tiles costs a few dozen round trips instead of one per tile
			int64_t t = (int64_t)tile - 1;
			bool done = false;
			while(!done)
			{
				uint32_t sv[8];
#pragma unroll
				for(int k = 0; k < 8; k++)
					sv[k] = t - k >= 0 ? status[(size_t)(t - k) * 256u + d] : (2u << 30);
#pragma unroll
				for(int k = 0; k < 8; k++)
				{
					if(done) break;
					uint32_t x = sv[k];
					while((x >> 30) == 0u) x = status[(size_t)(t - k) * 256u + d];     // not published yet: wait for this one
					excl += x & 0x3fffffffu;
					if((x >> 30) == 2u) done = true;
				}
				t -= 8;
			}
		}
		status[(size_t)tile * 256u + d] = (2u << 30) | (excl + run);
		dbase[d] = v.ghist[(kSortDigits + a.pass) * 256u + d] + excl;
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
#pragma unroll
	for(uint32_t r = 0; r < kSortItems; r++)
	{
		if(dr[r] != 0xffffffffu)
		{
			const uint32_t dig = dr[r] & 255u;
			const uint32_t lp = lpre[dig] + wcount[warp][dig] + (dr[r] >> 8);
			s_kv[lp] = make_uint2(key[r], val[r]);
			s_dig[lp] = (uint8_t)dig;
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
			if(a.last) olist[dst] = s_kv[idx].y;
			else okv[dst] = s_kv[idx];
		}
	}
}

static int run_passes(rn_scene *s, const SortView &v, bool with_priority_digit, cudaStream_t stream)
{
	const uint32_t grid_tiles = (s->n + kSortTile - 1) / kSortTile;
	uint32_t src = 0;
	const uint32_t n_pass = with_priority_digit ? 5u : 4u;
	sort_prefix_kernel<<<1, 32 * kSortDigits, 0, stream>>>(v);
	s->launches++;
	for(uint32_t p = 0; p < n_pass; p++)
	{
		PassArgs a;
		a.src = src;
		a.shift = p * 8u;
		a.last = p + 1 == n_pass ? 1u : 0u;
		a.pass = p;
		RN_CUDA(cudaMemsetAsync(v.status, 0, (size_t)grid_tiles * 256u * sizeof(uint32_t), stream));
		sort_onesweep_kernel<<<grid_tiles, kSortThreads, 0, stream>>>(v, a);
		s->launches++;
		src ^= 1u;
	}
	// no write-back pass: the last sweep wrote the entries straight into the list
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

} // namespace rn

using namespace rn;

extern "C" int rn_cull_sort(rn_scene_t *s, const float *camera_world_positions, uint32_t n_cameras, void *cuda_stream)
{
	RN_RANGE("rn_cull_sort");
	RN_REQUIRE(s && camera_world_positions, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_sort needs the lists of a cull");
	RN_REQUIRE(n_cameras == s->last_ncams, RN_ERR_INVALID, "the last cull had %u cameras, %u positions given", s->last_ncams, n_cameras);
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const uint32_t nb_max = (s->desc.max_nodes + kSortTile - 1) / kSortTile;
	RN_REQUIRE(s->desc.max_nodes < (1u << 30), RN_ERR_CAPACITY, "rn_cull_sort handles lists below 2^30 entries");
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
		// look-back status words of one pass + digit totals and their scans + the passes' tile tickets
		if(!rc) rc = dev_alloc(s, &s->sort_hist, (size_t)256 * nb_max + 2 * kSortDigits * 256 + 8);
		if(!rc) rc = dev_alloc(s, &s->sort_ranges, (size_t)RN_MAX_CAMERAS * 2);
		if(rc) return rc;
	}
	const uint32_t grid_flat = (s->n + 255) / 256;
	if(!s->sm_count)
		RN_CUDA(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, s->device));
	const uint32_t grid_build = std::max(1u, std::min(grid_flat, (uint32_t)s->sm_count * 8u));
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
		v.kv[0] = reinterpret_cast<uint2 *>(s->sort_scratch);
		v.kv[1] = reinterpret_cast<uint2 *>(s->sort_scratch) + cap;
		v.status = s->sort_hist;
		v.ghist = s->sort_hist + (size_t)256 * nb_max;
		v.ticket = v.ghist + 2 * kSortDigits * 256;
		v.nb_max = nb_max;
		v.rank_mode = s->list_rank_mode;
		v.rank_offset = s->list_rank_offset;
		v.transparent_group = s->transparent_group;
		v.first_sky_group = s->first_sky_group;
		memcpy(v.cam, camera_world_positions + 3 * c, sizeof(float) * 3);
		if(f & RN_CAMERA_SORT_FRONT_TO_BACK)
		{
			const bool with_prio = s->n_prio_groups > 1;
			sort_range_all_kernel<<<1, 1, 0, stream>>>(v);
			RN_CUDA(cudaMemsetAsync(v.ghist, 0, (2 * kSortDigits * 256 + 8) * sizeof(uint32_t), stream));
			sort_build_kernel<false><<<grid_build, 256, 0, stream>>>(v, with_prio ? 1u : 0u);
			s->launches += 2;
			int rc = run_passes(s, v, with_prio, stream);
			if(rc) return rc;
		}
		if((f & RN_CAMERA_SORT_TRANSPARENT_BACK_TO_FRONT) && s->transparent_group != 0xffffffffu)
		{
			sort_range_reset_kernel<<<1, 1, 0, stream>>>(v);
			sort_range_scan_kernel<<<grid_flat, 256, 0, stream>>>(v);
			sort_range_close_kernel<<<1, 1, 0, stream>>>(v);
			RN_CUDA(cudaMemsetAsync(v.ghist, 0, (2 * kSortDigits * 256 + 8) * sizeof(uint32_t), stream));
			sort_build_kernel<true><<<grid_build, 256, 0, stream>>>(v, 0u);
			s->launches += 4;
			int rc = run_passes(s, v, false, stream);
			if(rc) return rc;
		}
	}
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}
