// Visibility of the last cull as bitmasks in render-list (rank) order, and their expansion back into lists.
//
// A visibility list of SceneBasic::Render (Source/Scene/RNSceneBasic.cpp:576-593) is the ascending sequence of the
// render-list ranks that passed CanRender, so one bit per rank holds exactly the same information in 1/32 of a list
// entry: 2.1 MB per camera for 16.8 M nodes instead of 4 B per visible node (61 MB for the six views of the
// benchmark frame).  That is the form the multi-GPU exchange ships (shard.PipelinedMaskExchange): every rank
// all-gathers the masks of its rank range and expands the cameras it consumes with rn_masks_expand, which reproduces
// the reference list bit for bit (tests/test_gpu_masks.py).
//
// Both kernels are HBM-bound integer work: rank_mask_kernel reads 4 B (rank -> slot) + 1 B (mask byte) per node and
// writes V/8 B; the expansion reads the words once or twice and writes 4 B per visible node.
#include "rn_internal.h"
#include <algorithm>
#include <string.h>
#include <mutex>

namespace rn {

constexpr uint32_t kMaskThreads = 256;

// The cull kernels store the mask bytes in RANK order (rn_scene.cu), so word w of camera c is bit c of the 32 bytes
// [32 w, 32 w + 32): one thread per word loads its 32 bytes (two 128-bit loads; the bytes beyond the node count are
// kept zero) and gathers bit c of four bytes at a time into a nibble with one multiply.  Consecutive threads store
// consecutive words: coalesced 128-byte stores.  1 B read and V/8 B written per node.
//
// MULTICAST: `masks` is an NVLS multicast address (one virtual address mapped onto the same buffer of every GPU of
// the group): a multimem store is replicated by the NVSwitch to all of them, so the kernel that BUILDS the masks also
// performs the all-gather - every GPU's block lands in every GPU's copy while the warps are still computing.
template <uint32_t NCAMS, bool MULTICAST>
__global__ void __launch_bounds__(kMaskThreads) rank_mask_kernel(const uint8_t *__restrict__ vis, uint32_t words, uint32_t stride,
                                                                 uint32_t *__restrict__ masks)
{
	const uint32_t w = blockIdx.x * kMaskThreads + threadIdx.x;
	if(w >= words)
		return;
	const uint4 lo = __ldg(reinterpret_cast<const uint4 *>(vis) + 2u * w), hi = __ldg(reinterpret_cast<const uint4 *>(vis) + 2u * w + 1u);
	const uint32_t v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
	for(uint32_t c = 0; c < NCAMS; c++)
	{
		uint32_t word = 0;
#pragma unroll
		for(uint32_t g = 0; g < 8; g++)
		{
			// bits c, 8 + c, 16 + c, 24 + c of v[g] -> bits 28..31 of the product, shifted in from the top
			const uint32_t t = (v[g] >> c) & 0x01010101u;
			word = (word >> 4) | ((t * 0x10204080u) & 0xf0000000u);
		}
		uint32_t *dst = masks + (size_t)c * stride + w;
		if(MULTICAST)
			asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(dst), "r"(word) : "memory");
		else
			*dst = word;
	}
}

// The same words stored into the gathered buffers of several GPUs of ONE process through peer-mapped pointers (NVLink
// stores): the fallback of rn_group where no NVLS multicast object can be had.  dst.p[k] = row 0 of this GPU's block inside
// GPU k's buffer.
struct FanoutDst
{
	uint32_t *p[RN_MAX_GROUP];
	uint32_t n;
};

template <uint32_t NCAMS>
__global__ void __launch_bounds__(kMaskThreads) rank_mask_fanout_kernel(const uint8_t *__restrict__ vis, uint32_t words, uint32_t stride,
                                                                        const __grid_constant__ FanoutDst dst)
{
	const uint32_t w = blockIdx.x * kMaskThreads + threadIdx.x;
	if(w >= words)
		return;
	const uint4 lo = __ldg(reinterpret_cast<const uint4 *>(vis) + 2u * w), hi = __ldg(reinterpret_cast<const uint4 *>(vis) + 2u * w + 1u);
	const uint32_t v[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
#pragma unroll
	for(uint32_t c = 0; c < NCAMS; c++)
	{
		uint32_t word = 0;
#pragma unroll
		for(uint32_t g = 0; g < 8; g++)
		{
			const uint32_t t = (v[g] >> c) & 0x01010101u;
			word = (word >> 4) | ((t * 0x10204080u) & 0xf0000000u);
		}
		for(uint32_t k = 0; k < dst.n; k++)
			dst.p[k][(size_t)c * stride + w] = word;
	}
}

// ---- expansion: words -> ascending ranks -----------------------------------------------------------------------
constexpr uint32_t kExpandThreads = 256;
constexpr uint32_t kExpandWordsPerBlock = kExpandThreads * 4;   // 32 768 ranks per CTA

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *s_warp, uint32_t *total)
{
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	uint32_t x = v;
#pragma unroll
	for(int o = 1; o < 32; o <<= 1)
	{
		const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
		if(lane >= (uint32_t)o) x += y;
	}
	if(lane == 31) s_warp[warp] = x;
	__syncthreads();
	if(warp == 0)
	{
		uint32_t w = lane < kExpandThreads / 32 ? s_warp[lane] : 0u, z = w;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, z, o);
			if(lane >= (uint32_t)o) z += y;
		}
		if(lane < kExpandThreads / 32) s_warp[lane] = z - w;
		if(lane == 31) *total = z;
	}
	__syncthreads();
	return s_warp[warp] + x - v;
}

constexpr uint32_t kExpandMaxSegments = 32;

// The rank ranges of several GPUs as ONE virtual word array: CTA b works on words [1024 (b - first_block[s]), ...) of
// the segment s that holds it.
struct ExpandArgs
{
	const uint32_t *words[kExpandMaxSegments];
	uint32_t n_words[kExpandMaxSegments];
	uint32_t rank_base[kExpandMaxSegments];
	uint32_t first_block[kExpandMaxSegments + 1];
	uint32_t n_segments;
};

__device__ __forceinline__ uint32_t segment_of(const ExpandArgs &a, uint32_t block)
{
	uint32_t s = 0;
	while(s + 1 < a.n_segments && block >= a.first_block[s + 1]) s++;
	return s;
}

// pass 1: set bits per CTA
__global__ void __launch_bounds__(kExpandThreads) expand_count_kernel(const __grid_constant__ ExpandArgs a,
                                                                      uint32_t *__restrict__ block_totals)
{
	__shared__ uint32_t s_warp[32];
	const uint32_t seg = segment_of(a, blockIdx.x);
	const uint32_t *__restrict__ words = a.words[seg];
	const uint32_t n_words = a.n_words[seg];
	const uint32_t w0 = (blockIdx.x - a.first_block[seg]) * kExpandWordsPerBlock + threadIdx.x;
	uint32_t cnt = 0;
#pragma unroll
	for(int k = 0; k < 4; k++)
		if(w0 + k * kExpandThreads < n_words) cnt += __popc(__ldg(words + w0 + k * kExpandThreads));
	cnt = __reduce_add_sync(0xffffffffu, cnt);
	if((threadIdx.x & 31u) == 0) s_warp[threadIdx.x >> 5] = cnt;
	__syncthreads();
	if(threadIdx.x < 32)
	{
		uint32_t v = threadIdx.x < kExpandThreads / 32 ? s_warp[threadIdx.x] : 0u;
		v = __reduce_add_sync(0xffffffffu, v);
		if(threadIdx.x == 0) block_totals[blockIdx.x] = v;
	}
}

// pass 2 (one CTA): exclusive scan of the CTA totals on top of the running count of the output list
__global__ void __launch_bounds__(kExpandThreads) expand_scan_kernel(uint32_t *__restrict__ block_totals, uint32_t n_blocks,
                                                                     uint32_t *__restrict__ list_count)
{
	__shared__ uint32_t s_warp[32];
	__shared__ uint32_t s_total;
	uint32_t base = *list_count;
	for(uint32_t b0 = 0; b0 < n_blocks; b0 += kExpandThreads)
	{
		const uint32_t b = b0 + threadIdx.x;
		const uint32_t v = b < n_blocks ? block_totals[b] : 0u;
		const uint32_t ex = block_exclusive_scan(v, s_warp, &s_total);
		if(b < n_blocks) block_totals[b] = base + ex;
		base += s_total;
		__syncthreads();
	}
	if(threadIdx.x == 0) *list_count = base;
}

// pass 3: four rounds of 256 words; a thread unpacks its word into shared memory behind the entries of the threads
// before it, then the CTA copies the round's entries out with coalesced stores
__global__ void __launch_bounds__(kExpandThreads) expand_emit_kernel(const __grid_constant__ ExpandArgs a,
                                                                     const uint32_t *__restrict__ block_offsets,
                                                                     uint32_t *__restrict__ list, uint32_t capacity)
{
	__shared__ uint32_t s_warp[32];
	__shared__ uint32_t s_total;
	__shared__ uint32_t s_out[kExpandThreads * 32];
	const uint32_t seg = segment_of(a, blockIdx.x);
	const uint32_t *__restrict__ words = a.words[seg];
	const uint32_t n_words = a.n_words[seg];
	const uint32_t wblock = (blockIdx.x - a.first_block[seg]) * kExpandWordsPerBlock;
	uint32_t out = block_offsets[blockIdx.x];
	uint32_t w[4];
#pragma unroll
	for(int k = 0; k < 4; k++)
	{
		const uint32_t wi = wblock + k * kExpandThreads + threadIdx.x;
		w[k] = wi < n_words ? __ldg(words + wi) : 0u;
	}
#pragma unroll
	for(int k = 0; k < 4; k++)
	{
		uint32_t bits = w[k];
		uint32_t pos = block_exclusive_scan(__popc(bits), s_warp, &s_total);
		const uint32_t r0 = a.rank_base[seg] + (wblock + k * kExpandThreads + threadIdx.x) * 32u;
		while(bits)
		{
			s_out[pos++] = r0 + __ffs(bits) - 1u;
			bits &= bits - 1u;
		}
		__syncthreads();
		const uint32_t total = s_total;
		for(uint32_t j = threadIdx.x; j < total; j += kExpandThreads)
			if(out + j < capacity) __stcs(list + out + j, s_out[j]);
		out += total;
		__syncthreads();
	}
}

} // namespace rn

using namespace rn;

static inline uint32_t mask_words(const rn_scene *s) { return (s->n + 31u) / 32u; }

template <bool MULTICAST>
static int build_masks_launch(rn_scene *s, uint32_t *out, uint32_t stride, cudaStream_t stream)
{
	const uint32_t words = mask_words(s);
	const uint32_t grid = (words + kMaskThreads - 1) / kMaskThreads;
	switch(s->last_ncams)
	{
#define RN_MASK_CASE(V) case V: rank_mask_kernel<V, MULTICAST><<<grid, kMaskThreads, 0, stream>>>(s->vis, words, stride, out); break;
		RN_MASK_CASE(1) RN_MASK_CASE(2) RN_MASK_CASE(3) RN_MASK_CASE(4) RN_MASK_CASE(5) RN_MASK_CASE(6) RN_MASK_CASE(7)
		default: rank_mask_kernel<8, MULTICAST><<<grid, kMaskThreads, 0, stream>>>(s->vis, words, stride, out); break;
#undef RN_MASK_CASE
	}
	RN_CUDA(cudaGetLastError());
	s->launches++;
	return RN_OK;
}

static int build_masks(rn_scene *s, uint32_t *out, uint32_t stride, cudaStream_t stream, bool multicast = false)
{
	return multicast ? build_masks_launch<true>(s, out, stride, stream) : build_masks_launch<false>(s, out, stride, stream);
}

extern "C" int rn_cull_build_masks(rn_scene_t *s, void *cuda_stream)
{
	RN_RANGE("rn_cull_build_masks");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_build_masks needs a cull");
	RN_CUDA(cudaSetDevice(s->device));
	if(!s->rank_masks)
	{
		const size_t cap_words = ((size_t)s->desc.max_nodes + 31) / 32;
		int rc = dev_alloc(s, &s->rank_masks, cap_words * RN_MAX_CAMERAS);
		if(rc) return rc;
	}
	s->have_masks = false;
	if(s->n == 0)
		return RN_OK;
	int rc = build_masks(s, s->rank_masks, mask_words(s), static_cast<cudaStream_t>(cuda_stream));
	if(rc) return rc;
	s->have_masks = true;
	s->mask_ncams = s->last_ncams;
	return RN_OK;
}

extern "C" int rn_cull_build_masks_to(rn_scene_t *s, uint32_t *device_out, uint32_t row_stride_words, void *cuda_stream)
{
	RN_REQUIRE(s && device_out, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_build_masks_to needs a cull");
	RN_REQUIRE(row_stride_words >= mask_words(s), RN_ERR_INVALID, "row stride %u < %u words per camera", row_stride_words, mask_words(s));
	RN_CUDA(cudaSetDevice(s->device));
	if(s->n == 0)
		return RN_OK;
	return build_masks(s, device_out, row_stride_words, static_cast<cudaStream_t>(cuda_stream));
}

extern "C" int rn_cull_build_masks_multicast(rn_scene_t *s, uint32_t *multicast_out, uint32_t row_stride_words, void *cuda_stream)
{
	RN_REQUIRE(s && multicast_out, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_build_masks_multicast needs a cull");
	RN_REQUIRE(row_stride_words >= mask_words(s), RN_ERR_INVALID, "row stride %u < %u words per camera", row_stride_words, mask_words(s));
	RN_CUDA(cudaSetDevice(s->device));
	if(s->n == 0)
		return RN_OK;
	return build_masks(s, multicast_out, row_stride_words, static_cast<cudaStream_t>(cuda_stream), true);
}

namespace rn {
int build_masks_fanout(rn_scene *s, uint32_t *const *dsts, uint32_t n_dst, uint32_t stride, cudaStream_t stream)
{
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "the mask fan-out needs a cull");
	RN_REQUIRE(n_dst >= 1 && n_dst <= RN_MAX_GROUP, RN_ERR_INVALID, "1..%d destinations", RN_MAX_GROUP);
	if(s->n == 0)
		return RN_OK;
	FanoutDst d;
	memset(&d, 0, sizeof(d));
	d.n = n_dst;
	for(uint32_t k = 0; k < n_dst; k++) d.p[k] = dsts[k];
	const uint32_t words = mask_words(s);
	const uint32_t grid = (words + kMaskThreads - 1) / kMaskThreads;
	switch(s->last_ncams)
	{
#define RN_FAN_CASE(V) case V: rank_mask_fanout_kernel<V><<<grid, kMaskThreads, 0, stream>>>(s->vis, words, stride, d); break;
		RN_FAN_CASE(1) RN_FAN_CASE(2) RN_FAN_CASE(3) RN_FAN_CASE(4) RN_FAN_CASE(5) RN_FAN_CASE(6) RN_FAN_CASE(7)
		default: rank_mask_fanout_kernel<8><<<grid, kMaskThreads, 0, stream>>>(s->vis, words, stride, d); break;
#undef RN_FAN_CASE
	}
	RN_CUDA(cudaGetLastError());
	s->launches++;
	return RN_OK;
}
}

extern "C" int rn_cull_device_masks(rn_scene_t *s, const uint32_t **masks, uint32_t *words_per_camera, uint32_t *n_cameras)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	if(masks) *masks = s->have_masks ? s->rank_masks : nullptr;
	if(words_per_camera) *words_per_camera = mask_words(s);
	if(n_cameras) *n_cameras = s->have_masks ? s->mask_ncams : 0;
	return RN_OK;
}

extern "C" int rn_cull_read_masks_async(rn_scene_t *s, uint32_t n_cameras, uint32_t *out_words, void *cuda_stream)
{
	RN_REQUIRE(s && out_words, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->have_masks, RN_ERR_STATE, "no masks: run rn_cull_build_masks after the cull");
	RN_REQUIRE(n_cameras <= s->mask_ncams, RN_ERR_INVALID, "%u cameras asked, the masks hold %u", n_cameras, s->mask_ncams);
	RN_CUDA(cudaSetDevice(s->device));
	// the camera rows are contiguous on both sides: one copy
	RN_CUDA(cudaMemcpyAsync(out_words, s->rank_masks, (size_t)n_cameras * mask_words(s) * sizeof(uint32_t), cudaMemcpyDeviceToHost,
	                        static_cast<cudaStream_t>(cuda_stream)));
	return RN_OK;
}

extern "C" int rn_masks_expand(rn_scene_t *s, const rn_mask_segment *segments, uint32_t n_segments, uint32_t *device_list,
                               uint32_t capacity, uint32_t *device_count, void *cuda_stream)
{
	RN_RANGE("rn_masks_expand");
	RN_REQUIRE(s && (segments || n_segments == 0) && device_list && device_count, RN_ERR_INVALID, "NULL argument");
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	// at most 32 segments per launch set; longer inputs continue on the running device-side count
	for(uint32_t s0 = 0; s0 < n_segments; s0 += kExpandMaxSegments)
	{
		ExpandArgs a;
		memset(&a, 0, sizeof(a));
		uint32_t n_blocks = 0;
		for(uint32_t k = s0; k < std::min(n_segments, s0 + kExpandMaxSegments); k++)
		{
			if(segments[k].n_words == 0)
				continue;
			RN_REQUIRE(segments[k].device_words, RN_ERR_INVALID, "segment %u: NULL words", k);
			const uint32_t q = a.n_segments++;
			a.words[q] = segments[k].device_words;
			a.n_words[q] = segments[k].n_words;
			a.rank_base[q] = segments[k].rank_base;
			a.first_block[q] = n_blocks;
			n_blocks += (segments[k].n_words + kExpandWordsPerBlock - 1) / kExpandWordsPerBlock;
			a.first_block[q + 1] = n_blocks;
		}
		if(n_blocks == 0)
			continue;
		if(n_blocks > s->expand_blocks_cap)
		{
			// grows only; cudaFree synchronises, so no kernel of an earlier call still reads the old buffer
			if(s->expand_totals)
			{
				RN_CUDA(cudaFree(s->expand_totals));
				s->device_bytes -= (uint64_t)s->expand_blocks_cap * sizeof(uint32_t);
				s->expand_totals = nullptr;
				s->expand_blocks_cap = 0;
			}
			const uint32_t cap = std::max(n_blocks, 4096u);
			int rc = dev_alloc(s, &s->expand_totals, cap);
			if(rc) return rc;
			s->expand_blocks_cap = cap;
		}
		expand_count_kernel<<<n_blocks, kExpandThreads, 0, stream>>>(a, s->expand_totals);
		expand_scan_kernel<<<1, kExpandThreads, 0, stream>>>(s->expand_totals, n_blocks, device_count);
		expand_emit_kernel<<<n_blocks, kExpandThreads, 0, stream>>>(a, s->expand_totals, device_list, capacity);
		RN_CUDA(cudaGetLastError());
		s->launches += 3;
	}
	return RN_OK;
}

// ---- rn_masks_expand without any allocation ---------------------------------------------------------------------------
// The per-CTA totals of the three expand kernels live in a file-scope device array instead of the scene's cudaMalloc'd
// scratch.  Every GPU that loads librn_cuda holds these 64 KiB until the process exits.  The kernels are the ones above,
// unchanged: they take the scratch as a pointer.
namespace rn {
constexpr uint32_t kExpandStaticBlocks = RN_MASKS_EXPAND_STATIC_MAX_WORDS / kExpandWordsPerBlock;   // 16 384 CTAs
static_assert(RN_MASKS_EXPAND_STATIC_MAX_WORDS % kExpandWordsPerBlock == 0, "the limit is a whole number of CTAs");
static __device__ __align__(256) uint32_t g_expand_static_totals[kExpandStaticBlocks];

constexpr int kStaticMaxDevices = 64;
static std::mutex g_expand_static_lock;
static uint32_t *g_expand_static_ptr[kStaticMaxDevices];   // device address of the array above, per device; nullptr = not asked yet

// In host code `g_expand_static_totals` is only a stand-in for the symbol: the address a kernel may use comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int expand_static_scratch(int device, uint32_t **out)
{
	std::lock_guard<std::mutex> hold(g_expand_static_lock);
	if(!g_expand_static_ptr[device])
	{
		uint32_t *p = nullptr;
		RN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_expand_static_totals));
		RN_REQUIRE(p, RN_ERR_CUDA, "cudaGetSymbolAddress gave no address for the static expand scratch on device %d", device);
		g_expand_static_ptr[device] = p;
	}
	*out = g_expand_static_ptr[device];
	return RN_OK;
}
}

extern "C" int rn_masks_expand_static(int device, const rn_mask_segment *segments, uint32_t n_segments, uint32_t *device_list,
                                      uint32_t capacity, uint32_t *device_count, void *cuda_stream)
{
	RN_RANGE("rn_masks_expand_static");
	RN_REQUIRE((segments || n_segments == 0) && device_list && device_count, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_segments <= kExpandMaxSegments, RN_ERR_INVALID, "rn_masks_expand_static takes at most %u segments per call, got %u",
	           kExpandMaxSegments, n_segments);
	RN_REQUIRE(device < kStaticMaxDevices, RN_ERR_INVALID, "device %d: rn_masks_expand_static knows devices 0..%d", device, kStaticMaxDevices - 1);
	ExpandArgs a;
	memset(&a, 0, sizeof(a));
	uint32_t n_blocks = 0;
	for(uint32_t k = 0; k < n_segments; k++)
	{
		if(segments[k].n_words == 0)
			continue;
		RN_REQUIRE(segments[k].device_words, RN_ERR_INVALID, "segment %u: NULL words", k);
		const uint32_t q = a.n_segments++;
		a.words[q] = segments[k].device_words;
		a.n_words[q] = segments[k].n_words;
		a.rank_base[q] = segments[k].rank_base;
		a.first_block[q] = n_blocks;
		n_blocks += (uint32_t)(((uint64_t)segments[k].n_words + kExpandWordsPerBlock - 1) / kExpandWordsPerBlock);
		a.first_block[q + 1] = n_blocks;
		RN_REQUIRE(n_blocks <= kExpandStaticBlocks, RN_ERR_CAPACITY,
		           "rn_masks_expand_static: the segments need more than %u CTAs of 1024 words (limit RN_MASKS_EXPAND_STATIC_MAX_WORDS = %u mask words "
		           "= %llu ranks, every segment rounded up to 1024 words); use rn_masks_expand",
		           kExpandStaticBlocks, (unsigned)RN_MASKS_EXPAND_STATIC_MAX_WORDS, (unsigned long long)RN_MASKS_EXPAND_STATIC_MAX_WORDS * 32ull);
	}
	if(n_blocks == 0)
		return RN_OK;
	// the arguments are checked; from here on a device is needed
	if(device < 0)
		RN_CUDA(cudaGetDevice(&device));
	RN_REQUIRE(device < kStaticMaxDevices, RN_ERR_INVALID, "device %d: rn_masks_expand_static knows devices 0..%d", device, kStaticMaxDevices - 1);
	RN_CUDA(cudaSetDevice(device));
	uint32_t *totals = nullptr;
	int rc = expand_static_scratch(device, &totals);
	if(rc) return rc;
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	// nothing in the array has to start at zero: pass 1 writes entries [0, n_blocks), the only ones passes 2 and 3 read
	expand_count_kernel<<<n_blocks, kExpandThreads, 0, stream>>>(a, totals);
	expand_scan_kernel<<<1, kExpandThreads, 0, stream>>>(totals, n_blocks, device_count);
	expand_emit_kernel<<<n_blocks, kExpandThreads, 0, stream>>>(a, totals, device_list, capacity);
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}
