// librn_cuda - the pipelined frame: one GPU, the render list cut into rank ranges, results in host memory.
//
// What a C++ host (the Rayne shim) calls per frame when the visibility has to reach the CPU: the reference's
// SceneBasic::Render hands `sceneNodesToRender` to host code (Source/Scene/RNSceneBasic.cpp:576-640), so the end-to-end
// frame is  dirty rows -> device, transform + cull, visibility -> host,  and the last leg is PCIe-bound.  The render list is
// cut into a few rank ranges where no parent / child edge crosses (rn_plan_ranges; the cuts of SURVEY.md §8(e)); every
// range is an rn_scene_t of its own on its own prioritised stream, so the copy-out of range k runs under the kernels of
// range k + 1, and a ring of result buffers keeps frame i's copy-out under frame i + 1's kernels.  Nothing in the frame
// makes the host wait: the copy-out is a KERNEL (pack_kernel) that stores into mapped pinned host memory, which lets the
// device pick the form of every (range, camera) segment from the count the compaction just produced:
//   RN_VIS_LISTS     the list itself, 4 bytes per visible node (global render-list ranks, ascending)
//   RN_VIS_MASKS     one bit per rank of the range (bit b of word w = rank first + 32 w + b)
//   RN_VIS_ADAPTIVE  whichever of the two is smaller for that segment (a list wins below 1/32 visible)
// Both carry exactly the list of the reference's cull loop (ascending ranks = render-list order); rn_frame_list decodes
// either back into a flat list on the host.
#include "rn_internal.h"

#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

namespace rn {

struct PackView
{
	const uint32_t *lists[RN_MAX_CAMERAS];
	const uint32_t *list_counts;
	const uint32_t *masks;          // [camera][words], may be nullptr when form == RN_VIS_LISTS
	uint32_t words;                 // mask words per camera of this range
	uint32_t form;
	uint32_t n_cams;
	uint32_t region_words;          // capacity of one camera's region in the host buffer
	uint32_t *host_data;            // mapped pinned memory: [camera][region_words]
	uint32_t *host_header;          // mapped pinned memory: [camera][2] = form taken, visible count
};

constexpr uint32_t kPackThreads = 256;

// grid = (blocks, cameras).  Streaming stores of consecutive words by consecutive threads: full PCIe write combining.
__global__ void __launch_bounds__(kPackThreads) pack_kernel(const __grid_constant__ PackView v)
{
	const uint32_t c = blockIdx.y;
	const uint32_t count = v.list_counts[c];
	const bool as_list = v.form == RN_VIS_LISTS || (v.form == RN_VIS_ADAPTIVE && count <= v.words);
	const uint32_t total = min(as_list ? count : v.words, v.region_words);
	const uint32_t *__restrict__ src = as_list ? v.lists[c] : v.masks + (size_t)c * v.words;
	uint32_t *__restrict__ dst = v.host_data + (size_t)c * v.region_words;
	if((reinterpret_cast<uintptr_t>(src) & 15u) == 0)
	{
		// 16 bytes per thread, 512 contiguous bytes per warp and instruction: whole PCIe write bursts
		const uint32_t quads = total / 4u;
		const uint4 *__restrict__ s4 = reinterpret_cast<const uint4 *>(src);
		uint4 *__restrict__ d4 = reinterpret_cast<uint4 *>(dst);
		for(uint32_t i = blockIdx.x * kPackThreads + threadIdx.x; i < quads; i += gridDim.x * kPackThreads)
			__stcs(d4 + i, __ldg(s4 + i));
		if(blockIdx.x == 0 && threadIdx.x < total - quads * 4u)
			__stcs(dst + quads * 4u + threadIdx.x, __ldg(src + quads * 4u + threadIdx.x));
	}
	else
	{
		for(uint32_t i = blockIdx.x * kPackThreads + threadIdx.x; i < total; i += gridDim.x * kPackThreads)
			__stcs(dst + i, __ldg(src + i));
	}
	if(blockIdx.x == 0 && threadIdx.x == 0)
	{
		v.host_header[c * 2] = as_list ? (uint32_t)RN_VIS_LISTS : (uint32_t)RN_VIS_MASKS;
		v.host_header[c * 2 + 1] = count;
	}
}

struct PipeRange
{
	rn_scene *scene = nullptr;
	cudaStream_t stream = nullptr;
	uint32_t first = 0, count = 0;
};

struct PipeSlot
{
	uint32_t *data = nullptr;       // mapped pinned: per range, per camera a region
	uint32_t *header = nullptr;     // mapped pinned: [range][RN_MAX_CAMERAS][2]
	size_t data_words = 0;
	std::vector<size_t> range_ofs;  // word offset of range k's block in `data`
	std::vector<uint32_t> region;   // region words per camera of range k
	std::vector<cudaEvent_t> done;  // per range
	std::vector<rn_frame_segment> segments;
	uint32_t n_cams = 0;
	uint32_t form = 0;
	uint64_t frame = 0;
	bool busy = false;              // submitted, not yet released
	bool engine_copy = false;       // single-range mode: masks + raw counts came through the copy engine (header = 8 counts)
};

} // namespace rn

struct rn_pipeline
{
	rn_pipeline_desc desc{};
	int device = 0;
	uint32_t n = 0;
	std::vector<rn::PipeRange> ranges;
	std::vector<uint32_t> bounds;       // n_ranges + 1
	std::vector<rn::PipeSlot> slots;
	uint64_t submitted = 0, acquired = 0, released = 0;
	uint64_t h2d_bytes = 0;             // dirty rows staged since the last submit
	uint64_t pending_h2d = 0;
	// single-range mode: the copy-out runs on its own stream beside the next frame's kernels, the scene flips between two
	// sets of cull outputs
	cudaStream_t copy_stream = nullptr;
	cudaEvent_t cull_done = nullptr;
	cudaEvent_t pack_done[2] = {nullptr, nullptr};
	uint32_t out_set = 0;
};

using namespace rn;

// ---- rank ranges ----------------------------------------------------------------------------------------------------
extern "C" int rn_plan_ranges(const uint32_t *parent, uint32_t n, uint32_t n_ranges, uint32_t *bounds)
{
	RN_REQUIRE(parent && bounds, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_ranges >= 1, RN_ERR_INVALID, "n_ranges must be >= 1");
	// crossing[r] = number of parent / child edges that cross the boundary between ranks r - 1 and r
	std::vector<int32_t> cover((size_t)n + 2, 0);
	for(uint32_t i = 0; i < n; i++)
	{
		const uint32_t p = parent[i];
		if(p == RN_NO_PARENT) continue;
		RN_REQUIRE(p < n, RN_ERR_INVALID, "node %u has an invalid parent %u", i, p);
		const uint32_t lo = std::min(i, p), hi = std::max(i, p);
		cover[lo + 1] += 1;
		cover[hi + 1] -= 1;
	}
	std::vector<uint32_t> ok;
	int32_t run = 0;
	for(uint32_t r = 0; r <= n; r++)
	{
		run += cover[r];
		if(run == 0) ok.push_back(r);
	}
	RN_REQUIRE(ok.size() >= (size_t)n_ranges + 1, RN_ERR_INVALID, "the hierarchy cannot be split into %u edge-free rank ranges (only %zu independent pieces)",
	           n_ranges, ok.size() - 1);
	std::vector<size_t> idx(1, 0);
	for(uint32_t k = 1; k < n_ranges; k++)
	{
		const uint64_t target = (uint64_t)n * k / n_ranges;
		const size_t lo_i = idx.back() + 1, hi_i = ok.size() - 1 - (n_ranges - k);
		const size_t i = std::lower_bound(ok.begin(), ok.end(), (uint32_t)target) - ok.begin();
		// the admissible cut closest to the target (the one below it on a tie)
		auto dist = [&](size_t j) { return ok[j] >= target ? (uint64_t)ok[j] - target : target - (uint64_t)ok[j]; };
		size_t best = std::min(std::max(i, lo_i), hi_i);
		bool have = false;
		if(i >= 1 && i - 1 >= lo_i && i - 1 <= hi_i) { best = i - 1; have = true; }
		if(i >= lo_i && i <= hi_i && (!have || dist(i) < dist(best))) best = i;
		idx.push_back(best);
	}
	for(uint32_t k = 0; k < n_ranges; k++) bounds[k] = ok[idx[k]];
	bounds[n_ranges] = n;
	return RN_OK;
}

static void pipeline_free(rn_pipeline *p)
{
	cudaSetDevice(p->device);
	cudaDeviceSynchronize();
	for(PipeSlot &s : p->slots)
	{
		if(s.data) cudaFreeHost(s.data);
		if(s.header) cudaFreeHost(s.header);
		for(cudaEvent_t e : s.done) if(e) cudaEventDestroy(e);
	}
	for(PipeRange &r : p->ranges)
	{
		if(r.scene) rn_scene_destroy(r.scene);
		if(r.stream) cudaStreamDestroy(r.stream);
	}
	if(p->copy_stream) cudaStreamDestroy(p->copy_stream);
	if(p->cull_done) cudaEventDestroy(p->cull_done);
	for(cudaEvent_t e : p->pack_done) if(e) cudaEventDestroy(e);
	delete p;
}

extern "C" int rn_pipeline_create(rn_pipeline_t **out, const rn_pipeline_desc *desc, const rn_node_columns *cols, uint32_t count)
{
	RN_REQUIRE(out && desc && cols, RN_ERR_INVALID, "NULL argument");
	*out = nullptr;
	RN_REQUIRE(count > 0, RN_ERR_INVALID, "no nodes");
	RN_REQUIRE(cols->position && cols->scale && cols->rotation && cols->parent && cols->flags && cols->render_group && cols->bounding_sphere,
	           RN_ERR_INVALID, "rn_pipeline_create needs every node column");
	const uint32_t n_ranges = desc->n_ranges ? desc->n_ranges : 1;
	const uint32_t ring = desc->frames_in_flight ? desc->frames_in_flight : 2;
	RN_REQUIRE(n_ranges <= 64 && ring <= 8, RN_ERR_INVALID, "n_ranges <= 64 and frames_in_flight <= 8");
	rn_pipeline *p = new(std::nothrow) rn_pipeline();
	RN_REQUIRE(p, RN_ERR_NOMEM, "out of host memory");
	p->desc = *desc;
	p->desc.n_ranges = n_ranges;
	p->desc.frames_in_flight = ring;
	p->device = desc->device;
	p->n = count;
	p->bounds.resize(n_ranges + 1);
	int rc = rn_plan_ranges(cols->parent, count, n_ranges, p->bounds.data());
	if(rc) { delete p; return rc; }
	p->ranges.resize(n_ranges);
	for(uint32_t k = 0; k < n_ranges && !rc; k++)
	{
		PipeRange &r = p->ranges[k];
		r.first = p->bounds[k];
		r.count = p->bounds[k + 1] - p->bounds[k];
		rn_scene_desc sd;
		memset(&sd, 0, sizeof(sd));
		sd.device = desc->device;
		sd.max_nodes = std::max(r.count, 1u);
		sd.tile_nodes = desc->tile_nodes; sd.tile_inner_nodes = desc->tile_inner_nodes; sd.tile_unit_nodes = desc->tile_unit_nodes;
		sd.flags = desc->flags;
		rc = rn_scene_create(&r.scene, &sd);
		if(rc) break;
		// earlier ranges get more urgent streams: otherwise range 0's compaction and copy-out queue behind the already
		// launched transform CTAs of the later ranges and the first copy starts when nearly all kernels are done
		void *st = nullptr;
		rc = rn_stream_create(&st, (int)(n_ranges - 1 - k));
		if(rc) break;
		r.stream = static_cast<cudaStream_t>(st);
		if(r.count == 0) continue;
		std::vector<uint32_t> local(r.count);
		for(uint32_t i = 0; i < r.count; i++)
		{
			const uint32_t q = cols->parent[r.first + i];
			local[i] = q == RN_NO_PARENT ? RN_NO_PARENT : q - r.first;
		}
		rn_node_columns c;
		memset(&c, 0, sizeof(c));
		c.position = cols->position + (size_t)r.first * 3; c.scale = cols->scale + (size_t)r.first * 3;
		c.rotation = cols->rotation + (size_t)r.first * 4; c.parent = local.data();
		c.flags = cols->flags + r.first; c.render_group = cols->render_group + r.first;
		c.bounding_sphere = cols->bounding_sphere + (size_t)r.first * 4;
		if(cols->bounding_box) c.bounding_box = cols->bounding_box + (size_t)r.first * 6;
		if(cols->render_priority) c.render_priority = cols->render_priority + r.first;
		rc = rn_scene_upload(r.scene, &c, 0, r.count);
		if(!rc) rc = rn_scene_set_list_mode(r.scene, 1, desc->rank_base + r.first);
	}
	if(rc) { pipeline_free(p); return rc; }
	if(n_ranges == 1 && ring >= 2)
	{
		void *st = nullptr;
		rc = rn_stream_create(&st, 1);
		cudaError_t e = cudaSuccess;
		if(!rc)
		{
			p->copy_stream = static_cast<cudaStream_t>(st);
			e = cudaEventCreateWithFlags(&p->cull_done, cudaEventDisableTiming);
			if(e == cudaSuccess) e = cudaEventCreateWithFlags(&p->pack_done[0], cudaEventDisableTiming);
			if(e == cudaSuccess) e = cudaEventCreateWithFlags(&p->pack_done[1], cudaEventDisableTiming);
		}
		if(rc || e != cudaSuccess) { pipeline_free(p); return rc ? rc : cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__); }
	}
	p->slots.resize(ring);
	for(PipeSlot &s : p->slots)
	{
		s.done.assign(n_ranges, nullptr);
		for(uint32_t k = 0; k < n_ranges; k++)
		{
			cudaError_t e = cudaEventCreateWithFlags(&s.done[k], cudaEventDisableTiming);
			if(e != cudaSuccess) { pipeline_free(p); return cuda_fail(e, "cudaEventCreate", __FILE__, __LINE__); }
		}
	}
	*out = p;
	return RN_OK;
}

extern "C" void rn_pipeline_destroy(rn_pipeline_t *p)
{
	if(p) pipeline_free(p);
}

extern "C" int rn_pipeline_ranges(const rn_pipeline_t *p, uint32_t *bounds, uint32_t capacity, uint32_t *n_ranges)
{
	RN_REQUIRE(p && n_ranges, RN_ERR_INVALID, "NULL argument");
	*n_ranges = p->desc.n_ranges;
	if(bounds)
	{
		RN_REQUIRE(capacity >= p->bounds.size(), RN_ERR_CAPACITY, "bounds holds %u entries, %zu needed", capacity, p->bounds.size());
		memcpy(bounds, p->bounds.data(), sizeof(uint32_t) * p->bounds.size());
	}
	return RN_OK;
}

// SetPosition & co on a dirty set of ranks (0-based inside this pipeline's node columns): routed to the ranges, staged
// through each range's pinned buffer and enqueued on its stream - they precede the next rn_pipeline_submit.
extern "C" int rn_pipeline_update_nodes(rn_pipeline_t *p, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count)
{
	RN_REQUIRE(p && cols && (ranks || count == 0), RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(cols->parent == nullptr && cols->bounding_box == nullptr && cols->render_priority == nullptr, RN_ERR_INVALID,
	           "hierarchy, boxes and render priorities change through a new pipeline");
	const uint32_t nr = p->desc.n_ranges;
	std::vector<std::vector<uint32_t>> sel(nr);
	for(uint32_t i = 0; i < count; i++)
	{
		RN_REQUIRE(ranks[i] < p->n, RN_ERR_INVALID, "rank %u out of range (%u nodes)", ranks[i], p->n);
		const uint32_t k = (uint32_t)(std::upper_bound(p->bounds.begin(), p->bounds.end(), ranks[i]) - p->bounds.begin()) - 1;
		sel[k].push_back(i);
	}
	std::vector<float> pos, scale, rot, sphere;
	std::vector<uint32_t> flags, group, local;
	for(uint32_t k = 0; k < nr; k++)
	{
		const size_t m = sel[k].size();
		if(!m) continue;
		local.resize(m);
		auto gather = [&](const float *src, std::vector<float> &dst, int w) -> const float * {
			if(!src) return nullptr;
			dst.resize(m * w);
			for(size_t j = 0; j < m; j++) memcpy(&dst[j * w], src + (size_t)sel[k][j] * w, sizeof(float) * w);
			return dst.data();
		};
		auto gather_u = [&](const uint32_t *src, std::vector<uint32_t> &dst) -> const uint32_t * {
			if(!src) return nullptr;
			dst.resize(m);
			for(size_t j = 0; j < m; j++) dst[j] = src[sel[k][j]];
			return dst.data();
		};
		for(size_t j = 0; j < m; j++) local[j] = ranks[sel[k][j]] - p->ranges[k].first;
		rn_node_columns c;
		memset(&c, 0, sizeof(c));
		c.position = gather(cols->position, pos, 3); c.scale = gather(cols->scale, scale, 3); c.rotation = gather(cols->rotation, rot, 4);
		c.bounding_sphere = gather(cols->bounding_sphere, sphere, 4);
		c.flags = gather_u(cols->flags, flags); c.render_group = gather_u(cols->render_group, group);
		int rc = rn_scene_update_nodes_async(p->ranges[k].scene, &c, local.data(), (uint32_t)m, p->ranges[k].stream);
		if(rc) return rc;
		const size_t per = 1 + (c.position ? 3 : 0) + (c.scale ? 3 : 0) + (c.rotation ? 4 : 0) + (c.bounding_sphere ? 4 : 0) + (c.flags ? 1 : 0) + (c.render_group ? 1 : 0);
		p->pending_h2d += per * 4 * m;
	}
	return RN_OK;
}

static int slot_prepare(rn_pipeline *p, PipeSlot &s, uint32_t n_cams, uint32_t form)
{
	const uint32_t nr = p->desc.n_ranges;
	std::vector<uint32_t> region(nr);
	std::vector<size_t> ofs(nr);
	size_t total = 0;
	for(uint32_t k = 0; k < nr; k++)
	{
		const uint32_t words = (p->ranges[k].count + 31u) / 32u;
		// LISTS: a camera may see the whole range; MASKS / ADAPTIVE: never more than the mask
		const uint32_t need = form == RN_VIS_LISTS ? p->ranges[k].count : words;
		region[k] = (need + 3u) & ~3u;      // keep every region 16-byte aligned
		ofs[k] = total;
		total += (size_t)region[k] * n_cams;
	}
	if(total > s.data_words)
	{
		if(s.data) RN_CUDA(cudaFreeHost(s.data));
		s.data = nullptr;
		RN_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&s.data), std::max<size_t>(total, 4) * sizeof(uint32_t), cudaHostAllocMapped));
		s.data_words = total;
	}
	if(!s.header)
		RN_CUDA(cudaHostAlloc(reinterpret_cast<void **>(&s.header), (size_t)nr * RN_MAX_CAMERAS * 2 * sizeof(uint32_t), cudaHostAllocMapped));
	s.region.swap(region);
	s.range_ofs.swap(ofs);
	s.n_cams = n_cams;
	s.form = form;
	s.segments.assign((size_t)nr * n_cams, rn_frame_segment{});
	return RN_OK;
}

extern "C" int rn_pipeline_submit(rn_pipeline_t *p, const rn_camera_frustum *cameras, uint32_t n_cameras, rn_visibility_form form)
{
	RN_RANGE("rn_pipeline_submit");
	RN_REQUIRE(p && cameras, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_cameras >= 1 && n_cameras <= RN_MAX_CAMERAS, RN_ERR_INVALID, "n_cameras must be 1..%d", RN_MAX_CAMERAS);
	RN_REQUIRE(form == RN_VIS_LISTS || form == RN_VIS_MASKS || form == RN_VIS_ADAPTIVE, RN_ERR_INVALID, "unknown visibility form %d", (int)form);
	const uint32_t ring = p->desc.frames_in_flight;
	RN_REQUIRE(p->submitted - p->released < ring, RN_ERR_STATE, "%u frames in flight already: rn_pipeline_acquire / rn_pipeline_release the oldest first", ring);
	RN_CUDA(cudaSetDevice(p->device));
	PipeSlot &s = p->slots[p->submitted % ring];
	int rc = slot_prepare(p, s, n_cameras, (uint32_t)form);
	if(rc) return rc;
	const uint32_t nr = p->desc.n_ranges;
	const bool overlap = p->copy_stream != nullptr;      // single range: copy-out on its own stream, two output sets
	for(uint32_t k = 0; k < nr; k++)
	{
		PipeRange &r = p->ranges[k];
		cudaStream_t pack_stream = r.stream;
		if(r.count)
		{
			if(overlap && p->submitted >= 2)     // this output set was last read by the copy-out of the frame before the previous one
				RN_CUDA(cudaStreamWaitEvent(r.stream, p->pack_done[p->out_set], 0));
			rc = rn_transform_cull(r.scene, cameras, n_cameras, nullptr, r.stream);
			if(!rc && form != RN_VIS_LISTS) rc = rn_cull_build_masks(r.scene, r.stream);
			if(rc) return rc;
			PackView v;
			memset(&v, 0, sizeof(v));
			for(uint32_t c = 0; c < n_cameras; c++) v.lists[c] = r.scene->lists[c];
			v.list_counts = r.scene->list_counts;
			v.masks = form != RN_VIS_LISTS ? r.scene->rank_masks : nullptr;
			v.words = (r.count + 31u) / 32u;
			v.form = (uint32_t)form;
			v.n_cams = n_cameras;
			v.region_words = s.region[k];
			v.host_data = s.data + s.range_ofs[k];
			v.host_header = s.header + (size_t)k * RN_MAX_CAMERAS * 2;
			const uint32_t most = form == RN_VIS_LISTS ? r.count : v.words;
			const uint32_t blocks = std::max(1u, std::min((most + kPackThreads * 16 - 1) / (kPackThreads * 16), 96u));
			s.engine_copy = false;
			if(overlap)
			{
				RN_CUDA(cudaEventRecord(p->cull_done, r.stream));
				RN_CUDA(cudaStreamWaitEvent(p->copy_stream, p->cull_done, 0));
				pack_stream = p->copy_stream;
			}
			if(overlap && form != RN_VIS_LISTS)
			{
				// The persistent tile kernel of the next frame fills every SM, so a copy KERNEL would not get a CTA slot
				// until it retires; the masks have a fixed size, so the copy ENGINE moves them (and the counts) instead -
				// it runs truly beside the next frame.  (The adaptive encoding therefore means masks in this mode.)
				RN_CUDA(cudaMemcpy2DAsync(v.host_data, (size_t)v.region_words * 4, v.masks, (size_t)v.words * 4, (size_t)v.words * 4, n_cameras,
				                          cudaMemcpyDeviceToHost, p->copy_stream));
				RN_CUDA(cudaMemcpyAsync(v.host_header, v.list_counts, sizeof(uint32_t) * RN_MAX_CAMERAS, cudaMemcpyDeviceToHost, p->copy_stream));
				s.engine_copy = true;
			}
			else
			{
				pack_kernel<<<dim3(blocks, n_cameras), kPackThreads, 0, pack_stream>>>(v);
				RN_CUDA(cudaGetLastError());
				r.scene->launches++;
			}
			if(overlap)
			{
				RN_CUDA(cudaEventRecord(p->pack_done[p->out_set], p->copy_stream));
				rc = scene_flip_outputs(r.scene);          // the next frame writes the other set while this one is copied out
				if(rc) return rc;
				p->out_set ^= 1u;
			}
		}
		RN_CUDA(cudaEventRecord(s.done[k], pack_stream));
	}
	s.frame = p->submitted;
	s.busy = true;
	p->submitted++;
	p->h2d_bytes = p->pending_h2d;
	p->pending_h2d = 0;
	return RN_OK;
}

extern "C" int rn_pipeline_acquire(rn_pipeline_t *p, rn_frame_view *view)
{
	RN_RANGE("rn_pipeline_acquire");
	RN_REQUIRE(p && view, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(p->acquired < p->submitted, RN_ERR_STATE, "no submitted frame to acquire");
	RN_REQUIRE(p->acquired == p->released, RN_ERR_STATE, "release the acquired frame first");
	RN_CUDA(cudaSetDevice(p->device));
	PipeSlot &s = p->slots[p->acquired % p->desc.frames_in_flight];
	const uint32_t nr = p->desc.n_ranges;
	uint64_t d2h = 0;
	for(uint32_t k = 0; k < nr; k++)
	{
		RN_CUDA(cudaEventSynchronize(s.done[k]));
		for(uint32_t c = 0; c < s.n_cams; c++)
		{
			rn_frame_segment &g = s.segments[(size_t)k * s.n_cams + c];
			g.rank_first = p->desc.rank_base + p->ranges[k].first;
			g.rank_count = p->ranges[k].count;
			if(p->ranges[k].count == 0)
			{
				g.form = RN_VIS_LISTS; g.count = 0; g.data = nullptr;
				continue;
			}
			const uint32_t *h = s.header + ((size_t)k * RN_MAX_CAMERAS + c) * 2;
			g.form = s.engine_copy ? (uint32_t)RN_VIS_MASKS : h[0];
			g.count = s.engine_copy ? s.header[c] : h[1];
			g.data = s.data + s.range_ofs[k] + (size_t)c * s.region[k];
			d2h += 8 + 4ull * (g.form == RN_VIS_LISTS ? g.count : (g.rank_count + 31u) / 32u);
		}
	}
	view->frame = s.frame;
	view->n_cameras = s.n_cams;
	view->n_ranges = nr;
	view->segments = s.segments.data();
	view->d2h_bytes = d2h;
	view->h2d_bytes = p->h2d_bytes;
	p->acquired++;
	return RN_OK;
}

extern "C" int rn_pipeline_release(rn_pipeline_t *p)
{
	RN_REQUIRE(p, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(p->released < p->acquired, RN_ERR_STATE, "no acquired frame to release");
	p->slots[p->released % p->desc.frames_in_flight].busy = false;
	p->released++;
	return RN_OK;
}

// Host decode of one camera's visibility of an acquired frame into a flat, ascending list of global ranks - the
// `sceneNodesToRender` of the reference's cull loop.  List segments are copied, mask segments are walked bit by bit.
extern "C" int rn_frame_list(const rn_frame_view *view, uint32_t camera, uint32_t *out_ranks, uint32_t capacity, uint32_t *out_count)
{
	RN_REQUIRE(view && out_count, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(camera < view->n_cameras, RN_ERR_INVALID, "camera %u of %u", camera, view->n_cameras);
	uint64_t total = 0;
	for(uint32_t k = 0; k < view->n_ranges; k++)
		total += view->segments[(size_t)k * view->n_cameras + camera].count;
	*out_count = (uint32_t)total;
	if(!out_ranks)
		return RN_OK;
	RN_REQUIRE(total <= capacity, RN_ERR_CAPACITY, "the list has %llu entries, the buffer holds %u", (unsigned long long)total, capacity);
	uint32_t at = 0;
	for(uint32_t k = 0; k < view->n_ranges; k++)
	{
		const rn_frame_segment &g = view->segments[(size_t)k * view->n_cameras + camera];
		if(g.count == 0) continue;
		if(g.form == RN_VIS_LISTS)
		{
			memcpy(out_ranks + at, g.data, sizeof(uint32_t) * g.count);
			at += g.count;
		}
		else
		{
			const uint32_t words = (g.rank_count + 31u) / 32u;
			for(uint32_t w = 0; w < words; w++)
			{
				uint32_t bits = g.data[w];
				while(bits)
				{
					out_ranks[at++] = g.rank_first + w * 32u + (uint32_t)__builtin_ctz(bits);
					bits &= bits - 1u;
				}
			}
		}
	}
	return RN_OK;
}

extern "C" int rn_pipeline_scene(rn_pipeline_t *p, uint32_t range, rn_scene_t **scene, void **cuda_stream)
{
	RN_REQUIRE(p && range < p->desc.n_ranges, RN_ERR_INVALID, "bad range");
	if(scene) *scene = p->ranges[range].scene;
	if(cuda_stream) *cuda_stream = p->ranges[range].stream;
	return RN_OK;
}
