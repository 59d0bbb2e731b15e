// librn_cuda - several GPUs driven by ONE host process (Rayne is a single-process C++ engine).
//
// the render list is cut into one rank range per GPU where no parent / child edge crosses
// (rn_plan_ranges), every GPU propagates and culls its range against all cameras, and the path's only exchange is the
// visibility: every GPU ends up with the rank-order bitmasks of every range (12.6 MB per GPU and frame for 16.8 M nodes x 6
// views, instead of 4 bytes per visible node).  The exchange is fused into the kernel that produces the masks:
//   RN_GROUP_MULTICAST  the gathered buffer of every GPU is backed by one CUDA multicast object (NVLS); the mask kernel
//                       stores each word once with multimem.st and the NVSwitch replicates it into all copies
//   RN_GROUP_PEER       peer-mapped pointers (cudaDeviceEnablePeerAccess): the mask kernel stores each word into every
//                       GPU's buffer over NVLink
// No collective library is involved; ordering between the GPUs is done with CUDA events across devices (the host only
// enqueues, it never waits): a frame's masks are complete on GPU j's stream once that stream has waited for the mask
// kernels of all GPUs, and the next frame's stores wait for the previous frame's consumers the same way.
#include "rn_internal.h"

#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

namespace rn {

struct GroupMember
{
	int device = 0;
	rn_scene *scene = nullptr;
	cudaStream_t stream = nullptr;
	cudaEvent_t masks_done = nullptr;    // this GPU's mask kernel of the current frame has finished
	cudaEvent_t frame_end = nullptr;     // everything enqueued on this GPU's stream up to the next frame's start
	uint32_t first = 0, count = 0;
	uint32_t word_ofs = 0, words = 0;    // this GPU's block inside a camera row of the gathered buffer
	uint32_t *gathered = nullptr;        // this GPU's copy: [RN_MAX_CAMERAS][row_words]
	uint32_t *list_scratch = nullptr;    // rn_group_expand output when the caller passes none
	// multicast transport
	CUmemGenericAllocationHandle mem = 0;
	CUdeviceptr uc_va = 0;
};

// driver entry points, resolved through the runtime (librn_cuda does not link libcuda)
struct DriverApi
{
	CUresult (*MulticastCreate)(CUmemGenericAllocationHandle *, const CUmulticastObjectProp *) = nullptr;
	CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
	CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long) = nullptr;
	CUresult (*MulticastGetGranularity)(size_t *, const CUmulticastObjectProp *, CUmulticastGranularity_flags) = nullptr;
	CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
	CUresult (*MemCreate)(CUmemGenericAllocationHandle *, size_t, const CUmemAllocationProp *, unsigned long long) = nullptr;
	CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
	CUresult (*MemAddressReserve)(CUdeviceptr *, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
	CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
	CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
	CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
	CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc *, size_t) = nullptr;
	CUresult (*MemGetAllocationGranularity)(size_t *, const CUmemAllocationProp *, CUmemAllocationGranularity_flags) = nullptr;
	CUresult (*DeviceGet)(CUdevice *, int) = nullptr;
	CUresult (*DeviceGetAttribute)(int *, CUdevice_attribute, CUdevice) = nullptr;
	bool ok = false;
};

static bool load_driver(DriverApi &d)
{
	auto get = [](const char *name, void **fn) -> bool {
		cudaDriverEntryPointQueryResult q;
		*fn = nullptr;
		return cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess && *fn;
	};
	d.ok = get("cuMulticastCreate", (void **)&d.MulticastCreate) && get("cuMulticastAddDevice", (void **)&d.MulticastAddDevice) &&
	       get("cuMulticastBindMem", (void **)&d.MulticastBindMem) && get("cuMulticastGetGranularity", (void **)&d.MulticastGetGranularity) &&
	       get("cuMulticastUnbind", (void **)&d.MulticastUnbind) && get("cuMemCreate", (void **)&d.MemCreate) &&
	       get("cuMemRelease", (void **)&d.MemRelease) && get("cuMemAddressReserve", (void **)&d.MemAddressReserve) &&
	       get("cuMemAddressFree", (void **)&d.MemAddressFree) && get("cuMemMap", (void **)&d.MemMap) && get("cuMemUnmap", (void **)&d.MemUnmap) &&
	       get("cuMemSetAccess", (void **)&d.MemSetAccess) && get("cuMemGetAllocationGranularity", (void **)&d.MemGetAllocationGranularity) &&
	       get("cuDeviceGet", (void **)&d.DeviceGet) && get("cuDeviceGetAttribute", (void **)&d.DeviceGetAttribute);
	cudaGetLastError();
	return d.ok;
}

} // namespace rn

struct rn_group
{
	rn_group_desc desc{};
	uint32_t n = 0;                     // nodes
	uint32_t n_dev = 0;
	std::vector<rn::GroupMember> dev;
	std::vector<uint32_t> bounds;
	uint32_t row_words = 0;             // words of one camera row of the gathered buffer (all GPUs' blocks)
	uint32_t transport = 0;             // RN_GROUP_PEER / RN_GROUP_MULTICAST actually in use
	uint32_t last_ncams = 0;
	uint64_t frames = 0;
	rn::DriverApi drv;
	CUmemGenericAllocationHandle mc = 0;
	CUdeviceptr mc_va = 0;
	size_t mc_size = 0;
	char transport_note[256] = "";
};

using namespace rn;

static void group_free(rn_group *g)
{
	for(GroupMember &m : g->dev)
	{
		cudaSetDevice(m.device);
		cudaDeviceSynchronize();
	}
	if(g->transport == RN_GROUP_MULTICAST && g->drv.ok)
	{
		if(g->mc_va) { g->drv.MemUnmap(g->mc_va, g->mc_size); g->drv.MemAddressFree(g->mc_va, g->mc_size); }
		for(GroupMember &m : g->dev)
		{
			if(m.uc_va) { g->drv.MemUnmap(m.uc_va, g->mc_size); g->drv.MemAddressFree(m.uc_va, g->mc_size); }
			if(m.mem) g->drv.MemRelease(m.mem);
			m.gathered = nullptr;
		}
		if(g->mc) g->drv.MemRelease(g->mc);
	}
	for(GroupMember &m : g->dev)
	{
		cudaSetDevice(m.device);
		if(m.gathered) cudaFree(m.gathered);
		if(m.list_scratch) cudaFree(m.list_scratch);
		if(m.masks_done) cudaEventDestroy(m.masks_done);
		if(m.frame_end) cudaEventDestroy(m.frame_end);
		if(m.scene) rn_scene_destroy(m.scene);
		if(m.stream) cudaStreamDestroy(m.stream);
	}
	delete g;
}

// One multicast object over all GPUs of the group, each GPU's physical buffer bound to it and mapped twice: at a unicast
// address (what the GPU reads) and behind the group-wide multicast address (what the mask kernels store through).
static bool setup_multicast(rn_group *g, size_t bytes)
{
	DriverApi &d = g->drv;
	if(!load_driver(d)) { snprintf(g->transport_note, sizeof(g->transport_note), "driver without the multicast entry points"); return false; }
	const uint32_t n = g->n_dev;
	std::vector<CUdevice> cud(n);
	for(uint32_t k = 0; k < n; k++)
	{
		cudaSetDevice(g->dev[k].device);
		cudaFree(nullptr);                      // make sure the primary context exists
		int sup = 0;
		if(d.DeviceGet(&cud[k], g->dev[k].device) != CUDA_SUCCESS ||
		   d.DeviceGetAttribute(&sup, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cud[k]) != CUDA_SUCCESS || !sup)
		{
			snprintf(g->transport_note, sizeof(g->transport_note), "device %d does not support multicast objects", g->dev[k].device);
			return false;
		}
	}
	CUmulticastObjectProp mp;
	memset(&mp, 0, sizeof(mp));
	mp.numDevices = n;
	mp.size = bytes;
	mp.handleTypes = 0;
	size_t gran = 0;
	if(d.MulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) != CUDA_SUCCESS || !gran)
	{ snprintf(g->transport_note, sizeof(g->transport_note), "cuMulticastGetGranularity failed"); return false; }
	CUmemAllocationProp ap;
	memset(&ap, 0, sizeof(ap));
	ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
	ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
	size_t mgran = 0;
	ap.location.id = g->dev[0].device;
	d.MemGetAllocationGranularity(&mgran, &ap, CU_MEM_ALLOC_GRANULARITY_MINIMUM);
	gran = std::max(gran, mgran);
	const size_t size = (bytes + gran - 1) / gran * gran;
	mp.size = size;
	g->mc_size = size;
	CUresult r = d.MulticastCreate(&g->mc, &mp);
	if(r != CUDA_SUCCESS) { g->mc = 0; snprintf(g->transport_note, sizeof(g->transport_note), "cuMulticastCreate failed (%d)", (int)r); return false; }
	for(uint32_t k = 0; k < n; k++)
		if((r = d.MulticastAddDevice(g->mc, cud[k])) != CUDA_SUCCESS)
		{ snprintf(g->transport_note, sizeof(g->transport_note), "cuMulticastAddDevice failed (%d)", (int)r); return false; }
	std::vector<CUmemAccessDesc> acc(n);
	for(uint32_t k = 0; k < n; k++)
	{
		acc[k].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
		acc[k].location.id = g->dev[k].device;
		acc[k].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
	}
	for(uint32_t k = 0; k < n; k++)
	{
		GroupMember &m = g->dev[k];
		cudaSetDevice(m.device);
		ap.location.id = m.device;
		if((r = d.MemCreate(&m.mem, size, &ap, 0)) != CUDA_SUCCESS) { m.mem = 0; snprintf(g->transport_note, sizeof(g->transport_note), "cuMemCreate failed (%d)", (int)r); return false; }
		if((r = d.MulticastBindMem(g->mc, 0, m.mem, 0, size, 0)) != CUDA_SUCCESS)
		{ snprintf(g->transport_note, sizeof(g->transport_note), "cuMulticastBindMem failed (%d)", (int)r); return false; }
		if((r = d.MemAddressReserve(&m.uc_va, size, gran, 0, 0)) != CUDA_SUCCESS) { m.uc_va = 0; return false; }
		if((r = d.MemMap(m.uc_va, size, 0, m.mem, 0)) != CUDA_SUCCESS) return false;
		if((r = d.MemSetAccess(m.uc_va, size, &acc[k], 1)) != CUDA_SUCCESS) return false;
		m.gathered = reinterpret_cast<uint32_t *>(m.uc_va);
		if(cudaMemset(m.gathered, 0, bytes) != cudaSuccess) return false;
	}
	if((r = d.MemAddressReserve(&g->mc_va, size, gran, 0, 0)) != CUDA_SUCCESS) { g->mc_va = 0; return false; }
	if((r = d.MemMap(g->mc_va, size, 0, g->mc, 0)) != CUDA_SUCCESS) return false;
	if((r = d.MemSetAccess(g->mc_va, size, acc.data(), n)) != CUDA_SUCCESS) return false;
	snprintf(g->transport_note, sizeof(g->transport_note), "multicast object over %u devices, %zu bytes", n, size);
	return true;
}

static bool setup_peer(rn_group *g, size_t bytes)
{
	const uint32_t n = g->n_dev;
	for(uint32_t a = 0; a < n; a++)
	{
		cudaSetDevice(g->dev[a].device);
		for(uint32_t b = 0; b < n; b++)
		{
			if(a == b) continue;
			int can = 0;
			cudaDeviceCanAccessPeer(&can, g->dev[a].device, g->dev[b].device);
			if(!can) { snprintf(g->transport_note, sizeof(g->transport_note), "device %d cannot access device %d", g->dev[a].device, g->dev[b].device); return false; }
			cudaError_t e = cudaDeviceEnablePeerAccess(g->dev[b].device, 0);
			if(e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return false; }
			cudaGetLastError();
		}
	}
	for(uint32_t k = 0; k < n; k++)
	{
		cudaSetDevice(g->dev[k].device);
		if(cudaMalloc(reinterpret_cast<void **>(&g->dev[k].gathered), bytes) != cudaSuccess) return false;
		cudaMemset(g->dev[k].gathered, 0, bytes);
	}
	snprintf(g->transport_note, sizeof(g->transport_note), "peer-mapped stores over %u devices", n);
	return true;
}

extern "C" int rn_group_create(rn_group_t **out, const rn_group_desc *desc, const rn_node_columns *cols, uint32_t count)
{
	RN_REQUIRE(out && desc && cols, RN_ERR_INVALID, "NULL argument");
	*out = nullptr;
	RN_REQUIRE(desc->n_devices >= 1 && desc->n_devices <= RN_MAX_GROUP, RN_ERR_INVALID, "n_devices must be 1..%d", RN_MAX_GROUP);
	RN_REQUIRE(count > 0 && cols->parent, RN_ERR_INVALID, "rn_group_create needs the node columns");
	int ndev = 0;
	RN_CUDA(cudaGetDeviceCount(&ndev));
	for(uint32_t k = 0; k < desc->n_devices; k++)
	{
		RN_REQUIRE(desc->devices[k] >= 0 && desc->devices[k] < ndev, RN_ERR_CUDA, "CUDA device %d not available (%d devices)", desc->devices[k], ndev);
		for(uint32_t j = 0; j < k; j++)
			RN_REQUIRE(desc->devices[j] != desc->devices[k], RN_ERR_INVALID, "device %d listed twice", desc->devices[k]);
	}
	rn_group *g = new(std::nothrow) rn_group();
	RN_REQUIRE(g, RN_ERR_NOMEM, "out of host memory");
	g->desc = *desc;
	g->n = count;
	g->n_dev = desc->n_devices;
	g->bounds.resize(g->n_dev + 1);
	int rc = rn_plan_ranges(cols->parent, count, g->n_dev, g->bounds.data());
	if(rc) { delete g; return rc; }
	g->dev.resize(g->n_dev);
	uint32_t ofs = 0;
	for(uint32_t k = 0; k < g->n_dev && !rc; k++)
	{
		GroupMember &m = g->dev[k];
		m.device = desc->devices[k];
		m.first = g->bounds[k];
		m.count = g->bounds[k + 1] - g->bounds[k];
		m.words = (m.count + 31u) / 32u;
		m.word_ofs = ofs;
		ofs += (m.words + 3u) & ~3u;          // 16-byte aligned blocks
		cudaError_t e = cudaSetDevice(m.device);
		if(e != cudaSuccess) { rc = cuda_fail(e, "cudaSetDevice", __FILE__, __LINE__); break; }
		rn_scene_desc sd;
		memset(&sd, 0, sizeof(sd));
		sd.device = m.device;
		sd.max_nodes = std::max(m.count, 1u);
		sd.tile_nodes = desc->tile_nodes; sd.tile_inner_nodes = desc->tile_inner_nodes; sd.tile_unit_nodes = desc->tile_unit_nodes;
		sd.flags = desc->flags;
		rc = rn_scene_create(&m.scene, &sd);
		if(rc) break;
		if((e = cudaStreamCreateWithFlags(&m.stream, cudaStreamNonBlocking)) != cudaSuccess ||
		   (e = cudaEventCreateWithFlags(&m.masks_done, cudaEventDisableTiming)) != cudaSuccess ||
		   (e = cudaEventCreateWithFlags(&m.frame_end, cudaEventDisableTiming)) != cudaSuccess)
		{ rc = cuda_fail(e, "stream / event creation", __FILE__, __LINE__); break; }
		std::vector<uint32_t> local(m.count);
		for(uint32_t i = 0; i < m.count; i++)
		{
			const uint32_t q = cols->parent[m.first + i];
			local[i] = q == RN_NO_PARENT ? RN_NO_PARENT : q - m.first;
		}
		rn_node_columns c;
		memset(&c, 0, sizeof(c));
		c.position = cols->position + (size_t)m.first * 3; c.scale = cols->scale + (size_t)m.first * 3;
		c.rotation = cols->rotation + (size_t)m.first * 4; c.parent = local.data();
		c.flags = cols->flags + m.first; c.render_group = cols->render_group + m.first;
		c.bounding_sphere = cols->bounding_sphere + (size_t)m.first * 4;
		if(cols->bounding_box) c.bounding_box = cols->bounding_box + (size_t)m.first * 6;
		if(cols->render_priority) c.render_priority = cols->render_priority + m.first;
		if(m.count) rc = rn_scene_upload(m.scene, &c, 0, m.count);
		if(!rc) rc = rn_scene_set_list_mode(m.scene, 1, m.first);
	}
	if(rc) { group_free(g); return rc; }
	g->row_words = ofs;
	const size_t bytes = (size_t)RN_MAX_CAMERAS * g->row_words * sizeof(uint32_t);
	bool ok = false;
	if(g->n_dev > 1 && desc->transport != RN_GROUP_PEER)
	{
		ok = setup_multicast(g, bytes);
		if(ok) g->transport = RN_GROUP_MULTICAST;
		else if(desc->transport == RN_GROUP_MULTICAST)
		{
			set_error("rn_group_create: the multicast transport was requested but is not available: %s", g->transport_note);
			group_free(g);
			return RN_ERR_PEER;
		}
		else
		{
			// undo the partial multicast set-up before falling back
			char why[256];
			snprintf(why, sizeof(why), "%s", g->transport_note);
			if(g->mc_va) { g->drv.MemUnmap(g->mc_va, g->mc_size); g->drv.MemAddressFree(g->mc_va, g->mc_size); g->mc_va = 0; }
			for(GroupMember &m : g->dev)
			{
				if(m.uc_va) { g->drv.MemUnmap(m.uc_va, g->mc_size); g->drv.MemAddressFree(m.uc_va, g->mc_size); m.uc_va = 0; }
				if(m.mem) { g->drv.MemRelease(m.mem); m.mem = 0; }
				m.gathered = nullptr;
			}
			if(g->mc) { g->drv.MemRelease(g->mc); g->mc = 0; }
			cudaGetLastError();
			ok = setup_peer(g, bytes);
			if(ok)
			{
				g->transport = RN_GROUP_PEER;
				char note[256];
				snprintf(note, sizeof(note), "%.120s (multicast: %.100s)", g->transport_note, why);
				snprintf(g->transport_note, sizeof(g->transport_note), "%s", note);
			}
		}
	}
	else
	{
		ok = setup_peer(g, bytes);
		if(ok) g->transport = RN_GROUP_PEER;
	}
	if(!ok)
	{
		set_error("rn_group_create: no transport between the GPUs: %s", g->transport_note);
		group_free(g);
		return RN_ERR_PEER;
	}
	*out = g;
	return RN_OK;
}

extern "C" void rn_group_destroy(rn_group_t *g)
{
	if(g) group_free(g);
}

extern "C" int rn_group_info(const rn_group_t *g, uint32_t *transport, uint32_t *bounds, uint32_t capacity, uint32_t *row_words, const char **note)
{
	RN_REQUIRE(g, RN_ERR_INVALID, "NULL group");
	if(transport) *transport = g->transport;
	if(row_words) *row_words = g->row_words;
	if(note) *note = g->transport_note;
	if(bounds)
	{
		RN_REQUIRE(capacity >= g->bounds.size(), RN_ERR_CAPACITY, "bounds holds %u entries, %zu needed", capacity, g->bounds.size());
		memcpy(bounds, g->bounds.data(), sizeof(uint32_t) * g->bounds.size());
	}
	return RN_OK;
}

extern "C" int rn_group_update_nodes(rn_group_t *g, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count)
{
	RN_REQUIRE(g && cols && (ranks || count == 0), RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(cols->parent == nullptr && cols->bounding_box == nullptr && cols->render_priority == nullptr, RN_ERR_INVALID,
	           "hierarchy, boxes and render priorities change through a new group");
	std::vector<std::vector<uint32_t>> sel(g->n_dev);
	for(uint32_t i = 0; i < count; i++)
	{
		RN_REQUIRE(ranks[i] < g->n, RN_ERR_INVALID, "rank %u out of range (%u nodes)", ranks[i], g->n);
		const uint32_t k = (uint32_t)(std::upper_bound(g->bounds.begin(), g->bounds.end(), ranks[i]) - g->bounds.begin()) - 1;
		sel[k].push_back(i);
	}
	std::vector<float> pos, scale, rot, sphere;
	std::vector<uint32_t> flags, group, local;
	for(uint32_t k = 0; k < g->n_dev; k++)
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
		for(size_t j = 0; j < m; j++) local[j] = ranks[sel[k][j]] - g->dev[k].first;
		rn_node_columns c;
		memset(&c, 0, sizeof(c));
		c.position = gather(cols->position, pos, 3); c.scale = gather(cols->scale, scale, 3); c.rotation = gather(cols->rotation, rot, 4);
		c.bounding_sphere = gather(cols->bounding_sphere, sphere, 4);
		c.flags = gather_u(cols->flags, flags); c.render_group = gather_u(cols->render_group, group);
		int rc = rn_scene_update_nodes_async(g->dev[k].scene, &c, local.data(), (uint32_t)m, g->dev[k].stream);
		if(rc) return rc;
	}
	return RN_OK;
}

// One frame on every GPU, enqueued without waiting: transform + cull of the GPU's range, then the mask kernel that stores
// the range's visibility words into every GPU's gathered buffer.  On return every GPU's stream is ordered behind the mask
// kernels of ALL GPUs: whatever the caller enqueues next on rn_group_scene's stream (rn_group_expand, a renderer's draw
// preparation) sees the whole scene's visibility.
extern "C" int rn_group_frame(rn_group_t *g, const rn_camera_frustum *cameras, uint32_t n_cameras)
{
	RN_RANGE("rn_group_frame");
	RN_REQUIRE(g && cameras, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_cameras >= 1 && n_cameras <= RN_MAX_CAMERAS, RN_ERR_INVALID, "n_cameras must be 1..%d", RN_MAX_CAMERAS);
	const uint32_t n = g->n_dev;
	// the stores of this frame must not overtake the readers of the previous one on any GPU
	if(g->frames)
	{
		for(uint32_t k = 0; k < n; k++)
		{
			RN_CUDA(cudaSetDevice(g->dev[k].device));
			RN_CUDA(cudaEventRecord(g->dev[k].frame_end, g->dev[k].stream));
		}
		for(uint32_t k = 0; k < n; k++)
		{
			RN_CUDA(cudaSetDevice(g->dev[k].device));
			for(uint32_t j = 0; j < n; j++)
				if(j != k) RN_CUDA(cudaStreamWaitEvent(g->dev[k].stream, g->dev[j].frame_end, 0));
		}
	}
	for(uint32_t k = 0; k < n; k++)
	{
		GroupMember &m = g->dev[k];
		RN_CUDA(cudaSetDevice(m.device));
		if(m.count)
		{
			int rc = rn_transform_cull(m.scene, cameras, n_cameras, nullptr, m.stream);
			if(rc) return rc;
			if(g->transport == RN_GROUP_MULTICAST)
				rc = rn_cull_build_masks_multicast(m.scene, reinterpret_cast<uint32_t *>(g->mc_va) + m.word_ofs, g->row_words, m.stream);
			else
			{
				uint32_t *dsts[RN_MAX_GROUP];
				for(uint32_t j = 0; j < n; j++) dsts[j] = g->dev[j].gathered + m.word_ofs;
				rc = build_masks_fanout(m.scene, dsts, n, g->row_words, m.stream);
			}
			if(rc) return rc;
		}
		RN_CUDA(cudaEventRecord(m.masks_done, m.stream));
	}
	for(uint32_t k = 0; k < n; k++)
	{
		RN_CUDA(cudaSetDevice(g->dev[k].device));
		for(uint32_t j = 0; j < n; j++)
			if(j != k) RN_CUDA(cudaStreamWaitEvent(g->dev[k].stream, g->dev[j].masks_done, 0));
	}
	g->last_ncams = n_cameras;
	g->frames++;
	return RN_OK;
}

extern "C" int rn_group_synchronize(rn_group_t *g)
{
	RN_REQUIRE(g, RN_ERR_INVALID, "NULL group");
	for(GroupMember &m : g->dev)
	{
		RN_CUDA(cudaSetDevice(m.device));
		RN_CUDA(cudaStreamSynchronize(m.stream));
	}
	return RN_OK;
}

extern "C" int rn_group_member(rn_group_t *g, uint32_t member, rn_scene_t **scene, void **cuda_stream, const uint32_t **gathered_masks)
{
	RN_REQUIRE(g && member < g->n_dev, RN_ERR_INVALID, "bad member index");
	if(scene) *scene = g->dev[member].scene;
	if(cuda_stream) *cuda_stream = g->dev[member].stream;
	if(gathered_masks) *gathered_masks = g->dev[member].gathered;
	return RN_OK;
}

// The visibility list of `camera` over the WHOLE scene as ascending global ranks, built on GPU `member` from its copy of the
// gathered masks (rn_masks_expand over the GPUs' blocks in order).  device_list / device_count: memory on that GPU, or NULL
// to use (and get back through out_list / out_count) buffers owned by the group.
extern "C" int rn_group_expand(rn_group_t *g, uint32_t member, uint32_t camera, uint32_t *device_list, uint32_t capacity, uint32_t *device_count,
                               const uint32_t **out_list, const uint32_t **out_count)
{
	RN_RANGE("rn_group_expand");
	RN_REQUIRE(g && member < g->n_dev, RN_ERR_INVALID, "bad member index");
	RN_REQUIRE(camera < g->last_ncams, RN_ERR_STATE, "camera %u was not part of the last frame (%u cameras)", camera, g->last_ncams);
	GroupMember &m = g->dev[member];
	RN_CUDA(cudaSetDevice(m.device));
	if(!device_list || !device_count)
	{
		if(!m.list_scratch)
			RN_CUDA(cudaMalloc(reinterpret_cast<void **>(&m.list_scratch), ((size_t)g->n + 4) * sizeof(uint32_t)));
		device_list = m.list_scratch + 4;
		device_count = m.list_scratch;
		capacity = g->n;
	}
	RN_CUDA(cudaMemsetAsync(device_count, 0, sizeof(uint32_t), m.stream));
	rn_mask_segment seg[RN_MAX_GROUP];
	for(uint32_t j = 0; j < g->n_dev; j++)
	{
		seg[j].device_words = m.gathered + (size_t)camera * g->row_words + g->dev[j].word_ofs;
		seg[j].n_words = g->dev[j].words;
		seg[j].rank_base = g->dev[j].first;
	}
	int rc = rn_masks_expand(m.scene, seg, g->n_dev, device_list, capacity, device_count, m.stream);
	if(rc) return rc;
	if(out_list) *out_list = device_list;
	if(out_count) *out_count = device_count;
	return RN_OK;
}

// rn_group_expand + copy to host memory (waits for that GPU's stream): for tests and host-side consumers.
extern "C" int rn_group_read_list(rn_group_t *g, uint32_t member, uint32_t camera, uint32_t *out_ranks, uint32_t capacity, uint32_t *out_count)
{
	RN_REQUIRE(g && out_count, RN_ERR_INVALID, "NULL argument");
	const uint32_t *lst = nullptr, *cnt = nullptr;
	int rc = rn_group_expand(g, member, camera, nullptr, 0, nullptr, &lst, &cnt);
	if(rc) return rc;
	GroupMember &m = g->dev[member];
	uint32_t n = 0;
	RN_CUDA(cudaMemcpyAsync(&n, cnt, sizeof(uint32_t), cudaMemcpyDeviceToHost, m.stream));
	RN_CUDA(cudaStreamSynchronize(m.stream));
	*out_count = n;
	if(out_ranks)
	{
		RN_REQUIRE(n <= capacity, RN_ERR_CAPACITY, "the list has %u entries, the buffer holds %u", n, capacity);
		if(n)
		{
			RN_CUDA(cudaMemcpyAsync(out_ranks, lst, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, m.stream));
			RN_CUDA(cudaStreamSynchronize(m.stream));
		}
	}
	return RN_OK;
}
