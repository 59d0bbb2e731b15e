// librn_cuda - scene nodes: upload + tiling (host), world-transform propagation and frustum culling (sm_100a).
//
// Reference behaviour being replaced (file:line relative to the Rayne tree):
//   SceneNode::UpdateInternalData            Source/Scene/RNSceneNode.cpp:501-524   world position/rotation/scale
//   SceneNode::UpdateInternalTransformData   Source/Scene/RNSceneNode.cpp:526-546   world matrix = P.world * local
//   SceneNode::UpdateInternalBoundsData      Source/Scene/RNSceneNode.cpp:570-588   transformed bounding sphere
//   SceneNode::CanRenderUtil                 Source/Scene/RNSceneNode.cpp:405-417   hidden / group / NoCulling / frustum
//   Camera::InFrustum(Vector3, float)        Source/Scene/RNCamera.cpp:440-467      frustum sphere + 6 planes, strict >
//   SceneBasic::Render cull loop             Source/Scene/RNSceneBasic.cpp:576-593  list order = render-list order
//
// Data layout in HBM: nodes live in *slots*; the slot range of a tile equals its rank range, and
// inside a tile the nodes are sorted by sublevel (0 = spine nodes whose world transform was computed by the
// pre-pass, k >= 1 = depth k-1 inside a compact subtree).  One CTA processes one tile: the four input columns of
// the tile are staged into shared memory with four TMA bulk copies, the tile's levels are walked with the parents'
// world TRS + world matrix read from shared memory, results are written once.
//
// Compiled with -fmad=false: see rn_math.cuh.
#include "rn_internal.h"
#include "rn_math.cuh"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>

namespace rn {

// ---------------------------------------------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------------------------------------------
static thread_local char g_error[512] = "";

void set_error(const char *fmt, ...)
{
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(g_error, sizeof(g_error), fmt, ap);
	va_end(ap);
}

int cuda_fail(cudaError_t err, const char *what, const char *file, int line)
{
	set_error("CUDA error %d (%s) in %s at %s:%d", (int)err, cudaGetErrorString(err), what, file, line);
	return err == cudaErrorMemoryAllocation ? RN_ERR_NOMEM : RN_ERR_CUDA;
}

// ---------------------------------------------------------------------------------------------------------------
// device helpers: mbarrier + TMA bulk copy (cp.async.bulk, 1-D)
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init()
{
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar)
{
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
	             ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// pulls `bytes` (a multiple of 16) at `src` into L2 without a destination: cp.async.bulk.prefetch.L2 (SASS UBLKPF)
__device__ __forceinline__ void bulk_prefetch_l2(const void *src, uint32_t bytes)
{
	asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
	asm volatile(
	    "{\n"
	    ".reg .pred p;\n"
	    "RN_WAIT:\n"
	    "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
	    "@p bra RN_DONE;\n"
	    "bra RN_WAIT;\n"
	    "RN_DONE:\n"
	    "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}

// ---------------------------------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------------------------------

struct SceneView
{
	const float4 *c0, *c1, *c2, *c3;
	float4 *wpos, *wrot;
	float2 *wsyz;
	float4 *wmat;
	uint8_t *vis;              // camera bit mask per SLOT
	uint32_t *tile_counts;     // [tile][RN_MAX_CAMERAS] visible nodes of the last cull
	uint32_t *super_totals;    // [tile / kSuperTiles][RN_MAX_CAMERAS]
	uint32_t *tile_counter;    // dynamic tile hand-out of the persistent tile kernel (zeroed before every launch)
	const TileMeta *tiles;
	uint32_t n_tiles;
	// second ("late") launch of the tile kernel: tiles whose nodes hang under a node that another tile computes
	// (rn_scene_add_nodes); tile_list = absolute indices into tiles_abs, nullptr in the main launch
	const TileMeta *tiles_abs;
	const uint32_t *tile_list;
	uint32_t n_list, tile_head;
	uint32_t inner_cap;        // P: parents kept in shared memory per tile
	uint32_t warp_smem;        // bytes of shared memory per warp
	uint32_t n;
	uint32_t prefetch_spine_rows;   // tile kernel: L2 prefetch of the next tile's sublevel-0 rows (experiment, RN_TILE_PREFETCH_SPINE=1)
};

struct CullParams
{
	uint32_t n;
	uint32_t simple;                   // bit c: camera c has UseSimpleCulling
	// built on the host (fill_cull_params), copied into shared memory by every CTA:
	float4 test[RN_MAX_CAMERAS * 7];   // [c*7] = frustum sphere (centre xyz, radius); [c*7 + 1 + k] = plane k (normal xyz, d)
	uint32_t group_mask[64];           // 256 bytes: _renderGroup byte -> bit mask of the cameras whose render group intersects it
};

struct NodeWorld
{
	Vec3 wp;
	Quat wr;
	Vec3 ws;
	Affine M;
};

__device__ __forceinline__ float std_max(float a, float b) { return (a < b) ? b : a; } // std::max

// ---------------------------------------------------------------------------------------------------------------
// Culling.  The predicate of the reference, per node and camera (SceneNode::CanRenderUtil,
// Source/Scene/RNSceneNode.cpp:405-417; Camera::InFrustum, Source/Scene/RNCamera.cpp:440-473):
//     !Hidden && (group & cameraGroup) && (NoCulling || (!(dist(centre, p) > R + r) && (simple || all 6 planes: !(n.p - d > r))))
// cull_exact<C> evaluates it per lane in the reference's fp32 operation order.  In front of it sits a warp-level
// classifier: the 32 nodes of a pass are siblings / cousins, i.e. spatially close, so for most (camera, test) pairs
// ALL of them fall on the same side.  The warp reduces the bounding box of its sphere centres and the radius range,
// every lane classifies one (camera, plane) or (camera, frustum sphere) pair against that box with a rigorous
// rounding margin, and a camera whose outcome is decided for the whole warp never runs the per-node tests.
// Undecided cameras (the box straddles a plane, NaNs, negative radii) take the exact path, so the result is always
// bit-identical to the reference's.
// ---------------------------------------------------------------------------------------------------------------
struct CullTable
{
	float4 test[RN_MAX_CAMERAS * 7];   // [c*7] = frustum sphere (centre xyz, radius); [c*7 + 1 + k] = plane k (normal xyz, d)
	uint8_t group_mask[256];           // _renderGroup byte -> bit mask of the cameras whose render group intersects it
	// per-LANE constants of the classifier (the same for every warp of the CTA; read at cull time instead of living in
	// ten registers for the whole kernel)
	float4 lane_e0[32], lane_e1[32];
	uint32_t lane_kind[32];            // kind0 | kind1 << 8
};

// Per-lane constants of the classifier: round 0 = lane l < 30 tests plane l%6 of camera l/6 (cameras 0-4);
// round 1 = lane l < 18 tests plane l%6 of camera 5 + l/6, lanes 18-25 test the frustum sphere of camera l-18.
// kind: 0 = none, 1 = plane, 2 = plane of a simple-culling camera (always inside), 3 = sphere
// The predicate for ONE camera whose seven test vectors sit in registers (loaded warp-uniformly from the CTA's cull
// table): frustum-sphere pre-test, then the six planes, in the reference's fp32 operation order.
struct CamTests
{
	float4 sphere;      // frustum centre xyz, radius
	float4 plane[6];    // normal xyz, d
	bool simple;        // Camera::Flags::UseSimpleCulling
};

__device__ __forceinline__ bool cull_exact(const CamTests &cam, float px, float py, float pz, float radius)
{
	// _frustumCenter.GetDistance(position) > _frustumRadius + radius  (Source/Scene/RNCamera.cpp:444-445)
	const float dx = cam.sphere.x - px, dy = cam.sphere.y - py, dz = cam.sphere.z - pz;
	const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
	if(dist > cam.sphere.w + radius)
		return false;
	if(cam.simple)
		return true;
	bool inside = true;
#pragma unroll
	for(int k = 0; k < 6; k++)
	{
		// Plane::GetDistance: position.Dot(normal) - d   (Source/Math/RNPlane.h:99-102)
		const float d = px * cam.plane[k].x + py * cam.plane[k].y + pz * cam.plane[k].z - cam.plane[k].w;
		if(d > radius)
			inside = false;
	}
	return inside;
}

// Warp-wide fp32 min / max in one instruction (sm_100a: redux.sync.{min,max}.NaN.f32 -> CREDUX.F32.NAN).  A NaN in
// any lane makes the result NaN, which the `sane` guard of cull_node turns into "undecided".
__device__ __forceinline__ float warp_min(float f)
{
	float r;
	asm volatile("redux.sync.min.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(f));
	return r;
}
__device__ __forceinline__ float warp_max(float f)
{
	float r;
	asm volatile("redux.sync.max.NaN.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(f));
	return r;
}

// NaN-propagating two-input min / max (fminf / fmaxf would DROP a NaN, and the classifier's guard relies on seeing it)
__device__ __forceinline__ float min_nan(float a, float b)
{
	float r;
	asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
	return r;
}
__device__ __forceinline__ float max_nan(float a, float b)
{
	float r;
	asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
	return r;
}

struct WarpBox
{
	float lox, hix, loy, hiy, loz, hiz, rlo, rhi;
};

constexpr float kMargin = 1.9073486328125e-6f;   // 2^-19 = 32 ulp(1): covers the <= 8 roundings of either evaluation

// Classifies one plane against the box: bit 0 = every node is outside (culled), bit 1 = every node is inside.
// Reference value per node: d = fl(fl(fl(px nx + py ny) + pz nz) - dd), |d - exact| <= 4u * S with
// S = |px nx| + |py ny| + |pz nz| + |dd|; the bounds below carry at most 6u * S more.
__device__ __forceinline__ uint32_t classify_plane(const float4 &e, const WarpBox &b)
{
	const float a0 = e.x * b.lox, a1 = e.x * b.hix, b0 = e.y * b.loy, b1 = e.y * b.hiy, c0 = e.z * b.loz, c1 = e.z * b.hiz;
	const float lo = fminf(a0, a1) + fminf(b0, b1) + fminf(c0, c1) - e.w;
	const float hi = fmaxf(a0, a1) + fmaxf(b0, b1) + fmaxf(c0, c1) - e.w;
	const float sum = fabsf(e.x) * (fabsf(b.lox) + fabsf(b.hix)) + fabsf(e.y) * (fabsf(b.loy) + fabsf(b.hiy)) +
	                  fabsf(e.z) * (fabsf(b.loz) + fabsf(b.hiz)) + fabsf(e.w);
	const float m = kMargin * sum + 1e-30f;
	uint32_t r = 0;
	if(lo - m > b.rhi) r = 1;          // even the nearest point of the box is farther out than the largest radius
	if(hi + m <= b.rlo) r = 2;         // even the farthest point is within the smallest radius
	return r;
}

// Classifies the frustum-sphere pre-test: node culled iff sqrt(|c - p|^2) > R + r (relative error of either side < 5u).
__device__ __forceinline__ uint32_t classify_sphere(const float4 &e, const WarpBox &b)
{
	const float ux = b.lox - e.x, vx = e.x - b.hix, uy = b.loy - e.y, vy = e.y - b.hiy, uz = b.loz - e.z, vz = e.z - b.hiz;
	const float gx = fmaxf(fmaxf(ux, vx), 0.0f), gy = fmaxf(fmaxf(uy, vy), 0.0f), gz = fmaxf(fmaxf(uz, vz), 0.0f);
	const float Gx = fmaxf(fabsf(ux), fabsf(vx)), Gy = fmaxf(fabsf(uy), fabsf(vy)), Gz = fmaxf(fabsf(uz), fabsf(vz));
	const float near2 = gx * gx + gy * gy + gz * gz;
	const float far2 = Gx * Gx + Gy * Gy + Gz * Gz;
	const float tlo = e.w + b.rlo, thi = e.w + b.rhi;
	uint32_t r = 0;
	if(tlo >= 0.0f)
	{
		if(near2 * (1.0f - 2.0f * kMargin) > thi * thi * (1.0f + 2.0f * kMargin)) r = 1;
		if(far2 * (1.0f + 2.0f * kMargin) <= tlo * tlo * (1.0f - 2.0f * kMargin)) r = 2;
	}
	return r;
}

__device__ __forceinline__ uint32_t classify(uint32_t kind, const float4 &e, const WarpBox &b)
{
	if(kind == 1) return classify_plane(e, b);
	if(kind == 3) return classify_sphere(e, b);
	return kind == 2 ? 2u : 0u;
}

// What the cull needs of one node: the tested sphere (SceneNode::UpdateInternalBoundsData, sphere part), the flags word
// and where its mask byte goes.
struct CullItem
{
	float px, py, pz, radius;
	uint32_t fw;        // flags | render group << 8 | (rank - tile start) << 16
	uint32_t vidx;      // index of the node's mask byte (rank order)
	bool active;        // false: the lane carries a copy of a real node (keeps the warp converged) and stores nothing
};

// Sphere::operator*=(Vector3) then SetRotation (Source/Math/RNSphere.h:62-73); Camera::InFrustum(const Sphere&) tests
// position + offset (Source/Scene/RNCamera.cpp:469-473).
__device__ __forceinline__ void bound_sphere(const Vec3 &wp, const Quat &wr, const Vec3 &ws, const float4 &sph, CullItem &it)
{
	const float s = std_max(std_max(ws.x, ws.y), ws.z);
	it.radius = sph.w * s;
	it.px = wp.x; it.py = wp.y; it.pz = wp.z;
	// A zero offset stays (+-)0 through the scaling and the rotation and `position + offset` returns the position:
	// skipping the rotation is exact (up to the sign of a zero, which no comparison below can see).
	if(sph.x != 0.0f || sph.y != 0.0f || sph.z != 0.0f)
	{
		Vec3 off = {sph.x * s, sph.y * s, sph.z * s};
		off = quat_rotate(wr, off);
		it.px = wp.x + off.x; it.py = wp.y + off.y; it.pz = wp.z + off.z;
	}
}

// The predicate for all cameras and TWO nodes per lane (64 nodes per call: two consecutive passes of a tile are
// siblings / cousins, so one bounding box serves both and the warp-level work - 8 reductions, 2 classification rounds,
// the votes - is paid once per 64 nodes).  Converged call (all 32 lanes).  Returns the camera bit masks.
__device__ __forceinline__ void cull_pair(const CullParams &cp, const CullTable &tab, const CullItem &A, const CullItem &B,
                                          uint32_t &maskA, uint32_t &maskB)
{
	const uint32_t lane = threadIdx.x & 31u;
	__syncwarp();
	WarpBox b;
	b.lox = warp_min(min_nan(A.px, B.px)); b.hix = warp_max(max_nan(A.px, B.px));
	b.loy = warp_min(min_nan(A.py, B.py)); b.hiy = warp_max(max_nan(A.py, B.py));
	b.loz = warp_min(min_nan(A.pz, B.pz)); b.hiz = warp_max(max_nan(A.pz, B.pz));
	b.rlo = warp_min(min_nan(A.radius, B.radius)); b.rhi = warp_max(max_nan(A.radius, B.radius));
	// NaN / inf-inf anywhere makes `guard` NaN: then nothing is decided at warp level
	const float guard = (b.lox - b.hix) + (b.loy - b.hiy) + (b.loz - b.hiz) + (b.rlo - b.rhi);
	const bool sane = guard <= 0.0f;
	const uint32_t kinds = tab.lane_kind[lane];
	const uint32_t r0 = sane ? classify(kinds & 0xffu, tab.lane_e0[lane], b) : 0u;
	const uint32_t out0 = __ballot_sync(0xffffffffu, r0 == 1), in0 = __ballot_sync(0xffffffffu, r0 == 2);
	const uint32_t r1 = sane ? classify(kinds >> 8, tab.lane_e1[lane], b) : 0u;
	const uint32_t out1 = __ballot_sync(0xffffffffu, r1 == 1), in1 = __ballot_sync(0xffffffffu, r1 == 2);
	// lane c (< 8) folds the votes of camera c: culled for all 64 nodes / visible for all / undecided
	const uint32_t c = lane & 7u;
	const uint32_t sh = c < 5u ? 6u * c : 6u * (c - 5u);
	const uint32_t pout = ((c < 5u ? out0 : out1) >> sh) & 63u, pin = ((c < 5u ? in0 : in1) >> sh) & 63u;
	const uint32_t sout = (out1 >> (18u + c)) & 1u, sin = (in1 >> (18u + c)) & 1u;
	const bool culled = (sout | pout) != 0u;
	const uint32_t cams = (1u << cp.n) - 1u;
	const uint32_t all_out = __ballot_sync(0xffffffffu, culled) & cams;
	const uint32_t all_in = __ballot_sync(0xffffffffu, !culled && sin && pin == 63u) & cams;
	uint32_t fa = all_in, fb = all_in;   // per lane: cameras whose frustum test the node passes
	uint32_t undecided = cams & ~(all_out | all_in);
	while(undecided)
	{
		const uint32_t k = (uint32_t)__ffs((int)undecided) - 1u;
		undecided &= undecided - 1u;
		CamTests cam;
		cam.sphere = tab.test[k * 7u];
#pragma unroll
		for(int q = 0; q < 6; q++)
			cam.plane[q] = tab.test[k * 7u + 1u + q];
		cam.simple = (cp.simple >> k) & 1u;
		if(cull_exact(cam, A.px, A.py, A.pz, A.radius)) fa |= 1u << k;
		if(cull_exact(cam, B.px, B.py, B.pz, B.radius)) fb |= 1u << k;
	}
	__syncwarp();
	const uint32_t ga = tab.group_mask[(A.fw >> 8) & 0xffu], gb = tab.group_mask[(B.fw >> 8) & 0xffu];
	maskA = (!A.active || (A.fw & (RN_NODE_HIDDEN | kNodeRemoved))) ? 0u : ((A.fw & RN_NODE_NOCULLING) ? ga : (ga & fa));
	maskB = (!B.active || (B.fw & (RN_NODE_HIDDEN | kNodeRemoved))) ? 0u : ((B.fw & RN_NODE_NOCULLING) ? gb : (gb & fb));
}

// Fills the CTA's cull table incl. the lane constants (all threads of the CTA; contains a __syncthreads).
__device__ __forceinline__ void cull_setup(const CullParams &cp, CullTable &tab)
{
	for(uint32_t t = threadIdx.x; t < RN_MAX_CAMERAS * 7; t += blockDim.x)
		tab.test[t] = cp.test[t];
	for(uint32_t g = threadIdx.x; g < 64; g += blockDim.x)
		reinterpret_cast<uint32_t *>(tab.group_mask)[g] = cp.group_mask[g];
	const uint32_t simple = cp.simple;
	if(threadIdx.x < 32)
	{
		const uint32_t lane = threadIdx.x;
		uint32_t kind0 = 0, kind1 = 0;
		float4 e0 = make_float4(0.0f, 0.0f, 0.0f, 0.0f), e1 = e0;
		if(lane < 30)
		{
			const uint32_t c = lane / 6, k = lane % 6;
			if(c < cp.n) { kind0 = ((simple >> c) & 1u) ? 2u : 1u; e0 = cp.test[c * 7 + 1 + k]; }
		}
		if(lane < 18)
		{
			const uint32_t c = 5 + lane / 6, k = lane % 6;
			if(c < cp.n) { kind1 = ((simple >> c) & 1u) ? 2u : 1u; e1 = cp.test[c * 7 + 1 + k]; }
		}
		else if(lane < 26)
		{
			const uint32_t c = lane - 18;
			if(c < cp.n) { kind1 = 3u; e1 = cp.test[c * 7]; }
		}
		tab.lane_e0[lane] = e0;
		tab.lane_e1[lane] = e1;
		tab.lane_kind[lane] = kind0 | (kind1 << 8);
	}
	__syncthreads();
}

// Per-lane visible counters of a tile, one byte per camera (a tile has at most 255 passes).
struct VisCount
{
	uint32_t lo, hi;   // cameras 0-3, 4-7
};

__device__ __forceinline__ void vis_count_add(VisCount &vc, uint32_t mask)
{
	// bit i of a nibble -> byte i
	vc.lo += ((mask & 0xfu) * 0x00204081u) & 0x01010101u;
	vc.hi += (((mask >> 4) & 0xfu) * 0x00204081u) & 0x01010101u;
}

// Warp-reduces the counters and publishes them: tile_counts[tile][camera], super_totals[tile / 256][camera] += .
__device__ __forceinline__ void vis_count_store(const SceneView &sv, uint32_t tile, const VisCount &vc, uint32_t lane)
{
	uint32_t mine = 0;
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++)
	{
		const uint32_t v = __reduce_add_sync(0xffffffffu, ((c < 4 ? vc.lo : vc.hi) >> (8 * (c & 3))) & 0xffu);
		if(lane == (uint32_t)c)
			mine = v;
	}
	if(lane < RN_MAX_CAMERAS)
	{
		sv.tile_counts[tile * RN_MAX_CAMERAS + lane] = mine;
		if(mine)
			atomicAdd(&sv.super_totals[(tile / kSuperTiles) * RN_MAX_CAMERAS + lane], mine);
	}
}

template <bool FMA>
__device__ __forceinline__ void compute_world(const Vec3 &pos, const Quat &rot, const Vec3 &scale, bool hasParent,
                                              const NodeWorld &P, NodeWorld &out)
{
	const Affine L = local_matrix<FMA>(pos, rot, scale);
	if(hasParent)
	{
		// SceneNode::UpdateInternalData (RNSceneNode.cpp:509-511)
		const Vec3 rv = quat_rotate(P.wr, pos);
		out.wp.x = P.wp.x + P.ws.x * rv.x;
		out.wp.y = P.wp.y + P.ws.y * rv.y;
		out.wp.z = P.wp.z + P.ws.z * rv.z;
		out.wr = quat_mul(P.wr, rot);
		out.ws.x = P.ws.x * scale.x;
		out.ws.y = P.ws.y * scale.y;
		out.ws.z = P.ws.z * scale.z;
		// SceneNode::UpdateInternalTransformData (RNSceneNode.cpp:537)
		out.M = affine_mul<FMA>(P.M, L);
	}
	else
	{
		out.wp = pos;
		out.wr = rot;
		out.ws = scale;
		out.M = L;
	}
}

__device__ __forceinline__ void load_world_global(const SceneView &sv, uint32_t slot, NodeWorld &w)
{
	const float4 a = sv.wpos[slot];
	const float4 b = sv.wrot[slot];
	const float2 c = sv.wsyz[slot];
	w.wp = {a.x, a.y, a.z};
	w.wr = {b.x, b.y, b.z, b.w};
	w.ws = {a.w, c.x, c.y};
	const float4 m0 = sv.wmat[slot * 4u + 0], m1 = sv.wmat[slot * 4u + 1], m2 = sv.wmat[slot * 4u + 2],
	             m3 = sv.wmat[slot * 4u + 3];
	w.M.m[0] = m0.x; w.M.m[1] = m0.y; w.M.m[2] = m0.z;
	w.M.m[3] = m1.x; w.M.m[4] = m1.y; w.M.m[5] = m1.z;
	w.M.m[6] = m2.x; w.M.m[7] = m2.y; w.M.m[8] = m2.z;
	w.M.m[9] = m3.x; w.M.m[10] = m3.y; w.M.m[11] = m3.z;
}

__device__ __forceinline__ void store_world_global(const SceneView &sv, uint32_t slot, const NodeWorld &w)
{
	sv.wpos[slot] = make_float4(w.wp.x, w.wp.y, w.wp.z, w.ws.x);
	sv.wrot[slot] = make_float4(w.wr.x, w.wr.y, w.wr.z, w.wr.w);
	sv.wsyz[slot] = make_float2(w.ws.y, w.ws.z);
	sv.wmat[slot * 4u + 0] = make_float4(w.M.m[0], w.M.m[1], w.M.m[2], 0.0f);
	sv.wmat[slot * 4u + 1] = make_float4(w.M.m[3], w.M.m[4], w.M.m[5], 0.0f);
	sv.wmat[slot * 4u + 2] = make_float4(w.M.m[6], w.M.m[7], w.M.m[8], 0.0f);
	sv.wmat[slot * 4u + 3] = make_float4(w.M.m[9], w.M.m[10], w.M.m[11], 1.0f);
}

// Spine pre-pass: one launch per hierarchy depth of the spine (the few nodes whose subtree does not fit a tile).
template <bool FMA>
__global__ void __launch_bounds__(256) spine_kernel(SceneView sv, const uint32_t *__restrict__ slots, uint32_t count)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= count)
		return;
	const uint32_t slot = slots[i];
	const float4 a = sv.c0[slot], b = sv.c1[slot], c = sv.c2[slot];
	const Vec3 pos = {a.x, a.y, a.z};
	const Vec3 scale = {a.w, c.x, c.y};
	const Quat rot = {b.x, b.y, b.z, b.w};
	const uint32_t parent = __float_as_uint(c.z);
	NodeWorld P, W;
	const bool hasParent = parent != RN_NO_PARENT;
	if(hasParent)
		load_world_global(sv, parent, P);
	compute_world<FMA>(pos, rot, scale, hasParent, P, W);
	store_world_global(sv, slot, W);
}

// The same pre-pass in ONE launch for shallow spines (the usual case: the top few levels of big trees): every thread
// recomputes the world transforms of its node's ancestors from their local columns - the same operations on the same
// inputs, hence the same bits as the per-depth launches - and stores only its own node.  No dependency between threads,
// no barrier, one launch instead of one per depth (each costs ~5 us of latency, a fifth of a 1 M-node frame).
constexpr uint32_t kSpineChainMax = 12;

template <bool FMA>
__global__ void __launch_bounds__(256) spine_chain_kernel(SceneView sv, const uint32_t *__restrict__ slots, uint32_t count, uint32_t zero_words)
{
	// the frame's first kernel also clears the per-superblock totals and the tile hand-out counter in front of them (what a
	// separate memset did: one launch less per frame)
	if(blockIdx.x == 0)
		for(uint32_t w = threadIdx.x; w < zero_words; w += blockDim.x)
			sv.tile_counter[w] = 0;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= count)
		return;
	uint32_t chain[kSpineChainMax];
	uint32_t depth = 0, cur = slots[i];
	for(;;)
	{
		chain[depth++] = cur;
		const uint32_t parent = __float_as_uint(sv.c2[cur].z);
		if(parent == RN_NO_PARENT || depth == kSpineChainMax)
			break;
		cur = parent;
	}
	NodeWorld P, W;
	bool hasParent = false;
	for(uint32_t k = depth; k-- > 0;)
	{
		const uint32_t slot = chain[k];
		const float4 a = sv.c0[slot], b = sv.c1[slot], c = sv.c2[slot];
		const Vec3 pos = {a.x, a.y, a.z};
		const Vec3 scale = {a.w, c.x, c.y};
		const Quat rot = {b.x, b.y, b.z, b.w};
		compute_world<FMA>(pos, rot, scale, hasParent, P, W);
		P = W;
		hasParent = true;
	}
	store_world_global(sv, chain[0], W);
}

// Per-warp shared memory: the world TRS + matrix of the tile's INNER nodes (nodes with children), SoA (88 bytes
// per inner node), the tile descriptor, one staging buffer for the four input columns of a pass (4 x 32 x float4,
// filled by TMA bulk copies) and the mbarrier those copies complete on.
struct WarpSmem
{
	float4 *a, *b, *m0, *m1, *m2;   // (wp.xyz, ws.x) ; wr ; matrix floats 0-3, 4-7, 8-11
	float2 *syz;                    // ws.y, ws.z
	const TileMeta *meta;
	float4 *stage;                  // [4][32]
	uint64_t *bar;
};

constexpr uint32_t kStageBytes = 4u * 32u * 16u;

__host__ __device__ inline uint32_t warp_smem_bytes(uint32_t inner_cap)
{
	return inner_cap * 88u + (uint32_t)sizeof(TileMeta) + kStageBytes + 16u;
}

__device__ __forceinline__ WarpSmem carve_warp_smem(unsigned char *base, uint32_t P)
{
	WarpSmem s;
	float4 *f = reinterpret_cast<float4 *>(base);
	s.a = f; s.b = f + P; s.m0 = f + 2 * P; s.m1 = f + 3 * P; s.m2 = f + 4 * P;
	s.syz = reinterpret_cast<float2 *>(f + 5 * P);
	s.meta = reinterpret_cast<const TileMeta *>(s.syz + P);
	s.stage = reinterpret_cast<float4 *>(reinterpret_cast<unsigned char *>(s.syz + P) + sizeof(TileMeta));
	s.bar = reinterpret_cast<uint64_t *>(s.stage + 128);
	return s;
}

__device__ __forceinline__ void load_world_smem(const WarpSmem &s, uint32_t j, NodeWorld &w)
{
	const float4 a = s.a[j], b = s.b[j];
	const float2 c = s.syz[j];
	const float4 q0 = s.m0[j], q1 = s.m1[j], q2 = s.m2[j];
	w.wp = {a.x, a.y, a.z};
	w.wr = {b.x, b.y, b.z, b.w};
	w.ws = {a.w, c.x, c.y};
	w.M.m[0] = q0.x; w.M.m[1] = q0.y; w.M.m[2] = q0.z; w.M.m[3] = q0.w;
	w.M.m[4] = q1.x; w.M.m[5] = q1.y; w.M.m[6] = q1.z; w.M.m[7] = q1.w;
	w.M.m[8] = q2.x; w.M.m[9] = q2.y; w.M.m[10] = q2.z; w.M.m[11] = q2.w;
}

__device__ __forceinline__ void store_world_smem(const WarpSmem &s, uint32_t j, const NodeWorld &w)
{
	s.a[j] = make_float4(w.wp.x, w.wp.y, w.wp.z, w.ws.x);
	s.b[j] = make_float4(w.wr.x, w.wr.y, w.wr.z, w.wr.w);
	s.syz[j] = make_float2(w.ws.y, w.ws.z);
	s.m0[j] = make_float4(w.M.m[0], w.M.m[1], w.M.m[2], w.M.m[3]);
	s.m1[j] = make_float4(w.M.m[4], w.M.m[5], w.M.m[6], w.M.m[7]);
	s.m2[j] = make_float4(w.M.m[8], w.M.m[9], w.M.m[10], w.M.m[11]);
}

__device__ __forceinline__ void prefetch_l2(const void *p)
{
	asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// Old value of *p, *p += 1 - as atom.inc with a bound below all-ones, not atom.add: ptxas turns every atom.add with a
// uniform address (also one under `if(lane == 0)`, also as inline PTX, also an atom.inc bounded by 0xffffffff) into its
// warp-aggregated form (vote, one ATOMG, SHFL of the result), and the shuffle waits for the atomic's round trip on the
// spot.  The tile hand-out needs the value one tile later, so the latency can stay hidden behind the tile
// (profiles/r02fin2_tile_kernel_ncu.txt: 3 % of the kernel's warp time sat on that shuffle).  The counter counts tiles: it
// never comes near the bound, where atom.inc would wrap to zero.
__device__ __forceinline__ uint32_t atomic_inc_relaxed(uint32_t *p)
{
	uint32_t old;
	asm volatile("atom.global.relaxed.gpu.inc.u32 %0, [%1], 0xfffffffe;" : "=r"(old) : "l"(p) : "memory");
	return old;
}

// World-transform propagation (+ culling): PERSISTENT warps, one tile per warp at a time.  A tile is a contiguous slot
// range holding whole compact subtrees, sorted by sublevel (depth inside the subtree) with the inner nodes of every
// sublevel first.  The warp walks the sublevels in order, 32 nodes per pass: the four input columns of the NEXT pass
// arrive by TMA bulk copies while the current one computes (and the pass after that is pulled into L2 by bulk
// prefetches), parent world TRS + matrix come from the warp's shared memory (written by an earlier pass of the same
// warp, so only __syncwarp separates the levels - no block barrier, the warps of a CTA never wait for each other),
// the outputs are written once.  Every node is read once and written once; parents never come back from HBM.
// Sublevel 0 (spine nodes stored in this tile, transformed by the pre-pass) goes through the same loop with its world
// TRS loaded instead of computed.  The lanes of a pass stay converged (a lane without a node recomputes the last node of
// the level and stores nothing) so that the cull can use warp votes.  The cull of a pass is deferred by one pass and
// done for two passes at once (cull_pair).
// Tiles are handed out dynamically (first round by warp index, then an atomic counter); the descriptor of the next tile
// is fetched while the current one is processed and the TMA pipeline runs across the tile boundary, so a warp never
// sits through the load latency of a tile's first pass.
struct PassPos
{
	uint32_t lev, j0, je;
};

template <bool CULL, bool FMA, int CTAS>
__global__ void __launch_bounds__(kTileThreads, CTAS) tile_kernel(const __grid_constant__ SceneView sv, const __grid_constant__ CullParams cp)
{
	extern __shared__ __align__(16) unsigned char smem_raw[];
	__shared__ CullTable tab;
	__shared__ __align__(16) TileMeta s_meta[kWarpsPerCta][2];
	if(CULL)
		cull_setup(cp, tab);
	const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
	// work items: every tile in list order (main launch) or the late tiles (second launch)
	const bool late_launch = sv.tile_list != nullptr;
	const uint32_t n_work = late_launch ? sv.n_list : sv.n_tiles;
	auto meta_of = [&](uint32_t i) -> const TileMeta * { return late_launch ? sv.tiles_abs + sv.tile_list[i] : sv.tiles + i; };
	auto count_row = [&](uint32_t i) -> uint32_t { return late_launch ? sv.tile_list[i] - sv.tile_head : i; };
	uint32_t tile = blockIdx.x * kWarpsPerCta + warp;
	if(tile >= n_work)
		return;
	const WarpSmem s = carve_warp_smem(smem_raw + (size_t)warp * sv.warp_smem, sv.inner_cap);
	const uint32_t first_dynamic = gridDim.x * kWarpsPerCta;
	// the tile after this one: claimed now, its descriptor loaded during the first pass of the current tile
	uint32_t claim = 0;
	if(lane == 0)
		claim = atomic_inc_relaxed(sv.tile_counter);
	if(lane < sizeof(TileMeta) / 4)
		reinterpret_cast<uint32_t *>(&s_meta[warp][0])[lane] = reinterpret_cast<const uint32_t *>(meta_of(tile))[lane];
	if(lane == 0)
	{
		mbar_init(s.bar, 1);
		mbar_fence_init();
	}
	__syncwarp();
	uint32_t cur = 0;                 // which of the two descriptor slots holds the current tile
	uint32_t phase = 0;

	// The four input columns of pass p+1 are fetched by TMA bulk copies (completion on the warp's mbarrier) while pass p
	// computes; a pass starts by moving its 4 x float4 from the stage to registers.  The 32 slots behind it (the pass
	// after the next; the columns are padded, so running past the tile or the scene is harmless) are pulled into L2.
	// Lane k < 4 issues the copy of column k, lane 4 + k its prefetch: eight bulk operations for the price of two
	// instruction streams instead of eight executed by one lane.
	const uint32_t ncols = CULL ? 4u : 3u;
	const float4 *const my_col = (&sv.c0)[lane & 3u];       // c0..c3 are consecutive kernel parameters
	auto issue = [&](uint32_t slot0, uint32_t cnt) {
		const uint32_t bytes = cnt * 16u;
		// The stage was just read through the generic proxy and is about to be overwritten through the async proxy (TMA):
		// every lane orders its own reads before the copies with a proxy fence, then the warp meets, then the elected
		// lanes issue.  __syncwarp alone does NOT order the two proxies - without the fence a pass now and then computed
		// on the columns of the NEXT pass (seen on scenes whose tiles are mostly single-leaf units with parents in global
		// memory: the long parent loads keep the stage registers' producers in flight; profiles/r02_tile_proxy_fence.txt).
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
		__syncwarp();
		if(lane == 0)
			mbar_expect_tx(s.bar, bytes * ncols);
		if(lane < ncols)
			bulk_g2s(s.stage + 32u * lane, my_col + slot0, bytes, s.bar);
		else if(lane >= 4u && lane < 4u + ncols)
			bulk_prefetch_l2(my_col + slot0 + cnt, 512u);
	};
	// first pass of a tile: sublevel 0 only matters to the cull
	auto first_pass = [&](const TileMeta &m, PassPos &p) -> bool {
		p.lev = 0; p.j0 = 0; p.je = m.lvl_end[0];
		if(m.late && !late_launch)
			return false;                 // this tile belongs to the second launch
		if(!CULL || p.je == 0)
		{
			p.lev = 1;
			p.j0 = p.je;
			p.je = m.lvl_end[m.nlev > 1 ? 1 : 0];
		}
		return p.lev < m.nlev;
	};
	PassPos pos;
	bool have = first_pass(s_meta[warp][0], pos);
	if(have && pos.lev)
		issue(s_meta[warp][0].start + pos.j0, min(32u, pos.je - pos.j0));
	for(;;)
	{
		const TileMeta &meta = s_meta[warp][cur];
		const uint32_t start = meta.start, nlev = meta.nlev;
		// ---- the descriptor of the next tile: load issued now, parked in the other slot after the first pass ----
		claim = __shfl_sync(0xffffffffu, claim, 0);
		const uint32_t next_tile = first_dynamic + claim;
		const bool more = next_tile < n_work;
		uint32_t mreg = 0;
		if(more && lane < sizeof(TileMeta) / 4)
			mreg = __ldg(reinterpret_cast<const uint32_t *>(meta_of(next_tile)) + lane);
		if(more && lane == 0)
			claim = atomic_inc_relaxed(sv.tile_counter);
		bool next_ready = !more;
		auto park_next = [&]() {
			if(!next_ready)
			{
				if(lane < sizeof(TileMeta) / 4)
					reinterpret_cast<uint32_t *>(&s_meta[warp][cur ^ 1u])[lane] = mreg;
				__syncwarp();
				next_ready = true;
				// The spine nodes stored in the next tile (its sublevel 0): their world TRS and local rows are read with
				// plain loads at the start of that tile (the cull's sublevel-0 pass), and their world TRS + matrix are what
				// the unit roots of this tile and the next one read as parents (children precede their parent in the render
				// list, so a unit's spine parent sits in the unit's own tile or in the one after it).  Both reads sit at the
				// head of a dependency chain with the load latency exposed (11 % of the kernel's warp time,
				// profiles/r02fin2_tile_kernel_ncu.txt).  Experiment knob RN_TILE_PREFETCH_SPINE=1: pull them into L2 a tile
				// (or at least a pass) ahead - measured without effect (profiles/r02ab_*: 0.5198 ms with, 0.5182 ms without;
				// the latency that is exposed is not the DRAM's), so it is off by default.
				if(sv.prefetch_spine_rows)
				{
					const TileMeta &nm = s_meta[warp][cur ^ 1u];
					if(lane < nm.lvl_end[0])
					{
						const uint32_t ps = nm.start + lane;
						prefetch_l2(sv.wmat + 4u * (size_t)ps);
						prefetch_l2(sv.wpos + ps);
						prefetch_l2(sv.wrot + ps);
						prefetch_l2(sv.wsyz + ps);
						if(CULL)
						{
							prefetch_l2(sv.c2 + ps);
							prefetch_l2(sv.c3 + ps);
						}
					}
				}
			}
		};
		VisCount vc = {0u, 0u};
		if(have)
		{
			uint32_t lev = pos.lev, j0 = pos.j0, je = pos.je;
			uint32_t jb = j0;                                    // first slot offset of the current sublevel
			uint32_t ib = 0, ni = lev ? meta.inner_end[1] : 0;   // inner nodes: first index, count in this sublevel
			bool pending = false;
			CullItem held;
			for(;;)
			{
				const uint32_t n_cur = min(32u, je - j0);
				const bool active = lane < n_cur;
				const uint32_t l = active ? lane : n_cur - 1;
				const uint32_t j = j0 + l;
				const uint32_t slot = start + j;
				// next pass of this tile
				uint32_t nlv = lev, nj0 = j0 + 32, nje = je;
				if(nj0 >= je)
				{
					nlv = lev + 1;
					nj0 = je;
					nje = meta.lvl_end[nlv < nlev ? nlv : lev];
				}
				const bool has_next = nlv < nlev;
				// what the TMA engine fetches while this pass computes: the next pass of this tile, or the first pass
				// of the next tile
				uint32_t f_slot = start + nj0, f_cnt = min(32u, nje - nj0);
				bool fetch = has_next && nlv;
				if(!has_next && more)
				{
					park_next();
					const TileMeta &nm = s_meta[warp][cur ^ 1u];
					PassPos npos;
					fetch = first_pass(nm, npos) && npos.lev;
					f_slot = nm.start + npos.j0;
					f_cnt = min(32u, npos.je - npos.j0);
				}
				NodeWorld W;
				float4 d;
				uint32_t fw;
				if(!CULL || lev)
				{
					mbar_wait(s.bar, phase);
					phase ^= 1u;
					const float4 a = s.stage[l], b = s.stage[32 + l], c = s.stage[64 + l];
					if(CULL) d = s.stage[96 + l];
					if(fetch)
						issue(f_slot, f_cnt);      // proxy fence + __syncwarp inside
					const Vec3 pos3 = {a.x, a.y, a.z};
					const Vec3 scale = {a.w, c.x, c.y};
					const Quat rot = {b.x, b.y, b.z, b.w};
					const uint32_t parent = __float_as_uint(c.z);
					fw = __float_as_uint(c.w);
					NodeWorld P;
					const bool hasParent = parent != RN_NO_PARENT;
					if(hasParent)
					{
						if(parent & kParentInTile)
							load_world_smem(s, parent & 0xffffu, P);
						else
							load_world_global(sv, parent, P);
					}
					compute_world<FMA>(pos3, rot, scale, hasParent, P, W);
					if(active)
					{
						store_world_global(sv, slot, W);
						if(j - jb < ni)
							store_world_smem(s, ib + (j - jb), W);
					}
					__syncwarp();
				}
				else
				{
					// spine nodes stored inside this tile: world transform already computed by the pre-pass (CULL only)
					if(fetch)
						issue(f_slot, f_cnt);
					const float4 a = sv.wpos[slot], b = sv.wrot[slot];
					const float2 c = sv.wsyz[slot];
					fw = __float_as_uint(sv.c2[slot].w);
					d = sv.c3[slot];
					W.wp = {a.x, a.y, a.z};
					W.wr = {b.x, b.y, b.z, b.w};
					W.ws = {a.w, c.x, c.y};
				}
				park_next();      // after the first pass of the tile the descriptor load has long landed
				if(CULL)
				{
					CullItem cur_item;
					bound_sphere(W.wp, W.wr, W.ws, d, cur_item);
					cur_item.fw = fw;
					cur_item.vidx = start + ((fw >> kRankShift) & kRankMask);
					cur_item.active = active;
					if(!pending && has_next)
					{
						held = cur_item;
						pending = true;
					}
					else
					{
						if(!pending)
						{
							held = cur_item;
							held.active = false;
						}
						uint32_t ma, mb;
						cull_pair(cp, tab, held, cur_item, ma, mb);
						if(held.active) sv.vis[held.vidx] = (uint8_t)ma;
						if(cur_item.active) sv.vis[cur_item.vidx] = (uint8_t)mb;
						vis_count_add(vc, ma);
						vis_count_add(vc, mb);
						pending = false;
					}
				}
				if(!has_next)
					break;
				if(nlv != lev)
				{
					jb = nj0;
					ib = meta.inner_end[lev];
					ni = meta.inner_end[nlv] - ib;
				}
				lev = nlv; j0 = nj0; je = nje;
			}
		}
		else if(more)
		{
			// a tile without a pass (a tile of spine nodes in a propagate-only launch): nothing was fetched ahead
			park_next();
			PassPos npos;
			if(first_pass(s_meta[warp][cur ^ 1u], npos) && npos.lev)
				issue(s_meta[warp][cur ^ 1u].start + npos.j0, min(32u, npos.je - npos.j0));
		}
		if(CULL && !(meta.late && !late_launch))
			vis_count_store(sv, count_row(tile), vc, lane);
		if(!more)
			break;
		tile = next_tile;
		cur ^= 1u;
		have = first_pass(s_meta[warp][cur], pos);
	}
}

// rn_cull on stored world transforms: one warp per tile, 64 consecutive slots per pass.
__global__ void __launch_bounds__(256) cull_kernel(const __grid_constant__ SceneView sv, const __grid_constant__ CullParams cp)
{
	__shared__ CullTable tab;
	cull_setup(cp, tab);
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(tile >= sv.n_tiles)
		return;
	const uint32_t start = sv.tiles[tile].start, count = sv.tiles[tile].count;
	VisCount vc = {0u, 0u};
	for(uint32_t j0 = 0; j0 < count; j0 += 64)
	{
		CullItem it[2];
#pragma unroll
		for(int h = 0; h < 2; h++)
		{
			const uint32_t j = j0 + 32u * h + lane;
			it[h].active = j < count;
			const uint32_t slot = start + (it[h].active ? j : count - 1);
			const float4 a = sv.wpos[slot], b = sv.wrot[slot];
			const float2 c = sv.wsyz[slot];
			const uint32_t fw = __float_as_uint(sv.c2[slot].w);
			const Vec3 wp = {a.x, a.y, a.z};
			const Quat wr = {b.x, b.y, b.z, b.w};
			const Vec3 ws = {a.w, c.x, c.y};
			bound_sphere(wp, wr, ws, sv.c3[slot], it[h]);
			it[h].fw = fw;
			it[h].vidx = start + ((fw >> kRankShift) & kRankMask);
		}
		uint32_t ma, mb;
		cull_pair(cp, tab, it[0], it[1], ma, mb);
		if(it[0].active) sv.vis[it[0].vidx] = (uint8_t)ma;
		if(it[1].active) sv.vis[it[1].vidx] = (uint8_t)mb;
		vis_count_add(vc, ma);
		vis_count_add(vc, mb);
	}
	vis_count_store(sv, tile, vc, lane);
}

// ---------------------------------------------------------------------------------------------------------------
// Ordered stream compaction of the visibility masks into one list per camera.
// The cull kernels leave per-tile per-camera counts (a tile is a rank range AND a slot range, so its count does not
// depend on the order inside) and add them into per-superblock totals (256 tiles).  A CTA of the emit kernel owns 32
// consecutive tiles: its base offset = totals of the earlier superblocks + counts of the earlier tiles of its own
// superblock (two short coalesced sums, no look-back, no spinning), then every warp walks its tiles' RANKS in order
// and writes the slots of the visible nodes with ballot/popc ranking.  List order = render-list order, exactly.
// ---------------------------------------------------------------------------------------------------------------
constexpr uint32_t kEmitThreads = 256;
constexpr uint32_t kEmitTiles = 16;
constexpr uint32_t kEmitUnroll = 4;

struct EmitView
{
	const uint8_t *vis;
	const uint32_t *slot_of_rank;
	const TileMeta *tiles;
	const uint32_t *tile_counts;    // [tile][RN_MAX_CAMERAS]
	const uint32_t *super_totals;   // [superblock][RN_MAX_CAMERAS]
	uint32_t *lists[RN_MAX_CAMERAS];
	uint32_t *list_counts;
	uint32_t n_tiles, n_cams;
	uint32_t rank_mode;             // 0: lists hold slots; 1: lists hold rank + rank_offset (multi-GPU exchange)
	uint32_t rank_offset;
};

__global__ void __launch_bounds__(kEmitThreads, 6) emit_kernel(const __grid_constant__ EmitView ev)
{
	__shared__ uint32_t s_off[kEmitTiles][RN_MAX_CAMERAS];
	__shared__ uint32_t s_cnt[kEmitTiles][RN_MAX_CAMERAS];
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	const uint32_t first = blockIdx.x * kEmitTiles;
	const uint32_t sb = first / kSuperTiles;
	{
		const uint32_t c = warp; // 8 warps = RN_MAX_CAMERAS
		uint32_t acc = 0;
		for(uint32_t q = lane; q < sb; q += 32)
			acc += ev.super_totals[q * RN_MAX_CAMERAS + c];
		for(uint32_t t = sb * kSuperTiles + lane; t < first; t += 32)
			acc += ev.tile_counts[t * RN_MAX_CAMERAS + c];
		acc = __reduce_add_sync(0xffffffffu, acc);
		const uint32_t t = first + lane;
		const uint32_t cnt = (lane < kEmitTiles && t < ev.n_tiles) ? ev.tile_counts[t * RN_MAX_CAMERAS + c] : 0u;
		uint32_t x = cnt;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
			if(lane >= (uint32_t)o)
				x += y;
		}
		if(lane < kEmitTiles)
		{
			s_off[lane][c] = acc + x - cnt;
			s_cnt[lane][c] = cnt;
		}
		if(blockIdx.x == gridDim.x - 1 && lane == 31 && c < ev.n_cams)
			ev.list_counts[c] = acc + x;
	}
	__syncthreads();
	for(uint32_t k = warp; k < kEmitTiles; k += kEmitThreads / 32)
	{
		const uint32_t t = first + k;
		if(t >= ev.n_tiles)
			break;
		// cameras that see anything of this tile (most tiles are invisible to most cameras)
		uint32_t cams = 0;
#pragma unroll
		for(int c = 0; c < RN_MAX_CAMERAS; c++)
			if((uint32_t)c < ev.n_cams && s_cnt[k][c])
				cams |= 1u << c;
		if(!cams)
			continue;
		const uint32_t start = ev.tiles[t].start, count = ev.tiles[t].count;
		// A camera that sees EVERY node of the tile (the common case next to "sees nothing": tiles are subtrees, and a
		// subtree is usually inside or outside a frustum as a whole) needs no mask bytes and no ranking - its part of
		// the list is the tile's rank range itself, copied with coalesced loads and stores.
		uint32_t full = 0;
#pragma unroll
		for(int c = 0; c < RN_MAX_CAMERAS; c++)
			if(((cams >> c) & 1u) && s_cnt[k][c] == count)
				full |= 1u << c;
		if(full)
		{
			for(uint32_t j = lane; j < count; j += 32)
			{
				const uint32_t id = ev.rank_mode ? start + j + ev.rank_offset : __ldg(ev.slot_of_rank + start + j);
#pragma unroll
				for(int c = 0; c < RN_MAX_CAMERAS; c++)
					if((full >> c) & 1u)
						ev.lists[c][s_off[k][c] + j] = id;
			}
			cams &= ~full;
			if(!cams)
				continue;
		}
		uint32_t off[RN_MAX_CAMERAS];
#pragma unroll
		for(int c = 0; c < RN_MAX_CAMERAS; c++)
			off[c] = s_off[k][c];
		// 128 ranks per iteration, four consecutive ranks per lane.  The mask bytes are stored in RANK order, so a lane's
		// four bytes are one aligned 32-bit load (and its four slots one 128-bit load); bytes of the neighbouring tiles
		// that share the first / last word are masked off.
		const uint32_t end = start + count;
		for(uint32_t q0 = start & ~3u; q0 < end; q0 += 32 * kEmitUnroll)
		{
			const uint32_t q = q0 + lane * kEmitUnroll;
			uint32_t vv = 0;
			uint4 id = make_uint4(0u, 0u, 0u, 0u);
			if(q < end)
			{
				vv = __ldg(reinterpret_cast<const uint32_t *>(ev.vis + q));
				if(q < start) vv &= 0xffffffffu << (8u * (start - q));
				if(q + 4u > end) vv &= 0xffffffffu >> (8u * (q + 4u - end));
				if(ev.rank_mode)
					id = make_uint4(q + ev.rank_offset, q + 1u + ev.rank_offset, q + 2u + ev.rank_offset, q + 3u + ev.rank_offset);
				else if(vv)
					id = __ldg(reinterpret_cast<const uint4 *>(ev.slot_of_rank + q));
			}
			// per-lane visible counts of ALL cameras at once, one byte per camera (cameras 0-3 in ca, 4-7 in cb): the four
			// low (high) bits of a mask byte spread to the four bytes of a register by one multiply.  A lane counts at most
			// 4 and a warp at most 128 per camera, so the packed bytes never carry and ONE pair of warp scans ranks every
			// camera (instead of one scan per camera).
			uint32_t ca = 0, cb = 0;
#pragma unroll
			for(int u = 0; u < (int)kEmitUnroll; u++)
			{
				const uint32_t b = (vv >> (8 * u)) & 0xffu;
				ca += ((b & 0xfu) * 0x00204081u) & 0x01010101u;
				cb += ((b >> 4) * 0x00204081u) & 0x01010101u;
			}
			uint32_t xa = ca, xb = cb;
#pragma unroll
			for(int o = 1; o < 32; o <<= 1)
			{
				const uint32_t ya = __shfl_up_sync(0xffffffffu, xa, o), yb = __shfl_up_sync(0xffffffffu, xb, o);
				if(lane >= (uint32_t)o)
				{
					xa += ya;
					xb += yb;
				}
			}
			const uint32_t ta = __shfl_sync(0xffffffffu, xa, 31), tb = __shfl_sync(0xffffffffu, xb, 31);
			const uint32_t ea = xa - ca, eb = xb - cb;     // exclusive prefixes, still packed
#pragma unroll
			for(int c = 0; c < RN_MAX_CAMERAS; c++)
			{
				if((cams >> c) & 1u)
				{
					const uint32_t sh = 8u * (uint32_t)(c & 3);
					const uint32_t tot = ((c < 4 ? ta : tb) >> sh) & 0xffu;
					if(tot)
					{
						uint32_t *out = ev.lists[c];
						const uint32_t p0 = off[c] + (((c < 4 ? ea : eb) >> sh) & 0xffu);
						const uint32_t b0 = (vv >> c) & 1u, b1 = (vv >> (8 + c)) & 1u, b2 = (vv >> (16 + c)) & 1u, b3 = (vv >> (24 + c)) & 1u;
						const uint32_t p1 = p0 + b0, p2 = p1 + b1, p3 = p2 + b2;
						if(b0) out[p0] = id.x;
						if(b1) out[p1] = id.y;
						if(b2) out[p2] = id.z;
						if(b3) out[p3] = id.w;
						off[c] += tot;
					}
				}
			}
		}
	}
}

// Scatter update of local columns (SetPosition / SetScale / SetRotation / SetFlags ... on a dirty set).
struct UpdateView
{
	float4 *c0, *c1, *c2, *c3;
	const uint32_t *slots;
	const float *pos, *scale, *rot, *sphere;
	const uint32_t *flags, *group;
	uint32_t count;
};

__global__ void __launch_bounds__(256) update_kernel(UpdateView u)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= u.count)
		return;
	const uint32_t slot = u.slots[i];
	if(u.pos || u.scale)
	{
		float4 a = u.c0[slot];
		if(u.pos) { a.x = u.pos[i * 3]; a.y = u.pos[i * 3 + 1]; a.z = u.pos[i * 3 + 2]; }
		if(u.scale) a.w = u.scale[i * 3];
		u.c0[slot] = a;
	}
	if(u.rot)
		u.c1[slot] = make_float4(u.rot[i * 4], u.rot[i * 4 + 1], u.rot[i * 4 + 2], u.rot[i * 4 + 3]);
	if(u.scale || u.flags || u.group)
	{
		float4 c = u.c2[slot];
		if(u.scale) { c.x = u.scale[i * 3 + 1]; c.y = u.scale[i * 3 + 2]; }
		uint32_t fw = __float_as_uint(c.w);
		if(u.flags) fw = (fw & ~0x7fu) | (u.flags[i] & 0x7fu);      // bit 7 is the library's hole mark
		if(u.group) fw = (fw & ~0xff00u) | ((u.group[i] & 0xffu) << 8);
		c.w = __uint_as_float(fw);
		u.c2[slot] = c;
	}
	if(u.sphere)
		u.c3[slot] = make_float4(u.sphere[i * 4], u.sphere[i * 4 + 1], u.sphere[i * 4 + 2], u.sphere[i * 4 + 3]);
}

// ---------------------------------------------------------------------------------------------------------------
// host: tiling
// ---------------------------------------------------------------------------------------------------------------
struct Tiling
{
	std::vector<uint32_t> slot_of_rank, rank_of_slot, depth;
	std::vector<uint8_t> spine;       // by rank
	std::vector<uint8_t> sublevel;    // by rank: 0 spine, k>=1 depth inside unit + 1
	std::vector<uint8_t> is_inner;    // by rank: non-spine node with at least one child
	std::vector<uint16_t> inner_idx;  // by rank: index of an inner node among the inner nodes of its tile
	std::vector<TileMeta> tiles;
	std::vector<uint32_t> spine_slots;      // grouped by depth
	std::vector<uint32_t> spine_level_ofs;
	uint32_t max_depth = 0;
};

// Builds the tiling described in DESIGN.md §3.  O(n) time and memory; returns false on an invalid hierarchy.
static bool build_tiling(uint32_t n, const uint32_t *parent, uint32_t cap, uint32_t inner_cap, uint32_t unit_cap, Tiling &t)
{
	const uint32_t kMaxHeight = kMaxSublevels - 2; // sublevels 1..kMaxSublevels-1 hold depths 0..kMaxHeight
	t.depth.assign(n, 0);
	// children CSR
	std::vector<uint32_t> child_ofs(n + 1, 0);
	for(uint32_t i = 0; i < n; i++)
	{
		if(parent[i] != RN_NO_PARENT)
		{
			if(parent[i] >= n || parent[i] == i) { set_error("node %u has an invalid parent %u", i, parent[i]); return false; }
			child_ofs[parent[i] + 1]++;
		}
	}
	for(uint32_t i = 0; i < n; i++) child_ofs[i + 1] += child_ofs[i];
	std::vector<uint32_t> child(child_ofs[n]);
	{
		std::vector<uint32_t> fill(child_ofs.begin(), child_ofs.end() - 1);
		for(uint32_t i = 0; i < n; i++)
			if(parent[i] != RN_NO_PARENT) child[fill[parent[i]]++] = i;
	}
	// BFS order from the roots
	std::vector<uint32_t> bfs;
	bfs.reserve(n);
	for(uint32_t i = 0; i < n; i++)
		if(parent[i] == RN_NO_PARENT) bfs.push_back(i);
	for(size_t h = 0; h < bfs.size(); h++)
	{
		const uint32_t u = bfs[h];
		for(uint32_t k = child_ofs[u]; k < child_ofs[u + 1]; k++)
		{
			t.depth[child[k]] = t.depth[u] + 1;
			bfs.push_back(child[k]);
		}
	}
	if(bfs.size() != n) { set_error("the parent column contains a cycle"); return false; }

	// subtree extent / size / height, children before parents
	std::vector<uint32_t> lo(n), hi(n), size(n, 1), height(n, 0), inner(n, 0);
	for(uint32_t i = 0; i < n; i++)
	{
		lo[i] = hi[i] = i;
		inner[i] = child_ofs[i + 1] > child_ofs[i] ? 1u : 0u;
	}
	for(size_t h = n; h-- > 0;)
	{
		const uint32_t u = bfs[h];
		const uint32_t p = parent[u];
		if(p != RN_NO_PARENT)
		{
			lo[p] = std::min(lo[p], lo[u]);
			hi[p] = std::max(hi[p], hi[u]);
			size[p] += size[u];
			inner[p] += inner[u];
			height[p] = std::max(height[p], height[u] + 1);
		}
	}
	// covered = inside a compact subtree (contiguous rank range, fits a tile); the rest is the spine
	std::vector<uint32_t> unit_root(n, RN_NO_PARENT);
	t.spine.assign(n, 0);
	t.sublevel.assign(n, 0);
	t.max_depth = 0;
	for(size_t h = 0; h < n; h++)
	{
		const uint32_t u = bfs[h];
		const uint32_t p = parent[u];
		t.max_depth = std::max(t.max_depth, t.depth[u]);
		if(p != RN_NO_PARENT && unit_root[p] != RN_NO_PARENT)
			unit_root[u] = unit_root[p];
		else if(hi[u] - lo[u] + 1 == size[u] && size[u] <= unit_cap && inner[u] <= inner_cap && height[u] <= kMaxHeight)
			unit_root[u] = u;
		if(unit_root[u] == RN_NO_PARENT)
			t.spine[u] = 1;
		else
			t.sublevel[u] = (uint8_t)(t.depth[u] - t.depth[unit_root[u]] + 1);
	}
	// greedy packing of units (a spine rank or the rank range of a compact subtree) into tiles
	t.tiles.clear();
	t.slot_of_rank.assign(n, 0);
	t.rank_of_slot.assign(n, 0);
	t.is_inner.assign(n, 0);
	t.inner_idx.assign(n, 0);
	for(uint32_t i = 0; i < n; i++)
		t.is_inner[i] = (!t.spine[i] && child_ofs[i + 1] > child_ofs[i]) ? 1 : 0;
	uint32_t r = 0;
	while(r < n)
	{
		const uint32_t a = r;
		uint32_t b = r;
		uint32_t tile_inner = 0;
		while(b < n)
		{
			uint32_t ue, ui = 0;
			if(t.spine[b])
				ue = b + 1;
			else
			{
				const uint32_t root = unit_root[b];
				if(lo[root] != b) { set_error("internal tiling error at rank %u", b); return false; }
				ue = hi[root] + 1;
				ui = inner[root];
			}
			if(ue - a > cap || tile_inner + ui > inner_cap)
				break;
			tile_inner += ui;
			b = ue;
		}
		if(b == a) { set_error("internal tiling error: empty tile at rank %u", a); return false; }
		TileMeta tm;
		memset(&tm, 0, sizeof(tm));
		tm.start = a;
		tm.count = (uint16_t)(b - a);
		// counting sort of the tile's ranks by (sublevel, inner nodes first), stable in rank
		uint32_t hist[2 * kMaxSublevels + 1] = {0};
		uint32_t nlev = 1;
		for(uint32_t x = a; x < b; x++)
		{
			hist[2 * t.sublevel[x] + (t.is_inner[x] ? 0 : 1) + 1]++;
			nlev = std::max<uint32_t>(nlev, t.sublevel[x] + 1u);
		}
		uint32_t inner_run = 0;
		for(uint32_t k = 0; k < kMaxSublevels; k++)
		{
			inner_run += hist[2 * k + 1];
			tm.inner_end[k] = (uint16_t)inner_run;
		}
		for(uint32_t k = 0; k < 2 * kMaxSublevels; k++) hist[k + 1] += hist[k];
		for(uint32_t k = 0; k < kMaxSublevels; k++) tm.lvl_end[k] = (uint16_t)hist[2 * k + 2];
		tm.nlev = (uint16_t)nlev;
		uint32_t inner_fill[kMaxSublevels];
		for(uint32_t k = 0; k < kMaxSublevels; k++) inner_fill[k] = k ? tm.inner_end[k - 1] : 0;
		for(uint32_t x = a; x < b; x++)
		{
			const uint32_t slot = a + hist[2 * t.sublevel[x] + (t.is_inner[x] ? 0 : 1)]++;
			t.slot_of_rank[x] = slot;
			t.rank_of_slot[slot] = x;
			if(t.is_inner[x])
				t.inner_idx[x] = (uint16_t)inner_fill[t.sublevel[x]]++;
		}
		t.tiles.push_back(tm);
		r = b;
	}
	// spine slots grouped by depth
	std::vector<uint32_t> per_depth(t.max_depth + 2, 0);
	uint32_t spine_max_depth = 0;
	bool any_spine = false;
	for(uint32_t i = 0; i < n; i++)
		if(t.spine[i]) { per_depth[t.depth[i] + 1]++; spine_max_depth = std::max(spine_max_depth, t.depth[i]); any_spine = true; }
	const uint32_t levels = any_spine ? spine_max_depth + 1 : 0;
	t.spine_level_ofs.assign(levels + 1, 0);
	for(uint32_t d = 0; d < levels; d++) t.spine_level_ofs[d + 1] = t.spine_level_ofs[d] + per_depth[d + 1];
	t.spine_slots.assign(t.spine_level_ofs[levels], 0);
	{
		std::vector<uint32_t> fill(t.spine_level_ofs.begin(), t.spine_level_ofs.end());
		for(uint32_t slot = 0; slot < n; slot++)
		{
			const uint32_t rk = t.rank_of_slot[slot];
			if(t.spine[rk]) t.spine_slots[fill[t.depth[rk]]++] = slot;
		}
	}
	return true;
}

static void free_nodes(rn_scene *s)
{
	void *ptrs[] = {s->c0, s->c1, s->c2, s->c3, s->wpos, s->wrot, s->wsyz, s->wmat, s->d_slot_of_rank, s->vis, s->tiles_base,
	                s->tile_counts, s->super_totals ? s->super_totals - 8 : nullptr, s->list_counts, s->spine_slots, s->scratch_cols,
	                s->d_late_tiles};
	for(void *p : ptrs)
		if(p) cudaFree(p);
	for(int c = 0; c < RN_MAX_CAMERAS; c++)
	{
		if(s->lists[c]) { cudaFree(s->lists[c]); s->lists[c] = nullptr; }
		if(s->lists_alt[c]) { cudaFree(s->lists_alt[c]); s->lists_alt[c] = nullptr; }
	}
	if(s->list_counts_alt) { cudaFree(s->list_counts_alt); s->list_counts_alt = nullptr; }
	if(s->rank_masks_alt) { cudaFree(s->rank_masks_alt); s->rank_masks_alt = nullptr; }
	extras_free(s);
	occlusion_free(s);
	if(s->scratch_host) { cudaFreeHost(s->scratch_host); s->scratch_host = nullptr; }
	if(s->scratch_done) { cudaEventDestroy(s->scratch_done); s->scratch_done = nullptr; }
	for(int i = 0; i < rn_scene::kProfileSlots; i++)
		for(int k = 0; k < 2; k++)
			if(s->prof_ev[i][k]) { cudaEventDestroy(s->prof_ev[i][k]); s->prof_ev[i][k] = nullptr; }
	s->c0 = s->c1 = s->c2 = s->c3 = s->wpos = s->wrot = s->wmat = nullptr;
	s->wsyz = nullptr; s->d_slot_of_rank = nullptr; s->vis = nullptr; s->tiles = nullptr;
	s->tile_counts = s->super_totals = s->list_counts = s->spine_slots = nullptr;
	s->scratch_cols = nullptr;
	s->tiles_base = nullptr; s->d_late_tiles = nullptr; s->d_late_cap = 0;
}

static int tile_env(const char *name, int dflt);

static SceneView make_view(const rn_scene *s)
{
	SceneView v;
	v.c0 = s->c0; v.c1 = s->c1; v.c2 = s->c2; v.c3 = s->c3;
	v.wpos = s->wpos; v.wrot = s->wrot; v.wsyz = s->wsyz; v.wmat = s->wmat;
	v.vis = s->vis; v.tiles = s->tiles; v.tile_counts = s->tile_counts; v.super_totals = s->super_totals;
	v.tile_counter = s->super_totals - 1;       // the word in front of the totals (one memset clears both)
	v.tiles_abs = s->tiles_base; v.tile_list = nullptr; v.n_list = 0; v.tile_head = s->tile_head;
	v.n_tiles = s->n_tiles; v.inner_cap = s->inner_cap; v.warp_smem = warp_smem_bytes(s->inner_cap); v.n = s->n;
	static const int prefetch_spine = tile_env("RN_TILE_PREFETCH_SPINE", 0);   // measured: no gain (0.5198 vs 0.5182 ms), stays off
	v.prefetch_spine_rows = (uint32_t)prefetch_spine;
	return v;
}

static int ensure_lists(rn_scene *s, uint32_t ncams)
{
	for(uint32_t c = 0; c < ncams; c++)
	{
		if(!s->lists[c])
		{
			int rc = dev_alloc(s, &s->lists[c], s->desc.max_nodes);
			if(rc) return rc;
		}
	}
	return RN_OK;
}

// Swaps the two sets of cull outputs (visibility lists, their counts, the rank-order masks).  The second set is allocated
// on the first flip (the lists per camera on demand, by ensure_lists, once they are current).
int scene_flip_outputs(rn_scene *s)
{
	if(!s->list_counts_alt)
	{
		int rc = dev_alloc(s, &s->list_counts_alt, (size_t)RN_MAX_CAMERAS);
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->list_counts_alt, 0, sizeof(uint32_t) * RN_MAX_CAMERAS));
	}
	for(int c = 0; c < RN_MAX_CAMERAS; c++) std::swap(s->lists[c], s->lists_alt[c]);
	std::swap(s->list_counts, s->list_counts_alt);
	std::swap(s->rank_masks, s->rank_masks_alt);
	s->have_masks = false;
	return RN_OK;
}

int clear_super_totals(rn_scene *s, cudaStream_t stream);

// clear = true: the launch also zeroes the super totals + tile counter (callers that pass it skip clear_super_totals)
static int launch_spine(rn_scene *s, cudaStream_t stream, bool clear = false)
{
	const SceneView v = make_view(s);
	const size_t levels = s->spine_level_ofs.empty() ? 0 : s->spine_level_ofs.size() - 1;
	if(levels >= 1 && levels <= kSpineChainMax)
	{
		// shallow spine: all depths in one launch (every thread walks its own ancestor chain)
		const uint32_t cnt = s->spine_level_ofs[levels];
		const uint32_t zero = clear ? (uint32_t)(((s->n_tiles + kSuperTiles - 1) / kSuperTiles) * RN_MAX_CAMERAS + 1) : 0u;
		if(s->fma_matrices)
			spine_chain_kernel<true><<<(cnt + 255) / 256, 256, 0, stream>>>(v, s->spine_slots, cnt, zero);
		else
			spine_chain_kernel<false><<<(cnt + 255) / 256, 256, 0, stream>>>(v, s->spine_slots, cnt, zero);
		s->launches++;
		RN_CUDA(cudaGetLastError());
		return RN_OK;
	}
	if(clear)
	{
		int rc = clear_super_totals(s, stream);
		if(rc) return rc;
	}
	for(size_t d = 0; d + 1 < s->spine_level_ofs.size(); d++)
	{
		const uint32_t cnt = s->spine_level_ofs[d + 1] - s->spine_level_ofs[d];
		if(!cnt) continue;
		if(s->fma_matrices)
			spine_kernel<true><<<(cnt + 255) / 256, 256, 0, stream>>>(v, s->spine_slots + s->spine_level_ofs[d], cnt);
		else
			spine_kernel<false><<<(cnt + 255) / 256, 256, 0, stream>>>(v, s->spine_slots + s->spine_level_ofs[d], cnt);
		s->launches++;
	}
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

// Experiment knobs (environment, read once): RN_TILE_CTAS = 4 | 5 resident CTAs per SM the tile kernel is compiled for
// (5: 96 registers, 4: 128 - the default: no spills, measured fastest, profiles/r02_tile_sweep.jsonl); RN_TILE_PERSIST = CTAs per SM of the launch grid (0 = one warp per tile, no persistence).
static int tile_env(const char *name, int dflt)
{
	const char *v = getenv(name);
	return v && *v ? atoi(v) : dflt;
}

template <bool CULL, bool FMA, int CTAS>
static int launch_tiles_impl(rn_scene *s, const CullParams &cp, cudaStream_t stream, int persist)
{
	const size_t smem = (size_t)warp_smem_bytes(s->inner_cap) * kWarpsPerCta;
	// cudaFuncSetAttribute is per device (and per function): remembered in the scene's device slot, not in a static
	size_t &configured = s->tile_smem_configured[(CULL ? 1 : 0) + (FMA ? 2 : 0) + (CTAS == 4 ? 4 : 0)];
	if(configured < smem)
	{
		RN_CUDA(cudaFuncSetAttribute(tile_kernel<CULL, FMA, CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		RN_CUDA(cudaFuncSetAttribute(tile_kernel<CULL, FMA, CTAS>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
		configured = smem;
	}
	const bool prof = s->profile && s->prof_used < rn_scene::kProfileSlots;
	if(prof)
	{
		if(!s->prof_ev[s->prof_used][0])
		{
			RN_CUDA(cudaEventCreate(&s->prof_ev[s->prof_used][0]));
			RN_CUDA(cudaEventCreate(&s->prof_ev[s->prof_used][1]));
		}
		RN_CUDA(cudaEventRecord(s->prof_ev[s->prof_used][0], stream));
	}
	// persistent warps: as many CTAs as fit on the device at once, the rest of the tiles is handed out dynamically
	if(!s->sm_count)
		RN_CUDA(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, s->device));
	uint32_t grid = (s->n_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
	if(persist > 0)
		grid = std::min<uint32_t>(grid, (uint32_t)s->sm_count * (uint32_t)persist);
	tile_kernel<CULL, FMA, CTAS><<<grid, kTileThreads, smem, stream>>>(make_view(s), cp);
	s->launches++;
	if(!s->late_tiles.empty())
	{
		// tiles added under nodes that the launch above computes (rn_scene_add_nodes): now their parents are in memory
		SceneView lv = make_view(s);
		lv.tile_list = s->d_late_tiles;
		lv.n_list = (uint32_t)s->late_tiles.size();
		RN_CUDA(cudaMemsetAsync(s->super_totals - 1, 0, sizeof(uint32_t), stream));     // the hand-out counter only
		const uint32_t lgrid = std::min<uint32_t>((lv.n_list + kWarpsPerCta - 1) / kWarpsPerCta,
		                                          persist > 0 ? (uint32_t)s->sm_count * (uint32_t)persist : 0xffffffffu);
		tile_kernel<CULL, FMA, CTAS><<<lgrid, kTileThreads, smem, stream>>>(lv, cp);
		s->launches++;
	}
	if(prof)
	{
		RN_CUDA(cudaEventRecord(s->prof_ev[s->prof_used][1], stream));
		s->prof_used++;
	}
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

template <bool CULL>
static int launch_tiles(rn_scene *s, const CullParams &cp, cudaStream_t stream)
{
	static const int ctas = tile_env("RN_TILE_CTAS", 4);
	static const int persist = tile_env("RN_TILE_PERSIST", ctas);
	if(ctas == 4)
		return s->fma_matrices ? launch_tiles_impl<CULL, true, 4>(s, cp, stream, persist) : launch_tiles_impl<CULL, false, 4>(s, cp, stream, persist);
	return s->fma_matrices ? launch_tiles_impl<CULL, true, 5>(s, cp, stream, persist) : launch_tiles_impl<CULL, false, 5>(s, cp, stream, persist);
}

int clear_super_totals(rn_scene *s, cudaStream_t stream)
{
	// + the tile counter of the persistent tile kernel, which lives right in front of the totals
	const size_t supers = (s->n_tiles + kSuperTiles - 1) / kSuperTiles;
	RN_CUDA(cudaMemsetAsync(s->super_totals - 1, 0, (supers * RN_MAX_CAMERAS + 1) * sizeof(uint32_t), stream));
	return RN_OK;
}

static int launch_cull_flat(rn_scene *s, const CullParams &cp, cudaStream_t stream)
{
	const uint32_t grid = (s->n_tiles + 7) / 8;
	cull_kernel<<<grid, 256, 0, stream>>>(make_view(s), cp);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

int launch_compaction(rn_scene *s, uint32_t ncams, cudaStream_t stream)
{
	EmitView ev;
	ev.vis = s->vis;
	ev.slot_of_rank = s->d_slot_of_rank;
	ev.tiles = s->tiles;
	ev.tile_counts = s->tile_counts;
	ev.super_totals = s->super_totals;
	for(int c = 0; c < RN_MAX_CAMERAS; c++) ev.lists[c] = s->lists[c];
	ev.list_counts = s->list_counts;
	ev.n_tiles = s->n_tiles;
	ev.n_cams = ncams;
	ev.rank_mode = s->list_rank_mode;
	ev.rank_offset = s->list_rank_offset;
	emit_kernel<<<(s->n_tiles + kEmitTiles - 1) / kEmitTiles, kEmitThreads, 0, stream>>>(ev);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

static int fill_cull_params(const rn_camera_frustum *cams, uint32_t ncams, CullParams &cp)
{
	RN_REQUIRE(cams != nullptr, RN_ERR_INVALID, "cameras is NULL");
	RN_REQUIRE(ncams >= 1 && ncams <= RN_MAX_CAMERAS, RN_ERR_INVALID, "n_cameras must be 1..%d, got %u", RN_MAX_CAMERAS, ncams);
	memset(&cp, 0, sizeof(cp));
	cp.n = ncams;
	for(uint32_t c = 0; c < ncams; c++)
	{
		if(cams[c].flags & RN_CAMERA_USE_SIMPLE_CULLING) cp.simple |= 1u << c;
		cp.test[c * 7] = make_float4(cams[c].center[0], cams[c].center[1], cams[c].center[2], cams[c].radius);
		for(int k = 0; k < 6; k++)
			cp.test[c * 7 + 1 + k] = make_float4(cams[c].planes[k][0], cams[c].planes[k][1], cams[c].planes[k][2], cams[c].planes[k][3]);
	}
	uint8_t *gm = reinterpret_cast<uint8_t *>(cp.group_mask);
	for(uint32_t g = 0; g < 256; g++)
	{
		uint32_t m = 0;
		for(uint32_t c = 0; c < ncams; c++)
			if(g & cams[c].render_group & 0xffu) m |= 1u << c;
		gm[g] = (uint8_t)m;
	}
	return RN_OK;
}

static void fill_lists(rn_scene *s, uint32_t ncams, rn_vis_list *out)
{
	if(!out) return;
	for(uint32_t c = 0; c < ncams; c++)
	{
		out[c].slots = s->lists[c];
		out[c].count = s->list_counts + c;
		out[c].capacity = s->desc.max_nodes;
		out[c].reserved = 0;
	}
}

} // namespace rn

using namespace rn;

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
extern "C" const char *rn_last_error(void) { return g_error; }
extern "C" int rn_abi_version(void) { return RN_ABI_VERSION; }

extern "C" int rn_scene_create(rn_scene_t **out, const rn_scene_desc *desc)
{
	RN_REQUIRE(out && desc, RN_ERR_INVALID, "rn_scene_create: NULL argument");
	*out = nullptr;
	RN_REQUIRE(desc->max_nodes > 0, RN_ERR_INVALID, "max_nodes must be > 0");
	uint32_t cap = desc->tile_nodes ? desc->tile_nodes : 512;
	RN_REQUIRE(cap >= 64 && cap <= 1024 && (cap % 16) == 0, RN_ERR_INVALID, "tile_nodes must be a multiple of 16 in 64..1024, got %u", cap);
	RN_REQUIRE(desc->max_nodes < 0x80000000u, RN_ERR_INVALID, "max_nodes must be below 2^31");
	uint32_t inner_cap = desc->tile_inner_nodes ? desc->tile_inner_nodes : 128;
	RN_REQUIRE(inner_cap >= 16 && inner_cap <= 1024 && (inner_cap % 2) == 0 &&
	           (size_t)warp_smem_bytes(inner_cap) * kWarpsPerCta <= 227u * 1024u,
	           RN_ERR_INVALID, "tile_inner_nodes must be an even number in 16..%u, got %u", (227u * 1024u / kWarpsPerCta - 64u - kStageBytes - 16u) / 88u, inner_cap);
	// 128: several small subtrees per tile fill the 32-wide passes of every level (6 units of 85 nodes in a tile of the
	// benchmark's 4-ary trees: 94 % of the lanes busy instead of 82 %); measured best on the benchmark scene together with
	// 128 inner nodes, the most that still lets four CTAs share an SM (profiles/r02ae_tiling.jsonl)
	const uint32_t unit_cap = desc->tile_unit_nodes ? desc->tile_unit_nodes : std::min(cap, 128u);
	RN_REQUIRE(unit_cap >= 8 && unit_cap <= cap, RN_ERR_INVALID, "tile_unit_nodes must be in 8..tile_nodes, got %u", unit_cap);
	RN_REQUIRE(desc->max_particles_per_emitter < 16383, RN_ERR_INVALID,
	           "max_particles_per_emitter must be < 16383 (reference: RNParticleEmitter.cpp:115)");
	int ndev = 0;
	RN_CUDA(cudaGetDeviceCount(&ndev));
	RN_REQUIRE(desc->device >= 0 && desc->device < ndev, RN_ERR_CUDA, "CUDA device %d not available (%d devices)", desc->device, ndev);
	RN_CUDA(cudaSetDevice(desc->device));
	cudaDeviceProp prop;
	RN_CUDA(cudaGetDeviceProperties(&prop, desc->device));
	RN_REQUIRE(prop.major == 10, RN_ERR_CUDA, "device %d is sm_%d%d; librn_cuda is built for sm_100a only", desc->device, prop.major, prop.minor);
	rn_scene *s = new(std::nothrow) rn_scene();
	RN_REQUIRE(s, RN_ERR_NOMEM, "out of host memory");
	s->desc = *desc;
	s->device = desc->device;
	s->tile_cap = cap;
	s->inner_cap = inner_cap;
	s->unit_cap = unit_cap;
	s->fma_matrices = !(desc->flags & RN_SCENE_STRICT_MATRICES);
	const size_t n = desc->max_nodes;
	int rc = RN_OK;
	if(!rc) rc = dev_alloc(s, &s->c0, n + 64);   // + the reach of the L2 prefetch
	if(!rc) rc = dev_alloc(s, &s->c1, n + 64);   // + the reach of the L2 prefetch
	if(!rc) rc = dev_alloc(s, &s->c2, n + 64);   // + the reach of the L2 prefetch
	if(!rc) rc = dev_alloc(s, &s->c3, n + 64);   // + the reach of the L2 prefetch
	if(!rc) rc = dev_alloc(s, &s->wpos, n);
	if(!rc) rc = dev_alloc(s, &s->wrot, n);
	if(!rc) rc = dev_alloc(s, &s->wsyz, n);
	if(!rc) rc = dev_alloc(s, &s->wmat, n * 4);
	if(!rc) rc = dev_alloc(s, &s->d_slot_of_rank, n + 4);                  // read four at a time by the compaction
	const size_t vis_bytes = ((n + 31) / 32) * 32 + 32;                    // mask bytes in rank order, read 4 / 32 at a time
	if(!rc) rc = dev_alloc(s, &s->vis, vis_bytes);

	if(!rc) rc = dev_alloc(s, &s->list_counts, (size_t)RN_MAX_CAMERAS);
	if(!rc) rc = dev_alloc(s, &s->spine_slots, n);
	if(rc)
	{
		free_nodes(s);
		delete s;
		return rc;
	}
	cudaMemset(s->list_counts, 0, sizeof(uint32_t) * RN_MAX_CAMERAS);
	cudaMemset(s->vis, 0, vis_bytes);
	cudaMemset(s->d_slot_of_rank, 0, sizeof(uint32_t) * (n + 4));
	*out = s;
	return RN_OK;
}

extern "C" void rn_scene_destroy(rn_scene_t *s)
{
	if(!s) return;
	cudaSetDevice(s->device);
	cudaDeviceSynchronize();
	particles_destroy(s);
	free_nodes(s);
	delete s;
}

extern "C" int rn_scene_node_count(const rn_scene_t *s, uint32_t *out_count)
{
	RN_REQUIRE(s && out_count, RN_ERR_INVALID, "NULL argument");
	*out_count = s->n;
	return RN_OK;
}

static int upload_full(rn_scene *s, const rn_node_columns *cols, uint32_t count)
{
	RN_REQUIRE(cols->position && cols->scale && cols->rotation && cols->parent && cols->flags && cols->render_group && cols->bounding_sphere,
	           RN_ERR_INVALID, "a full upload needs every column");
	RN_REQUIRE(count <= s->desc.max_nodes, RN_ERR_CAPACITY, "%u nodes exceed max_nodes %u", count, s->desc.max_nodes);
	if(s->desc.tile_nodes == 0)
	{
		// tile size left to the library: large scenes take 512-node tiles (fewest passes per node); a small scene is cut
		// into more, smaller tiles so that its few thousand nodes still spread over the GPU's warps - a tile is walked by
		// ONE warp, pass after pass (10 000 nodes: 46 us per frame with 512-node tiles, 26 us with 64-node tiles; 1 M nodes:
		// 76 us with 512, 91 us with 128)
		const uint32_t want = ((count / 600u) + 15u) & ~15u;
		s->tile_cap = std::min(512u, std::max(64u, want));
		if(s->desc.tile_unit_nodes == 0)
			s->unit_cap = std::min(s->tile_cap, 128u);
	}
	Tiling t;
	if(!build_tiling(count, cols->parent, s->tile_cap, s->inner_cap, s->unit_cap, t))
		return RN_ERR_INVALID;
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());

	// pack the columns in slot order
	std::vector<float4> h0(count), h1(count), h2(count), h3(count);
	for(const TileMeta &tm : t.tiles)
	{
		for(uint32_t j = 0; j < tm.count; j++)
		{
			const uint32_t slot = tm.start + j;
			const uint32_t r = t.rank_of_slot[slot];
			const float *p = cols->position + (size_t)r * 3, *sc = cols->scale + (size_t)r * 3, *q = cols->rotation + (size_t)r * 4,
			            *sp = cols->bounding_sphere + (size_t)r * 4;
			const uint32_t pr = cols->parent[r];
			// parent reference: a parent inside the same compact subtree lives in the warp's shared memory
			// (index among the tile's inner nodes); a spine parent is read from global memory by slot
			uint32_t pbits = RN_NO_PARENT;
			if(pr != RN_NO_PARENT)
				pbits = (!t.spine[r] && !t.spine[pr]) ? (kParentInTile | t.inner_idx[pr]) : t.slot_of_rank[pr];
			// flags word: SceneNode flags | render group << 8 | (rank - tile start) << 16 (where the node's mask byte goes)
			const uint32_t fw = (cols->flags[r] & 0xffu) | ((cols->render_group[r] & 0xffu) << 8) | ((r - tm.start) << kRankShift);
			float pf, ff;
			memcpy(&pf, &pbits, 4);
			memcpy(&ff, &fw, 4);
			h0[slot] = make_float4(p[0], p[1], p[2], sc[0]);
			h1[slot] = make_float4(q[0], q[1], q[2], q[3]);
			h2[slot] = make_float4(sc[1], sc[2], pf, ff);
			h3[slot] = make_float4(sp[0], sp[1], sp[2], sp[3]);
		}
	}
	RN_CUDA(cudaMemcpy(s->c0, h0.data(), sizeof(float4) * count, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c1, h1.data(), sizeof(float4) * count, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c2, h2.data(), sizeof(float4) * count, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c3, h3.data(), sizeof(float4) * count, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->d_slot_of_rank, t.slot_of_rank.data(), sizeof(uint32_t) * count, cudaMemcpyHostToDevice));
	// mask bytes of ranks beyond the new node count must read as "invisible" (rn_cull_build_masks reads whole words)
	RN_CUDA(cudaMemset(s->vis, 0, (((size_t)s->desc.max_nodes + 31) / 32) * 32 + 32));
	if(!t.spine_slots.empty())
		RN_CUDA(cudaMemcpy(s->spine_slots, t.spine_slots.data(), sizeof(uint32_t) * t.spine_slots.size(), cudaMemcpyHostToDevice));

	// optional columns: local AABB (A4 / C5a) and render priority (A11)
	if(cols->bounding_box)
	{
		std::vector<float4> hb0(count);
		std::vector<float2> hb1(count);
		for(uint32_t slot = 0; slot < count; slot++)
		{
			const float *b = cols->bounding_box + (size_t)t.rank_of_slot[slot] * 6;
			hb0[slot] = make_float4(b[0], b[1], b[2], b[3]);
			hb1[slot] = make_float2(b[4], b[5]);
		}
		int brc = RN_OK;
		if(!s->box0) brc = dev_alloc(s, &s->box0, (size_t)s->desc.max_nodes);
		if(!brc && !s->box1) brc = dev_alloc(s, &s->box1, (size_t)s->desc.max_nodes);
		if(brc) return brc;
		RN_CUDA(cudaMemcpy(s->box0, hb0.data(), sizeof(float4) * count, cudaMemcpyHostToDevice));
		RN_CUDA(cudaMemcpy(s->box1, hb1.data(), sizeof(float2) * count, cudaMemcpyHostToDevice));
	}
	else
	{
		if(s->box0) { cudaFree(s->box0); s->box0 = nullptr; }
		if(s->box1) { cudaFree(s->box1); s->box1 = nullptr; }
	}
	s->n_prio_groups = 1;
	s->transparent_group = s->first_sky_group = 0xffffffffu;
	if(cols->render_priority)
	{
		std::vector<int32_t> distinct(cols->render_priority, cols->render_priority + count);
		std::sort(distinct.begin(), distinct.end());
		distinct.erase(std::unique(distinct.begin(), distinct.end()), distinct.end());
		RN_REQUIRE(distinct.size() <= 256, RN_ERR_CAPACITY, "%zu distinct render priorities; rn_cull_sort supports 256", distinct.size());
		std::vector<uint8_t> hg(count);
		for(uint32_t slot = 0; slot < count; slot++)
		{
			const int32_t p = cols->render_priority[t.rank_of_slot[slot]];
			hg[slot] = (uint8_t)(std::lower_bound(distinct.begin(), distinct.end(), p) - distinct.begin());
		}
		s->n_prio_groups = (uint32_t)distinct.size();
		for(size_t g = 0; g < distinct.size(); g++)
		{
			if(distinct[g] == RN_RENDER_TRANSPARENT) s->transparent_group = (uint32_t)g;
			if(distinct[g] >= RN_RENDER_SKY && s->first_sky_group == 0xffffffffu) s->first_sky_group = (uint32_t)g;
		}
		if(!s->prio_group)
		{
			int prc = dev_alloc(s, &s->prio_group, (size_t)s->desc.max_nodes);
			if(prc) return prc;
		}
		RN_CUDA(cudaMemcpy(s->prio_group, hg.data(), count, cudaMemcpyHostToDevice));
	}
	else if(s->prio_group) { cudaFree(s->prio_group); s->prio_group = nullptr; }
	s->have_inverse = s->have_bounds = false;
	// a full upload re-tiles the scene: volume assignments (slot order) of an earlier upload are void
	if(s->node_volume) { cudaFree(s->node_volume); s->node_volume = nullptr; }
	s->volumes_on = false;
	if(s->lod_node_model) { cudaFree(s->lod_node_model); s->lod_node_model = nullptr; }   // the same goes for the node -> model table
	// ... and for the occlusion pass's per-slot mesh table and frame counters (a node that re-enters a scene starts at 0)
	if(s->occ_mesh_of_slot) { cudaFree(s->occ_mesh_of_slot); s->occ_mesh_of_slot = nullptr; }
	if(s->occ_counter) { cudaFree(s->occ_counter); s->occ_counter = nullptr; }

	// per-tile arrays, with room in front of and behind the tiles of this upload: rn_scene_add_nodes puts new tiles at the
	// head or at the tail of the list without moving a byte (every spare 32 slots of max_nodes can become one tile)
	if(s->tiles_base) { cudaFree(s->tiles_base); s->tiles_base = nullptr; s->tiles = nullptr; }
	if(s->tile_counts) { cudaFree(s->tile_counts); s->tile_counts = nullptr; }
	if(s->super_totals) { cudaFree(s->super_totals - 8); s->super_totals = nullptr; }
	const size_t nt = t.tiles.size();
	const size_t spare_tiles = std::min<size_t>(((size_t)s->desc.max_nodes - count) / 32, (size_t)1 << 20);
	const size_t room = spare_tiles ? spare_tiles + 16 : 0;
	s->tile_capacity = (uint32_t)(nt + 2 * room);
	s->tile_head = (uint32_t)room;
	int rc = dev_alloc(s, &s->tiles_base, std::max<size_t>(s->tile_capacity, 1));
	if(!rc) rc = dev_alloc(s, &s->tile_counts, std::max<size_t>(s->tile_capacity, 1) * RN_MAX_CAMERAS);
	uint32_t *totals_alloc = nullptr;
	if(!rc) rc = dev_alloc(s, &totals_alloc, 8 + (((size_t)s->tile_capacity + kSuperTiles - 1) / kSuperTiles + 1) * RN_MAX_CAMERAS);
	if(rc) return rc;
	s->super_totals = totals_alloc + 8;          // the word in front of it is the persistent kernel's tile counter
	s->tiles = s->tiles_base + s->tile_head;
	if(nt) RN_CUDA(cudaMemcpy(s->tiles, t.tiles.data(), sizeof(TileMeta) * nt, cudaMemcpyHostToDevice));

	s->n = count;
	s->n_tiles = (uint32_t)nt;
	s->max_depth = t.max_depth;
	s->spine_nodes = (uint32_t)t.spine_slots.size();
	s->spine_level_ofs = t.spine_level_ofs;
	// host mirrors of what the incremental edits need
	s->slot_high = count;
	s->live_nodes = count;
	s->slot_state.assign(s->desc.max_nodes, 0);
	for(uint32_t r = 0; r < count; r++)
		s->slot_state[t.slot_of_rank[r]] = (uint8_t)(8u | (t.spine[r] ? 1u : 0u));
	s->prio_of_slot.clear();
	s->live_by_prio.clear();
	if(cols->render_priority)
	{
		s->prio_of_slot.assign(s->desc.max_nodes, 0);
		std::vector<int32_t> sorted(cols->render_priority, cols->render_priority + count);
		for(uint32_t r = 0; r < count; r++) s->prio_of_slot[t.slot_of_rank[r]] = cols->render_priority[r];
		std::sort(sorted.begin(), sorted.end());
		for(size_t i = 0; i < sorted.size();)
		{
			size_t j = i;
			while(j < sorted.size() && sorted[j] == sorted[i]) j++;
			s->live_by_prio.push_back({sorted[i], (uint32_t)(j - i)});
			i = j;
		}
	}
	else if(count)
		s->live_by_prio.push_back({0, count});
	s->late_tiles.clear();
	s->slot_of_rank.swap(t.slot_of_rank);
	s->rank_of_slot.swap(t.rank_of_slot);
	s->slot_of_rank.resize(s->desc.max_nodes, 0);
	s->rank_of_slot.resize(s->desc.max_nodes, 0);
	s->have_world = false;
	s->last_ncams = 0;
	s->have_masks = false;
	return RN_OK;
}

// wait = true (the two synchronous entry points): the update is complete on return, so a following rn_transform_* on ANY
// stream (rn_stream_create hands out non-blocking streams, which do not order against the legacy default stream) sees it.
static int update_scattered(rn_scene *s, const rn_node_columns *cols, const uint32_t *ranks, uint32_t first, uint32_t count,
                            cudaStream_t stream = 0, bool wait = true)
{
	RN_REQUIRE(cols->parent == nullptr, RN_ERR_INVALID, "the hierarchy can only change through a full upload");
	RN_REQUIRE(cols->bounding_box == nullptr && cols->render_priority == nullptr, RN_ERR_INVALID,
	           "bounding_box / render_priority can only be set through a full upload");
	if(count == 0) return RN_OK;
	RN_CUDA(cudaSetDevice(s->device));
	// staging layout (words): slots | pos | scale | rot | sphere | flags | group ; packed into ONE pinned buffer
	// and moved with ONE async copy on `stream` (the legacy default stream for the two original entry points),
	// followed by the scatter kernel on the same stream.
	const size_t per = 1 + 3 + 3 + 4 + 4 + 1 + 1;
	if(s->scratch_cap < count)
	{
		if(s->scratch_done) RN_CUDA(cudaEventSynchronize(s->scratch_done));
		if(s->scratch_cols) { cudaFree(s->scratch_cols); s->scratch_cols = nullptr; }
		if(s->scratch_host) { cudaFreeHost(s->scratch_host); s->scratch_host = nullptr; }
		const uint32_t cap = std::max<uint32_t>(count, 1024);
		int rc = dev_alloc(s, &s->scratch_cols, per * cap);
		if(rc) return rc;
		RN_CUDA(cudaMallocHost(reinterpret_cast<void **>(&s->scratch_host), per * cap * sizeof(float)));
		s->scratch_cap = cap;
		if(!s->scratch_done) RN_CUDA(cudaEventCreateWithFlags(&s->scratch_done, cudaEventDisableTiming));
	}
	else if(s->scratch_done)
		RN_CUDA(cudaEventSynchronize(s->scratch_done)); // the previous copy out of the pinned buffer has finished
	float *h = s->scratch_host;
	float *d = s->scratch_cols;
	UpdateView u;
	memset(&u, 0, sizeof(u));
	u.c0 = s->c0; u.c1 = s->c1; u.c2 = s->c2; u.c3 = s->c3;
	u.count = count;
	uint32_t *hs = reinterpret_cast<uint32_t *>(h);
	for(uint32_t i = 0; i < count; i++)
	{
		const uint32_t r = ranks ? ranks[i] : first + i;
		RN_REQUIRE(r < s->n, RN_ERR_INVALID, "rank %u out of range (%u nodes)", r, s->n);
		hs[i] = s->slot_of_rank[r];
	}
	size_t ofs = count;
	u.slots = reinterpret_cast<const uint32_t *>(d);
	auto stage = [&](const void *src, size_t words) -> const float * {
		const float *dev = nullptr;
		if(src)
		{
			memcpy(h + ofs, src, words * sizeof(float));
			dev = d + ofs;
			ofs += words;
		}
		return dev;
	};
	u.pos = stage(cols->position, (size_t)count * 3);
	u.scale = stage(cols->scale, (size_t)count * 3);
	u.rot = stage(cols->rotation, (size_t)count * 4);
	u.sphere = stage(cols->bounding_sphere, (size_t)count * 4);
	u.flags = reinterpret_cast<const uint32_t *>(stage(cols->flags, count));
	u.group = reinterpret_cast<const uint32_t *>(stage(cols->render_group, count));
	RN_CUDA(cudaMemcpyAsync(d, h, ofs * sizeof(float), cudaMemcpyHostToDevice, stream));
	update_kernel<<<(count + 255) / 256, 256, 0, stream>>>(u);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	// both staging buffers (pinned host AND device) are free again once the scatter kernel has read them: the next
	// update - on whatever stream - waits for this event before it overwrites them
	RN_CUDA(cudaEventRecord(s->scratch_done, stream));
	if(wait)
		RN_CUDA(cudaStreamSynchronize(stream));
	s->have_world = false;
	s->have_inverse = s->have_bounds = false;
	return RN_OK;
}

extern "C" int rn_scene_upload(rn_scene_t *s, const rn_node_columns *cols, uint32_t first, uint32_t count)
{
	RN_RANGE("rn_scene_upload");
	RN_REQUIRE(s && cols, RN_ERR_INVALID, "NULL argument");
	if(cols->parent)
	{
		RN_REQUIRE(first == 0, RN_ERR_INVALID, "a full upload (parent column given) must start at rank 0");
		return upload_full(s, cols, count);
	}
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "range [%u, %u) exceeds the %u uploaded nodes", first, first + count, s->n);
	return update_scattered(s, cols, nullptr, first, count);
}

extern "C" int rn_scene_update_nodes(rn_scene_t *s, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count)
{
	RN_RANGE("rn_scene_update_nodes");
	RN_REQUIRE(s && cols && (ranks || count == 0), RN_ERR_INVALID, "NULL argument");
	return update_scattered(s, cols, ranks, 0, count);
}

extern "C" int rn_scene_update_nodes_async(rn_scene_t *s, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count,
                                           void *cuda_stream)
{
	RN_RANGE("rn_scene_update_nodes_async");
	RN_REQUIRE(s && cols && (ranks || count == 0), RN_ERR_INVALID, "NULL argument");
	return update_scattered(s, cols, ranks, 0, count, static_cast<cudaStream_t>(cuda_stream), false);
}

// ---------------------------------------------------------------------------------------------------------------
// Incremental edits: SceneBasic::AddNode / RemoveNode without re-tiling the scene.
//
// New nodes become NEW tiles.  A tile's slot range (and with it the range of its mask bytes) may sit anywhere in the
// node arrays - only the ORDER of the tile array is the order of the render list - so a batch is tiled on its own, its
// tiles take the next free slots behind the high-water mark and their descriptors are written in front of (or behind) the
// existing ones, into the room the upload left on both sides.  Nothing that exists moves; lists (slot mode) keep naming
// every node by the slot it always had.  Where the reference puts the batch (SceneBasic::AddRenderNode,
// Source/Scene/RNSceneBasic.cpp:652-676): at the very front if its priority does not exceed the head's - each PushFront
// lands before the previous one, so the batch appears in REVERSE call order -, at the very back if it is not below the
// tail's; a position in the middle of the list is left to a full upload (RN_ERR_UNSUPPORTED).
// A new node may hang under an existing one (AddChild on a node that is in the scene).  If that parent is computed by a
// tile (not by the spine pre-pass) the new tile is marked `late` and runs in a second launch of the tile kernel.
// Removed nodes become holes: bit 7 of their flags word makes every cull treat them as hidden.
// ---------------------------------------------------------------------------------------------------------------
static bool live_head_tail(const rn_scene *s, int32_t &head, int32_t &tail)
{
	bool any = false;
	for(const auto &g : s->live_by_prio)
	{
		if(!g.second) continue;
		if(!any) head = g.first;
		tail = g.first;
		any = true;
	}
	return any;
}

static void live_add(rn_scene *s, int32_t prio, int64_t delta)
{
	for(auto &g : s->live_by_prio)
		if(g.first == prio) { g.second = (uint32_t)((int64_t)g.second + delta); return; }
	s->live_by_prio.push_back({prio, (uint32_t)delta});
	std::sort(s->live_by_prio.begin(), s->live_by_prio.end());
}

__global__ void __launch_bounds__(256) remove_kernel(float4 *c2, const uint32_t *slots, uint32_t count)
{
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= count) return;
	float4 c = c2[slots[i]];
	c.w = __uint_as_float(__float_as_uint(c.w) | kNodeRemoved);
	c2[slots[i]] = c;
}

extern "C" int rn_scene_add_nodes(rn_scene_t *s, const rn_node_columns *cols, uint32_t count, uint32_t *out_slots)
{
	RN_RANGE("rn_scene_add_nodes");
	RN_REQUIRE(s && cols && out_slots, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(cols->position && cols->scale && cols->rotation && cols->parent && cols->flags && cols->render_group && cols->bounding_sphere,
	           RN_ERR_INVALID, "rn_scene_add_nodes needs every node column");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "rn_scene_add_nodes extends an uploaded scene");
	RN_REQUIRE(!s->list_rank_mode, RN_ERR_STATE, "incremental edits need slot lists (rn_scene_set_list_mode(scene, 0, 0)): ranks are not stable under insertion");
	if(count == 0) return RN_OK;
	// ---- where does the reference put this batch? one priority per batch ----
	const int32_t prio = cols->render_priority ? cols->render_priority[0] : 0;
	for(uint32_t i = 0; i < count; i++)
		if((cols->render_priority ? cols->render_priority[i] : 0) != prio)
		{ set_error("rn_scene_add_nodes: one render priority per batch (split the batch)"); return RN_ERR_UNSUPPORTED; }
	int32_t head = 0, tail = 0;
	const bool any = live_head_tail(s, head, tail);
	const bool front = !any || prio <= head;
	if(!front && prio < tail) { set_error("rn_scene_add_nodes: priority %d belongs in the middle of the render list", prio); return RN_ERR_UNSUPPORTED; }
	uint32_t group = 0;
	if(s->prio_group)
	{
		// dense priority index of the sort: the priority must be one the upload already knew
		bool found = false;
		std::vector<int32_t> distinct;
		for(uint32_t slot = 0; slot < s->slot_high; slot++) if(s->slot_state[slot] & 8u) distinct.push_back(s->prio_of_slot[slot]);
		std::sort(distinct.begin(), distinct.end());
		distinct.erase(std::unique(distinct.begin(), distinct.end()), distinct.end());
		for(size_t g = 0; g < distinct.size(); g++) if(distinct[g] == prio) { group = (uint32_t)g; found = true; }
		if(!found || distinct.size() != s->n_prio_groups) { set_error("rn_scene_add_nodes: a render priority the scene has not seen yet"); return RN_ERR_UNSUPPORTED; }
	}
	else if(prio != 0) { set_error("rn_scene_add_nodes: the scene was uploaded without render priorities"); return RN_ERR_UNSUPPORTED; }
	// ---- the batch in list order (front: every PushFront lands before the previous one) ----
	std::vector<uint32_t> seq_of(count);              // list position -> call-sequence index
	for(uint32_t q = 0; q < count; q++) seq_of[q] = front ? count - 1 - q : q;
	std::vector<uint32_t> pos_of(count);
	for(uint32_t q = 0; q < count; q++) pos_of[seq_of[q]] = q;
	std::vector<uint32_t> lparent(count);             // parent inside the batch (list position) or none, for the tiling
	std::vector<uint32_t> ext(count, RN_NO_PARENT);   // slot of an existing parent
	for(uint32_t q = 0; q < count; q++)
	{
		const uint32_t pr = cols->parent[seq_of[q]];
		lparent[q] = RN_NO_PARENT;
		if(pr == RN_NO_PARENT) continue;
		if(pr & RN_PARENT_IS_SLOT)
		{
			const uint32_t slot = pr & ~RN_PARENT_IS_SLOT;
			RN_REQUIRE(slot < s->slot_high && (s->slot_state[slot] & 8u) && !(s->slot_state[slot] & 4u), RN_ERR_INVALID, "node %u: parent slot %u is not a live node", seq_of[q], slot);
			if(s->slot_state[slot] & 2u) { set_error("rn_scene_add_nodes: the parent (slot %u) was itself added under a tile-computed node", slot); return RN_ERR_UNSUPPORTED; }
			ext[q] = slot;
		}
		else
		{
			RN_REQUIRE(pr < count && pr != seq_of[q], RN_ERR_INVALID, "node %u has an invalid parent %u", seq_of[q], pr);
			lparent[q] = pos_of[pr];
		}
	}
	Tiling t;
	if(!build_tiling(count, lparent.data(), s->tile_cap, s->inner_cap, s->unit_cap, t))
		return RN_ERR_INVALID;
	if(!t.spine_slots.empty()) { set_error("rn_scene_add_nodes: a subtree of the batch does not fit a tile (%zu spine nodes)", t.spine_slots.size()); return RN_ERR_UNSUPPORTED; }
	// ---- room? ----
	const uint32_t k_tiles = (uint32_t)t.tiles.size();
	uint64_t need = 0;
	for(const TileMeta &tm : t.tiles) need += (tm.count + 31u) & ~31u;
	const uint32_t base = (s->slot_high + 31u) & ~31u;
	RN_REQUIRE((uint64_t)base + need <= s->desc.max_nodes, RN_ERR_CAPACITY, "%u more nodes need %llu slots, %u are free (max_nodes %u)", count,
	           (unsigned long long)need, s->desc.max_nodes - std::min(base, s->desc.max_nodes), s->desc.max_nodes);
	if(front) RN_REQUIRE(s->tile_head >= k_tiles, RN_ERR_CAPACITY, "no room in front of the tile list (%u tiles)", k_tiles);
	else RN_REQUIRE(s->tile_head + s->n_tiles + k_tiles <= s->tile_capacity, RN_ERR_CAPACITY, "no room behind the tile list (%u tiles)", k_tiles);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	// ---- pack the new slots [base, base + need) ----
	const size_t span = (size_t)need;
	std::vector<float4> h0(span, make_float4(0, 0, 0, 0)), h1(span, make_float4(0, 0, 0, 1)), h2(span), h3(span, make_float4(0, 0, 0, 0));
	{
		float none, holefw;
		const uint32_t nopar = RN_NO_PARENT, hole = kNodeRemoved | RN_NODE_HIDDEN;
		memcpy(&none, &nopar, 4); memcpy(&holefw, &hole, 4);
		for(size_t i = 0; i < span; i++) h2[i] = make_float4(0, 0, none, holefw);    // padding slots are never part of a pass
	}
	std::vector<uint32_t> sor(span);
	for(size_t i = 0; i < span; i++) sor[i] = base + (uint32_t)i;
	std::vector<float4> hb0; std::vector<float2> hb1;
	if(s->box0) { hb0.assign(span, make_float4(-1, -1, -1, 1)); hb1.assign(span, make_float2(1, 1)); }
	std::vector<TileMeta> metas(k_tiles);
	uint32_t at = base;
	for(uint32_t k = 0; k < k_tiles; k++)
	{
		TileMeta tm = t.tiles[k];
		const uint32_t old_start = tm.start;          // position of the tile's first node in the batch's list order
		bool late = false;
		for(uint32_t j = 0; j < tm.count; j++)
		{
			const uint32_t bslot = old_start + j;     // slot inside the batch's own tiling
			const uint32_t q = t.rank_of_slot[bslot]; // list position
			const uint32_t i = seq_of[q];             // the caller's index
			const uint32_t slot = at + j;
			const float *p = cols->position + (size_t)i * 3, *sc = cols->scale + (size_t)i * 3, *qr = cols->rotation + (size_t)i * 4,
			            *sp = cols->bounding_sphere + (size_t)i * 4;
			uint32_t pbits = RN_NO_PARENT;
			if(lparent[q] != RN_NO_PARENT)
				pbits = kParentInTile | t.inner_idx[lparent[q]];
			else if(ext[q] != RN_NO_PARENT)
			{
				pbits = ext[q];
				if(!(s->slot_state[ext[q]] & 1u)) late = true;        // computed by a tile, not by the pre-pass
			}
			const uint32_t fw = (cols->flags[i] & 0x7fu) | ((cols->render_group[i] & 0xffu) << 8) | ((q - old_start) << kRankShift);
			float pf, ff;
			memcpy(&pf, &pbits, 4); memcpy(&ff, &fw, 4);
			const size_t o = slot - base;
			h0[o] = make_float4(p[0], p[1], p[2], sc[0]);
			h1[o] = make_float4(qr[0], qr[1], qr[2], qr[3]);
			h2[o] = make_float4(sc[1], sc[2], pf, ff);
			h3[o] = make_float4(sp[0], sp[1], sp[2], sp[3]);
			if(s->box0 && cols->bounding_box)
			{
				const float *b = cols->bounding_box + (size_t)i * 6;
				hb0[o] = make_float4(b[0], b[1], b[2], b[3]);
				hb1[o] = make_float2(b[4], b[5]);
			}
			// the tile's rank range IS its slot range: the mask byte / rank -> slot entry of list position q - old_start
			sor[(at - base) + (q - old_start)] = slot;
			out_slots[i] = slot;
		}
		tm.start = at;
		tm.late = late ? 1u : 0u;
		metas[k] = tm;
		at += (tm.count + 31u) & ~31u;
	}
	RN_CUDA(cudaMemcpy(s->c0 + base, h0.data(), sizeof(float4) * span, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c1 + base, h1.data(), sizeof(float4) * span, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c2 + base, h2.data(), sizeof(float4) * span, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->c3 + base, h3.data(), sizeof(float4) * span, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(s->d_slot_of_rank + base, sor.data(), sizeof(uint32_t) * span, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemset(s->vis + base, 0, span));
	if(s->box0)
	{
		RN_CUDA(cudaMemcpy(s->box0 + base, hb0.data(), sizeof(float4) * span, cudaMemcpyHostToDevice));
		RN_CUDA(cudaMemcpy(s->box1 + base, hb1.data(), sizeof(float2) * span, cudaMemcpyHostToDevice));
	}
	if(s->prio_group) RN_CUDA(cudaMemset(s->prio_group + base, (int)group, span));
	if(s->node_volume) RN_CUDA(cudaMemset(s->node_volume + base, 0, span));
	if(s->lod_node_model) RN_CUDA(cudaMemset(s->lod_node_model + base, 0xff, span * sizeof(uint16_t)));
	if(s->occ_mesh_of_slot) RN_CUDA(cudaMemset(s->occ_mesh_of_slot + base, 0xff, span * sizeof(int32_t)));
	if(s->occ_counter) RN_CUDA(cudaMemset(s->occ_counter + base, 0, span * sizeof(uint16_t)));
	// ---- the tile descriptors: in front of the list or behind it ----
	const uint32_t first_abs = front ? s->tile_head - k_tiles : s->tile_head + s->n_tiles;
	RN_CUDA(cudaMemcpy(s->tiles_base + first_abs, metas.data(), sizeof(TileMeta) * k_tiles, cudaMemcpyHostToDevice));
	if(front) { s->tile_head -= k_tiles; s->tiles = s->tiles_base + s->tile_head; }
	s->n_tiles += k_tiles;
	bool new_late = false;
	for(uint32_t k = 0; k < k_tiles; k++)
		if(metas[k].late) { s->late_tiles.push_back(first_abs + k); new_late = true; }
	if(new_late)
	{
		if(s->late_tiles.size() > s->d_late_cap)
		{
			if(s->d_late_tiles) cudaFree(s->d_late_tiles);
			s->d_late_cap = (uint32_t)std::max<size_t>(s->late_tiles.size() * 2, 256);
			int rc = dev_alloc(s, &s->d_late_tiles, s->d_late_cap);
			if(rc) return rc;
		}
		RN_CUDA(cudaMemcpy(s->d_late_tiles, s->late_tiles.data(), sizeof(uint32_t) * s->late_tiles.size(), cudaMemcpyHostToDevice));
	}
	// ---- host mirrors ----
	for(uint32_t k = 0; k < k_tiles; k++)
		for(uint32_t j = 0; j < metas[k].count; j++)
		{
			const uint32_t slot = metas[k].start + j;
			s->slot_state[slot] = (uint8_t)(8u | (metas[k].late ? 2u : 0u));
			if(!s->prio_of_slot.empty()) s->prio_of_slot[slot] = prio;
		}
	for(size_t i = 0; i < span; i++)
	{
		s->slot_of_rank[base + i] = sor[i];
		s->rank_of_slot[sor[i]] = base + (uint32_t)i;
	}
	s->slot_high = base + (uint32_t)need;
	s->n = s->slot_high;
	s->live_nodes += count;
	live_add(s, prio, (int64_t)count);
	s->have_world = false;
	s->have_inverse = s->have_bounds = s->have_masks = false;
	s->last_ncams = 0;
	s->edits_fast++;
	s->edits_nodes += count;
	return RN_OK;
}

extern "C" int rn_scene_remove_nodes(rn_scene_t *s, const uint32_t *slots, uint32_t count)
{
	RN_RANGE("rn_scene_remove_nodes");
	RN_REQUIRE(s && (slots || count == 0), RN_ERR_INVALID, "NULL argument");
	if(count == 0) return RN_OK;
	for(uint32_t i = 0; i < count; i++)
		RN_REQUIRE(slots[i] < s->slot_high && (s->slot_state[slots[i]] & 8u) && !(s->slot_state[slots[i]] & 4u), RN_ERR_INVALID,
		           "slot %u is not a live node", slots[i]);
	RN_CUDA(cudaSetDevice(s->device));
	uint32_t *d = nullptr;
	RN_CUDA(cudaMalloc(reinterpret_cast<void **>(&d), sizeof(uint32_t) * count));
	cudaError_t e = cudaMemcpy(d, slots, sizeof(uint32_t) * count, cudaMemcpyHostToDevice);
	if(e == cudaSuccess)
	{
		remove_kernel<<<(count + 255) / 256, 256>>>(s->c2, d, count);
		e = cudaDeviceSynchronize();
	}
	cudaFree(d);
	if(e != cudaSuccess) return cuda_fail(e, "rn_scene_remove_nodes", __FILE__, __LINE__);
	s->launches++;
	for(uint32_t i = 0; i < count; i++)
	{
		s->slot_state[slots[i]] |= 4u;
		live_add(s, s->prio_of_slot.empty() ? 0 : s->prio_of_slot[slots[i]], -1);
	}
	s->live_nodes -= count;
	s->edits_fast++;
	s->edits_nodes += count;
	return RN_OK;
}

extern "C" int rn_scene_update_nodes_by_slot(rn_scene_t *s, const rn_node_columns *cols, const uint32_t *slots, uint32_t count)
{
	RN_REQUIRE(s && cols && (slots || count == 0), RN_ERR_INVALID, "NULL argument");
	std::vector<uint32_t> ranks(count);
	for(uint32_t i = 0; i < count; i++)
	{
		RN_REQUIRE(slots[i] < s->slot_high && (s->slot_state[slots[i]] & 8u), RN_ERR_INVALID, "slot %u is not a node", slots[i]);
		ranks[i] = s->rank_of_slot[slots[i]];
	}
	return update_scattered(s, cols, ranks.data(), 0, count);
}

extern "C" int rn_scene_edit_stats(const rn_scene_t *s, uint32_t *live_nodes, uint32_t *slots_in_use, uint32_t *late_tiles, uint64_t *edits)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	if(live_nodes) *live_nodes = s->live_nodes;
	if(slots_in_use) *slots_in_use = s->slot_high;
	if(late_tiles) *late_tiles = (uint32_t)s->late_tiles.size();
	if(edits) *edits = s->edits_fast;
	return RN_OK;
}

extern "C" int rn_transform_propagate(rn_scene_t *s, void *cuda_stream)
{
	RN_RANGE("rn_transform_propagate");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "no nodes uploaded");
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	int rc = launch_spine(s, stream, true);      // also zeroes the tile counter of the persistent kernel
	if(rc) return rc;
	CullParams cp;
	memset(&cp, 0, sizeof(cp));
	rc = launch_tiles<false>(s, cp, stream);
	if(rc) return rc;
	s->have_world = true;
	s->have_bounds = false;
	return RN_OK;
}

static bool wants_aabb(const rn_camera_frustum *cams, uint32_t n)
{
	for(uint32_t c = 0; c < n; c++)
		if(cams[c].flags & RN_CAMERA_TEST_AABB) return true;
	return false;
}

// rn_cull on stored transforms; with RN_CAMERA_TEST_AABB cameras the mask is post-filtered before the compaction
static int cull_stored(rn_scene *s, const rn_camera_frustum *cameras, uint32_t n_cameras, const CullParams &cp, rn_vis_list *out_lists,
                       cudaStream_t stream)
{
	int rc = clear_super_totals(s, stream);
	if(!rc) rc = launch_cull_flat(s, cp, stream);
	if(!rc && wants_aabb(cameras, n_cameras))
	{
		rc = clear_super_totals(s, stream);
		if(!rc) rc = aabb_filter_launch(s, cameras, n_cameras, stream);
	}
	if(!rc && s->volumes_on)
	{
		rc = clear_super_totals(s, stream);
		if(!rc) rc = volume_filter_launch(s, n_cameras, stream);
	}
	if(!rc) rc = launch_compaction(s, n_cameras, stream);
	if(rc) return rc;
	s->last_ncams = n_cameras;
	s->have_masks = false;
	for(uint32_t c = 0; c < n_cameras; c++) s->last_cam_flags[c] = cameras[c].flags;
	fill_lists(s, n_cameras, out_lists);
	return RN_OK;
}

extern "C" int rn_cull(rn_scene_t *s, const rn_camera_frustum *cameras, uint32_t n_cameras, rn_vis_list *out_lists, void *cuda_stream)
{
	RN_RANGE("rn_cull");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "no nodes uploaded");
	RN_REQUIRE(s->have_world, RN_ERR_STATE, "rn_cull needs world transforms: call rn_transform_propagate first (or rn_transform_cull)");
	CullParams cp;
	int rc = fill_cull_params(cameras, n_cameras, cp);
	if(rc) return rc;
	RN_CUDA(cudaSetDevice(s->device));
	rc = ensure_lists(s, n_cameras);
	if(rc) return rc;
	return cull_stored(s, cameras, n_cameras, cp, out_lists, static_cast<cudaStream_t>(cuda_stream));
}

extern "C" int rn_transform_cull(rn_scene_t *s, const rn_camera_frustum *cameras, uint32_t n_cameras, rn_vis_list *out_lists,
                                 void *cuda_stream)
{
	RN_RANGE("rn_transform_cull");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(s->n > 0, RN_ERR_STATE, "no nodes uploaded");
	CullParams cp;
	int rc = fill_cull_params(cameras, n_cameras, cp);
	if(rc) return rc;
	RN_CUDA(cudaSetDevice(s->device));
	rc = ensure_lists(s, n_cameras);
	if(rc) return rc;
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	if(wants_aabb(cameras, n_cameras))
	{
		// extension path: the AABB filter works on stored transforms
		rc = launch_spine(s, stream, true);
		if(!rc) rc = launch_tiles<false>(s, cp, stream);
		if(rc) return rc;
		s->have_world = true;
		s->have_bounds = false;
		return cull_stored(s, cameras, n_cameras, cp, out_lists, stream);
	}
	rc = launch_spine(s, stream, true);
	if(!rc) rc = launch_tiles<true>(s, cp, stream);
	if(!rc && s->volumes_on)
	{
		rc = clear_super_totals(s, stream);
		if(!rc) rc = volume_filter_launch(s, n_cameras, stream);
	}
	if(!rc) rc = launch_compaction(s, n_cameras, stream);
	if(rc) return rc;
	s->have_world = true;
	s->have_bounds = false;
	s->last_ncams = n_cameras;
	s->have_masks = false;
	for(uint32_t c = 0; c < n_cameras; c++) s->last_cam_flags[c] = cameras[c].flags;
	fill_lists(s, n_cameras, out_lists);
	return RN_OK;
}

extern "C" int rn_scene_set_list_mode(rn_scene_t *s, int ranks, uint32_t rank_offset)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	s->list_rank_mode = ranks ? 1u : 0u;
	s->list_rank_offset = ranks ? rank_offset : 0u;
	return RN_OK;
}

extern "C" int rn_cull_read_counts(rn_scene_t *s, uint32_t *out_counts, uint32_t n_cameras, void *cuda_stream)
{
	RN_REQUIRE(s && out_counts, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_cameras <= s->last_ncams, RN_ERR_STATE, "the last cull had %u cameras, %u requested", s->last_ncams, n_cameras);
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	RN_CUDA(cudaMemcpyAsync(out_counts, s->list_counts, sizeof(uint32_t) * n_cameras, cudaMemcpyDeviceToHost, stream));
	RN_CUDA(cudaStreamSynchronize(stream));
	return RN_OK;
}

extern "C" int rn_cull_read_list(rn_scene_t *s, uint32_t camera, uint32_t *out_slots, uint32_t capacity, uint32_t *out_count,
                                 void *cuda_stream)
{
	RN_REQUIRE(s && out_count, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(camera < s->last_ncams, RN_ERR_STATE, "camera %u was not part of the last cull (%u cameras)", camera, s->last_ncams);
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	uint32_t cnt = 0;
	RN_CUDA(cudaMemcpyAsync(&cnt, s->list_counts + camera, sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
	RN_CUDA(cudaStreamSynchronize(stream));
	*out_count = cnt;
	if(out_slots)
	{
		RN_REQUIRE(cnt <= capacity, RN_ERR_CAPACITY, "list of camera %u has %u entries, buffer holds %u", camera, cnt, capacity);
		if(cnt)
		{
			RN_CUDA(cudaMemcpyAsync(out_slots, s->lists[camera], sizeof(uint32_t) * cnt, cudaMemcpyDeviceToHost, stream));
			RN_CUDA(cudaStreamSynchronize(stream));
		}
	}
	return RN_OK;
}

static int read_lists(rn_scene *s, uint32_t n_cameras, uint32_t *const *out_slots, const uint32_t *capacities, uint32_t *out_counts,
                      cudaStream_t stream, bool wait)
{
	RN_REQUIRE(s && out_slots && capacities && out_counts, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(n_cameras <= s->last_ncams, RN_ERR_STATE, "the last cull had %u cameras, %u requested", s->last_ncams, n_cameras);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaMemcpyAsync(out_counts, s->list_counts, sizeof(uint32_t) * n_cameras, cudaMemcpyDeviceToHost, stream));
	RN_CUDA(cudaStreamSynchronize(stream));
	for(uint32_t c = 0; c < n_cameras; c++)
	{
		RN_REQUIRE(out_counts[c] <= capacities[c], RN_ERR_CAPACITY, "list of camera %u has %u entries, buffer holds %u", c, out_counts[c], capacities[c]);
		if(out_counts[c] && out_slots[c])
			RN_CUDA(cudaMemcpyAsync(out_slots[c], s->lists[c], sizeof(uint32_t) * out_counts[c], cudaMemcpyDeviceToHost, stream));
	}
	if(wait)
		RN_CUDA(cudaStreamSynchronize(stream));
	return RN_OK;
}

extern "C" int rn_cull_read_lists(rn_scene_t *s, uint32_t n_cameras, uint32_t *const *out_slots, const uint32_t *capacities,
                                  uint32_t *out_counts, void *cuda_stream)
{
	RN_RANGE("rn_cull_read_lists");
	return read_lists(s, n_cameras, out_slots, capacities, out_counts, static_cast<cudaStream_t>(cuda_stream), true);
}

extern "C" int rn_cull_read_lists_async(rn_scene_t *s, uint32_t n_cameras, uint32_t *const *out_slots, const uint32_t *capacities,
                                        uint32_t *out_counts, void *cuda_stream)
{
	return read_lists(s, n_cameras, out_slots, capacities, out_counts, static_cast<cudaStream_t>(cuda_stream), false);
}

extern "C" int rn_stream_create(void **out_stream, int urgency)
{
	RN_REQUIRE(out_stream, RN_ERR_INVALID, "NULL argument");
	int least = 0, greatest = 0;   // numerically: greatest priority <= least priority
	RN_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
	int prio = least - (urgency < 0 ? 0 : urgency);
	if(prio < greatest) prio = greatest;
	cudaStream_t st = nullptr;
	RN_CUDA(cudaStreamCreateWithPriority(&st, cudaStreamNonBlocking, prio));
	*out_stream = st;
	return RN_OK;
}

extern "C" int rn_stream_destroy(void *cuda_stream)
{
	if(cuda_stream) RN_CUDA(cudaStreamDestroy(static_cast<cudaStream_t>(cuda_stream)));
	return RN_OK;
}

extern "C" int rn_stream_synchronize(void *cuda_stream)
{
	RN_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(cuda_stream)));
	return RN_OK;
}

extern "C" int rn_scene_download(rn_scene_t *s, rn_download_what what, uint32_t first, uint32_t count, void *dst)
{
	RN_REQUIRE(s && dst, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "range [%u, %u) exceeds the %u uploaded nodes", first, first + count, s->n);
	if(count == 0) return RN_OK;
	if(what == RN_DL_SLOT_OF_RANK)
	{
		memcpy(dst, s->slot_of_rank.data() + first, sizeof(uint32_t) * count);
		return RN_OK;
	}
	if(what == RN_DL_RANK_OF_SLOT)
	{
		memcpy(dst, s->rank_of_slot.data() + first, sizeof(uint32_t) * count);
		return RN_OK;
	}
	if(what == RN_DL_INVERSE_WORLD_MATRIX)
		RN_REQUIRE(s->have_inverse, RN_ERR_STATE, "no inverse world matrices yet: call rn_transform_inverse first");
	else if(what == RN_DL_BOUNDING_SPHERE || what == RN_DL_BOUNDING_BOX)
		RN_REQUIRE(s->have_bounds, RN_ERR_STATE, "no transformed bounds yet: call rn_bounds_update after the propagate");
	else
		RN_REQUIRE(s->have_world, RN_ERR_STATE, "no world transforms yet: call rn_transform_propagate / rn_transform_cull first");
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	uint32_t slo = 0xffffffffu, shi = 0;
	for(uint32_t r = first; r < first + count; r++)
	{
		slo = std::min(slo, s->slot_of_rank[r]);
		shi = std::max(shi, s->slot_of_rank[r]);
	}
	const size_t span = (size_t)shi - slo + 1;
	float *out = static_cast<float *>(dst);
	switch(what)
	{
		case RN_DL_WORLD_POSITION:
		case RN_DL_WORLD_SCALE:
		{
			std::vector<float4> a(span);
			std::vector<float2> b(span);
			RN_CUDA(cudaMemcpy(a.data(), s->wpos + slo, sizeof(float4) * span, cudaMemcpyDeviceToHost));
			RN_CUDA(cudaMemcpy(b.data(), s->wsyz + slo, sizeof(float2) * span, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < count; i++)
			{
				const size_t k = s->slot_of_rank[first + i] - slo;
				if(what == RN_DL_WORLD_POSITION) { out[i * 3] = a[k].x; out[i * 3 + 1] = a[k].y; out[i * 3 + 2] = a[k].z; }
				else { out[i * 3] = a[k].w; out[i * 3 + 1] = b[k].x; out[i * 3 + 2] = b[k].y; }
			}
			break;
		}
		case RN_DL_WORLD_ROTATION:
		{
			std::vector<float4> a(span);
			RN_CUDA(cudaMemcpy(a.data(), s->wrot + slo, sizeof(float4) * span, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < count; i++)
			{
				const size_t k = s->slot_of_rank[first + i] - slo;
				out[i * 4] = a[k].x; out[i * 4 + 1] = a[k].y; out[i * 4 + 2] = a[k].z; out[i * 4 + 3] = a[k].w;
			}
			break;
		}
		case RN_DL_BOUNDING_SPHERE:
		{
			std::vector<float4> a(span);
			RN_CUDA(cudaMemcpy(a.data(), s->bsphere + slo, sizeof(float4) * span, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < count; i++)
				memcpy(out + (size_t)i * 4, &a[s->slot_of_rank[first + i] - slo], 16);
			break;
		}
		case RN_DL_BOUNDING_BOX:
		{
			std::vector<float2> a(span * 3);
			RN_CUDA(cudaMemcpy(a.data(), s->baabb + (size_t)slo * 3, sizeof(float2) * span * 3, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < count; i++)
				memcpy(out + (size_t)i * 6, &a[(s->slot_of_rank[first + i] - slo) * 3], 24);
			break;
		}
		case RN_DL_WORLD_MATRIX:
		case RN_DL_INVERSE_WORLD_MATRIX:
		{
			std::vector<float4> a(span * 4);
			RN_CUDA(cudaMemcpy(a.data(), (what == RN_DL_WORLD_MATRIX ? s->wmat : s->winv) + (size_t)slo * 4, sizeof(float4) * span * 4, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < count; i++)
			{
				const size_t k = s->slot_of_rank[first + i] - slo;
				memcpy(out + (size_t)i * 16, &a[k * 4], 64);
			}
			break;
		}
		default:
			RN_REQUIRE(false, RN_ERR_INVALID, "unknown download selector %d", (int)what);
	}
	return RN_OK;
}

extern "C" int rn_scene_device_outputs(rn_scene_t *s, const float **world_matrix, const float **world_pos_sx, const float **world_rot,
                                       const float **world_scale_yz)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	if(world_matrix) *world_matrix = reinterpret_cast<const float *>(s->wmat);
	if(world_pos_sx) *world_pos_sx = reinterpret_cast<const float *>(s->wpos);
	if(world_rot) *world_rot = reinterpret_cast<const float *>(s->wrot);
	if(world_scale_yz) *world_scale_yz = reinterpret_cast<const float *>(s->wsyz);
	return RN_OK;
}

extern "C" int rn_scene_get_stats(const rn_scene_t *s, rn_scene_stats *out)
{
	RN_REQUIRE(s && out, RN_ERR_INVALID, "NULL argument");
	out->nodes = s->n;
	out->tiles = s->n_tiles;
	out->spine_nodes = s->spine_nodes;
	out->spine_levels = s->spine_level_ofs.empty() ? 0 : (uint32_t)s->spine_level_ofs.size() - 1;
	out->max_depth = s->max_depth;
	out->tile_nodes = s->tile_cap;
	out->device_bytes = s->device_bytes;
	out->kernel_launches = s->launches;
	return RN_OK;
}

extern "C" int rn_scene_profile(rn_scene_t *s, int enable)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	s->profile = enable != 0;
	s->prof_used = 0;
	return RN_OK;
}

extern "C" int rn_scene_profile_read(rn_scene_t *s, float *tile_ms_sum, uint32_t *tile_launches)
{
	RN_REQUIRE(s && tile_ms_sum && tile_launches, RN_ERR_INVALID, "NULL argument");
	RN_CUDA(cudaSetDevice(s->device));
	float sum = 0.0f;
	for(int i = 0; i < s->prof_used; i++)
	{
		RN_CUDA(cudaEventSynchronize(s->prof_ev[i][1]));
		float ms = 0.0f;
		RN_CUDA(cudaEventElapsedTime(&ms, s->prof_ev[i][0], s->prof_ev[i][1]));
		sum += ms;
	}
	*tile_ms_sum = sum;
	*tile_launches = (uint32_t)s->prof_used;
	s->prof_used = 0;
	return RN_OK;
}
