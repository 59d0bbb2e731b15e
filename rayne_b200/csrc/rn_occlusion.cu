// librn_cuda - software occlusion culling of SceneBasic::Render on the device (sm_100a).
//
// Reference behaviour being replaced (Source/Scene/RNSceneBasic.cpp):
//   :445-457  occluders of a camera = nodes with Flags::Occluder that pass CanRender
//   :463-483  keep the 30 with the largest radius^2 / max(squared distance, 1), then order them front to back
//   :485-489  clear the 40 x 40 depth buffer (reverse Z: 1 = near, cleared to 0), matViewProj = projection *
//             WithTranslation(jitter) * view, jitter drawn from the shared random generator
//   :498-529  per occluder, in that order: TestBoundingBox against the buffer so far, the occludedFrameCounter
//             hysteresis, RasterizeMesh of LOD 0 / mesh 0 of an Entity that passed
//   :531-573  every node that passes CanRender: priorities >= RenderSky and visible occluders go through, the others
//             take TestBoundingBox + the same hysteresis (submitted while the counter is below 50)
//   :152-208  RasterizeClipSpaceTriangle, :210-318 RasterizeMesh (near-plane clipping), :320-405 TestBoundingBox
//
// Shape on the GPU, per camera and in the reference's camera order (the counters are shared between the cameras of a
// frame, so the cameras cannot run side by side):
//   occ_candidates_kernel  one warp per tile: the occluders of the camera (flag + visibility bit of the cull) with their
//                          score and distance, appended to a candidate array
//   occ_select_kernel      one CTA: 30 rounds of "largest key below the previous one" (key = score bits, then rank: the
//                          order a stable descending sort gives), then the front-to-back insertion sort of those <= 30;
//                          a camera that found occluders takes the next pre-drawn jitter and builds matViewProj
//   occ_raster_kernel      one CTA, depth buffer in shared memory: per occluder the box test (rect cells over all
//                          threads), the counter, then its triangles - one warp per triangle, lanes over the pixels of
//                          the bounding rect, atomicMax on the float bits (all depths written are > 0, so the integer
//                          order is the float order and the result does not depend on the triangle order)
//   occ_test_kernel        one warp per tile: box test + hysteresis of every visible node, clears the camera bit of the
//                          mask byte, recounts the tile totals for the compaction (like the AABB / volume filters)
// followed by one emit_kernel for all cameras.  Every float operation is in the reference's order (this TU is compiled
// with -fmad=false like the rest); tests/golden/occlusion_basic.npz holds the reference's own lists, counters, depth
// buffers and jitter for 60 frames, and tests/test_gpu_occlusion.py compares this file's output with them bit for bit.
#include "rn_internal.h"

#include <cstring>
#include <vector>

namespace rn {

constexpr int kOccW = 40, kOccH = 40, kOccCells = kOccW * kOccH;
constexpr uint32_t kOccMax = 30;          // occluders kept per camera (:470)
constexpr uint32_t kOccHold = 50;         // frames a node stays submitted after it was last seen (:506,563)
constexpr uint32_t kOccCounterMax = 1000; // (:502,559)

struct OccCand
{
	float score, dist;
	uint32_t rank, slot;
};

// device-resident state of one camera's pass
struct OccSel
{
	uint32_t n_candidates;
	uint32_t n_occ;                 // <= 30; 0 = the camera takes the plain path (no occluder passed CanRender)
	uint32_t slot[kOccMax];
	uint32_t visible[kOccMax];      // "visibleOccluders" membership (:506-514)
	uint32_t pad[2];
	float vp[16];                   // projection * WithTranslation(jitter) * view of this pass
	float offset[4];                // the jitter it took (xyz)
	uint32_t pad2[12];
};
static_assert(sizeof(OccSel) == 384, "OccSel layout");

struct OccView
{
	const float4 *c2;
	const float4 *wpos, *wmat;
	const float4 *bsphere;
	const float2 *baabb;
	uint8_t *vis;
	const uint8_t *prio_group;
	uint32_t first_sky_group;
	const TileMeta *tiles;
	uint32_t n_tiles;
	uint32_t *tile_counts, *super_totals;
	uint16_t *counter;              // SceneBasicInfo::occludedFrameCounter per slot
	const int32_t *mesh_of_slot;    // may be nullptr (no Entity occluder in the scene)
	const uint32_t *tri_ofs;
	const float *tris;
	uint32_t n_meshes;
	OccCand *cand;
	uint32_t cand_cap;
	OccSel *sel;
	float *depth;                   // [kOccCells] of this camera
	uint32_t camera;
	float view[16], projection[16];
	float jitter[RN_MAX_CAMERAS * 3];   // the frame's pre-drawn camera offsets, in draw order
	uint32_t *draws;                    // how many of them the cameras before this one took
	float pixel[2];
	float cam[3];
};

// a * b, both column-major 4 x 4, summed left to right (Matrix::operator*, Source/Math/RNMatrix.h:90-142)
__device__ void occ_mat_mul(const float *a, const float *b, float *o)
{
	for(int c = 0; c < 4; c++)
		for(int r = 0; r < 4; r++)
			o[c * 4 + r] = a[r] * b[c * 4] + a[4 + r] * b[c * 4 + 1] + a[8 + r] * b[c * 4 + 2] + a[12 + r] * b[c * 4 + 3];
}

__device__ __forceinline__ float smax(float a, float b) { return (a < b) ? b : a; }   // std::max
__device__ __forceinline__ float smin(float a, float b) { return (b < a) ? b : a; }   // std::min

struct V4 { float x, y, z, w; };

__device__ __forceinline__ V4 mat_vec(const float *m, const V4 &v)
{
	V4 r;
	r.x = m[0] * v.x + m[4] * v.y + m[8] * v.z + m[12] * v.w;
	r.y = m[1] * v.x + m[5] * v.y + m[9] * v.z + m[13] * v.w;
	r.z = m[2] * v.x + m[6] * v.y + m[10] * v.z + m[14] * v.w;
	r.w = m[3] * v.x + m[7] * v.y + m[11] * v.z + m[15] * v.w;
	return r;
}

// ---- candidates ---------------------------------------------------------------------------------------------------
__global__ void occ_reset_kernel(OccSel *sel)
{
	if(threadIdx.x == 0) { sel->n_candidates = 0; sel->n_occ = 0; }
}

__global__ void __launch_bounds__(256) occ_candidates_kernel(const __grid_constant__ OccView v)
{
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(tile >= v.n_tiles)
		return;
	const uint32_t start = v.tiles[tile].start, count = v.tiles[tile].count;
	for(uint32_t j0 = 0; j0 < count; j0 += 32)
	{
		const uint32_t j = j0 + lane;
		if(j >= count)
			continue;
		const uint32_t slot = start + j;
		const uint32_t fw = __float_as_uint(v.c2[slot].w);
		if(!(fw & RN_NODE_OCCLUDER))
			continue;
		const uint32_t rank = start + ((fw >> kRankShift) & kRankMask);
		if(!((v.vis[rank] >> v.camera) & 1u))
			continue;
		// a->GetWorldPosition().GetSquaredDistance(cameraWorldPosition); radius^2 / max(distance, 1)   (:466-470)
		const float4 p = v.wpos[slot];
		const float dx = p.x - v.cam[0], dy = p.y - v.cam[1], dz = p.z - v.cam[2];
		const float d2 = dx * dx + dy * dy + dz * dz;
		const float r = v.bsphere[slot].w;
		OccCand c;
		c.dist = d2;
		c.score = r * r / smax(d2, 1.0f);
		c.rank = rank;
		c.slot = slot;
		const uint32_t at = atomicAdd(&v.sel->n_candidates, 1u);
		if(at < v.cand_cap)
			v.cand[at] = c;
	}
}

// key of the stable descending sort by score: larger score first, then smaller rank
__device__ __forceinline__ unsigned long long occ_key(const OccCand &c)
{
	return ((unsigned long long)__float_as_uint(c.score) << 32) | (unsigned long long)(0xffffffffu - c.rank);
}

__global__ void __launch_bounds__(1024) occ_select_kernel(const __grid_constant__ OccView v)
{
	__shared__ unsigned long long s_best[32];
	__shared__ uint32_t s_idx[32];
	__shared__ OccCand s_pick[kOccMax];
	__shared__ unsigned long long s_last;
	__shared__ uint32_t s_n;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	const uint32_t m = min(v.sel->n_candidates, v.cand_cap);
	if(tid == 0) { s_last = ~0ull; s_n = 0; }
	__syncthreads();
	const uint32_t rounds = min(m, kOccMax);
	for(uint32_t r = 0; r < rounds; r++)
	{
		const unsigned long long below = s_last;
		unsigned long long best = 0;
		uint32_t idx = 0xffffffffu;
		for(uint32_t i = tid; i < m; i += blockDim.x)
		{
			const unsigned long long k = occ_key(v.cand[i]);
			if(k < below && (idx == 0xffffffffu || k > best)) { best = k; idx = i; }
		}
#pragma unroll
		for(int o = 16; o > 0; o >>= 1)
		{
			const unsigned long long ob = __shfl_xor_sync(0xffffffffu, best, o);
			const uint32_t oi = __shfl_xor_sync(0xffffffffu, idx, o);
			if(oi != 0xffffffffu && (idx == 0xffffffffu || ob > best)) { best = ob; idx = oi; }
		}
		if(lane == 0) { s_best[warp] = best; s_idx[warp] = idx; }
		__syncthreads();
		if(warp == 0)
		{
			best = s_best[lane]; idx = s_idx[lane];
#pragma unroll
			for(int o = 16; o > 0; o >>= 1)
			{
				const unsigned long long ob = __shfl_xor_sync(0xffffffffu, best, o);
				const uint32_t oi = __shfl_xor_sync(0xffffffffu, idx, o);
				if(oi != 0xffffffffu && (idx == 0xffffffffu || ob > best)) { best = ob; idx = oi; }
			}
			if(lane == 0 && idx != 0xffffffffu)
			{
				s_pick[s_n] = v.cand[idx];
				s_n = s_n + 1;
				s_last = best;
			}
		}
		__syncthreads();
	}
	if(tid == 0)
	{
		// "Sort remaining occluders front to back" (:473-477): ascending distance, ties keep the order of the first sort
		const uint32_t n = s_n;
		for(uint32_t i = 1; i < n; i++)
		{
			const OccCand c = s_pick[i];
			uint32_t j = i;
			while(j > 0 && s_pick[j - 1].dist > c.dist) { s_pick[j] = s_pick[j - 1]; j--; }
			s_pick[j] = c;
		}
		for(uint32_t i = 0; i < n; i++) { v.sel->slot[i] = s_pick[i].slot; v.sel->visible[i] = 0; }
		v.sel->n_occ = n;
		if(n)
		{
			// this camera takes the occlusion path: it consumes the next jitter draw (:488) and builds
			// matViewProj = projection * Matrix::WithTranslation(randomCameraOffset) * view   (:489)
			const uint32_t k = min(atomicAdd(v.draws, 1u), (uint32_t)RN_MAX_CAMERAS - 1u);
			const float ox = v.jitter[k * 3], oy = v.jitter[k * 3 + 1], oz = v.jitter[k * 3 + 2];
			const float tr[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, ox, oy, oz, 1};
			float pt[16], vp[16];
			occ_mat_mul(v.projection, tr, pt);
			occ_mat_mul(pt, v.view, vp);
			for(int i = 0; i < 16; i++) v.sel->vp[i] = vp[i];
			v.sel->offset[0] = ox; v.sel->offset[1] = oy; v.sel->offset[2] = oz;
		}
	}
}

// ---- TestBoundingBox (:320-405) -------------------------------------------------------------------------------------
// Projects the eight corners; returns 1 = visible for sure (a corner in front of the near plane), 0 = culled for sure
// (smaller than a pixel), 2 = compare maxz with the cells of the rect [x0, x1] x [y0, y1].
struct OccRect
{
	uint32_t x0, x1, y0, y1;
	float maxz;
};

__device__ __forceinline__ int occ_project_box(const float *vp, const float4 &wp, const float2 &b0, const float2 &b1, const float2 &b2,
                                               const float *pixel, OccRect &rc)
{
	const float mn[3] = {b0.x, b0.y, b1.x}, mx[3] = {b1.y, b2.x, b2.y};
	float maxx = 0, maxy = 0, maxz = 0, minx = 0, miny = 0;
#pragma unroll
	for(int i = 0; i < 8; i++)
	{
		// corner order of the reference (:326-333): max max max, max max min, max min min, max min max, then the same with min x
		const float ex = (i < 4) ? mx[0] : mn[0];
		const float ey = ((i & 3) < 2) ? mx[1] : mn[1];
		const float ez = ((i & 3) == 0 || (i & 3) == 3) ? mx[2] : mn[2];
		V4 c = {wp.x + ex, wp.y + ey, wp.z + ez, 1.0f + 0.0f};
		c = mat_vec(vp, c);
		if(c.z > c.w)
			return 1;
		const float w = c.w;
		c.x /= w; c.y /= w; c.z /= w;
		c.x = c.x * 0.5f + 0.5f;
		c.y = c.y * 0.5f + 0.5f;
		if(i == 0)
		{
			maxx = minx = c.x; maxy = miny = c.y; maxz = c.z;
		}
		else
		{
			maxx = smax(maxx, c.x); maxy = smax(maxy, c.y); maxz = smax(maxz, c.z);
			minx = smin(minx, c.x); miny = smin(miny, c.y);
		}
	}
	if((maxx - minx) < pixel[0] && (maxy - miny) < pixel[1])
		return 0;
	minx *= (float)kOccW; miny *= (float)kOccH;
	maxx *= (float)kOccW; maxy *= (float)kOccH;
	rc.x0 = (uint32_t)(uint16_t)smax(smin(floorf(minx) - 1.0f, (float)(kOccW - 1)), 0.0f);
	rc.x1 = (uint32_t)(uint16_t)smax(smin(ceilf(maxx) + 1.0f, (float)(kOccW - 1)), 0.0f);
	rc.y0 = (uint32_t)(uint16_t)smax(smin(floorf(miny) - 1.0f, (float)(kOccH - 1)), 0.0f);
	rc.y1 = (uint32_t)(uint16_t)smax(smin(ceilf(maxy) + 1.0f, (float)(kOccH - 1)), 0.0f);
	rc.maxz = maxz;
	return 2;
}

__device__ __forceinline__ bool occ_cell_passes(const float *depth, uint32_t x, uint32_t y, float maxz)
{
	return maxz > depth[kOccW * (kOccH - (int)y - 1) + (int)x];
}

// ---- rasteriser (:152-318) ------------------------------------------------------------------------------------------
__device__ __forceinline__ float occ_edge(float ax, float ay, float bx, float by, float cx, float cy)
{
	return (cx - ax) * (by - ay) - (cy - ay) * (bx - ax);   // edgeFunction, :27-30
}

__device__ __forceinline__ V4 v4_add(const V4 &a, const V4 &b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
__device__ __forceinline__ V4 v4_sub(const V4 &a, const V4 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w}; }
__device__ __forceinline__ V4 v4_scale(const V4 &a, float n) { return {a.x * n, a.y * n, a.z * n, a.w * n}; }

// RasterizeClipSpaceTriangle: the whole warp works on one triangle, lanes take the pixels of its bounding rect.
__device__ void occ_raster_triangle(int *depth_bits, V4 A, V4 B, V4 C, uint32_t lane)
{
	V4 *pts[3] = {&A, &B, &C};
#pragma unroll
	for(int k = 0; k < 3; k++)
	{
		V4 &p = *pts[k];
		const float w = p.w;
		p.x /= w; p.y /= w; p.z /= w; p.w /= w;
		p.x = p.x * 0.5f + 0.5f;
		p.x *= (float)kOccW;
		p.y = p.y * 0.5f + 0.5f;
		p.y *= (float)kOccH;
	}
	const float area = occ_edge(A.x, A.y, B.x, B.y, C.x, C.y);
	if(area < 0)
		return;
	const uint32_t minX = (uint32_t)(uint16_t)smin(smax(smin(A.x, smin(B.x, C.x)), 0.0f), (float)(kOccW - 1));
	const uint32_t minY = (uint32_t)(uint16_t)smin(smax(smin(A.y, smin(B.y, C.y)), 0.0f), (float)(kOccH - 1));
	const uint32_t maxX = (uint32_t)(uint16_t)smin(smax(smax(A.x, smax(B.x, C.x)), 0.0f), (float)(kOccW - 1));
	const uint32_t maxY = (uint32_t)(uint16_t)smin(smax(smax(A.y, smax(B.y, C.y)), 0.0f), (float)(kOccH - 1));
	if(maxX < minX || maxY < minY)
		return;
	const uint32_t w = maxX - minX + 1, cells = w * (maxY - minY + 1);
	for(uint32_t q = lane; q < cells; q += 32)
	{
		const uint32_t x = minX + q % w, y = minY + q / w;
		const float px = (float)x, py = (float)y;
		float w0 = occ_edge(B.x, B.y, C.x, C.y, px, py);
		float w1 = occ_edge(C.x, C.y, A.x, A.y, px, py);
		float w2 = occ_edge(A.x, A.y, B.x, B.y, px, py);
		if(w0 >= 0 && w1 >= 0 && w2 >= 0)
		{
			w0 /= area; w1 /= area; w2 /= area;
			const float depth = w0 * A.z + w1 * B.z + w2 * C.z - 0.000001f;
			// "if(depth <= 1) if(depth > cell) cell = depth": cells start at 0 and only grow, so only depths > 0 are ever
			// stored and for those the integer order of the bits is the float order
			if(depth <= 1.0f && depth > 0.0f)
				atomicMax(&depth_bits[kOccW * (kOccH - (int)y - 1) + (int)x], __float_as_int(depth));
		}
	}
}

// RasterizeMesh for one triangle (posA, posB, posC as the mesh iterator yields them): transform, trivial rejects,
// near-plane clipping into one or two triangles.
__device__ void occ_raster_mesh_triangle(int *depth_bits, const float *mvp, const float *t, uint32_t lane)
{
	const V4 pa = {t[0], t[1], t[2], 1.0f}, pb = {t[3], t[4], t[5], 1.0f}, pc = {t[6], t[7], t[8], 1.0f};
	V4 A = mat_vec(mvp, pb);     // the reference swaps A and B (:227-229)
	V4 B = mat_vec(mvp, pa);
	V4 C = mat_vec(mvp, pc);
	if(A.z < -A.w && B.z < -B.w && C.z < -C.w) return;
	if(A.z > A.w && B.z > B.w && C.z > C.w) return;
	const float dA = (A.z - (A.w)), dB = (B.z - (B.w)), dC = (C.z - (C.w));
	if(A.z > A.w && B.z > B.w)
	{
		const float tA = dA / (dA - dC), tB = dB / (dB - dC);
		A = v4_add(A, v4_scale(v4_sub(C, A), tA));
		B = v4_add(B, v4_scale(v4_sub(C, B), tB));
	}
	else if(A.z > A.w && C.z > C.w)
	{
		const float tA = dA / (dA - dB), tC = dC / (dC - dB);
		A = v4_add(A, v4_scale(v4_sub(B, A), tA));
		C = v4_add(C, v4_scale(v4_sub(B, C), tC));
	}
	else if(B.z > B.w && C.z > C.w)
	{
		const float tB = dB / (dB - dA), tC = dC / (dC - dA);
		B = v4_add(B, v4_scale(v4_sub(A, B), tB));
		C = v4_add(C, v4_scale(v4_sub(A, C), tC));
	}
	else if(A.z > A.w)
	{
		const float tB = dA / (dA - dB), tC = dA / (dA - dC);
		const V4 AB = v4_add(A, v4_scale(v4_sub(B, A), tB));
		const V4 AC = v4_add(A, v4_scale(v4_sub(C, A), tC));
		A = AB;
		occ_raster_triangle(depth_bits, C, AC, AB, lane);
	}
	else if(B.z > B.w)
	{
		const float tA = dB / (dB - dA), tC = dB / (dB - dC);
		const V4 BA = v4_add(B, v4_scale(v4_sub(A, B), tA));
		const V4 BC = v4_add(B, v4_scale(v4_sub(C, B), tC));
		B = BA;
		occ_raster_triangle(depth_bits, C, BA, BC, lane);
	}
	else if(C.z > C.w)
	{
		const float tA = dC / (dC - dA), tB = dC / (dC - dB);
		const V4 CA = v4_add(C, v4_scale(v4_sub(A, C), tA));
		const V4 CB = v4_add(C, v4_scale(v4_sub(B, C), tB));
		C = CA;
		occ_raster_triangle(depth_bits, B, CB, CA, lane);
	}
	occ_raster_triangle(depth_bits, A, B, C, lane);
}

constexpr uint32_t kRasterThreads = 512;

__global__ void __launch_bounds__(kRasterThreads) occ_raster_kernel(const __grid_constant__ OccView v)
{
	__shared__ int s_depth[kOccCells];
	__shared__ float s_mvp[16];
	__shared__ float s_vp[16];
	__shared__ OccRect s_rect;
	__shared__ int s_state;      // result of occ_project_box for the current occluder
	__shared__ int s_raster;
	const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	for(uint32_t i = tid; i < (uint32_t)kOccCells; i += blockDim.x)
		s_depth[i] = 0;                                     // std::fill(..., 0.0f)
	const uint32_t n_occ = v.sel->n_occ;
	if(tid < 16)
		s_vp[tid] = v.sel->vp[tid];
	__syncthreads();
	const float *depth_f = reinterpret_cast<const float *>(s_depth);
	for(uint32_t k = 0; k < n_occ; k++)
	{
		const uint32_t slot = v.sel->slot[k];
		if(tid == 0)
		{
			OccRect rc = {0, 0, 0, 0, 0.0f};
			s_state = occ_project_box(s_vp, v.wpos[slot], v.baabb[slot * 3u], v.baabb[slot * 3u + 1], v.baabb[slot * 3u + 2], v.pixel, rc);
			s_rect = rc;
		}
		__syncthreads();
		int res = s_state;
		if(res == 2)
		{
			const OccRect rc = s_rect;
			const uint32_t w = rc.x1 - rc.x0 + 1, cells = (rc.x1 >= rc.x0 && rc.y1 >= rc.y0) ? w * (rc.y1 - rc.y0 + 1) : 0u;
			int any = 0;
			for(uint32_t q = tid; q < cells; q += blockDim.x)
				if(occ_cell_passes(depth_f, rc.x0 + q % w, rc.y0 + q / w, rc.maxz)) any = 1;
			res = __syncthreads_or(any) ? 1 : 0;
		}
		if(tid == 0)
		{
			// the hysteresis of :502-514
			uint32_t cnt = v.counter[slot];
			if(!res && cnt < kOccCounterMax) cnt += 1;
			if(res || cnt < kOccHold)
			{
				if(res) cnt = 0;
				v.sel->visible[k] = 1;
			}
			v.counter[slot] = (uint16_t)cnt;
			const int mesh = (res && v.mesh_of_slot) ? v.mesh_of_slot[slot] : -1;
			s_raster = mesh;
			if(mesh >= 0)
			{
				// matModelViewProj = matViewProj * node->GetWorldTransform(): the full 4 x 4 product (Matrix::operator*)
				float wm[16];
				const float4 c0 = v.wmat[slot * 4u], c1 = v.wmat[slot * 4u + 1], c2 = v.wmat[slot * 4u + 2], c3 = v.wmat[slot * 4u + 3];
				wm[0] = c0.x; wm[1] = c0.y; wm[2] = c0.z; wm[3] = c0.w;
				wm[4] = c1.x; wm[5] = c1.y; wm[6] = c1.z; wm[7] = c1.w;
				wm[8] = c2.x; wm[9] = c2.y; wm[10] = c2.z; wm[11] = c2.w;
				wm[12] = c3.x; wm[13] = c3.y; wm[14] = c3.z; wm[15] = c3.w;
				for(int c = 0; c < 4; c++)
					for(int r = 0; r < 4; r++)
						s_mvp[c * 4 + r] = s_vp[r] * wm[c * 4] + s_vp[4 + r] * wm[c * 4 + 1] + s_vp[8 + r] * wm[c * 4 + 2] + s_vp[12 + r] * wm[c * 4 + 3];
			}
		}
		__syncthreads();
		const int mesh = s_raster;
		if(mesh >= 0 && (uint32_t)mesh < v.n_meshes)
		{
			const uint32_t t0 = v.tri_ofs[mesh], t1 = v.tri_ofs[mesh + 1];
			for(uint32_t t = t0 + warp; t < t1; t += kRasterThreads / 32)
				occ_raster_mesh_triangle(s_depth, s_mvp, v.tris + (size_t)t * 9, lane);
		}
		__syncthreads();
	}
	for(uint32_t i = tid; i < (uint32_t)kOccCells; i += blockDim.x)
		v.depth[i] = depth_f[i];
}

// ---- per-node test (:531-573) ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) occ_test_kernel(const __grid_constant__ OccView v)
{
	__shared__ float s_depth[kOccCells];
	__shared__ float s_vp[16];
	__shared__ uint32_t s_vis_slot[kOccMax];
	__shared__ uint32_t s_nvis;
	const uint32_t lane = threadIdx.x & 31u;
	const uint32_t n_occ = v.sel->n_occ;
	if(n_occ)
	{
		if(threadIdx.x < 16)
			s_vp[threadIdx.x] = v.sel->vp[threadIdx.x];
		for(uint32_t i = threadIdx.x; i < (uint32_t)kOccCells; i += blockDim.x)
			s_depth[i] = v.depth[i];
		if(threadIdx.x == 0)
		{
			uint32_t nv = 0;
			for(uint32_t k = 0; k < n_occ; k++)
				if(v.sel->visible[k]) s_vis_slot[nv++] = v.sel->slot[k];
			s_nvis = nv;
		}
	}
	__syncthreads();
	const uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if(tile >= v.n_tiles)
		return;
	const uint32_t start = v.tiles[tile].start, count = v.tiles[tile].count;
	const uint32_t bit = 1u << v.camera;
	uint32_t total[RN_MAX_CAMERAS];
#pragma unroll
	for(int c = 0; c < RN_MAX_CAMERAS; c++) total[c] = 0;
	for(uint32_t j0 = 0; j0 < count; j0 += 32)
	{
		const bool active = j0 + lane < count;
		const uint32_t slot = start + j0 + lane;
		const uint32_t fw = active ? __float_as_uint(v.c2[slot].w) : 0u;
		const uint32_t vidx = start + ((fw >> kRankShift) & kRankMask);
		uint32_t mask = active ? v.vis[vidx] : 0u;
		if(n_occ && (mask & bit))
		{
			bool tested = true;
			if(v.prio_group && (uint32_t)v.prio_group[slot] >= v.first_sky_group)
				tested = false;                              // RenderSky and above are never tested (:537-549)
			if(tested && (fw & RN_NODE_OCCLUDER))
			{
				const uint32_t nv = s_nvis;
				for(uint32_t k = 0; k < nv; k++)
					if(s_vis_slot[k] == slot) tested = false;    // a visible occluder (:552-556)
			}
			if(tested)
			{
				OccRect rc = {0, 0, 0, 0, 0.0f};
				int res = occ_project_box(s_vp, v.wpos[slot], v.baabb[slot * 3u], v.baabb[slot * 3u + 1], v.baabb[slot * 3u + 2], v.pixel, rc);
				if(res == 2)
				{
					res = 0;
					for(uint32_t y = rc.y0; y <= rc.y1 && !res; y++)
						for(uint32_t x = rc.x0; x <= rc.x1; x++)
							if(occ_cell_passes(s_depth, x, y, rc.maxz)) { res = 1; break; }
				}
				uint32_t cnt = v.counter[slot];
				if(!res && cnt < kOccCounterMax) cnt += 1;
				if(res || cnt < kOccHold)
				{
					if(res) cnt = 0;
				}
				else
				{
					mask &= ~bit;
					v.vis[vidx] = (uint8_t)mask;
				}
				v.counter[slot] = (uint16_t)cnt;
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

void occlusion_free(rn_scene *s)
{
	void *ptrs[] = {s->occ_counter, s->occ_mesh_of_slot, s->occ_tri_ofs, s->occ_tris, s->occ_cand, s->occ_sel, s->occ_depth};
	for(void *p : ptrs)
		if(p) cudaFree(p);
	s->occ_counter = nullptr; s->occ_mesh_of_slot = nullptr; s->occ_tri_ofs = nullptr; s->occ_tris = nullptr;
	s->occ_cand = nullptr; s->occ_sel = nullptr; s->occ_depth = nullptr;
	s->occ_meshes = 0;
}

static int ensure_counters(rn_scene *s)
{
	if(!s->occ_counter)
	{
		int rc = dev_alloc(s, &s->occ_counter, (size_t)s->desc.max_nodes);
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->occ_counter, 0, sizeof(uint16_t) * s->desc.max_nodes));
	}
	return RN_OK;
}

} // namespace rn

using namespace rn;

extern "C" int rn_occlusion_set_meshes(rn_scene_t *s, const uint32_t *triangle_offsets, const float *triangles, uint32_t n_meshes)
{
	RN_REQUIRE(s && (n_meshes == 0 || (triangle_offsets && triangles)), RN_ERR_INVALID, "NULL argument");
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	if(s->occ_tri_ofs) { cudaFree(s->occ_tri_ofs); s->occ_tri_ofs = nullptr; }
	if(s->occ_tris) { cudaFree(s->occ_tris); s->occ_tris = nullptr; }
	s->occ_meshes = 0;
	if(n_meshes == 0)
		return RN_OK;
	RN_REQUIRE(triangle_offsets[0] == 0, RN_ERR_INVALID, "triangle_offsets[0] must be 0");
	for(uint32_t m = 0; m < n_meshes; m++)
		RN_REQUIRE(triangle_offsets[m + 1] >= triangle_offsets[m], RN_ERR_INVALID, "triangle_offsets must ascend");
	const size_t nt = triangle_offsets[n_meshes];
	int rc = dev_alloc(s, &s->occ_tri_ofs, (size_t)n_meshes + 1);
	if(!rc) rc = dev_alloc(s, &s->occ_tris, nt * 9);
	if(rc) return rc;
	RN_CUDA(cudaMemcpy(s->occ_tri_ofs, triangle_offsets, sizeof(uint32_t) * (n_meshes + 1), cudaMemcpyHostToDevice));
	if(nt) RN_CUDA(cudaMemcpy(s->occ_tris, triangles, sizeof(float) * nt * 9, cudaMemcpyHostToDevice));
	s->occ_meshes = n_meshes;
	return RN_OK;
}

extern "C" int rn_occlusion_set_node_meshes(rn_scene_t *s, const int32_t *mesh_of_rank, uint32_t first, uint32_t count)
{
	RN_REQUIRE(s && mesh_of_rank, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "range [%u, %u) exceeds the %u uploaded nodes", first, first + count, s->n);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	if(!s->occ_mesh_of_slot)
	{
		int rc = dev_alloc(s, &s->occ_mesh_of_slot, (size_t)s->desc.max_nodes);
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->occ_mesh_of_slot, 0xff, sizeof(int32_t) * s->desc.max_nodes));   // -1: not an Entity
	}
	std::vector<int32_t> host(s->n);
	RN_CUDA(cudaMemcpy(host.data(), s->occ_mesh_of_slot, sizeof(int32_t) * s->n, cudaMemcpyDeviceToHost));
	for(uint32_t i = 0; i < count; i++)
		host[s->slot_of_rank[first + i]] = mesh_of_rank[i];
	RN_CUDA(cudaMemcpy(s->occ_mesh_of_slot, host.data(), sizeof(int32_t) * s->n, cudaMemcpyHostToDevice));
	return RN_OK;
}

extern "C" int rn_occlusion_reset_counters(rn_scene_t *s)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_CUDA(cudaSetDevice(s->device));
	int rc = ensure_counters(s);
	if(rc) return rc;
	RN_CUDA(cudaDeviceSynchronize());
	RN_CUDA(cudaMemset(s->occ_counter, 0, sizeof(uint16_t) * s->desc.max_nodes));
	return RN_OK;
}

extern "C" int rn_cull_occlusion(rn_scene_t *s, const rn_occlusion_camera *cameras, const uint32_t *camera_order, uint32_t n_cameras,
                                 const float *jitter, void *cuda_stream)
{
	RN_RANGE("rn_cull_occlusion");
	RN_REQUIRE(s && cameras && jitter, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(s->last_ncams > 0, RN_ERR_STATE, "rn_cull_occlusion needs the masks of a cull (rn_transform_cull / rn_cull)");
	RN_REQUIRE(n_cameras == s->last_ncams, RN_ERR_INVALID, "the last cull had %u cameras, %u given", s->last_ncams, n_cameras);
	RN_REQUIRE(s->have_world, RN_ERR_STATE, "no world transforms");
	RN_REQUIRE(s->have_bounds, RN_ERR_STATE, "rn_cull_occlusion reads the transformed bounds: call rn_bounds_update after the propagate");
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	int rc = ensure_counters(s);
	if(rc) return rc;
	if(!s->occ_sel)
	{
		rc = dev_alloc(s, &s->occ_sel, (size_t)RN_MAX_CAMERAS * sizeof(OccSel) / sizeof(uint32_t) + 1);   // + the draw counter
		if(!rc) rc = dev_alloc(s, &s->occ_depth, (size_t)RN_MAX_CAMERAS * kOccCells);
		if(!rc) rc = dev_alloc(s, &s->occ_cand, (size_t)s->desc.max_nodes * sizeof(OccCand) / sizeof(uint32_t));
		if(rc) return rc;
		RN_CUDA(cudaMemset(s->occ_sel, 0, RN_MAX_CAMERAS * sizeof(OccSel)));
		RN_CUDA(cudaMemset(s->occ_depth, 0, sizeof(float) * RN_MAX_CAMERAS * kOccCells));
	}
	bool seen[RN_MAX_CAMERAS] = {};
	const uint32_t grid_tiles = (s->n_tiles + 7) / 8;
	uint32_t *draws = s->occ_sel + (size_t)RN_MAX_CAMERAS * sizeof(OccSel) / sizeof(uint32_t);
	RN_CUDA(cudaMemsetAsync(draws, 0, sizeof(uint32_t), stream));
	for(uint32_t i = 0; i < n_cameras; i++)
	{
		const uint32_t c = camera_order ? camera_order[i] : i;
		RN_REQUIRE(c < n_cameras && !seen[c], RN_ERR_INVALID, "camera_order must be a permutation of 0..%u", n_cameras - 1);
		seen[c] = true;
		OccSel *sel = reinterpret_cast<OccSel *>(s->occ_sel) + c;
		if(!cameras[c].enabled)
		{
			occ_reset_kernel<<<1, 32, 0, stream>>>(sel);    // "used" reads as 0
			s->launches++;
			continue;
		}
		OccView v;
		memset(&v, 0, sizeof(v));
		v.c2 = s->c2; v.wpos = s->wpos; v.wmat = s->wmat; v.bsphere = s->bsphere; v.baabb = s->baabb;
		v.vis = s->vis; v.prio_group = s->prio_group; v.first_sky_group = s->first_sky_group;
		v.tiles = s->tiles; v.n_tiles = s->n_tiles; v.tile_counts = s->tile_counts; v.super_totals = s->super_totals;
		v.counter = s->occ_counter; v.mesh_of_slot = s->occ_meshes ? s->occ_mesh_of_slot : nullptr;
		v.tri_ofs = s->occ_tri_ofs; v.tris = s->occ_tris; v.n_meshes = s->occ_meshes;
		v.cand = reinterpret_cast<OccCand *>(s->occ_cand); v.cand_cap = s->desc.max_nodes;
		v.sel = sel;
		v.depth = s->occ_depth + (size_t)c * kOccCells;
		v.camera = c;
		memcpy(v.view, cameras[c].view, sizeof(v.view));
		memcpy(v.projection, cameras[c].projection, sizeof(v.projection));
		memcpy(v.jitter, jitter, sizeof(float) * 3 * n_cameras);
		v.draws = draws;
		v.pixel[0] = cameras[c].pixel_size[0]; v.pixel[1] = cameras[c].pixel_size[1];
		memcpy(v.cam, cameras[c].world_position, sizeof(v.cam));
		occ_reset_kernel<<<1, 32, 0, stream>>>(sel);
		occ_candidates_kernel<<<grid_tiles, 256, 0, stream>>>(v);
		occ_select_kernel<<<1, 1024, 0, stream>>>(v);
		occ_raster_kernel<<<1, kRasterThreads, 0, stream>>>(v);
		rc = clear_super_totals(s, stream);
		if(rc) return rc;
		occ_test_kernel<<<grid_tiles, 256, 0, stream>>>(v);
		s->launches += 5;
		RN_CUDA(cudaGetLastError());
	}
	rc = launch_compaction(s, n_cameras, stream);
	if(rc) return rc;
	s->have_masks = false;
	return RN_OK;
}

extern "C" int rn_occlusion_download(rn_scene_t *s, uint32_t camera, uint32_t *used, float *jitter_taken, uint32_t *draws_consumed,
                                     float *depth_buffer, uint32_t *counters_by_rank, uint32_t first, uint32_t count)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	RN_REQUIRE(camera < RN_MAX_CAMERAS, RN_ERR_INVALID, "camera %u", camera);
	RN_REQUIRE(s->occ_sel && s->occ_counter, RN_ERR_STATE, "rn_cull_occlusion has not run");
	RN_REQUIRE((uint64_t)first + count <= s->n, RN_ERR_INVALID, "range [%u, %u) exceeds the %u uploaded nodes", first, first + count, s->n);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	if(used || jitter_taken)
	{
		OccSel sel;
		RN_CUDA(cudaMemcpy(&sel, reinterpret_cast<OccSel *>(s->occ_sel) + camera, sizeof(sel), cudaMemcpyDeviceToHost));
		if(used) *used = sel.n_occ;
		if(jitter_taken) memcpy(jitter_taken, sel.offset, sizeof(float) * 3);
	}
	if(draws_consumed)
		RN_CUDA(cudaMemcpy(draws_consumed, s->occ_sel + (size_t)RN_MAX_CAMERAS * sizeof(OccSel) / sizeof(uint32_t), sizeof(uint32_t),
		                   cudaMemcpyDeviceToHost));
	if(depth_buffer)
		RN_CUDA(cudaMemcpy(depth_buffer, s->occ_depth + (size_t)camera * kOccCells, sizeof(float) * kOccCells, cudaMemcpyDeviceToHost));
	if(counters_by_rank && count)
	{
		std::vector<uint16_t> host(s->n);
		RN_CUDA(cudaMemcpy(host.data(), s->occ_counter, sizeof(uint16_t) * s->n, cudaMemcpyDeviceToHost));
		for(uint32_t i = 0; i < count; i++)
			counters_by_rank[i] = host[s->slot_of_rank[first + i]];
	}
	return RN_OK;
}
