// librn_cuda - particle emitters (sm_100a).
//
// Reference behaviour being replaced (file:line relative to the Rayne tree):
//   GenericParticle::Update                    Source/Scene/RNParticle.cpp:31-53     age, interpolate, integrate
//   Interpolator<T>::GetValue (Linear)         Source/Math/RNInterpolation.h:101-103 start + (difference * t / duration)
//   ParticleEmitter::UpdateParticles           Source/Scene/RNParticleEmitter.cpp:180-211 ordered kill, spawn count
//   GenericParticleEmitter::CreateParticle     Source/Scene/RNParticleEmitter.cpp:416-454 11 RNG draws per particle; both
//                                              branches of GetIsLocal() (:428-437)
//   ParticleEmitter::SetMaxParticlesSoft       Source/Scene/RNParticleEmitter.cpp:128-132,204-205 (unsigned wrap kept)
//   Random::MersenneTwister                    Source/Math/RNRandom.cpp:313-358      _offset = 0 after Seed, 31-bit outputs
//   Generator::GetRandomFloatRange             Source/Math/RNRandom.cpp:213-222      float(double(k) * 2^-31 * range + min)
//
// Layout in HBM: every emitter owns a fixed segment of `cap` particle slots in SoA float4 columns; the live
// particles of an emitter are the first `count` slots of its segment, in the reference's vector order.  A step
// reads generation g and writes the compacted survivors + the newly spawned particles to generation g^1.
// One CTA per emitter.  Compiled with -fmad=false (no contraction, fp32 and fp64).
#include "rn_internal.h"

#include <cfloat>
#include <cstring>
#include <new>

namespace rn {

constexpr uint32_t kMTN = 624;
constexpr uint32_t kMTM = 397;
constexpr uint32_t kDrawsPerParticle = 11;
constexpr uint32_t kParticleThreads = 256;
constexpr uint32_t kSpawnBatch = 256;

struct EmitterState
{
	rn_emitter_desc d;
	float time;        // ParticleEmitter::_time
	uint32_t count;    // _particles.size()
	uint32_t mt_index; // MersenneTwister::_offset
	uint32_t overflow; // sticky: a spawn was cut at the segment capacity (only reachable through the soft-maximum wrap)
};

struct ParticleBuffers
{
	float4 *a;  // position xyz, time
	float4 *b;  // velocity xyz, lifespan
	float4 *c;  // duration, size start, size end, rotation start
	float *d;   // rotation end
	float4 *g;  // GenericParticle::gravity xyz - only allocated when some emitter is non-local (else the emitter's gravity)
};

struct ParticleStore
{
	uint32_t n_emitters = 0;
	uint32_t cap = 0;
	uint32_t gen = 0;
	ParticleBuffers buf[2] = {};
	float4 *color = nullptr;   // derived: colour rgba
	float4 *misc = nullptr;    // derived: rotation, size x, size y, distance key (after sort)
	uint32_t *sorted = nullptr;
	EmitterState *emitters = nullptr;
	uint32_t *mt = nullptr;    // [n_emitters][624]
	// render mesh (rn_particles_build_mesh), allocated on first use: per emitter cap*4 vertices, cap*6 indices
	float *mesh_pos = nullptr;      // 3 floats per vertex
	float4 *mesh_color = nullptr;
	float2 *mesh_uv0 = nullptr, *mesh_uv1 = nullptr;
	uint16_t *mesh_idx = nullptr;
	float *mesh_scale = nullptr;    // per emitter: GetWorldScale().x
	uint32_t *mesh_flags = nullptr; // per emitter: RN_EMITTER_*
	bool have_mesh = false;
	// vertex stage of Shaders/Particles.hlsl (rn_particles_expand_billboards): clip position + colour per vertex
	float4 *clip_pos = nullptr, *clip_color = nullptr;
	float *clip_mats = nullptr;     // per emitter: inverse model-view [16], model-view-projection [16]
	bool have_clip = false;
	size_t sort_smem_configured = 0;
};

struct ParticleView
{
	ParticleBuffers in, out;
	float4 *color, *misc;
	EmitterState *emitters;
	uint32_t *mt;
	uint32_t cap;
	float delta;
};

struct PState
{
	float px, py, pz, time;
	float vx, vy, vz, life;
	float dur, s0, s1, r0, r1;
	float gx, gy, gz;   // GenericParticle::gravity
};

// GenericParticle::Update (RNParticle.cpp:43-53) for one particle; writes the derived fields.
__device__ __forceinline__ void particle_update(PState &p, float delta, const rn_emitter_desc &e, float4 &color, float4 &misc)
{
	p.time += delta;
	p.life -= delta;
	const float t = p.time, dur = p.dur;
	// Interpolator Linear: start + ((difference * time) / duration), difference = end - start cached at spawn
	const float rotation = p.r0 + ((p.r1 - p.r0) * t) / dur;
	color.x = e.start_color[0] + ((e.end_color[0] - e.start_color[0]) * t) / dur;
	color.y = e.start_color[1] + ((e.end_color[1] - e.start_color[1]) * t) / dur;
	color.z = e.start_color[2] + ((e.end_color[2] - e.start_color[2]) * t) / dur;
	color.w = e.start_color[3] + ((e.end_color[3] - e.start_color[3]) * t) / dur;
	const float size = p.s0 + ((p.s1 - p.s0) * t) / dur;
	misc = make_float4(rotation, size, size, 0.0f);
	// velocity += gravity * delta; position += velocity * delta
	p.vx += p.gx * delta; p.vy += p.gy * delta; p.vz += p.gz * delta;
	p.px += p.vx * delta; p.py += p.vy * delta; p.pz += p.vz * delta;
}

__device__ __forceinline__ void store_particle(const ParticleBuffers &o, size_t i, const PState &p)
{
	o.a[i] = make_float4(p.px, p.py, p.pz, p.time);
	o.b[i] = make_float4(p.vx, p.vy, p.vz, p.life);
	o.c[i] = make_float4(p.dur, p.s0, p.s1, p.r0);
	o.d[i] = p.r1;
}

// Generator::GetRandomFloatRange (RNRandom.cpp:213-222): double(k) * 2^-31 * range + min, one rounding to float.
__device__ __forceinline__ float random_range(uint32_t k31, float lo, float hi)
{
	const float range = hi - lo;
	const double u = (double)(int32_t)k31 * (1.0 / 2147483648.0);
	return (float)(u * (double)range + (double)lo);
}

__device__ __forceinline__ float fmin_std(float a, float b) { return (b < a) ? b : a; } // std::min
__device__ __forceinline__ float fmax_std(float a, float b) { return (a < b) ? b : a; } // std::max

// MT19937 block regeneration (RNRandom.cpp:330-347), restructured into three data-parallel phases + the last word.
__device__ void mt_twist(uint32_t *mt, uint32_t tid)
{
	const uint32_t U = 0x80000000u, L = 0x7fffffffu, A = 0x9908b0dfu;
	const uint32_t bounds[4] = {0, kMTN - kMTM, 2 * (kMTN - kMTM), kMTN - 1};
#pragma unroll
	for(int ph = 0; ph < 3; ph++)
	{
		uint32_t v[1];
		const uint32_t i = bounds[ph] + tid;
		const bool on = i < bounds[ph + 1];
		if(on)
		{
			const uint32_t y = (mt[i] & U) | (mt[i + 1] & L);
			const uint32_t src = (i + kMTM) % kMTN;
			v[0] = mt[src] ^ (y >> 1) ^ ((y & 1u) ? A : 0u);
		}
		__syncthreads();
		if(on) mt[i] = v[0];
		__syncthreads();
	}
	if(tid == 0)
	{
		const uint32_t y = (mt[kMTN - 1] & U) | (mt[0] & L);
		mt[kMTN - 1] = mt[kMTM - 1] ^ (y >> 1) ^ ((y & 1u) ? A : 0u);
	}
	__syncthreads();
}

__global__ void __launch_bounds__(kParticleThreads) particles_step_kernel(ParticleView pv)
{
	__shared__ uint32_t s_mt[kMTN];
	__shared__ uint32_t s_draws[kSpawnBatch * kDrawsPerParticle];
	__shared__ uint32_t s_wtot[kParticleThreads / 32];
	__shared__ rn_emitter_desc s_e;
	__shared__ uint32_t s_spawn, s_index;

	const uint32_t e = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
	EmitterState *es = pv.emitters + e;
	if(tid < sizeof(rn_emitter_desc) / 4)
		reinterpret_cast<uint32_t *>(&s_e)[tid] = reinterpret_cast<const uint32_t *>(&es->d)[tid];
	const uint32_t count = es->count;
	const float etime = es->time;
	const uint32_t mt_index = es->mt_index;
	__syncthreads();
	const size_t seg = (size_t)e * pv.cap;
	const float delta = pv.delta;
	// non-local emitters carry a per-particle gravity column (CTA-uniform branch: local emitters never touch it)
	const bool nonlocal = (s_e.flags & RN_EMITTER_DESC_NONLOCAL) != 0u;

	// ---- update + order-preserving kill (RNParticleEmitter.cpp:182-196) ----
	uint32_t out = 0;
	for(uint32_t base = 0; base < count; base += kParticleThreads)
	{
		const uint32_t i = base + tid;
		PState p;
		float4 color, misc;
		bool alive = false;
		if(i < count)
		{
			const float4 a = pv.in.a[seg + i], b = pv.in.b[seg + i], c = pv.in.c[seg + i];
			p.px = a.x; p.py = a.y; p.pz = a.z; p.time = a.w;
			p.vx = b.x; p.vy = b.y; p.vz = b.z; p.life = b.w;
			p.dur = c.x; p.s0 = c.y; p.s1 = c.z; p.r0 = c.w;
			p.r1 = pv.in.d[seg + i];
			p.gx = s_e.gravity[0]; p.gy = s_e.gravity[1]; p.gz = s_e.gravity[2];
			if(nonlocal)
			{
				const float4 g = pv.in.g[seg + i];
				p.gx = g.x; p.gy = g.y; p.gz = g.z;
			}
			particle_update(p, delta, s_e, color, misc);
			alive = !(p.life <= 0.0f);
		}
		const uint32_t m = __ballot_sync(0xffffffffu, alive);
		if(lane == 0) s_wtot[warp] = __popc(m);
		__syncthreads();
		uint32_t pre = 0, tot = 0;
#pragma unroll
		for(uint32_t w = 0; w < kParticleThreads / 32; w++)
		{
			const uint32_t x = s_wtot[w];
			pre += w < warp ? x : 0;
			tot += x;
		}
		if(alive)
		{
			const size_t o = seg + out + pre + __popc(m & ((1u << lane) - 1u));
			store_particle(pv.out, o, p);
			if(nonlocal) pv.out.g[o] = make_float4(p.gx, p.gy, p.gz, 0.0f);
			pv.color[o] = color;
			pv.misc[o] = misc;
		}
		out += tot;
		__syncthreads();
	}

	// ---- spawn count (RNParticleEmitter.cpp:198-210) ----
	if(tid == 0)
	{
		uint32_t spawn = 0;
		float newtime = etime;
		if(!(s_e.spawn_rate < FLT_EPSILON))
		{
			const float t = etime + delta;
			spawn = (uint32_t)floorf(t / s_e.spawn_rate);
			newtime = fmodf(t, s_e.spawn_rate);
			// if(_maxParticles > 0 && _maxParticlesSoft > 0) spawn = min(_maxParticlesSoft - size, spawn)   (:204-205)
			const uint32_t soft = (s_e.flags & RN_EMITTER_DESC_SOFT_MAX) ? s_e.max_particles_soft : s_e.max_particles;
			if(spawn > 0 && s_e.max_particles > 0 && soft > 0)
			{
				const uint32_t room = soft - out; // unsigned, like the reference: wraps when out > soft
				spawn = room < spawn ? room : spawn;
			}
			// the segment holds pv.cap particles; only the wrap above can ask for more
			if(spawn > pv.cap - out)
			{
				spawn = pv.cap - out;
				es->overflow = 1u;
			}
		}
		s_spawn = spawn;
		es->time = newtime;
	}
	__syncthreads();
	const uint32_t spawn = s_spawn;
	if(spawn == 0)
	{
		if(tid == 0) es->count = out;
		return;
	}

	// ---- spawn (SpawnParticles -> CreateParticle, RNParticleEmitter.cpp:156-171,416-454) ----
	uint32_t *gmt = pv.mt + (size_t)e * kMTN;
	for(uint32_t k = tid; k < kMTN; k += kParticleThreads) s_mt[k] = gmt[k];
	if(tid == 0) s_index = mt_index;
	__syncthreads();
	for(uint32_t done = 0; done < spawn; done += kSpawnBatch)
	{
		const uint32_t batch = min(kSpawnBatch, spawn - done);
		const uint32_t need = batch * kDrawsPerParticle;
		uint32_t produced = 0;
		while(produced < need)
		{
			uint32_t index = s_index;
			__syncthreads();
			if(index >= kMTN)
			{
				mt_twist(s_mt, tid);
				index = 0;
			}
			const uint32_t take = min(kMTN - index, need - produced);
			for(uint32_t k = tid; k < take; k += kParticleThreads)
			{
				uint32_t y = s_mt[index + k];
				y ^= (y >> 11);
				y ^= (y << 7) & 0x9d2c5680u;
				y ^= (y << 15) & 0xefc60000u;
				y ^= (y >> 18);
				s_draws[produced + k] = y & 0x7fffffffu;
			}
			produced += take;
			__syncthreads();
			if(tid == 0) s_index = index + take;
			__syncthreads();
		}
		if(tid < batch)
		{
			const uint32_t *dr = s_draws + tid * kDrawsPerParticle;
			const rn_emitter_desc &d = s_e;
			PState p;
			// GetRandomVector3Range swaps min/max per component (RNRandom.cpp:250-252)
			p.px = random_range(dr[0], fmin_std(d.position_min[0], d.position_max[0]), fmax_std(d.position_min[0], d.position_max[0]));
			p.py = random_range(dr[1], fmin_std(d.position_min[1], d.position_max[1]), fmax_std(d.position_min[1], d.position_max[1]));
			p.pz = random_range(dr[2], fmin_std(d.position_min[2], d.position_max[2]), fmax_std(d.position_min[2], d.position_max[2]));
			const float lifespan = random_range(dr[3], d.life_span[0], d.life_span[1]);
			p.life = lifespan;
			p.vx = d.velocity[0] + random_range(dr[4], fmin_std(d.velocity_min[0], d.velocity_max[0]), fmax_std(d.velocity_min[0], d.velocity_max[0]));
			p.vy = d.velocity[1] + random_range(dr[5], fmin_std(d.velocity_min[1], d.velocity_max[1]), fmax_std(d.velocity_min[1], d.velocity_max[1]));
			p.vz = d.velocity[2] + random_range(dr[6], fmin_std(d.velocity_min[2], d.velocity_max[2]), fmax_std(d.velocity_min[2], d.velocity_max[2]));
			p.gx = d.gravity[0]; p.gy = d.gravity[1]; p.gz = d.gravity[2];
			float sizeScale = 1.0f;
			if(nonlocal)
			{
				// RNParticleEmitter.cpp:428-437: sizeScale = max(max(scale.x, scale.y), scale.z); position *= scale;
				// position += GetWorldPosition(); gravity *= scale; velocity *= scale
				const float sx = d.world_scale[0], sy = d.world_scale[1], sz = d.world_scale[2];
				sizeScale = fmax_std(fmax_std(sx, sy), sz);
				p.px *= sx; p.py *= sy; p.pz *= sz;
				p.px += d.world_position[0]; p.py += d.world_position[1]; p.pz += d.world_position[2];
				p.gx *= sx; p.gy *= sy; p.gz *= sz;
				p.vx *= sx; p.vy *= sy; p.vz *= sz;
			}
			p.s0 = random_range(dr[7], d.start_size[0], d.start_size[1]) * sizeScale;
			p.s1 = random_range(dr[8], d.end_size[0], d.end_size[1]) * sizeScale;
			p.r0 = random_range(dr[9], d.start_rotation[0], d.start_rotation[1]) * 0.01745329251994329577f;
			p.r1 = random_range(dr[10], d.end_rotation[0], d.end_rotation[1]) * 0.01745329251994329577f;
			p.dur = lifespan;
			p.time = 0.0f;
			float4 color, misc;
			particle_update(p, 0.0f, d, color, misc); // particle->Update(0.0f), RNParticleEmitter.cpp:451
			const size_t o = seg + out + done + tid;
			store_particle(pv.out, o, p);
			if(nonlocal) pv.out.g[o] = make_float4(p.gx, p.gy, p.gz, 0.0f);
			pv.color[o] = color;
			pv.misc[o] = misc;
		}
		__syncthreads();
	}
	for(uint32_t k = tid; k < kMTN; k += kParticleThreads) gmt[k] = s_mt[k];
	if(tid == 0)
	{
		es->mt_index = s_index;
		es->count = out + spawn;
	}
}

// Extension P7: per-emitter ascending distance sort (stable).  One CTA per emitter, everything in shared memory: an LSD
// radix sort (8-bit digits) of (fp32 distance bits, particle index) - distances are >= 0, so the unsigned order of the
// bits is the float order, and a stable sort leaves equal distances in index order (what the (distance, index) key of a
// comparison sort gives).  Per pass: warp-private digit histograms of the warp's contiguous chunk (shared-memory
// atomics), prefix over digits and warps, then every warp walks its chunk again in order and places its elements
// (match.any rank inside a round of 32 + the warp's running digit offsets).  A pass whose digit is the same for every
// particle of the emitter (typical for the top byte) moves nothing.  4 passes x ~5 barriers instead of the 78 barrier
// stages of the bitonic network this replaces (258 -> 148 us for 1 000 emitters x 4 000).
constexpr uint32_t kPSortThreads = 512;
constexpr uint32_t kPSortWarps = kPSortThreads / 32;

__global__ void __launch_bounds__(kPSortThreads) particles_sort_kernel(ParticleBuffers cur, float4 *misc, uint32_t *sorted,
                                                                     const EmitterState *emitters, uint32_t cap, float cx, float cy, float cz,
                                                                     uint32_t padded)
{
	extern __shared__ __align__(16) unsigned char s_raw[];
	__shared__ uint32_t s_scan[kPSortWarps];
	uint32_t *keys0 = reinterpret_cast<uint32_t *>(s_raw), *keys1 = keys0 + padded;
	uint32_t *wcount = keys1 + padded;                                   // [kPSortWarps][256]
	uint16_t *idx0 = reinterpret_cast<uint16_t *>(wcount + kPSortWarps * 256), *idx1 = idx0 + padded;
	const uint32_t e = blockIdx.x, tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
	const uint32_t n = emitters[e].count;
	const size_t seg = (size_t)e * cap;
	for(uint32_t i = tid; i < n; i += kPSortThreads)
	{
		const float4 a = cur.a[seg + i];
		// Vector3::GetDistance (RNVector.h:537-541): length(position - camera)
		const float dx = a.x - cx, dy = a.y - cy, dz = a.z - cz;
		const float dist = sqrtf(dx * dx + dy * dy + dz * dz);
		reinterpret_cast<float *>(misc + seg + i)[3] = dist;      // misc.w = the distance key; the other three floats stay
		keys0[i] = __float_as_uint(dist);
		idx0[i] = (uint16_t)i;
	}
	// the warp's contiguous chunk, a multiple of 32 long
	const uint32_t chunk = (((n + kPSortWarps - 1) / kPSortWarps) + 31u) & ~31u;
	const uint32_t cbeg = min(n, warp * chunk), cend = min(n, cbeg + chunk);
	uint32_t *ks = keys0, *kd = keys1;
	uint16_t *is = idx0, *id = idx1;
	for(uint32_t shift = 0; shift < 32; shift += 8)
	{
		for(uint32_t i = tid; i < kPSortWarps * 256; i += kPSortThreads) wcount[i] = 0;
		__syncthreads();
		for(uint32_t i = cbeg + lane; i < cend; i += 32)
			atomicAdd(&wcount[warp * 256 + ((ks[i] >> shift) & 255u)], 1u);
		__syncthreads();
		// thread d < 256: exclusive prefix of digit d over the warps, then over the digits
		uint32_t run = 0;
		if(tid < 256)
		{
			for(uint32_t w = 0; w < kPSortWarps; w++)
			{
				const uint32_t t = wcount[w * 256 + tid];
				wcount[w * 256 + tid] = run;
				run += t;
			}
		}
		uint32_t sc = run;
#pragma unroll
		for(int o = 1; o < 32; o <<= 1)
		{
			const uint32_t y = __shfl_up_sync(0xffffffffu, sc, o);
			if(lane >= (uint32_t)o) sc += y;
		}
		if(lane == 31) s_scan[warp] = sc;
		// barrier + vote: a digit that every particle carries makes the pass the identity (uniform decision)
		if(__syncthreads_or(tid < 256 && n != 0 && run == n))
			continue;
		if(tid < 256)
		{
			uint32_t before = 0;
			for(uint32_t w = 0; w < warp; w++) before += s_scan[w];
			const uint32_t dbase = before + sc - run;
			for(uint32_t w = 0; w < kPSortWarps; w++) wcount[w * 256 + tid] += dbase;
		}
		__syncthreads();
		for(uint32_t i0 = cbeg; i0 < cend; i0 += 32)
		{
			const uint32_t i = i0 + lane;
			const bool valid = i < cend;
			const uint32_t key = valid ? ks[i] : 0u;
			const uint32_t dig = valid ? (key >> shift) & 255u : 0xffffu;
			const uint32_t peers = __match_any_sync(0xffffffffu, dig);
			const uint32_t leader = __ffs(peers) - 1u;
			uint32_t base = 0;
			if(valid && lane == leader)
			{
				base = wcount[warp * 256 + dig];
				wcount[warp * 256 + dig] = base + __popc(peers);
			}
			base = __shfl_sync(0xffffffffu, base, leader);
			if(valid)
			{
				const uint32_t dst = base + __popc(peers & ((1u << lane) - 1u));
				kd[dst] = key;
				id[dst] = is[i];
			}
			__syncwarp();
		}
		__syncthreads();
		uint32_t *tk = ks; ks = kd; kd = tk;
		uint16_t *ti = is; is = id; id = ti;
	}
	__syncthreads();
	for(uint32_t i = tid; i < n; i += kPSortThreads)
		sorted[seg + i] = is[i];
}

// ParticleEmitter::UpdateMesh (Source/Scene/RNParticleEmitter.cpp:213-336).  One thread per particle slot of the
// emitter's mesh: quad w holds particle w (or stop-1-w when rendered inversed); its six indices are
// particle*4 + {0,1,2,2,1,3} (the reference indexes with the particle index, :283-288); slots >= stop get degenerate
// indices slot*4 (:322-330) and their vertices are left alone.  188 bytes written per live particle, 48 read.
struct MeshView
{
	const float4 *a, *color, *misc;
	const EmitterState *emitters;
	const float *scale;
	const uint32_t *flags;
	float *pos;
	float4 *col;
	float2 *uv0, *uv1;
	uint16_t *idx;
	uint32_t cap;
};

__global__ void __launch_bounds__(128) particles_mesh_kernel(MeshView mv)
{
	const uint32_t e = blockIdx.y;
	const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t maxp = mv.emitters[e].d.max_particles;
	if(w >= maxp)
		return;
	const uint32_t count = mv.emitters[e].count;
	const uint32_t stop = count < maxp ? count : maxp;
	const size_t seg = (size_t)e * mv.cap;
	uint32_t *idx32 = reinterpret_cast<uint32_t *>(mv.idx + (seg + w) * 6);
	if(w >= stop)
	{
		const uint32_t v = (w * 4u) & 0xffffu;
		idx32[0] = idx32[1] = idx32[2] = v | (v << 16);
		return;
	}
	const uint32_t flags = mv.flags[e];
	const uint32_t pi = (flags & RN_EMITTER_RENDERED_INVERSED) ? stop - 1u - w : w;
	const float4 a = mv.a[seg + pi], c = mv.color[seg + pi], m = mv.misc[seg + pi];
	const float scale = mv.scale[e];
	// Vector2 halfSize = particle->size / 2.0f * scale
	const float hx = m.y / 2.0f * scale, hy = m.z / 2.0f * scale;
	float tx = hx, ty = hy, bx = hx, by = -hy;
	if(flags & RN_EMITTER_CAN_ROLL)
	{
		const float cs = cosf(m.x), sn = sinf(m.x);
		tx = cs * hx - sn * hy; ty = sn * hx + cs * hy;
		bx = cs * hx + sn * hy; by = sn * hx - cs * hy;
	}
	float4 *pos4 = reinterpret_cast<float4 *>(mv.pos + (seg + w) * 12);
	pos4[0] = make_float4(a.x, a.y, a.z, a.x);
	pos4[1] = make_float4(a.y, a.z, a.x, a.y);
	pos4[2] = make_float4(a.z, a.x, a.y, a.z);
	float4 *col = mv.col + (seg + w) * 4;
	col[0] = c; col[1] = c; col[2] = c; col[3] = c;
	float4 *uv0 = reinterpret_cast<float4 *>(mv.uv0 + (seg + w) * 4);
	uv0[0] = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
	uv0[1] = make_float4(0.0f, 1.0f, 1.0f, 1.0f);
	float4 *uv1 = reinterpret_cast<float4 *>(mv.uv1 + (seg + w) * 4);
	uv1[0] = make_float4(-tx, -ty, bx, by);
	uv1[1] = make_float4(-bx, -by, tx, ty);
	const uint32_t b = pi * 4u;
	idx32[0] = ((b + 0u) & 0xffffu) | (((b + 1u) & 0xffffu) << 16);
	idx32[1] = ((b + 2u) & 0xffffu) | (((b + 2u) & 0xffffu) << 16);
	idx32[2] = ((b + 1u) & 0xffffu) | (((b + 3u) & 0xffffu) << 16);
}

// particles_vertex of Shaders/Particles.hlsl:52-71 for every vertex of the live quads:
//   clip = modelViewProjection * (float4(position, 1) + inverseModelView * float4(texCoords2, 0, 0)),  colour * ambient.
// Column-major m[16] as Rayne uploads it; products summed left to right, the two zero components of texCoords2 dropped.
struct BillboardView
{
	const float *pos;
	const float4 *col;
	const float2 *uv1;
	const EmitterState *emitters;
	const float *mats;     // [emitter][32]
	float4 *clip, *out_col;
	float4 ambient;
	uint32_t cap;
};

__global__ void __launch_bounds__(128) particles_billboard_kernel(BillboardView bv)
{
	const uint32_t e = blockIdx.y;
	const uint32_t vtx = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t maxp = bv.emitters[e].d.max_particles;
	const uint32_t count = bv.emitters[e].count;
	const uint32_t stop = count < maxp ? count : maxp;
	if(vtx >= stop * 4u)
		return;
	const size_t v = (size_t)e * bv.cap * 4 + vtx;
	const float *im = bv.mats + (size_t)e * 32, *mvp = im + 16;
	const float px = bv.pos[v * 3 + 0], py = bv.pos[v * 3 + 1], pz = bv.pos[v * 3 + 2];
	const float2 t = bv.uv1[v];
	const float x = px + (__ldg(im + 0) * t.x + __ldg(im + 4) * t.y);
	const float y = py + (__ldg(im + 1) * t.x + __ldg(im + 5) * t.y);
	const float z = pz + (__ldg(im + 2) * t.x + __ldg(im + 6) * t.y);
	const float w = 1.0f + (__ldg(im + 3) * t.x + __ldg(im + 7) * t.y);
	float4 c;
	c.x = __ldg(mvp + 0) * x + __ldg(mvp + 4) * y + __ldg(mvp + 8) * z + __ldg(mvp + 12) * w;
	c.y = __ldg(mvp + 1) * x + __ldg(mvp + 5) * y + __ldg(mvp + 9) * z + __ldg(mvp + 13) * w;
	c.z = __ldg(mvp + 2) * x + __ldg(mvp + 6) * y + __ldg(mvp + 10) * z + __ldg(mvp + 14) * w;
	c.w = __ldg(mvp + 3) * x + __ldg(mvp + 7) * y + __ldg(mvp + 11) * z + __ldg(mvp + 15) * w;
	bv.clip[v] = c;
	const float4 k = bv.col[v];
	bv.out_col[v] = make_float4(k.x * bv.ambient.x, k.y * bv.ambient.y, k.z * bv.ambient.z, k.w * bv.ambient.w);
}

static void free_store(ParticleStore *p)
{
	if(!p) return;
	void *mesh[] = {p->mesh_pos, p->mesh_color, p->mesh_uv0, p->mesh_uv1, p->mesh_idx, p->mesh_scale, p->mesh_flags,
	                p->clip_pos, p->clip_color, p->clip_mats};
	for(void *q : mesh)
		if(q) cudaFree(q);
	for(int g = 0; g < 2; g++)
	{
		if(p->buf[g].a) cudaFree(p->buf[g].a);
		if(p->buf[g].b) cudaFree(p->buf[g].b);
		if(p->buf[g].c) cudaFree(p->buf[g].c);
		if(p->buf[g].d) cudaFree(p->buf[g].d);
		if(p->buf[g].g) cudaFree(p->buf[g].g);
	}
	if(p->color) cudaFree(p->color);
	if(p->misc) cudaFree(p->misc);
	if(p->sorted) cudaFree(p->sorted);
	if(p->emitters) cudaFree(p->emitters);
	if(p->mt) cudaFree(p->mt);
	delete p;
}

int particles_destroy(rn_scene *scene)
{
	free_store(scene->particles);
	scene->particles = nullptr;
	return RN_OK;
}

template <class T>
static int palloc(rn_scene *s, T **p, size_t count)
{
	*p = nullptr;
	if(count == 0) count = 1;
	RN_CUDA(cudaMalloc(reinterpret_cast<void **>(p), count * sizeof(T)));
	s->device_bytes += count * sizeof(T);
	return RN_OK;
}

} // namespace rn

using namespace rn;

extern "C" int rn_particles_configure(rn_scene_t *s, const rn_emitter_desc *emitters, uint32_t count)
{
	RN_REQUIRE(s && emitters, RN_ERR_INVALID, "NULL argument");
	RN_REQUIRE(count > 0 && count <= s->desc.max_emitters, RN_ERR_CAPACITY, "%u emitters exceed max_emitters %u", count, s->desc.max_emitters);
	const uint32_t cap = s->desc.max_particles_per_emitter;
	RN_REQUIRE(cap > 0, RN_ERR_CAPACITY, "the scene was created without particle storage");
	for(uint32_t i = 0; i < count; i++)
	{
		RN_REQUIRE(emitters[i].max_particles > 0 && emitters[i].max_particles <= cap, RN_ERR_CAPACITY,
		           "emitter %u: max_particles %u must be in 1..%u", i, emitters[i].max_particles, cap);
	}
	bool any_nonlocal = false;
	for(uint32_t i = 0; i < count; i++)
		any_nonlocal = any_nonlocal || (emitters[i].flags & RN_EMITTER_DESC_NONLOCAL);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	particles_destroy(s);
	ParticleStore *p = new(std::nothrow) ParticleStore();
	RN_REQUIRE(p, RN_ERR_NOMEM, "out of host memory");
	s->particles = p;
	p->n_emitters = count;
	p->cap = cap;
	const size_t total = (size_t)s->desc.max_emitters * cap;
	int rc = RN_OK;
	for(int g = 0; g < 2 && !rc; g++)
	{
		if(!rc) rc = palloc(s, &p->buf[g].a, total);
		if(!rc) rc = palloc(s, &p->buf[g].b, total);
		if(!rc) rc = palloc(s, &p->buf[g].c, total);
		if(!rc) rc = palloc(s, &p->buf[g].d, total);
		if(!rc && any_nonlocal) rc = palloc(s, &p->buf[g].g, total);
	}
	if(!rc) rc = palloc(s, &p->color, total);
	if(!rc) rc = palloc(s, &p->misc, total);
	if(!rc) rc = palloc(s, &p->sorted, total);
	if(!rc) rc = palloc(s, &p->emitters, (size_t)count);
	if(!rc) rc = palloc(s, &p->mt, (size_t)count * kMTN);
	if(rc)
	{
		particles_destroy(s);
		return rc;
	}
	// MersenneTwister::Seed (RNRandom.cpp:313-323): _offset = 0, so the first 624 outputs are the tempered seed words
	std::vector<EmitterState> es(count);
	std::vector<uint32_t> mt((size_t)count * kMTN);
	for(uint32_t i = 0; i < count; i++)
	{
		memset(&es[i], 0, sizeof(EmitterState));
		es[i].d = emitters[i];
		uint32_t *m = &mt[(size_t)i * kMTN];
		m[0] = emitters[i].seed;
		for(uint32_t k = 1; k < kMTN; k++)
			m[k] = 1812433253u * (m[k - 1] ^ (m[k - 1] >> 30)) + k;
	}
	RN_CUDA(cudaMemcpy(p->emitters, es.data(), sizeof(EmitterState) * count, cudaMemcpyHostToDevice));
	RN_CUDA(cudaMemcpy(p->mt, mt.data(), sizeof(uint32_t) * mt.size(), cudaMemcpyHostToDevice));
	return RN_OK;
}

extern "C" int rn_particles_step(rn_scene_t *s, float delta, void *cuda_stream)
{
	RN_RANGE("rn_particles_step");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && p->n_emitters, RN_ERR_STATE, "rn_particles_configure has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	ParticleView pv;
	pv.in = p->buf[p->gen];
	pv.out = p->buf[p->gen ^ 1];
	pv.color = p->color;
	pv.misc = p->misc;
	pv.emitters = p->emitters;
	pv.mt = p->mt;
	pv.cap = p->cap;
	pv.delta = delta;
	particles_step_kernel<<<p->n_emitters, kParticleThreads, 0, static_cast<cudaStream_t>(cuda_stream)>>>(pv);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	p->gen ^= 1;
	return RN_OK;
}

extern "C" int rn_particles_sort(rn_scene_t *s, const float cam[3], void *cuda_stream)
{
	RN_RANGE("rn_particles_sort");
	RN_REQUIRE(s && cam, RN_ERR_INVALID, "NULL argument");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && p->n_emitters, RN_ERR_STATE, "rn_particles_configure has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	const uint32_t padded = (p->cap + 31u) & ~31u;
	// two key buffers (4 B), two index buffers (2 B) and the warps' digit counters
	const size_t smem = (size_t)padded * 12 + (size_t)kPSortWarps * 256 * sizeof(uint32_t);
	RN_REQUIRE(smem <= 227u * 1024u, RN_ERR_CAPACITY, "rn_particles_sort: %u particles per emitter need %zu bytes of shared memory", p->cap, smem);
	if(p->sort_smem_configured < smem)
	{
		RN_CUDA(cudaFuncSetAttribute(particles_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		p->sort_smem_configured = smem;
	}
	particles_sort_kernel<<<p->n_emitters, 512, smem, static_cast<cudaStream_t>(cuda_stream)>>>(p->buf[p->gen], p->misc, p->sorted, p->emitters,
	                                                                                          p->cap, cam[0], cam[1], cam[2], padded);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	return RN_OK;
}

extern "C" int rn_particles_counts(rn_scene_t *s, uint32_t *out_counts, float *out_times, uint32_t n_emitters)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && n_emitters <= p->n_emitters, RN_ERR_STATE, "%u emitters requested, %u configured", n_emitters, p ? p->n_emitters : 0);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	std::vector<EmitterState> es(n_emitters);
	RN_CUDA(cudaMemcpy(es.data(), p->emitters, sizeof(EmitterState) * n_emitters, cudaMemcpyDeviceToHost));
	uint32_t overflowed = 0xffffffffu;
	for(uint32_t i = 0; i < n_emitters; i++)
	{
		if(out_counts) out_counts[i] = es[i].count;
		if(out_times) out_times[i] = es[i].time;
		if(es[i].overflow && overflowed == 0xffffffffu) overflowed = i;
	}
	RN_REQUIRE(overflowed == 0xffffffffu, RN_ERR_CAPACITY,
	           "emitter %u asked for more particles than max_particles_per_emitter (%u): its live count exceeded the soft maximum, "
	           "so the reference's unsigned `soft - size` wrapped (RNParticleEmitter.cpp:204-205)", overflowed, p->cap);
	return RN_OK;
}

// host-side edit of a few fields of the emitter records (tiny, synchronous)
template <class F>
static int edit_emitters(rn_scene *s, uint32_t first, uint32_t count, F &&f)
{
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && (uint64_t)first + count <= p->n_emitters, RN_ERR_INVALID, "emitters [%u, %u) out of range", first, first + count);
	if(count == 0) return RN_OK;
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	std::vector<EmitterState> es(count);
	RN_CUDA(cudaMemcpy(es.data(), p->emitters + first, sizeof(EmitterState) * count, cudaMemcpyDeviceToHost));
	for(uint32_t i = 0; i < count; i++) f(es[i], i);
	RN_CUDA(cudaMemcpy(p->emitters + first, es.data(), sizeof(EmitterState) * count, cudaMemcpyHostToDevice));
	return RN_OK;
}

extern "C" int rn_particles_set_soft_max(rn_scene_t *s, uint32_t first, uint32_t count, const uint32_t *soft)
{
	RN_REQUIRE(s && soft, RN_ERR_INVALID, "NULL argument");
	return edit_emitters(s, first, count, [&](EmitterState &e, uint32_t i) {
		e.d.flags |= RN_EMITTER_DESC_SOFT_MAX;
		e.d.max_particles_soft = soft[i];
	});
}

extern "C" int rn_particles_set_emitter_transforms(rn_scene_t *s, uint32_t first, uint32_t count, const float *world_position,
                                                   const float *world_scale)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	return edit_emitters(s, first, count, [&](EmitterState &e, uint32_t i) {
		if(world_position) memcpy(e.d.world_position, world_position + (size_t)i * 3, 12);
		if(world_scale) memcpy(e.d.world_scale, world_scale + (size_t)i * 3, 12);
	});
}

extern "C" int rn_particles_download(rn_scene_t *s, uint32_t emitter, rn_particle_field field, void *dst, uint32_t capacity,
                                     uint32_t *out_count)
{
	RN_REQUIRE(s && out_count, RN_ERR_INVALID, "NULL argument");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && emitter < p->n_emitters, RN_ERR_INVALID, "emitter %u out of range", emitter);
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	EmitterState es;
	RN_CUDA(cudaMemcpy(&es, p->emitters + emitter, sizeof(es), cudaMemcpyDeviceToHost));
	const uint32_t n = es.count;
	*out_count = n;
	if(!dst || n == 0) return RN_OK;
	RN_REQUIRE(n <= capacity, RN_ERR_CAPACITY, "emitter %u has %u particles, buffer holds %u", emitter, n, capacity);
	const size_t seg = (size_t)emitter * p->cap;
	const ParticleBuffers &b = p->buf[p->gen];
	std::vector<float4> t0(n), t1(n), t2(n);
	std::vector<float> t3(n);
	float *out = static_cast<float *>(dst);
	switch(field)
	{
		case RN_PF_POSITION:
		case RN_PF_VELOCITY:
		case RN_PF_TIMING:
		case RN_PF_RANGES:
			RN_CUDA(cudaMemcpy(t0.data(), b.a + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
			RN_CUDA(cudaMemcpy(t1.data(), b.b + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
			RN_CUDA(cudaMemcpy(t2.data(), b.c + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
			RN_CUDA(cudaMemcpy(t3.data(), b.d + seg, sizeof(float) * n, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < n; i++)
			{
				if(field == RN_PF_POSITION) { out[i * 3] = t0[i].x; out[i * 3 + 1] = t0[i].y; out[i * 3 + 2] = t0[i].z; }
				else if(field == RN_PF_VELOCITY) { out[i * 3] = t1[i].x; out[i * 3 + 1] = t1[i].y; out[i * 3 + 2] = t1[i].z; }
				else if(field == RN_PF_TIMING) { out[i * 3] = t0[i].w; out[i * 3 + 1] = t1[i].w; out[i * 3 + 2] = t2[i].x; }
				else { out[i * 4] = t2[i].y; out[i * 4 + 1] = t2[i].z; out[i * 4 + 2] = t2[i].w; out[i * 4 + 3] = t3[i]; }
			}
			break;
		case RN_PF_DERIVED:
			RN_CUDA(cudaMemcpy(t0.data(), p->misc + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
			RN_CUDA(cudaMemcpy(t1.data(), p->color + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
			for(uint32_t i = 0; i < n; i++)
			{
				out[i * 7] = t0[i].x; out[i * 7 + 1] = t0[i].y; out[i * 7 + 2] = t0[i].z;
				out[i * 7 + 3] = t1[i].x; out[i * 7 + 4] = t1[i].y; out[i * 7 + 5] = t1[i].z; out[i * 7 + 6] = t1[i].w;
			}
			break;
		case RN_PF_SORTED:
			RN_CUDA(cudaMemcpy(dst, p->sorted + seg, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
			break;
		case RN_PF_GRAVITY:
			if(b.g && (es.d.flags & RN_EMITTER_DESC_NONLOCAL))
			{
				RN_CUDA(cudaMemcpy(t0.data(), b.g + seg, sizeof(float4) * n, cudaMemcpyDeviceToHost));
				for(uint32_t i = 0; i < n; i++) { out[i * 3] = t0[i].x; out[i * 3 + 1] = t0[i].y; out[i * 3 + 2] = t0[i].z; }
			}
			else
				for(uint32_t i = 0; i < n; i++) { out[i * 3] = es.d.gravity[0]; out[i * 3 + 1] = es.d.gravity[1]; out[i * 3 + 2] = es.d.gravity[2]; }
			break;
		default:
			RN_REQUIRE(false, RN_ERR_INVALID, "unknown particle field %d", (int)field);
	}
	return RN_OK;
}

extern "C" int rn_particles_build_mesh(rn_scene_t *s, const float *world_scale_x, const uint32_t *flags, rn_particle_mesh *out,
                                       void *cuda_stream)
{
	RN_RANGE("rn_particles_build_mesh");
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && p->n_emitters, RN_ERR_STATE, "rn_particles_configure has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const size_t total = (size_t)s->desc.max_emitters * p->cap;
	if(!p->mesh_pos)
	{
		int rc = palloc(s, &p->mesh_pos, total * 12);
		if(!rc) rc = palloc(s, &p->mesh_color, total * 4);
		if(!rc) rc = palloc(s, &p->mesh_uv0, total * 4);
		if(!rc) rc = palloc(s, &p->mesh_uv1, total * 4);
		if(!rc) rc = palloc(s, &p->mesh_idx, total * 6);
		if(!rc) rc = palloc(s, &p->mesh_scale, (size_t)p->n_emitters);
		if(!rc) rc = palloc(s, &p->mesh_flags, (size_t)p->n_emitters);
		if(rc) return rc;
		// the reference's mesh buffers start zeroed (calloc'ed staging); unused vertices stay that way
		RN_CUDA(cudaMemsetAsync(p->mesh_pos, 0, total * 12 * sizeof(float), stream));
		RN_CUDA(cudaMemsetAsync(p->mesh_color, 0, total * 4 * sizeof(float4), stream));
		RN_CUDA(cudaMemsetAsync(p->mesh_uv0, 0, total * 4 * sizeof(float2), stream));
		RN_CUDA(cudaMemsetAsync(p->mesh_uv1, 0, total * 4 * sizeof(float2), stream));
		RN_CUDA(cudaMemsetAsync(p->mesh_idx, 0, total * 6 * sizeof(uint16_t), stream));
	}
	{
		std::vector<float> sc(p->n_emitters, 1.0f);
		std::vector<uint32_t> fl(p->n_emitters, 0u);
		if(world_scale_x) memcpy(sc.data(), world_scale_x, sizeof(float) * p->n_emitters);
		if(flags) memcpy(fl.data(), flags, sizeof(uint32_t) * p->n_emitters);
		// small synchronous parameter copies (pageable host memory): n_emitters * 8 bytes
		RN_CUDA(cudaMemcpyAsync(p->mesh_scale, sc.data(), sizeof(float) * p->n_emitters, cudaMemcpyHostToDevice, stream));
		RN_CUDA(cudaMemcpyAsync(p->mesh_flags, fl.data(), sizeof(uint32_t) * p->n_emitters, cudaMemcpyHostToDevice, stream));
		RN_CUDA(cudaStreamSynchronize(stream));
	}
	MeshView mv;
	mv.a = p->buf[p->gen].a; mv.color = p->color; mv.misc = p->misc;
	mv.emitters = p->emitters; mv.scale = p->mesh_scale; mv.flags = p->mesh_flags;
	mv.pos = p->mesh_pos; mv.col = p->mesh_color; mv.uv0 = p->mesh_uv0; mv.uv1 = p->mesh_uv1; mv.idx = p->mesh_idx;
	mv.cap = p->cap;
	const dim3 grid((p->cap + 127) / 128, p->n_emitters);
	particles_mesh_kernel<<<grid, 128, 0, stream>>>(mv);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	p->have_mesh = true;
	if(out)
	{
		out->position = p->mesh_pos; out->color = reinterpret_cast<const float *>(p->mesh_color);
		out->uv0 = reinterpret_cast<const float *>(p->mesh_uv0); out->uv1 = reinterpret_cast<const float *>(p->mesh_uv1);
		out->indices = p->mesh_idx;
		out->vertex_stride_per_emitter = p->cap * 4;
		out->index_stride_per_emitter = p->cap * 6;
	}
	return RN_OK;
}

extern "C" int rn_particles_download_mesh(rn_scene_t *s, uint32_t emitter, float *position, float *color, float *uv0, float *uv1,
                                          uint16_t *indices)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && emitter < p->n_emitters, RN_ERR_INVALID, "emitter %u out of range", emitter);
	RN_REQUIRE(p->have_mesh, RN_ERR_STATE, "rn_particles_build_mesh has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	EmitterState es;
	RN_CUDA(cudaMemcpy(&es, p->emitters + emitter, sizeof(es), cudaMemcpyDeviceToHost));
	const size_t seg = (size_t)emitter * p->cap, nv = (size_t)es.d.max_particles * 4, ni = (size_t)es.d.max_particles * 6;
	if(position) RN_CUDA(cudaMemcpy(position, p->mesh_pos + seg * 12, nv * 3 * sizeof(float), cudaMemcpyDeviceToHost));
	if(color) RN_CUDA(cudaMemcpy(color, p->mesh_color + seg * 4, nv * sizeof(float4), cudaMemcpyDeviceToHost));
	if(uv0) RN_CUDA(cudaMemcpy(uv0, p->mesh_uv0 + seg * 4, nv * sizeof(float2), cudaMemcpyDeviceToHost));
	if(uv1) RN_CUDA(cudaMemcpy(uv1, p->mesh_uv1 + seg * 4, nv * sizeof(float2), cudaMemcpyDeviceToHost));
	if(indices) RN_CUDA(cudaMemcpy(indices, p->mesh_idx + seg * 6, ni * sizeof(uint16_t), cudaMemcpyDeviceToHost));
	return RN_OK;
}

extern "C" int rn_particles_expand_billboards(rn_scene_t *s, const float *inverse_model_view, const float *model_view_projection,
                                              const float *ambient_color, rn_particle_clip *out, void *cuda_stream)
{
	RN_REQUIRE(s && inverse_model_view && model_view_projection && ambient_color, RN_ERR_INVALID, "NULL argument");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && p->n_emitters, RN_ERR_STATE, "rn_particles_configure has not been called");
	RN_REQUIRE(p->have_mesh, RN_ERR_STATE, "rn_particles_build_mesh has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
	const size_t total = (size_t)s->desc.max_emitters * p->cap * 4;
	if(!p->clip_pos)
	{
		int rc = palloc(s, &p->clip_pos, total);
		if(!rc) rc = palloc(s, &p->clip_color, total);
		if(!rc) rc = palloc(s, &p->clip_mats, (size_t)s->desc.max_emitters * 32);
		if(rc) return rc;
		RN_CUDA(cudaMemsetAsync(p->clip_pos, 0, total * sizeof(float4), stream));
		RN_CUDA(cudaMemsetAsync(p->clip_color, 0, total * sizeof(float4), stream));
	}
	{
		// per emitter 2 x 64 bytes of matrices, interleaved; staged copy (pageable host memory)
		std::vector<float> m((size_t)p->n_emitters * 32);
		for(uint32_t e = 0; e < p->n_emitters; e++)
		{
			memcpy(&m[(size_t)e * 32], inverse_model_view + (size_t)e * 16, 64);
			memcpy(&m[(size_t)e * 32 + 16], model_view_projection + (size_t)e * 16, 64);
		}
		RN_CUDA(cudaMemcpyAsync(p->clip_mats, m.data(), m.size() * sizeof(float), cudaMemcpyHostToDevice, stream));
		RN_CUDA(cudaStreamSynchronize(stream));
	}
	BillboardView bv;
	bv.pos = p->mesh_pos; bv.col = p->mesh_color; bv.uv1 = p->mesh_uv1; bv.emitters = p->emitters; bv.mats = p->clip_mats;
	bv.clip = p->clip_pos; bv.out_col = p->clip_color;
	bv.ambient = make_float4(ambient_color[0], ambient_color[1], ambient_color[2], ambient_color[3]);
	bv.cap = p->cap;
	const dim3 grid((p->cap * 4 + 127) / 128, p->n_emitters);
	particles_billboard_kernel<<<grid, 128, 0, stream>>>(bv);
	s->launches++;
	RN_CUDA(cudaGetLastError());
	p->have_clip = true;
	if(out)
	{
		out->clip_position = reinterpret_cast<const float *>(p->clip_pos);
		out->color = reinterpret_cast<const float *>(p->clip_color);
		out->vertex_stride_per_emitter = p->cap * 4;
		out->reserved = 0;
	}
	return RN_OK;
}

extern "C" int rn_particles_download_clip(rn_scene_t *s, uint32_t emitter, float *clip_position, float *color)
{
	RN_REQUIRE(s, RN_ERR_INVALID, "NULL scene");
	ParticleStore *p = s->particles;
	RN_REQUIRE(p && emitter < p->n_emitters, RN_ERR_INVALID, "emitter %u out of range", emitter);
	RN_REQUIRE(p->have_clip, RN_ERR_STATE, "rn_particles_expand_billboards has not been called");
	RN_CUDA(cudaSetDevice(s->device));
	RN_CUDA(cudaDeviceSynchronize());
	EmitterState es;
	RN_CUDA(cudaMemcpy(&es, p->emitters + emitter, sizeof(es), cudaMemcpyDeviceToHost));
	const size_t seg = (size_t)emitter * p->cap * 4, nv = (size_t)es.d.max_particles * 4;
	if(clip_position) RN_CUDA(cudaMemcpy(clip_position, p->clip_pos + seg, nv * sizeof(float4), cudaMemcpyDeviceToHost));
	if(color) RN_CUDA(cudaMemcpy(color, p->clip_color + seg, nv * sizeof(float4), cudaMemcpyDeviceToHost));
	return RN_OK;
}
