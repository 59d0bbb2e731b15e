// TEST INFRASTRUCTURE — not part of the product.
//
// Headless driver that links the UNMODIFIED reference classes (RN::SceneNode, RN::Camera, RN::SceneBasic,
// RN::GenericParticleEmitter) built by oracle/build_ref.sh from /root/reference and runs them on a scene /
// emitter description read from a file.  Its outputs are (1) the parity goldens the C restatement
// (oracle/rn_oracle.c) and the CUDA path are checked against and (2) the timed CPU baseline ("reference" kind).
//
//   ref_harness scene     <in.bin> <out.bin>      run frames, dump render order / world TRS / matrices / lists
//   ref_harness particles <in.bin> <out.bin>      run K emitter steps, dump every live particle
//   ref_harness bench-scene     <in.bin> <frames> <warmup>   print JSON timing line (ms per frame)
//   ref_harness bench-particles <in.bin> <steps>  <warmup>   print JSON timing line
//   ref_harness cascades  <in.bin> <out.bin>      Light::UpdateShadows + Camera::MakeShadowSplit for a directional light
//   ref_harness occlusion <scene.bin> <occ.bin> <out.bin>   SceneBasic::Render with Flags::Occluder entities (software occlusion culling)
//
// File formats are documented in oracle/refio.py (the only reader/writer on the Python side).
//
// Build notes (see build_ref.sh): compiled with -fno-access-control so that protected Scene::Update/Render,
// private WorkQueue::InitializeQueues and ParticleEmitter::_particles are reachable; Base/RNBase.cpp is
// #included so the file-static RN::__functionPool can be created (reference: Source/Base/RNBase.cpp:164,174).

#include "Rayne.h"
#include "Base/RNBase.cpp"

#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <chrono>
#include <algorithm>
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

using namespace RN;

#include "ref_null_renderer.h"

namespace
{
	// A plain cull-only node: the base SceneNode::CanRender returns false (Source/Scene/RNSceneNode.cpp:419-422)
	// and Entity needs a Model, so this does what Entity::CanRender does minus the model check
	// (Source/Scene/RNEntity.cpp:154-159 -> SceneNode::CanRenderUtil).
	struct Recorder
	{
		std::vector<std::vector<uint32_t>> lists; // per camera (harness camera index)
		std::vector<Camera *> cameras;
		bool touchMatrices = false;
		float sink = 0.0f;
	};
	static Recorder *g_recorder = nullptr;

	class CullNode : public SceneNode
	{
	public:
		uint32_t id = 0;
		bool CanRender(Renderer *renderer, Camera *camera) const override { return CanRenderUtil(renderer, camera); }
		void Render(Renderer *renderer, Camera *camera) const override
		{
			SceneNode::Render(renderer, camera);
			if(g_recorder)
			{
				for(size_t i = 0; i < g_recorder->cameras.size(); i++)
				{
					if(g_recorder->cameras[i] == camera)
					{
						g_recorder->lists[i].push_back(id);
						break;
					}
				}
				if(g_recorder->touchMatrices)
				{
					// what Drawable::Update does for a visible node (Source/Rendering/RNRenderer.h:65-76)
					Matrix m = GetWorldTransform();
					g_recorder->sink += m.m[12];
				}
			}
		}
	};

	class HarnessScene : public SceneBasic
	{
	public:
		HarnessScene() {}
	};

	class HarnessVisScene : public SceneWithVisibilityLists
	{
	public:
		HarnessVisScene() {}
	};

	// ---- tiny binary IO ----------------------------------------------------------------------------------
	struct Reader
	{
		FILE *f;
		explicit Reader(const char *path) { f = fopen(path, "rb"); if(!f) { perror(path); exit(2); } }
		~Reader() { fclose(f); }
		template<class T> T get() { T v; if(fread(&v, sizeof(T), 1, f) != 1) { fprintf(stderr, "short read\n"); exit(2); } return v; }
		template<class T> std::vector<T> vec(size_t n) { std::vector<T> v(n); if(n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); } return v; }
	};
	struct Writer
	{
		FILE *f;
		explicit Writer(const char *path) { f = fopen(path, "wb"); if(!f) { perror(path); exit(2); } }
		~Writer() { fclose(f); }
		template<class T> void put(const T &v) { fwrite(&v, sizeof(T), 1, f); }
		template<class T> void vec(const std::vector<T> &v) { if(!v.empty()) fwrite(v.data(), sizeof(T), v.size(), f); }
	};

	struct CameraDesc
	{
		uint32_t flags;       // Camera::Flags bits to OR into Defaults (Orthogonal, UseSimpleCulling, ...)
		uint32_t renderGroup;
		float pos[3], rot[4], scale[3];
		float fov, aspect, clipNear, clipFar;
		float ortho[4];       // left, right, bottom, top
		float planeOffsets[4];// top, bottom, left, right (Camera::SetFrustumPlaneOffset order)
		uint32_t hasProjection;
		float projection[16];
	};

	struct SceneInput
	{
		uint32_t nNodes, nCameras, nFrames, options;
		float rootTranslate[3];
		std::vector<int32_t> parent;
		std::vector<float> pos, scale, rot;
		std::vector<uint32_t> flags, group;
		std::vector<int32_t> priority;
		std::vector<float> sphere;
		std::vector<float> box;   // optional (kOptBox): minExtend xyz, maxExtend xyz per node
		// optional (kOptVolumes): AxisAlignedBoxVolumes (boundsMin xyz, boundsMax xyz) and the volume of every node
		// (0 = stays in the default volume, k = listed in volume k)
		std::vector<float> volumeBounds;
		std::vector<uint32_t> volumeOfNode;
		std::vector<CameraDesc> cameras;
	};

	// kOptInstances: the dump gains, per camera, the view and projection*view matrices and, for camera 0, view*world
	//          and projView*world of every node (the matrix uniforms a renderer derives per drawable).
	// kOptVisLists: drive SceneWithVisibilityLists (default volume only) instead of SceneBasic.
	// kOptBox: the input carries a local AABB column (SceneNode::SetBoundingBox).
	enum : uint32_t { kOptInverse = 1u << 0, kOptTouchMatrices = 1u << 1, kOptVisLists = 1u << 2, kOptBox = 1u << 3, kOptInstances = 1u << 4,
	                  kOptVolumes = 1u << 5 };   // kOptVolumes (with kOptVisLists): SceneWithVisibilityLists::AddVolume

	SceneInput ReadScene(const char *path)
	{
		Reader r(path);
		SceneInput in;
		uint32_t magic = r.get<uint32_t>();
		uint32_t version = r.get<uint32_t>();
		if(magic != 0x49534e52u || version != 1) { fprintf(stderr, "bad scene file\n"); exit(2); }
		in.nNodes = r.get<uint32_t>(); in.nCameras = r.get<uint32_t>(); in.nFrames = r.get<uint32_t>(); in.options = r.get<uint32_t>();
		for(int i = 0; i < 3; i++) in.rootTranslate[i] = r.get<float>();
		size_t n = in.nNodes;
		in.parent = r.vec<int32_t>(n);
		in.pos = r.vec<float>(n * 3); in.scale = r.vec<float>(n * 3); in.rot = r.vec<float>(n * 4);
		in.flags = r.vec<uint32_t>(n); in.group = r.vec<uint32_t>(n); in.priority = r.vec<int32_t>(n);
		in.sphere = r.vec<float>(n * 4);
		if(in.options & kOptBox) in.box = r.vec<float>(n * 6);
		if(in.options & kOptVolumes)
		{
			uint32_t nVolumes = r.get<uint32_t>();
			in.volumeBounds = r.vec<float>((size_t)nVolumes * 6);
			in.volumeOfNode = r.vec<uint32_t>(n);
		}
		in.cameras = r.vec<CameraDesc>(in.nCameras);
		return in;
	}

	struct BuiltScene
	{
		Scene *scene;
		HarnessScene *basic = nullptr;
		HarnessVisScene *vis = nullptr;
		std::vector<CullNode *> nodes;
		std::vector<CullNode *> roots;
		std::vector<Camera *> cameras;
	};

	BuiltScene BuildScene(const SceneInput &in)
	{
		BuiltScene b;
		if(in.options & kOptVisLists) b.scene = b.vis = new HarnessVisScene();
		else b.scene = b.basic = new HarnessScene();
		b.nodes.resize(in.nNodes);
		for(uint32_t i = 0; i < in.nNodes; i++)
		{
			CullNode *node = new CullNode();
			node->id = i;
			node->SetPosition(Vector3(in.pos[i*3], in.pos[i*3+1], in.pos[i*3+2]));
			node->SetScale(Vector3(in.scale[i*3], in.scale[i*3+1], in.scale[i*3+2]));
			node->SetRotation(Quaternion(in.rot[i*4], in.rot[i*4+1], in.rot[i*4+2], in.rot[i*4+3]));
			node->SetFlags(in.flags[i]);
			node->SetRenderGroup(static_cast<uint8>(in.group[i]));
			if(in.priority[i] != SceneNode::RenderNormal) node->SetRenderPriority(in.priority[i]);
			if(in.options & kOptBox)
				node->SetBoundingBox(AABB(Vector3(in.box[i*6], in.box[i*6+1], in.box[i*6+2]), Vector3(in.box[i*6+3], in.box[i*6+4], in.box[i*6+5])), false);
			Sphere s(Vector3(in.sphere[i*4], in.sphere[i*4+1], in.sphere[i*4+2]), in.sphere[i*4+3]);
			node->SetBoundingSphere(s);
			b.nodes[i] = node;
		}
		// the trees are completed BEFORE their roots enter the scene (keeps every tree contiguous in the render list)
		for(uint32_t i = 0; i < in.nNodes; i++)
		{
			if(in.parent[i] >= 0) b.nodes[in.parent[i]]->AddChild(b.nodes[i]);
			else b.roots.push_back(b.nodes[i]);
		}
		for(CullNode *root : b.roots) b.scene->AddNode(root);
		if(b.vis && (in.options & kOptVolumes))
		{
			// the reference flow: nodes enter the scene (default volume), AddVolume then moves the listed ones out of it
			// (Source/Scene/RNSceneWithVisibilityLists.cpp:181-207); a volume lists its nodes in creation order here
			for(size_t k = 0; k < in.volumeBounds.size() / 6; k++)
			{
				auto *volume = new SceneWithVisibilityLists::AxisAlignedBoxVolume();
				volume->boundsMin = Vector3(in.volumeBounds[k*6], in.volumeBounds[k*6+1], in.volumeBounds[k*6+2]);
				volume->boundsMax = Vector3(in.volumeBounds[k*6+3], in.volumeBounds[k*6+4], in.volumeBounds[k*6+5]);
				for(uint32_t i = 0; i < in.nNodes; i++)
					if(in.volumeOfNode[i] == k + 1) volume->nodes.push_back(b.nodes[i]);
				b.vis->AddVolume(volume);
			}
		}

		for(uint32_t c = 0; c < in.nCameras; c++)
		{
			const CameraDesc &d = in.cameras[c];
			Camera *camera = new Camera();
			camera->SetFlags(camera->GetFlags() | d.flags);
			camera->SetRenderGroup(static_cast<uint8>(d.renderGroup));
			camera->SetPosition(Vector3(d.pos[0], d.pos[1], d.pos[2]));
			camera->SetRotation(Quaternion(d.rot[0], d.rot[1], d.rot[2], d.rot[3]));
			camera->SetScale(Vector3(d.scale[0], d.scale[1], d.scale[2]));
			camera->SetFOV(d.fov);
			camera->SetAspectRatio(d.aspect);
			camera->SetClipNear(d.clipNear);
			camera->SetClipFar(d.clipFar);
			if(d.flags & Camera::Flags::Orthogonal)
				camera->SetOrthogonalFrustum(d.ortho[3], d.ortho[2], d.ortho[0], d.ortho[1]);
			camera->SetFrustumPlaneOffset(d.planeOffsets[0], d.planeOffsets[1], d.planeOffsets[2], d.planeOffsets[3]);
			if(d.hasProjection)
			{
				Matrix m;
				for(int i = 0; i < 16; i++) m.m[i] = d.projection[i];
				camera->SetProjectionMatrix(m);
			}
			b.scene->AddNode(camera);
			b.cameras.push_back(camera);
		}
		return b;
	}

	void RunFrame(const SceneInput &in, BuiltScene &b, Renderer *renderer, bool translate)
	{
		if(translate)
		{
			Vector3 t(in.rootTranslate[0], in.rootTranslate[1], in.rootTranslate[2]);
			for(CullNode *root : b.roots) root->Translate(t);
		}
		b.scene->Update(1.0f / 60.0f);
		b.scene->Render(renderer);
	}

	int CmdScene(const char *inPath, const char *outPath, Renderer *renderer)
	{
		SceneInput in = ReadScene(inPath);
		BuiltScene b = BuildScene(in);
		Recorder rec;
		rec.cameras = b.cameras;
		rec.lists.resize(b.cameras.size());
		rec.touchMatrices = (in.options & kOptTouchMatrices) != 0;
		g_recorder = &rec;
		uint32_t frames = in.nFrames ? in.nFrames : 1;
		for(uint32_t f = 0; f < frames; f++)
		{
			for(auto &l : rec.lists) l.clear();
			RunFrame(in, b, renderer, f > 0);
		}
		g_recorder = nullptr;

		Writer w(outPath);
		w.put<uint32_t>(0x4f534e52u); w.put<uint32_t>(2);
		w.put<uint32_t>(in.nNodes); w.put<uint32_t>(in.nCameras);
		// render-list order as creation ids: SceneBasic::_renderNodes head to tail, or the default volume of
		// SceneWithVisibilityLists (cameras are nodes of that volume too; they are not CullNodes and are skipped)
		std::vector<uint32_t> order;
		if(b.basic)
		{
			for(IntrusiveList<SceneNode>::Member *m = b.basic->_renderNodes.GetHead(); m; m = m->GetNext())
				order.push_back(static_cast<CullNode *>(m->Get())->id);
		}
		else
		{
			for(SceneNode *node : b.vis->_defaultVolume->nodes)
				if(!node->IsKindOfClass(Camera::GetMetaClass())) order.push_back(static_cast<CullNode *>(node)->id);
		}
		w.put<uint32_t>(static_cast<uint32_t>(order.size()));
		w.vec(order);
		std::vector<float> wpos(in.nNodes * 3), wrot(in.nNodes * 4), wscale(in.nNodes * 3), wmat((size_t)in.nNodes * 16), sph((size_t)in.nNodes * 4);
		std::vector<float> winv, aabb((size_t)in.nNodes * 6);
		if(in.options & kOptInverse) winv.resize((size_t)in.nNodes * 16);
		for(uint32_t i = 0; i < in.nNodes; i++)
		{
			CullNode *node = b.nodes[i];
			Vector3 p = node->GetWorldPosition(); Quaternion q = node->GetWorldRotation(); Vector3 s = node->GetWorldScale();
			wpos[i*3] = p.x; wpos[i*3+1] = p.y; wpos[i*3+2] = p.z;
			wrot[i*4] = q.x; wrot[i*4+1] = q.y; wrot[i*4+2] = q.z; wrot[i*4+3] = q.w;
			wscale[i*3] = s.x; wscale[i*3+1] = s.y; wscale[i*3+2] = s.z;
			Matrix m = node->GetWorldTransform();
			for(int k = 0; k < 16; k++) wmat[(size_t)i*16 + k] = m.m[k];
			if(in.options & kOptInverse)
			{
				Matrix inv = node->GetInverseWorldTransform();
				for(int k = 0; k < 16; k++) winv[(size_t)i*16 + k] = inv.m[k];
			}
			Sphere bs = node->GetBoundingSphere();
			Vector3 c = bs.position + bs.offset;
			sph[i*4] = c.x; sph[i*4+1] = c.y; sph[i*4+2] = c.z; sph[i*4+3] = bs.radius;
			AABB bb = node->GetBoundingBox();   // position == world position; extents relative to it
			aabb[(size_t)i*6+0] = bb.minExtend.x; aabb[(size_t)i*6+1] = bb.minExtend.y; aabb[(size_t)i*6+2] = bb.minExtend.z;
			aabb[(size_t)i*6+3] = bb.maxExtend.x; aabb[(size_t)i*6+4] = bb.maxExtend.y; aabb[(size_t)i*6+5] = bb.maxExtend.z;
		}
		w.vec(wpos); w.vec(wrot); w.vec(wscale); w.vec(wmat); w.vec(sph);
		w.put<uint32_t>((in.options & kOptInverse) ? 1 : 0);
		w.vec(winv);
		w.vec(aabb);
		for(size_t c = 0; c < b.cameras.size(); c++)
		{
			Camera *cam = b.cameras[c];
			float fr[28];
			fr[0] = cam->_frustumCenter.x; fr[1] = cam->_frustumCenter.y; fr[2] = cam->_frustumCenter.z; fr[3] = cam->_frustumRadius;
			const Plane *planes[6] = { &cam->frustums._frustumLeft, &cam->frustums._frustumRight, &cam->frustums._frustumTop,
			                           &cam->frustums._frustumBottom, &cam->frustums._frustumNear, &cam->frustums._frustumFar };
			for(int k = 0; k < 6; k++)
			{
				fr[4 + k*4 + 0] = planes[k]->_normal.x; fr[4 + k*4 + 1] = planes[k]->_normal.y;
				fr[4 + k*4 + 2] = planes[k]->_normal.z; fr[4 + k*4 + 3] = planes[k]->_d;
			}
			for(int k = 0; k < 28; k++) w.put<float>(fr[k]);
			w.put<uint32_t>(static_cast<uint32_t>(cam->GetFlags()));
			w.put<uint32_t>(static_cast<uint32_t>(cam->GetRenderGroup()));
			w.put<uint32_t>(static_cast<uint32_t>(rec.lists[c].size()));
			w.vec(rec.lists[c]);
		}
		if(in.options & kOptInstances)
		{
			// the matrix products of a renderer's uniform upload (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168)
			// evaluated with the reference's own Matrix class.  (LOD selection, the other per-drawable step of
			// Entity::Render, is compiled out in this reference: RN_MODEL_LOD_DISABLED 1, Source/Rendering/RNModel.h:9.)
			for(size_t c = 0; c < b.cameras.size(); c++)
			{
				Camera *cam = b.cameras[c];
				Matrix view = cam->GetViewMatrix();
				Matrix projView = cam->GetProjectionMatrix() * view;
				for(int k = 0; k < 16; k++) w.put<float>(view.m[k]);
				for(int k = 0; k < 16; k++) w.put<float>(projView.m[k]);
				if(c == 0)
				{
					std::vector<float> mv((size_t)in.nNodes * 16), mvp((size_t)in.nNodes * 16);
					for(uint32_t i = 0; i < in.nNodes; i++)
					{
						Matrix model = b.nodes[i]->GetWorldTransform();
						Matrix a = view * model, p = projView * model;
						for(int k = 0; k < 16; k++) { mv[(size_t)i*16 + k] = a.m[k]; mvp[(size_t)i*16 + k] = p.m[k]; }
					}
					w.vec(mv); w.vec(mvp);
				}
			}
		}
		return 0;
	}

	int CmdBenchScene(const char *inPath, int frames, int warmup, Renderer *renderer)
	{
		SceneInput in = ReadScene(inPath);
		BuiltScene b = BuildScene(in);
		Recorder rec;
		rec.cameras = b.cameras;
		rec.lists.resize(b.cameras.size());
		rec.touchMatrices = true;
		g_recorder = &rec;
		for(int f = 0; f < warmup; f++) { for(auto &l : rec.lists) l.clear(); RunFrame(in, b, renderer, true); }
		double best = 1e30, total = 0.0;
		size_t visible = 0;
		for(int f = 0; f < frames; f++)
		{
			for(auto &l : rec.lists) l.clear();
			auto t0 = std::chrono::steady_clock::now();
			RunFrame(in, b, renderer, true);
			auto t1 = std::chrono::steady_clock::now();
			double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
			best = std::min(best, ms); total += ms;
			visible = 0; for(auto &l : rec.lists) visible += l.size();
		}
		g_recorder = nullptr;
		printf("{\"kind\": \"reference\", \"what\": \"scene\", \"nodes\": %u, \"cameras\": %u, \"frames\": %d, \"ms_best\": %.6f, \"ms_mean\": %.6f, \"visible\": %zu, \"sink\": %g}\n",
		       in.nNodes, in.nCameras, frames, best, total / frames, visible, (double)rec.sink);
		return 0;
	}

	// ---- particles -----------------------------------------------------------------------------------------
	struct EmitterDesc
	{
		uint32_t seed, maxParticles;
		float spawnRate;
		float lifeSpan[2];
		float startColor[4], endColor[4];
		float startSize[2], endSize[2], startRotation[2], endRotation[2];
		float gravity[3], velocity[3], velMin[3], velMax[3], posMin[3], posMax[3];
		uint32_t flags;                // bit 0: SetIsLocal(false); bit 1: SetMaxParticlesSoft(maxParticlesSoft)
		uint32_t maxParticlesSoft;
		float worldPosition[3], worldScale[3];
	};

	// options: the emitter switches UpdateMesh looks at (Source/Scene/RNParticleEmitter.cpp:229,235) + "dump the mesh"
	enum : uint32_t { kPOptRoll = 1u << 0, kPOptInversed = 1u << 1, kPOptMesh = 1u << 2 };

	struct ParticleInput
	{
		uint32_t nEmitters, nSteps;
		float delta;
		uint32_t options;
		uint32_t softAfterSteps;       // SetMaxParticlesSoft is applied after this many steps (0 = at creation)
		std::vector<EmitterDesc> emitters;
	};

	ParticleInput ReadParticles(const char *path)
	{
		Reader r(path);
		ParticleInput in;
		uint32_t magic = r.get<uint32_t>(); uint32_t version = r.get<uint32_t>();
		if(magic != 0x49504e52u || version != 3) { fprintf(stderr, "bad particle file\n"); exit(2); }
		in.nEmitters = r.get<uint32_t>(); in.nSteps = r.get<uint32_t>(); in.delta = r.get<float>(); in.options = r.get<uint32_t>();
		in.softAfterSteps = r.get<uint32_t>();
		in.emitters = r.vec<EmitterDesc>(in.nEmitters);
		return in;
	}

	std::vector<GenericParticleEmitter *> BuildEmitters(const ParticleInput &in, HarnessScene *scene)
	{
		std::vector<GenericParticleEmitter *> out;
		for(const EmitterDesc &d : in.emitters)
		{
			GenericParticleEmitter *e = new GenericParticleEmitter();
			e->GetGenerator()->Seed(d.seed);
			e->SetMaxParticles(d.maxParticles);
			e->SetSpawnRate(d.spawnRate);
			e->SetLifeSpan(Vector2(d.lifeSpan[0], d.lifeSpan[1]));
			e->SetStartColor(Color(d.startColor[0], d.startColor[1], d.startColor[2], d.startColor[3]));
			e->SetEndColor(Color(d.endColor[0], d.endColor[1], d.endColor[2], d.endColor[3]));
			e->SetStartSize(Vector2(d.startSize[0], d.startSize[1]));
			e->SetEndSize(Vector2(d.endSize[0], d.endSize[1]));
			e->SetStartRotation(Vector2(d.startRotation[0], d.startRotation[1]));
			e->SetEndRotation(Vector2(d.endRotation[0], d.endRotation[1]));
			e->SetGravity(Vector3(d.gravity[0], d.gravity[1], d.gravity[2]));
			e->SetVelocity(Vector3(d.velocity[0], d.velocity[1], d.velocity[2]));
			e->SetVelocityRandomizeMin(Vector3(d.velMin[0], d.velMin[1], d.velMin[2]));
			e->SetVelocityRandomizeMax(Vector3(d.velMax[0], d.velMax[1], d.velMax[2]));
			e->SetPositionRandomizeMin(Vector3(d.posMin[0], d.posMin[1], d.posMin[2]));
			e->SetPositionRandomizeMax(Vector3(d.posMax[0], d.posMax[1], d.posMax[2]));
			e->SetCanRollParticles((in.options & kPOptRoll) != 0);
			e->SetIsRenderedInversed((in.options & kPOptInversed) != 0);
			if(d.flags & 1u)
			{
				e->SetIsLocal(false);
				e->SetWorldPosition(Vector3(d.worldPosition[0], d.worldPosition[1], d.worldPosition[2]));
				e->SetWorldScale(Vector3(d.worldScale[0], d.worldScale[1], d.worldScale[2]));
			}
			if((d.flags & 2u) && in.softAfterSteps == 0) e->SetMaxParticlesSoft(d.maxParticlesSoft);
			scene->AddNode(e);
			out.push_back(e);
		}
		return out;
	}

	int CmdParticles(const char *inPath, const char *outPath)
	{
		ParticleInput in = ReadParticles(inPath);
		HarnessScene *scene = new HarnessScene();
		std::vector<GenericParticleEmitter *> emitters = BuildEmitters(in, scene);
		for(uint32_t s = 0; s < in.nSteps; s++)
		{
			if(in.softAfterSteps && s == in.softAfterSteps)
				for(size_t k = 0; k < emitters.size(); k++)
					if(in.emitters[k].flags & 2u) emitters[k]->SetMaxParticlesSoft(in.emitters[k].maxParticlesSoft);
			scene->Update(in.delta);
		}

		Writer w(outPath);
		w.put<uint32_t>(0x4f504e52u); w.put<uint32_t>(2); w.put<uint32_t>(in.nEmitters);
		for(GenericParticleEmitter *e : emitters)
		{
			uint32_t count = static_cast<uint32_t>(e->_particles.size());
			w.put<uint32_t>(count);
			w.put<float>(e->_time);
			// per particle 24 floats: position3 velocity3 time lifespan duration sizeStart sizeEnd rotStart rotEnd | rotation size2 color4 | gravity3
			std::vector<float> rec((size_t)count * 24);
			for(uint32_t i = 0; i < count; i++)
			{
				GenericParticle *p = static_cast<GenericParticle *>(e->_particles[i]);
				float *o = &rec[(size_t)i * 24];
				o[0] = p->position.x; o[1] = p->position.y; o[2] = p->position.z;
				o[3] = p->velocity.x; o[4] = p->velocity.y; o[5] = p->velocity.z;
				o[6] = p->time; o[7] = p->lifespan; o[8] = p->sizeInterpolator._duration;
				o[9] = p->sizeInterpolator._startValue.x; o[10] = p->sizeInterpolator._endValue.x;
				o[11] = p->rotationInterpolator._startValue; o[12] = p->rotationInterpolator._endValue;
				o[13] = p->rotation; o[14] = p->size.x; o[15] = p->size.y;
				o[16] = p->color.r; o[17] = p->color.g; o[18] = p->color.b; o[19] = p->color.a;
				o[20] = p->gravity.x; o[21] = p->gravity.y; o[22] = p->gravity.z; o[23] = 0.0f;
			}
			w.vec(rec);
			w.put<uint32_t>((in.options & kPOptMesh) ? e->_maxParticles : 0u);
			if(in.options & kPOptMesh)
			{
				// the mesh ParticleEmitter::UpdateMesh built in the last Update (RNParticleEmitter.cpp:213-336,345),
				// read back through the Mesh's own chunk iterators (layout independent)
				const uint32_t nv = e->_maxParticles * 4, ni = e->_maxParticles * 6;
				Mesh::Chunk chunk = e->_mesh->GetChunk();
				Mesh::ElementIterator<Vector3> vi = chunk.GetIterator<Vector3>(Mesh::VertexAttribute::Feature::Vertices);
				Mesh::ElementIterator<Color> ci = chunk.GetIterator<Color>(Mesh::VertexAttribute::Feature::Color0);
				Mesh::ElementIterator<Vector2> ti = chunk.GetIterator<Vector2>(Mesh::VertexAttribute::Feature::UVCoords0);
				Mesh::ElementIterator<Vector2> si = chunk.GetIterator<Vector2>(Mesh::VertexAttribute::Feature::UVCoords1);
				Mesh::ElementIterator<RN::uint16> ii = chunk.GetIterator<RN::uint16>(Mesh::VertexAttribute::Feature::Indices);
				std::vector<float> vp((size_t)nv * 3), vc((size_t)nv * 4), vt((size_t)nv * 2), vs((size_t)nv * 2);
				std::vector<uint16_t> idx(ni);
				for(uint32_t v = 0; v < nv; v++)
				{
					Vector3 a = *vi++; Color c = *ci++; Vector2 t = *ti++; Vector2 z = *si++;
					vp[v*3] = a.x; vp[v*3+1] = a.y; vp[v*3+2] = a.z;
					vc[v*4] = c.r; vc[v*4+1] = c.g; vc[v*4+2] = c.b; vc[v*4+3] = c.a;
					vt[v*2] = t.x; vt[v*2+1] = t.y; vs[v*2] = z.x; vs[v*2+1] = z.y;
				}
				for(uint32_t k = 0; k < ni; k++) idx[k] = *ii++;
				w.vec(vp); w.vec(vc); w.vec(vt); w.vec(vs); w.vec(idx);
			}
		}
		return 0;
	}

	// ---- shadow cascades ---------------------------------------------------------------------------------
	// The REAL Light::UpdateShadows (Source/Scene/RNLight.cpp:475-529) and Camera::MakeShadowSplit
	// (Source/Scene/RNCamera.cpp:240-287) for a directional light.  Light::ActivateDirectionalShadows wants depth
	// textures and framebuffers from the renderer; the harness fills the members it would have set (shadow cameras as
	// it configures them, :314-326, _shadowTarget, _shadowParameter) directly and gives every shadow camera's render
	// pass the frame its framebuffer would have had (resolution x resolution).
	// in : "RNCI" v1: u32 magic, version; CameraDesc target; f32 lightRot[4]; u32 nSplits; f32 maxDistance[nSplits];
	//      f32 resolution, clipNear, clipFar, directionalShadowDistance, distanceBlendFactor, maxShadowDist
	// out: "RNCO" v1: u32 magic, version, nSplits; per split: f32 pos[3], rot[4], ortho[4] (l, r, b, t), clipNear, clipFar,
	//      u32 flags, f32 frustum[28], f32 projview[16] (_shadowCameraMatrices)
	int CmdCascades(const char *inPath, const char *outPath)
	{
		Reader r(inPath);
		uint32_t magic = r.get<uint32_t>(); uint32_t version = r.get<uint32_t>();
		if(magic != 0x49434e52u || version != 1) { fprintf(stderr, "bad cascade file\n"); return 2; }
		CameraDesc d = r.get<CameraDesc>();
		float lq[4]; for(int i = 0; i < 4; i++) lq[i] = r.get<float>();
		uint32_t nSplits = r.get<uint32_t>();
		std::vector<float> maxDistance = r.vec<float>(nSplits);
		float resolution = r.get<float>(), clipNear = r.get<float>(), clipFar = r.get<float>();
		float shadowDistance = r.get<float>(), blend = r.get<float>(), maxShadowDist = r.get<float>();

		HarnessScene *scene = new HarnessScene();
		Camera *target = new Camera();
		target->SetFlags(target->GetFlags() | d.flags);
		target->SetPosition(Vector3(d.pos[0], d.pos[1], d.pos[2]));
		target->SetRotation(Quaternion(d.rot[0], d.rot[1], d.rot[2], d.rot[3]));
		target->SetScale(Vector3(d.scale[0], d.scale[1], d.scale[2]));
		target->SetFOV(d.fov);
		target->SetAspectRatio(d.aspect);
		target->SetClipNear(d.clipNear);
		target->SetClipFar(d.clipFar);
		if(d.hasProjection)
		{
			Matrix m;
			for(int i = 0; i < 16; i++) m.m[i] = d.projection[i];
			target->SetProjectionMatrix(m);
		}
		scene->AddNode(target);

		Light *light = new Light(Light::Type::DirectionalLight);
		light->SetWorldRotation(Quaternion(lq[0], lq[1], lq[2], lq[3]));
		scene->AddNode(light);

		ShadowParameter param(target, static_cast<size_t>(resolution));
		param.splits.clear();
		for(uint32_t i = 0; i < nSplits; i++) param.splits.push_back(ShadowSplit(1, 0, 2.0f, 512.0f, maxDistance[i]));
		param.clipNear = clipNear; param.clipFar = clipFar;
		param.directionalShadowDistance = shadowDistance; param.distanceBlendFactor = blend; param.maxShadowDist = maxShadowDist;
		light->_shadowParameter = param;
		light->_shadowTarget = target->Retain();
		std::vector<Camera *> cams;
		for(uint32_t i = 0; i < nSplits; i++)
		{
			light->_shadowCameraMatrices.push_back(Matrix());
			Camera *tempcam = new Camera();
			tempcam->GetRenderPass()->SetFrame(Rect(0.0f, 0.0f, resolution, resolution));
			tempcam->SetFlags(Camera::Flags::Orthogonal | Camera::Flags::RenderEarly);
			tempcam->SetLODCamera(target);
			tempcam->SetClipNear(param.clipNear);
			tempcam->SetClipFar(param.clipFar);
			scene->AddNode(tempcam);
			light->_shadowDepthCameras.AddObject(tempcam);
			cams.push_back(tempcam);
		}
		light->UpdateShadows();

		Writer w(outPath);
		w.put<uint32_t>(0x4f434e52u); w.put<uint32_t>(1); w.put<uint32_t>(nSplits);
		for(uint32_t i = 0; i < nSplits; i++)
		{
			Camera *cam = cams[i];
			cam->PostUpdate();
			cam->UpdateFrustum();
			Vector3 p = cam->GetWorldPosition(); Quaternion q = cam->GetWorldRotation();
			w.put<float>(p.x); w.put<float>(p.y); w.put<float>(p.z);
			w.put<float>(q.x); w.put<float>(q.y); w.put<float>(q.z); w.put<float>(q.w);
			w.put<float>(cam->_orthoLeft); w.put<float>(cam->_orthoRight); w.put<float>(cam->_orthoBottom); w.put<float>(cam->_orthoTop);
			w.put<float>(cam->GetClipNear()); w.put<float>(cam->GetClipFar());
			w.put<uint32_t>(static_cast<uint32_t>(cam->GetFlags()));
			float fr[28];
			fr[0] = cam->_frustumCenter.x; fr[1] = cam->_frustumCenter.y; fr[2] = cam->_frustumCenter.z; fr[3] = cam->_frustumRadius;
			const Plane *planes[6] = { &cam->frustums._frustumLeft, &cam->frustums._frustumRight, &cam->frustums._frustumTop,
			                           &cam->frustums._frustumBottom, &cam->frustums._frustumNear, &cam->frustums._frustumFar };
			for(int k = 0; k < 6; k++)
			{
				fr[4 + k*4 + 0] = planes[k]->_normal.x; fr[4 + k*4 + 1] = planes[k]->_normal.y;
				fr[4 + k*4 + 2] = planes[k]->_normal.z; fr[4 + k*4 + 3] = planes[k]->_d;
			}
			for(int k = 0; k < 28; k++) w.put<float>(fr[k]);
			for(int k = 0; k < 16; k++) w.put<float>(light->_shadowCameraMatrices[i].m[k]);
		}
		return 0;
	}

	// ---- software occlusion culling (SceneBasic::Render with Flags::Occluder nodes, RNSceneBasic.cpp:445-573) ----------
	// Occluders are real Entities (Model + Mesh built by the reference's Mesh::WithColoredCube / WithSphereMesh); every node
	// records its submission like CullNode does.  The jitter of :488 comes from the shared generator, seeded here; a second
	// generator with the same seed is drawn in lock step so that the offsets the reference used can be written out.
	struct OccNodeBase { uint32_t id = 0; };
	static void RecordSubmission(uint32_t id, Camera *camera)
	{
		if(!g_recorder) return;
		for(size_t i = 0; i < g_recorder->cameras.size(); i++)
			if(g_recorder->cameras[i] == camera) { g_recorder->lists[i].push_back(id); break; }
	}
	class OccEntity : public Entity
	{
	public:
		explicit OccEntity(Model *model) : Entity(model) {}
		uint32_t id = 0;
		void Render(Renderer *, Camera *camera) const override { RecordSubmission(id, camera); }
	};
	class OccPlain : public SceneNode
	{
	public:
		uint32_t id = 0;
		bool CanRender(Renderer *renderer, Camera *camera) const override { return CanRenderUtil(renderer, camera); }
		void Render(Renderer *, Camera *camera) const override { RecordSubmission(id, camera); }
	};

	int CmdOcclusion(const char *scenePath, const char *occPath, const char *outPath, Renderer *renderer)
	{
		SceneInput in = ReadScene(scenePath);
		Reader r(occPath);
		uint32_t magic = r.get<uint32_t>(), version = r.get<uint32_t>();
		if(magic != 0x494f4e52u || version != 1) { fprintf(stderr, "bad occlusion file\n"); return 2; }
		uint32_t seed = r.get<uint32_t>();
		float frameW = r.get<float>(), frameH = r.get<float>();
		float camMove[3]; for(int i = 0; i < 3; i++) camMove[i] = r.get<float>();
		uint32_t nMeshes = r.get<uint32_t>();
		struct MeshDesc { uint32_t kind; float p[4]; };
		std::vector<MeshDesc> meshDescs = r.vec<MeshDesc>(nMeshes);
		std::vector<int32_t> meshOfNode = r.vec<int32_t>(in.nNodes);

		std::vector<Mesh *> meshes;
		std::vector<Model *> models;
		Material *material = Material::WithShaders(renderer->GetDefaultShader(Shader::Type::Vertex, nullptr, Shader::UsageHint::Default),
		                                           renderer->GetDefaultShader(Shader::Type::Fragment, nullptr, Shader::UsageHint::Default));
		for(const MeshDesc &md : meshDescs)
		{
			Mesh *mesh = Mesh::WithColoredCube(Vector3(md.p[0], md.p[1], md.p[2]), Color(1.0f, 1.0f, 1.0f, 1.0f));
			mesh->Retain();
			meshes.push_back(mesh);
			models.push_back(new Model(mesh, material));
		}
		// custom indexed meshes (ids continue after the cubes): positions + uint16 indices through the reference's Mesh class
		uint32_t nCustom = r.get<uint32_t>();
		for(uint32_t k = 0; k < nCustom; k++)
		{
			uint32_t nv = r.get<uint32_t>(), ni = r.get<uint32_t>();
			std::vector<float> verts = r.vec<float>((size_t)nv * 3);
			std::vector<uint16_t> idx = r.vec<uint16_t>(ni);
			Mesh *mesh = new Mesh({Mesh::VertexAttribute(Mesh::VertexAttribute::Feature::Vertices, PrimitiveType::Vector3),
			                       Mesh::VertexAttribute(Mesh::VertexAttribute::Feature::Indices, PrimitiveType::Uint16)}, nv, ni);
			mesh->BeginChanges();
			Mesh::Chunk chunk = mesh->GetChunk();
			Mesh::ElementIterator<Vector3> vit = chunk.GetIterator<Vector3>(Mesh::VertexAttribute::Feature::Vertices);
			for(uint32_t v = 0; v < nv; v++) *vit++ = Vector3(verts[v*3], verts[v*3+1], verts[v*3+2]);
			Mesh::ElementIterator<uint16> iit = chunk.GetIterator<uint16>(Mesh::VertexAttribute::Feature::Indices);
			for(uint32_t v = 0; v < ni; v++) *iit++ = idx[v];
			mesh->changedVertices = true;
			mesh->changedIndices = true;
			mesh->EndChanges();
			meshes.push_back(mesh);
			models.push_back(new Model(mesh, material));
		}

		HarnessScene *scene = new HarnessScene();
		std::vector<SceneNode *> nodes(in.nNodes);
		std::vector<SceneNode *> roots;
		for(uint32_t i = 0; i < in.nNodes; i++)
		{
			SceneNode *node;
			if(meshOfNode[i] >= 0) { OccEntity *e = new OccEntity(models[meshOfNode[i]]); e->id = i; node = e; }
			else { OccPlain *pn = new OccPlain(); pn->id = i; node = pn; }
			node->SetPosition(Vector3(in.pos[i*3], in.pos[i*3+1], in.pos[i*3+2]));
			node->SetScale(Vector3(in.scale[i*3], in.scale[i*3+1], in.scale[i*3+2]));
			node->SetRotation(Quaternion(in.rot[i*4], in.rot[i*4+1], in.rot[i*4+2], in.rot[i*4+3]));
			node->SetFlags(in.flags[i]);
			node->SetRenderGroup(static_cast<uint8>(in.group[i]));
			if(in.priority[i] != SceneNode::RenderNormal) node->SetRenderPriority(in.priority[i]);
			// the bounds of the input file win over the model's (the same columns go to the oracle and to the GPU)
			if(in.options & kOptBox)
				node->SetBoundingBox(AABB(Vector3(in.box[i*6], in.box[i*6+1], in.box[i*6+2]), Vector3(in.box[i*6+3], in.box[i*6+4], in.box[i*6+5])), false);
			else
				node->SetBoundingBox(AABB(Vector3(-1.0f), Vector3(1.0f)), false);
			node->SetBoundingSphere(Sphere(Vector3(in.sphere[i*4], in.sphere[i*4+1], in.sphere[i*4+2]), in.sphere[i*4+3]));
			nodes[i] = node;
		}
		for(uint32_t i = 0; i < in.nNodes; i++)
		{
			if(in.parent[i] >= 0) nodes[in.parent[i]]->AddChild(nodes[i]);
			else roots.push_back(nodes[i]);
		}
		for(SceneNode *root : roots) scene->AddNode(root);
		std::vector<Camera *> cameras;
		for(uint32_t c = 0; c < in.nCameras; c++)
		{
			const CameraDesc &d = in.cameras[c];
			Camera *camera = new Camera();
			camera->SetFlags(camera->GetFlags() | d.flags);
			camera->SetRenderGroup(static_cast<uint8>(d.renderGroup));
			camera->SetPosition(Vector3(d.pos[0], d.pos[1], d.pos[2]));
			camera->SetRotation(Quaternion(d.rot[0], d.rot[1], d.rot[2], d.rot[3]));
			camera->SetScale(Vector3(d.scale[0], d.scale[1], d.scale[2]));
			camera->SetFOV(d.fov);
			camera->SetAspectRatio(d.aspect);
			camera->SetClipNear(d.clipNear);
			camera->SetClipFar(d.clipFar);
			if(d.flags & Camera::Flags::Orthogonal)
				camera->SetOrthogonalFrustum(d.ortho[3], d.ortho[2], d.ortho[0], d.ortho[1]);
			camera->SetFrustumPlaneOffset(d.planeOffsets[0], d.planeOffsets[1], d.planeOffsets[2], d.planeOffsets[3]);
			camera->GetRenderPass()->SetFrame(Rect(0.0f, 0.0f, frameW, frameH));
			scene->AddNode(camera);
			cameras.push_back(camera);
		}

		RandomNumberGenerator *shared = RandomNumberGenerator::GetSharedGenerator();
		if(!shared) { fprintf(stderr, "no shared random number generator\n"); return 2; }
		shared->Seed(seed);
		RandomNumberGenerator *mirror = new RandomNumberGenerator(RandomNumberGenerator::Type::MersenneTwister);
		mirror->Seed(seed);

		Recorder rec;
		rec.cameras = cameras;
		rec.lists.resize(cameras.size());
		g_recorder = &rec;
		uint32_t frames = in.nFrames ? in.nFrames : 1;

		Writer w(outPath);
		w.put<uint32_t>(0x4f4f4e52u); w.put<uint32_t>(1);
		w.put<uint32_t>(in.nNodes); w.put<uint32_t>(in.nCameras); w.put<uint32_t>(frames);
		std::vector<uint32_t> order;
		for(IntrusiveList<SceneNode>::Member *m = scene->_renderNodes.GetHead(); m; m = m->GetNext())
		{
			SceneNode *node = m->Get();
			OccEntity *e = node->Downcast<OccEntity>();
			order.push_back(e ? e->id : static_cast<OccPlain *>(node)->id);
		}
		w.put<uint32_t>(static_cast<uint32_t>(order.size()));
		w.vec(order);
		// the triangle soups exactly as RasterizeMesh's iterator walk yields them (posA, posB, posC per triangle, :216-224)
		w.put<uint32_t>(static_cast<uint32_t>(meshes.size()));
		for(Mesh *mesh : meshes)
		{
			Mesh::Chunk chunk = mesh->GetTrianglesChunk();
			Mesh::ElementIterator<Vector3> iterator = chunk.GetIterator<Vector3>(Mesh::VertexAttribute::Feature::Vertices);
			size_t triangleCount = mesh->GetIndicesCount() / 3;
			w.put<uint32_t>(static_cast<uint32_t>(triangleCount));
			for(size_t i = 0; i < triangleCount; i++)
			{
				const Vector3 &posA = *iterator++;
				const Vector3 &posB = *iterator++;
				const Vector3 &posC = *iterator;
				if(i < triangleCount - 1) iterator++;
				w.put<float>(posA.x); w.put<float>(posA.y); w.put<float>(posA.z);
				w.put<float>(posB.x); w.put<float>(posB.y); w.put<float>(posB.z);
				w.put<float>(posC.x); w.put<float>(posC.y); w.put<float>(posC.z);
			}
		}
		for(uint32_t f = 0; f < frames; f++)
		{
			for(auto &l : rec.lists) l.clear();
			if(f > 0)
			{
				Vector3 t(in.rootTranslate[0], in.rootTranslate[1], in.rootTranslate[2]);
				for(SceneNode *root : roots) root->Translate(t);
				for(Camera *camera : cameras) camera->Translate(Vector3(camMove[0], camMove[1], camMove[2]));
			}
			scene->Update(1.0f / 60.0f);
			// which cameras will take the occlusion path, in the order SceneBasic::Render visits them (:420-439): one jitter
			// draw each
			for(Camera *camera : cameras) camera->PostUpdate();
			std::vector<uint32_t> used(cameras.size(), 0);
			std::vector<Vector3> offsets(cameras.size());
			for(int prio = 0; prio < 3; prio++)
			{
				for(IntrusiveList<Camera>::Member *cm = scene->_cameras.GetHead(); cm; cm = cm->GetNext())
				{
					Camera *camera = cm->Get();
					if(camera->GetFlags() & Camera::Flags::NoRender || (prio == 0 && !(camera->GetFlags() & Camera::Flags::RenderEarly)) ||
					   (prio == 1 && (camera->GetFlags() & Camera::Flags::RenderEarly || camera->GetFlags() & Camera::Flags::RenderLate)) ||
					   (prio == 2 && !(camera->GetFlags() & Camera::Flags::RenderLate))) continue;
					bool any = false;
					for(SceneNode *node : nodes)
						if(node->HasFlags(SceneNode::Flags::Occluder) && node->CanRender(renderer, camera)) { any = true; break; }
					if(!any) continue;
					size_t ci = std::find(cameras.begin(), cameras.end(), camera) - cameras.begin();
					used[ci] = 1;
					offsets[ci] = mirror->GetRandomVector3Range(Vector3(-0.15f, -0.15f, 0.0f), Vector3(0.15f, 0.15f, 0.0f));
				}
			}
			scene->Render(renderer);
			for(size_t c = 0; c < cameras.size(); c++)
			{
				Camera *cam = cameras[c];
				w.put<uint32_t>(used[c]);
				w.put<float>(offsets[c].x); w.put<float>(offsets[c].y); w.put<float>(offsets[c].z);
				Matrix view = cam->GetViewMatrix(), proj = cam->GetProjectionMatrix();
				for(int k = 0; k < 16; k++) w.put<float>(view.m[k]);
				for(int k = 0; k < 16; k++) w.put<float>(proj.m[k]);
				Vector3 cp = cam->GetWorldPosition();
				w.put<float>(cp.x); w.put<float>(cp.y); w.put<float>(cp.z);
				float fr[28];
				fr[0] = cam->_frustumCenter.x; fr[1] = cam->_frustumCenter.y; fr[2] = cam->_frustumCenter.z; fr[3] = cam->_frustumRadius;
				const Plane *planes[6] = { &cam->frustums._frustumLeft, &cam->frustums._frustumRight, &cam->frustums._frustumTop,
				                           &cam->frustums._frustumBottom, &cam->frustums._frustumNear, &cam->frustums._frustumFar };
				for(int k = 0; k < 6; k++)
				{
					fr[4 + k*4 + 0] = planes[k]->_normal.x; fr[4 + k*4 + 1] = planes[k]->_normal.y;
					fr[4 + k*4 + 2] = planes[k]->_normal.z; fr[4 + k*4 + 3] = planes[k]->_d;
				}
				for(int k = 0; k < 28; k++) w.put<float>(fr[k]);
				w.put<uint32_t>(static_cast<uint32_t>(cam->GetFlags()));
				w.put<uint32_t>(static_cast<uint32_t>(cam->GetRenderGroup()));
				w.put<uint32_t>(static_cast<uint32_t>(rec.lists[c].size()));
				w.vec(rec.lists[c]);
			}
			// SceneBasicInfo::occludedFrameCounter of every node and the depth buffer the LAST camera left
			std::vector<uint32_t> counters(in.nNodes);
			for(uint32_t i = 0; i < in.nNodes; i++)
				counters[i] = static_cast<uint32_t>(static_cast<SceneBasicInfo *>(nodes[i]->GetSceneInfo())->occludedFrameCounter);
			w.vec(counters);
			for(int k = 0; k < 40 * 40; k++) w.put<float>(scene->_occlusionDepthBuffer[k]);
		}
		g_recorder = nullptr;
		// the shared generator must have been drawn exactly as often as the mirror
		Vector3 a = shared->GetRandomVector3Range(Vector3(0.0f), Vector3(1.0f)), b = mirror->GetRandomVector3Range(Vector3(0.0f), Vector3(1.0f));
		if(a.x != b.x || a.y != b.y || a.z != b.z) { fprintf(stderr, "jitter draws out of step with the reference\n"); return 3; }
		return 0;
	}

	int CmdBenchParticles(const char *inPath, int steps, int warmup)
	{
		ParticleInput in = ReadParticles(inPath);
		HarnessScene *scene = new HarnessScene();
		std::vector<GenericParticleEmitter *> emitters = BuildEmitters(in, scene);
		for(int s = 0; s < warmup; s++) scene->Update(in.delta);
		double best = 1e30, total = 0.0;
		for(int s = 0; s < steps; s++)
		{
			auto t0 = std::chrono::steady_clock::now();
			scene->Update(in.delta);
			auto t1 = std::chrono::steady_clock::now();
			double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
			best = std::min(best, ms); total += ms;
		}
		size_t live = 0;
		for(GenericParticleEmitter *e : emitters) live += e->_particles.size();
		printf("{\"kind\": \"reference\", \"what\": \"particles\", \"emitters\": %u, \"live\": %zu, \"steps\": %d, \"ms_best\": %.6f, \"ms_mean\": %.6f}\n",
		       in.nEmitters, live, steps, best, total / steps);
		return 0;
	}
}

static void OnSegv(int)
{
	void *frames[48];
	int n = backtrace(frames, 48);
	backtrace_symbols_fd(frames, n, 2);
	_Exit(139);
}

int main(int argc, char **argv)
{
	signal(SIGSEGV, OnSegv);
	if(argc < 3)
	{
		fprintf(stderr, "usage: %s scene|particles|bench-scene|bench-particles|cascades <in> ...\n", argv[0]);
		return 2;
	}
	// Minimal bootstrap: no Kernel, no manifest, no files (reference: RNBase.cpp:164-176, RNWorkQueue.cpp:65-71).
	RN::__functionPool = new MemoryPool();
	AutoreleasePool *pool = new AutoreleasePool(); // before anything that autoreleases (RNCSTR in InitializeQueues)
	WorkQueue::InitializeQueues();
	NullRenderer *renderer = new NullRenderer();
	renderer->Activate();

	std::string cmd = argv[1];
	int rc = 2;
	if(cmd == "scene" && argc >= 4) rc = CmdScene(argv[2], argv[3], renderer);
	else if(cmd == "particles" && argc >= 4) rc = CmdParticles(argv[2], argv[3]);
	else if(cmd == "bench-scene" && argc >= 5) rc = CmdBenchScene(argv[2], atoi(argv[3]), atoi(argv[4]), renderer);
	else if(cmd == "bench-particles" && argc >= 5) rc = CmdBenchParticles(argv[2], atoi(argv[3]), atoi(argv[4]));
	else if(cmd == "cascades" && argc >= 4) rc = CmdCascades(argv[2], argv[3]);
	else if(cmd == "occlusion" && argc >= 5) rc = CmdOcclusion(argv[2], argv[3], argv[4], renderer);
	else fprintf(stderr, "bad arguments\n");
	fflush(stdout);
	(void)pool;
	_Exit(rc); // skip static destructors of the half-booted engine
}
