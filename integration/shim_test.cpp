// TEST INFRASTRUCTURE - drives the same seeded scene through RN::SceneBasic (the unmodified reference) and through
// RN::SceneCUDA (integration/RNSceneCUDA.cpp -> librn_cuda.so) for a scripted sequence of frames - setters, AddNode,
// RemoveNode, AddChild (reparent), SetRenderPriority - and compares, per frame, what the two scenes submit per camera
// (node ids in order) and every node's world transform.  Prints one JSON line, exit code 0 = identical.
//
//   shim_test [roots] [frames]
//
// Built by oracle/build_ref.sh (target `shim`) against oracle/_ref/librayne_strict.a; needs a CUDA device at run time.
#include "Rayne.h"
#include "Base/RNBase.cpp"

#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <vector>
#include <algorithm>
#include <execinfo.h>
#include <signal.h>
#include <unistd.h>

using namespace RN;

#include "../oracle/ref_null_renderer.h"
#include "RNSceneCUDA.h"

namespace
{
	struct Recorder
	{
		std::vector<Camera *> cameras;
		std::vector<std::vector<uint32_t>> lists;
	};
	Recorder *g_recorder = nullptr;

	class TestNode : public SceneNode
	{
	public:
		uint32_t id = 0;
		bool CanRender(Renderer *renderer, Camera *camera) const override { return CanRenderUtil(renderer, camera); }
		void Render(Renderer *, Camera *camera) const override
		{
			if(!g_recorder) return;
			for(size_t i = 0; i < g_recorder->cameras.size(); i++)
				if(g_recorder->cameras[i] == camera) { g_recorder->lists[i].push_back(id); break; }
		}
	};

	struct Lcg
	{
		uint32_t x;
		explicit Lcg(uint32_t seed) : x(seed) {}
		float next() { x = x * 1664525u + 1013904223u; return static_cast<float>(x >> 8) / 16777216.0f; }
		float range(float lo, float hi) { return lo + (hi - lo) * next(); }
		uint32_t below(uint32_t n) { return static_cast<uint32_t>(next() * n) % n; }
	};

	// one of the two twin worlds
	struct World
	{
		Scene *scene = nullptr;
		std::vector<TestNode *> nodes;     // by id
		std::vector<TestNode *> roots;
		std::vector<Camera *> cameras;
		Recorder rec;
	};

	TestNode *MakeNode(World &w, Lcg &rng, int level)
	{
		static const float edge[] = {300.0f, 40.0f, 10.0f, 3.0f, 1.0f};
		TestNode *node = new TestNode();
		node->id = static_cast<uint32_t>(w.nodes.size());
		const float e = edge[level < 5 ? level : 4];
		node->SetPosition(Vector3(rng.range(-e, e), rng.range(-e, e), rng.range(-e, e)));
		node->SetRotation(Quaternion(Vector3(rng.range(0.0f, 360.0f), rng.range(0.0f, 360.0f), rng.range(0.0f, 360.0f))));
		const float s = rng.range(0.8f, 1.25f);
		node->SetScale(rng.next() < 0.2f ? Vector3(s, rng.range(0.7f, 1.3f), rng.range(0.7f, 1.3f)) : Vector3(s, s, s));
		const float f = rng.next();
		if(f < 0.03f) node->SetFlags(SceneNode::Flags::Hidden);
		else if(f < 0.06f) node->SetFlags(SceneNode::Flags::NoCulling);
		if(rng.next() < 0.1f) node->SetRenderGroup(static_cast<uint8>(1 + rng.below(3)));
		const float pr = rng.next();
		if(pr < 0.05f) node->SetRenderPriority(SceneNode::RenderTransparent);
		else if(pr < 0.08f) node->SetRenderPriority(SceneNode::RenderEarly);
		else if(pr < 0.10f) node->SetRenderPriority(SceneNode::RenderLate);   // (RenderSky and above compare equal under SortFrontToBack:
		                                                                        //  std::sort leaves their order unspecified, nothing to compare)
		if(rng.next() < 0.3f)
			node->SetBoundingSphere(Sphere(Vector3(rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f), rng.range(-1.0f, 1.0f)), rng.range(0.3f, 4.0f)));
		w.nodes.push_back(node);
		return node;
	}

	TestNode *MakeTree(World &w, Lcg &rng, int level, int depth)
	{
		TestNode *node = MakeNode(w, rng, level);
		if(level + 1 < depth)
		{
			const uint32_t kids = 1 + rng.below(4);
			for(uint32_t k = 0; k < kids; k++)
				node->AddChild(MakeTree(w, rng, level + 1, depth));
		}
		return node;
	}

	void BuildWorld(World &w, uint32_t nRoots, uint32_t seed)
	{
		Lcg rng(seed);
		for(uint32_t r = 0; r < nRoots; r++)
		{
			TestNode *root = MakeTree(w, rng, 0, 2 + static_cast<int>(rng.below(4)));   // complete before it enters the scene
			w.roots.push_back(root);
			w.scene->AddNode(root);
		}
		const float camPos[4][3] = {{0, 0, 0}, {40, 10, -20}, {-100, 30, 50}, {5, 5, 5}};
		for(int c = 0; c < 4; c++)
		{
			Camera *camera = new Camera();
			camera->SetPosition(Vector3(camPos[c][0], camPos[c][1], camPos[c][2]));
			camera->SetRotation(Quaternion(Vector3(30.0f * c, -10.0f * c, 0.0f)));
			camera->SetAspectRatio(16.0f / 9.0f);      // without it UpdateProjection asks the (absent) window for its size
			if(c == 1) { camera->SetFOV(95.0f); camera->SetClipFar(250.0f); camera->SetRenderGroup(3); }
			if(c == 2) camera->SetFlags(camera->GetFlags() | Camera::Flags::UseSimpleCulling | Camera::Flags::RenderEarly);
			if(c == 3) camera->SetFlags(camera->GetFlags() | Camera::Flags::SortFrontToBack | Camera::Flags::RenderLate);
			w.scene->AddNode(camera);
			w.cameras.push_back(camera);
		}
		w.rec.cameras = w.cameras;
		w.rec.lists.resize(w.cameras.size());
	}

	// the scripted edits of frame f, applied identically to both worlds (same ids, same generator state)
	void Mutate(World &w, uint32_t f, uint32_t seed)
	{
		Lcg rng(seed * 977u + f);
		const uint32_t n = static_cast<uint32_t>(w.nodes.size());
		for(TestNode *root : w.roots)
			if(root->GetSceneInfo()) root->Translate(Vector3(0.25f, 0.0f, -0.125f));
		for(uint32_t k = 0; k < n / 20 + 1; k++)
		{
			TestNode *node = w.nodes[rng.below(n)];
			switch(rng.below(5))
			{
				case 0: node->SetRotation(Quaternion(Vector3(rng.range(0.0f, 360.0f), rng.range(0.0f, 360.0f), 0.0f))); break;
				case 1: node->SetScale(Vector3(rng.range(0.5f, 1.5f))); break;
				case 2: node->SetFlags(node->GetFlags() ^ SceneNode::Flags::Hidden); break;
				case 3: node->SetPosition(node->GetPosition() + Vector3(rng.range(-2.0f, 2.0f), 0.0f, rng.range(-2.0f, 2.0f))); break;
				case 4: node->SetBoundingSphere(Sphere(Vector3(0.0f), rng.range(0.2f, 6.0f))); break;
			}
		}
		if(f % 8 == 1)
		{
			// a new tree, complete before AddNode
			TestNode *root = MakeTree(w, rng, 0, 3);
			w.roots.push_back(root);
			w.scene->AddNode(root);
		}
		if(f % 8 == 3)
		{
			// a root (with its subtree) leaves the scene
			TestNode *root = w.roots[rng.below(static_cast<uint32_t>(w.roots.size()))];
			if(root->GetSceneInfo() && !root->_scheduledForRemovalFromScene) w.scene->RemoveNode(root);
		}
		if(f % 8 == 5)
		{
			// reparent: a non-root node moves under another in-scene node that is not one of its descendants
			for(int tries = 0; tries < 20; tries++)
			{
				TestNode *a = w.nodes[rng.below(n)], *b = w.nodes[rng.below(n)];
				if(a == b || !a->GetSceneInfo() || !b->GetSceneInfo() || !a->GetParent()) continue;
				bool cycle = false;
				for(SceneNode *p = b; p; p = p->GetParent()) if(p == a) cycle = true;
				if(cycle) continue;
				b->AddChild(a);
				break;
			}
			// a fresh child attached to an in-scene parent enters the scene through AddChild
			TestNode *parent = w.nodes[rng.below(n)];
			if(parent->GetSceneInfo()) parent->AddChild(MakeNode(w, rng, 3));
		}
		if(f % 8 == 7)
		{
			TestNode *node = w.nodes[rng.below(n)];
			if(node->GetSceneInfo() && !node->_scheduledForRemovalFromScene)
				node->SetRenderPriority(node->GetRenderPriority() == SceneNode::RenderNormal ? SceneNode::RenderLate : SceneNode::RenderNormal);
		}
		for(Camera *camera : w.cameras) camera->Translate(Vector3(0.5f, 0.0f, 0.25f));
	}

	void RunFrame(World &w, Renderer *renderer)
	{
		for(auto &l : w.rec.lists) l.clear();
		g_recorder = &w.rec;
		w.scene->Update(1.0f / 60.0f);
		w.scene->Render(renderer);
		g_recorder = nullptr;
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
	const uint32_t nRoots = argc > 1 ? static_cast<uint32_t>(atoi(argv[1])) : 60;
	const uint32_t frames = argc > 2 ? static_cast<uint32_t>(atoi(argv[2])) : 13;
	RN::__functionPool = new MemoryPool();
	AutoreleasePool *pool = new AutoreleasePool();
	WorkQueue::InitializeQueues();
	NullRenderer *renderer = new NullRenderer();
	renderer->Activate();

	World ref, gpu;
	ref.scene = new SceneBasic();
	SceneCUDA *cuda = nullptr;
	try
	{
		cuda = new SceneCUDA(1u << 20, 0, true);
	}
	catch(Exception &e)
	{
		printf("{\"ok\": false, \"error\": \"%s\"}\n", e.GetReason().c_str());
		fflush(stdout);
		_Exit(3);
	}
	gpu.scene = cuda;
	BuildWorld(ref, nRoots, 2024);
	BuildWorld(gpu, nRoots, 2024);

	uint64_t entries = 0, listMismatch = 0, matrixMismatch = 0, matricesCompared = 0;
	int firstBadFrame = -1;
	for(uint32_t f = 0; f < frames; f++)
	{
		if(f > 0) { Mutate(ref, f, 5); Mutate(gpu, f, 5); }
		RunFrame(ref, renderer);
		RunFrame(gpu, renderer);
		bool bad = ref.nodes.size() != gpu.nodes.size();
		for(size_t c = 0; c < ref.cameras.size(); c++)
		{
			entries += ref.rec.lists[c].size();
			if(ref.cameras[c]->GetFlags() & Camera::Flags::SortFrontToBack)
			{
				// std::sort leaves entries with equal (priority, squared distance) in unspecified order: order such runs by id
				// in both lists before comparing
				const Vector3 cp = ref.cameras[c]->GetWorldPosition();
				auto canon = [&](std::vector<uint32_t> &l) {
					auto key = [&](uint32_t id, int32 &prio, float &d) {
						prio = ref.nodes[id]->GetRenderPriority();
						d = ref.nodes[id]->GetWorldPosition().GetSquaredDistance(cp);
					};
					size_t a = 0;
					while(a < l.size())
					{
						int32 pa; float da; key(l[a], pa, da);
						size_t b = a + 1;
						while(b < l.size()) { int32 pb; float db; key(l[b], pb, db); if(pb != pa || db != da) break; b++; }
						std::sort(l.begin() + a, l.begin() + b);
						a = b;
					}
				};
				canon(ref.rec.lists[c]);
				canon(gpu.rec.lists[c]);
			}
			if(ref.rec.lists[c] != gpu.rec.lists[c])
			{
				listMismatch++; bad = true;
				if(listMismatch <= 3)
				{
					size_t k = 0;
					while(k < ref.rec.lists[c].size() && k < gpu.rec.lists[c].size() && ref.rec.lists[c][k] == gpu.rec.lists[c][k]) k++;
					fprintf(stderr, "frame %u camera %zu: %zu vs %zu entries, first difference at %zu\n", f, c, ref.rec.lists[c].size(), gpu.rec.lists[c].size(), k);
				}
			}
		}
		for(size_t i = 0; i < ref.nodes.size() && i < gpu.nodes.size(); i++)
		{
			if(!ref.nodes[i]->GetSceneInfo()) continue;         // not in the scene (any more)
			if(!gpu.nodes[i]->GetSceneInfo()) { matrixMismatch++; bad = true; continue; }
			const Matrix a = ref.nodes[i]->GetWorldTransform();
			const Matrix b = cuda->GetWorldTransformOf(gpu.nodes[i]);
			matricesCompared++;
			for(int k = 0; k < 16; k++)
				if(!(a.m[k] == b.m[k])) { matrixMismatch++; bad = true; break; }   // == : +0 and -0 compare equal
		}
		if(bad && firstBadFrame < 0) firstBadFrame = static_cast<int>(f);
	}
	const bool ok = listMismatch == 0 && matrixMismatch == 0;
	printf("{\"ok\": %s, \"frames\": %u, \"nodes_created\": %zu, \"nodes_in_scene\": %zu, \"cameras\": %zu, \"list_entries_compared\": %llu, "
	       "\"matrices_compared\": %llu, \"list_mismatches\": %llu, \"matrix_mismatches\": %llu, \"first_bad_frame\": %d, "
	       "\"full_uploads\": %llu, \"incremental_edits\": %llu, \"dirty_rows_uploaded\": %llu}\n",
	       ok ? "true" : "false", frames, ref.nodes.size(), cuda->GetNodeCount(), ref.cameras.size(), (unsigned long long)entries,
	       (unsigned long long)matricesCompared, (unsigned long long)listMismatch, (unsigned long long)matrixMismatch, firstBadFrame,
	       (unsigned long long)cuda->GetFullUploads(), (unsigned long long)cuda->GetIncrementalEdits(), (unsigned long long)cuda->GetDirtyRowsUploaded());
	fflush(stdout);
	(void)pool;
	_Exit(ok ? 0 : 1);
}
