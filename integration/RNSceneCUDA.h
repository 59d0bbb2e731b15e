// RN::SceneCUDA - the Rayne-side binding of librn_cuda: a RN::Scene subclass (the extension point Rayne documents itself,
// Source/Scene/RNScene.h:30-54) that keeps RN::SceneNode / RN::Camera objects as the front end and replaces the two CPU loops
// of RN::SceneBasic (Source/Scene/RNSceneBasic.cpp:48-118 Update, :407-650 Render) by calls into the C ABI of
// include/rn_cuda.h.  This is the file a Rayne maintainer adds as Modules/CUDAScene; it is compiled here against the
// reference's own headers and library (oracle/build_ref.sh shim) and exercised by integration/shim_test.cpp, which drives
// the same scene through RN::SceneBasic and through this class and compares what they submit.
//
// Not part of librn_cuda.so (which has no Rayne dependency).
#pragma once
#include "Rayne.h"

extern "C" {
#include "rn_cuda.h"
}

#include <mutex>
#include <vector>

namespace RN
{
	class SceneCUDA;

	// per-node record, like SceneBasicInfo (Source/Scene/RNSceneBasic.h:60-68)
	class SceneCUDAInfo : public SceneInfo
	{
	public:
		SceneCUDAInfo(Scene *scene) : SceneInfo(scene), rank(0), slot(0xffffffffu), dirty(false) {}
		uint32 rank;     // row of the node on the device (rn_scene_download addresses rows)
		uint32 slot;     // the node's device slot: what list entries name; 0xffffffff = not on the device yet
		bool dirty;      // local state changed since the last upload
	};

	// Setter hook: SceneNode::DidUpdate forwards every ChangeSet to the node's attachments (Source/Scene/RNSceneNode.cpp:
	// 492-497), which is the public way to observe SetPosition / SetRotation / SetScale / SetFlags / AddChild on a plain
	// SceneNode.  One instance per node, added by SceneCUDA::AddNode.
	class SceneCUDAHook : public SceneNodeAttachment
	{
	public:
		SceneCUDAHook(SceneCUDA *scene) : _scene(scene) {}
	protected:
		void DidUpdate(SceneNode::ChangeSet changeSet) override;
	private:
		SceneCUDA *_scene;
	};

	class SceneCUDA : public Scene
	{
	public:
		friend class SceneCUDAHook;

		// strictMatrices: world matrices bit-exact with Rayne's strict-FP build (RN_SCENE_STRICT_MATRICES); the default
		// uses fused multiply-adds on the matrix chain (within 1e-5)
		SceneCUDA(uint32 maxNodes, int device = 0, bool strictMatrices = false);
		~SceneCUDA() override;

		void AddNode(SceneNode *node) override;
		void RemoveNode(SceneNode *node) override;

		// what SceneNode::GetWorldTransform / GetWorldPosition would return, read back from the device
		Matrix GetWorldTransformOf(SceneNode *node);
		Vector3 GetWorldPositionOf(SceneNode *node);

		size_t GetNodeCount() const { return _nodeCount; }
		rn_scene_t *GetHandle() const { return _scene; }
		uint64 GetFullUploads() const { return _fullUploads; }
		uint64 GetIncrementalEdits() const { return _incrementalEdits; }
		uint64 GetDirtyRowsUploaded() const { return _dirtyRows; }

	protected:
		void Update(float delta) override;
		void Render(Renderer *renderer) override;

	private:
		void FlushDeletionQueue();
		void AddRenderNode(SceneNode *node);
		void MarkDirty(SceneNode *node, SceneNode::ChangeSet changeSet);
		void UploadAll();
		bool ApplyEdits();
		void UploadDirty();
		void FillRow(SceneNode *node, size_t row);
		static void Check(int status);

		rn_scene_t *_scene;
		uint32 _maxNodes;
		IntrusiveList<SceneNode> _renderNodes;   // the render list, kept exactly as SceneBasic keeps it (priority insert rules)
		size_t _nodeCount;
		std::vector<SceneNode *> _toRemove;
		std::vector<SceneNode *> _dirty;
		IntrusiveList<Camera> _cameras;          // PushFront like SceneBasic (:719)
		std::vector<SceneNode *> _nodeOfSlot;    // device slot -> node (what a list entry refers to)
		std::vector<uint32> _slotOfRank;
		bool _structureDirty;                    // something happened that only a full upload can express
		struct Edit { bool add; SceneNode *node; uint32 slot; };
		std::vector<Edit> _edits;                // AddRenderNode / removal events since the last upload, in order
		std::mutex _lock;
		uint64 _fullUploads, _dirtyRows, _incrementalEdits;

		// host staging of the node columns (rank order)
		std::vector<float> _pos, _scale, _rot, _sphere, _box;
		std::vector<uint32> _parent, _flags, _group;
		std::vector<int32> _prio;
		std::vector<uint32> _listScratch[RN_MAX_CAMERAS];
	};
}
