// RN::SceneCUDA - see RNSceneCUDA.h.  Reference behaviour mirrored here, call by call:
//   SceneBasic::AddNode / RemoveNode / FlushDeletionQueue   Source/Scene/RNSceneBasic.cpp:683-746, :120-150
//   SceneBasic::Update                                      :48-118   (node->Update for non-static nodes)
//   SceneBasic::Render                                      :407-650  (camera order, cull loop, sorts, SubmitCamera)
#include "RNSceneCUDA.h"

#include <algorithm>
#include <cstring>

namespace RN
{
	void SceneCUDAHook::DidUpdate(SceneNode::ChangeSet changeSet)
	{
		_scene->MarkDirty(GetParent(), changeSet);
	}

	void SceneCUDA::Check(int status)
	{
		// the C ABI never throws; the binding maps a failure to Rayne's exception type
		if(status != RN_OK)
			throw InconsistencyException(RNSTR("librn_cuda: " << rn_last_error()));
	}

	SceneCUDA::SceneCUDA(uint32 maxNodes, int device, bool strictMatrices) :
		_scene(nullptr), _maxNodes(maxNodes), _nodeCount(0), _structureDirty(true), _fullUploads(0), _dirtyRows(0), _incrementalEdits(0)
	{
		rn_scene_desc desc;
		memset(&desc, 0, sizeof(desc));
		desc.device = device;
		desc.max_nodes = maxNodes;
		desc.flags = strictMatrices ? RN_SCENE_STRICT_MATRICES : 0;
		Check(rn_scene_create(&_scene, &desc));
	}

	SceneCUDA::~SceneCUDA()
	{
		rn_scene_destroy(_scene);
	}

	// Same contract as SceneBasic::AddNode: the node is retained, gets its scene info (which adds the children FIRST,
	// SceneNode::UpdateSceneInfo, Source/Scene/RNSceneNode.cpp:270-294), then enters the render list.
	void SceneCUDA::AddNode(SceneNode *node)
	{
if(node->GetSceneInfo() && node->GetSceneInfo()->GetScene() == this && node->_scheduledForRemovalFromScene)
		{
			// re-added before the deletion queue was flushed (what SetRenderPriority does on purpose, RNSceneNode.cpp:214-219):
			// SceneBasic takes the node out of the list, lets its children leave and re-enter the scene through the scene info
			// and inserts the node again (RNSceneBasic.cpp:686-707) - the whole subtree moves to where fresh inserts put it
			{
				std::lock_guard<std::mutex> guard(_lock);
				_toRemove.erase(std::remove(_toRemove.begin(), _toRemove.end(), node), _toRemove.end());
			}
			const bool isCamera = node->IsKindOfClass(Camera::GetMetaClass());
			SceneInfo *info = node->GetSceneInfo();
			info->Retain();
			if(!isCamera)
				_renderNodes.Erase(node->_sceneRenderEntry);
			node->UpdateSceneInfo(nullptr);
			node->UpdateSceneInfo(info);
			if(!isCamera)
				AddRenderNode(node);
			info->Release();
			node->_scheduledForRemovalFromScene = false;
			_structureDirty = true;
			return;
		}
		RN_ASSERT(node->GetSceneInfo() == nullptr, "AddNode() must be called on a Node not owned by a scene");
		node->Retain();
		SceneCUDAInfo *info = new SceneCUDAInfo(this);
		node->UpdateSceneInfo(info);
		info->Release();
		std::lock_guard<std::mutex> guard(_lock);
		if(node->IsKindOfClass(Camera::GetMetaClass()))
		{
			_cameras.PushFront(static_cast<Camera *>(node)->_cameraSceneEntry);
		}
		else
		{
			AddRenderNode(node);
			_nodeCount++;
			SceneCUDAHook *hook = new SceneCUDAHook(this);
			node->AddAttachment(hook);
			hook->Release();
			_edits.push_back(Edit{true, node, 0});      // applied by the next Update / Render (rn_scene_add_nodes)
		}
	}

	// Where a node enters the render list (SceneBasic::AddRenderNode, RNSceneBasic.cpp:652-676): the list stays sorted by
	// render priority; a node whose priority does not exceed the head's goes to the very front, one whose priority is not
	// below the tail's to the very back, anything else behind the last entry whose priority is not larger.
	void SceneCUDA::AddRenderNode(SceneNode *node)
	{
		const int32 priority = node->GetRenderPriority();
		IntrusiveList<SceneNode>::Member *head = _renderNodes.GetHead(), *tail = _renderNodes.GetTail();
		if(!head || priority <= head->Get()->GetRenderPriority())
			_renderNodes.PushFront(node->_sceneRenderEntry);
		else if(priority >= tail->Get()->GetRenderPriority())
			_renderNodes.PushBack(node->_sceneRenderEntry);
		else
		{
			IntrusiveList<SceneNode>::Member *at = tail;
			while(at->Get()->GetRenderPriority() > priority)
				at = at->GetPrevious();
			_renderNodes.InsertAfter(node->_sceneRenderEntry, at);
		}
	}

	void SceneCUDA::RemoveNode(SceneNode *node)
	{
		RN_ASSERT(node->GetSceneInfo() && node->GetSceneInfo()->GetScene() == this && node->_scheduledForRemovalFromScene == false,
		          "RemoveNode() must be called on a Node owned by the scene");
		std::lock_guard<std::mutex> guard(_lock);
		_toRemove.push_back(node);
		node->_scheduledForRemovalFromScene = true;
	}

	void SceneCUDA::FlushDeletionQueue()
	{
		// UpdateSceneInfo(nullptr) calls RemoveNode for every child (RNSceneNode.cpp:274-281), which queues more work: run
		// until the queue stays empty (SceneBasic's Enumerate sees the entries appended while it runs)
		for(;;)
		{
			std::vector<SceneNode *> queue;
			{
				std::lock_guard<std::mutex> guard(_lock);
				queue.swap(_toRemove);
			}
			if(queue.empty())
				break;
			for(SceneNode *node : queue)
			{
				if(node->IsKindOfClass(Camera::GetMetaClass()))
					_cameras.Erase(static_cast<Camera *>(node)->_cameraSceneEntry);
				else
				{
					_renderNodes.Erase(node->_sceneRenderEntry);
					_nodeCount--;
					const uint32 slot = static_cast<SceneCUDAInfo *>(node->GetSceneInfo())->slot;
					if(slot != 0xffffffffu)
						_edits.push_back(Edit{false, node, slot});
					else      // never reached the device: forget its pending addition
						_edits.erase(std::remove_if(_edits.begin(), _edits.end(), [node](const Edit &e) { return e.add && e.node == node; }), _edits.end());
				}
				_dirty.erase(std::remove(_dirty.begin(), _dirty.end(), node), _dirty.end());
				node->UpdateSceneInfo(nullptr);
				node->_scheduledForRemovalFromScene = false;
				node->Autorelease();
			}
		}
	}

	void SceneCUDA::MarkDirty(SceneNode *node, SceneNode::ChangeSet changeSet)
	{
		if(!node)
			return;
		SceneCUDAInfo *info = static_cast<SceneCUDAInfo *>(node->GetSceneInfo());
		if(!info || info->GetScene() != this)
			return;
		if(changeSet & (SceneNode::ChangeSet::Parent | SceneNode::ChangeSet::RenderPriority))
			_structureDirty = true;          // the hierarchy / the render list changed: next Update uploads everything
		if((changeSet & (SceneNode::ChangeSet::Position | SceneNode::ChangeSet::Flags | SceneNode::ChangeSet::Generic)) && !info->dirty)
		{
			// SceneNode::DidUpdate(Position) also walks the whole subtree (RNSceneNode.cpp:463-467) and so reaches the hooks
			// of all descendants; their LOCAL state did not change, but telling the two apart needs the old values - the
			// rows are uploaded again (40 bytes each), the device recomputes every world transform each frame anyway
			info->dirty = true;
			_dirty.push_back(node);
		}
	}

	void SceneCUDA::FillRow(SceneNode *node, size_t r)
	{
		const Vector3 p = node->GetPosition(), s = node->GetScale();
		const Quaternion q = node->GetRotation();
		_pos[r * 3] = p.x; _pos[r * 3 + 1] = p.y; _pos[r * 3 + 2] = p.z;
		_scale[r * 3] = s.x; _scale[r * 3 + 1] = s.y; _scale[r * 3 + 2] = s.z;
		_rot[r * 4] = q.x; _rot[r * 4 + 1] = q.y; _rot[r * 4 + 2] = q.z; _rot[r * 4 + 3] = q.w;
		_flags[r] = static_cast<uint32>(node->GetFlags());
		_group[r] = node->GetRenderGroup();
		// the LOCAL bounding volumes (SetBoundingSphere / SetBoundingBox); the world ones are derived on the device
		const Sphere &bs = node->_boundingSphere;
		_sphere[r * 4] = bs.offset.x; _sphere[r * 4 + 1] = bs.offset.y; _sphere[r * 4 + 2] = bs.offset.z; _sphere[r * 4 + 3] = bs.radius;
		const AABB &bb = node->_boundingBox;
		_box[r * 6] = bb.minExtend.x; _box[r * 6 + 1] = bb.minExtend.y; _box[r * 6 + 2] = bb.minExtend.z;
		_box[r * 6 + 3] = bb.maxExtend.x; _box[r * 6 + 4] = bb.maxExtend.y; _box[r * 6 + 5] = bb.maxExtend.z;
	}

	void SceneCUDA::UploadAll()
	{
		const size_t n = _nodeCount;
		RN_ASSERT(n <= _maxNodes, "more nodes than the SceneCUDA was created for");
		// rank = position in the render list, head to tail
		std::vector<SceneNode *> byRank;
		byRank.reserve(n);
		for(IntrusiveList<SceneNode>::Member *m = _renderNodes.GetHead(); m; m = m->GetNext())
		{
			static_cast<SceneCUDAInfo *>(m->Get()->GetSceneInfo())->rank = static_cast<uint32>(byRank.size());
			byRank.push_back(m->Get());
		}
		_pos.resize(n * 3); _scale.resize(n * 3); _rot.resize(n * 4); _sphere.resize(n * 4); _box.resize(n * 6);
		_parent.resize(n); _flags.resize(n); _group.resize(n); _prio.resize(n);
		for(size_t r = 0; r < n; r++)
		{
			SceneNode *node = byRank[r];
			FillRow(node, r);
			SceneNode *p = node->GetParent();
			SceneCUDAInfo *pinfo = p ? static_cast<SceneCUDAInfo *>(p->GetSceneInfo()) : nullptr;
			// a parent that is not a render node of this scene (a camera, a node outside the scene) cannot be propagated on
			// the device: such children would need their parent's world transform fed in as their own root transform
			RN_ASSERT(!p || (pinfo && pinfo->GetScene() == this && !p->IsKindOfClass(Camera::GetMetaClass())), "SceneCUDA: parent outside the render list");
			_parent[r] = p ? pinfo->rank : RN_NO_PARENT;
			_prio[r] = node->GetRenderPriority();
			static_cast<SceneCUDAInfo *>(node->GetSceneInfo())->dirty = false;
		}
		_dirty.clear();
		if(n)
		{
			rn_node_columns cols;
			memset(&cols, 0, sizeof(cols));
			cols.position = _pos.data(); cols.scale = _scale.data(); cols.rotation = _rot.data(); cols.parent = _parent.data();
			cols.flags = _flags.data(); cols.render_group = _group.data(); cols.bounding_sphere = _sphere.data();
			cols.bounding_box = _box.data(); cols.render_priority = _prio.data();
			Check(rn_scene_upload(_scene, &cols, 0, static_cast<uint32_t>(n)));
			_slotOfRank.resize(n);
			Check(rn_scene_download(_scene, RN_DL_SLOT_OF_RANK, 0, static_cast<uint32_t>(n), _slotOfRank.data()));
			_nodeOfSlot.assign(n, nullptr);
			for(size_t r = 0; r < n; r++)
			{
				_nodeOfSlot[_slotOfRank[r]] = byRank[r];
				static_cast<SceneCUDAInfo *>(byRank[r]->GetSceneInfo())->slot = _slotOfRank[r];
			}
		}
		_edits.clear();
		_structureDirty = false;
		_fullUploads++;
	}

	// The AddNode / RemoveNode events since the last upload, replayed in order through the incremental entry points: runs
	// of additions with one render priority become one rn_scene_add_nodes (the library places the run where
	// AddRenderNode's rules put it), runs of removals one rn_scene_remove_nodes.  Returns false when the library wants a
	// full upload instead (RN_ERR_UNSUPPORTED / RN_ERR_CAPACITY) - the render list kept here is authoritative either way.
	bool SceneCUDA::ApplyEdits()
	{
		size_t at = 0;
		while(at < _edits.size())
		{
			size_t end = at + 1;
			if(!_edits[at].add)
			{
				std::vector<uint32> slots;
				for(end = at; end < _edits.size() && !_edits[end].add; end++)
					slots.push_back(_edits[end].slot);
				if(!slots.empty() && rn_scene_remove_nodes(_scene, slots.data(), static_cast<uint32_t>(slots.size())) != RN_OK)
					return false;
				for(uint32 slot : slots) _nodeOfSlot[slot] = nullptr;
				_incrementalEdits++;
				at = end;
				continue;
			}
			const int32 priority = _edits[at].node->GetRenderPriority();
			while(end < _edits.size() && _edits[end].add && _edits[end].node->GetRenderPriority() == priority) end++;
			const size_t m = end - at;
			_pos.resize(std::max(_pos.size(), m * 3)); _scale.resize(std::max(_scale.size(), m * 3)); _rot.resize(std::max(_rot.size(), m * 4));
			_sphere.resize(std::max(_sphere.size(), m * 4)); _box.resize(std::max(_box.size(), m * 6));
			_flags.resize(std::max(_flags.size(), m)); _group.resize(std::max(_group.size(), m));
			_parent.resize(std::max(_parent.size(), m)); _prio.resize(std::max(_prio.size(), m));
			// batch index of every node of the run, to resolve parents inside it
			for(size_t i = 0; i < m; i++)
				static_cast<SceneCUDAInfo *>(_edits[at + i].node->GetSceneInfo())->rank = static_cast<uint32>(i);
			for(size_t i = 0; i < m; i++)
			{
				SceneNode *node = _edits[at + i].node;
				if(!node->GetSceneInfo() || node->GetSceneInfo()->GetScene() != this) return false;
				FillRow(node, i);
				_prio[i] = priority;
				SceneNode *p = node->GetParent();
				if(!p)
					_parent[i] = RN_NO_PARENT;
				else
				{
					SceneCUDAInfo *pinfo = static_cast<SceneCUDAInfo *>(p->GetSceneInfo());
					if(!pinfo || pinfo->GetScene() != this || p->IsKindOfClass(Camera::GetMetaClass())) return false;
					if(pinfo->slot != 0xffffffffu) _parent[i] = RN_PARENT_IS_SLOT | pinfo->slot;
					else if(pinfo->rank < m && _edits[at + pinfo->rank].node == p) _parent[i] = pinfo->rank;
					else return false;
				}
			}
			rn_node_columns cols;
			memset(&cols, 0, sizeof(cols));
			cols.position = _pos.data(); cols.scale = _scale.data(); cols.rotation = _rot.data(); cols.parent = _parent.data();
			cols.flags = _flags.data(); cols.render_group = _group.data(); cols.bounding_sphere = _sphere.data();
			cols.bounding_box = _box.data(); cols.render_priority = _prio.data();
			std::vector<uint32> slots(m);
			if(rn_scene_add_nodes(_scene, &cols, static_cast<uint32_t>(m), slots.data()) != RN_OK)
				return false;
			for(size_t i = 0; i < m; i++)
			{
				SceneNode *node = _edits[at + i].node;
				SceneCUDAInfo *info = static_cast<SceneCUDAInfo *>(node->GetSceneInfo());
				info->slot = slots[i];
				info->dirty = false;
				uint32 row = 0;
				Check(rn_scene_download(_scene, RN_DL_RANK_OF_SLOT, slots[i], 1, &row));
				info->rank = row;
				if(_nodeOfSlot.size() <= slots[i]) _nodeOfSlot.resize(slots[i] + 1, nullptr);
				_nodeOfSlot[slots[i]] = node;
				_dirty.erase(std::remove(_dirty.begin(), _dirty.end(), node), _dirty.end());
			}
			_incrementalEdits++;
			at = end;
		}
		_edits.clear();
		return true;
	}

	void SceneCUDA::UploadDirty()
	{
		const size_t m = _dirty.size();
		if(!m)
			return;
		std::vector<uint32> slots(m);
		_pos.resize(std::max(_pos.size(), m * 3)); _scale.resize(std::max(_scale.size(), m * 3)); _rot.resize(std::max(_rot.size(), m * 4));
		_sphere.resize(std::max(_sphere.size(), m * 4)); _box.resize(std::max(_box.size(), m * 6));
		_flags.resize(std::max(_flags.size(), m)); _group.resize(std::max(_group.size(), m));
		for(size_t i = 0; i < m; i++)
		{
			SceneCUDAInfo *info = static_cast<SceneCUDAInfo *>(_dirty[i]->GetSceneInfo());
			slots[i] = info->slot;
			info->dirty = false;
			FillRow(_dirty[i], i);
		}
		rn_node_columns cols;
		memset(&cols, 0, sizeof(cols));
		cols.position = _pos.data(); cols.scale = _scale.data(); cols.rotation = _rot.data();
		cols.flags = _flags.data(); cols.render_group = _group.data(); cols.bounding_sphere = _sphere.data();
		Check(rn_scene_update_nodes_by_slot(_scene, &cols, slots.data(), static_cast<uint32_t>(m)));
		_dirtyRows += m;
		_dirty.clear();
	}

	void SceneCUDA::Update(float delta)
	{
		WillUpdate(delta);
		FlushDeletionQueue();
		// SceneBasic::Update: node->Update(delta) for every node that is not Static (attachments, subclasses); the order
		// between nodes of one update priority is not observable for plain nodes
		for(IntrusiveList<SceneNode>::Member *m = _renderNodes.GetHead(); m;)
		{
			SceneNode *node = m->Get();
			m = m->GetNext();
			UpdateNode(node, delta);
		}
		for(IntrusiveList<Camera>::Member *m = _cameras.GetHead(); m; m = m->GetNext())
			UpdateNode(m->Get(), delta);
		FlushDeletionQueue();
		if(_structureDirty || _fullUploads == 0 || !ApplyEdits())
			UploadAll();
		else
			UploadDirty();
		Scene::Update(delta);
		DidUpdate(delta);
	}

	void SceneCUDA::Render(Renderer *renderer)
	{
		WillRender(renderer);
		if(_structureDirty || _fullUploads == 0 || !ApplyEdits())
			UploadAll();        // AddNode / reparenting between Update and Render
		else
			UploadDirty();
		for(IntrusiveList<Camera>::Member *m = _cameras.GetHead(); m; m = m->GetNext())
			m->Get()->PostUpdate();
		// the cameras in SceneBasic::Render's order (:420-439)
		std::vector<Camera *> cams;
		for(int cameraPriority = 0; cameraPriority < 3; cameraPriority++)
		{
			for(IntrusiveList<Camera>::Member *m = _cameras.GetHead(); m; m = m->GetNext())
			{
				Camera *camera = m->Get();
				if(camera->GetFlags() & Camera::Flags::NoRender || (cameraPriority == 0 && !(camera->GetFlags() & Camera::Flags::RenderEarly)) ||
				   (cameraPriority == 1 && (camera->GetFlags() & Camera::Flags::RenderEarly || camera->GetFlags() & Camera::Flags::RenderLate)) ||
				   (cameraPriority == 2 && !(camera->GetFlags() & Camera::Flags::RenderLate)))
					continue;
				if(camera->GetIsMultiviewCamera())
					continue;
				cams.push_back(camera);
			}
		}
		const size_t n = _nodeCount;
		for(size_t c0 = 0; c0 < cams.size() && n; c0 += RN_MAX_CAMERAS)
		{
			const uint32 v = static_cast<uint32>(std::min<size_t>(RN_MAX_CAMERAS, cams.size() - c0));
			rn_camera_frustum frusta[RN_MAX_CAMERAS];
			float positions[RN_MAX_CAMERAS * 3];
			for(uint32 c = 0; c < v; c++)
			{
				// the 28 floats Camera::UpdateFrustum just computed (RNCamera.cpp:342-384), the flags and the render group
				Camera *cam = cams[c0 + c];
				rn_camera_frustum &f = frusta[c];
				const Vector3 &center = cam->GetFrustumCenter();
				f.center[0] = center.x; f.center[1] = center.y; f.center[2] = center.z;
				f.radius = cam->GetFrustumRadius();
				const Plane *planes[6] = {&cam->frustums._frustumLeft, &cam->frustums._frustumRight, &cam->frustums._frustumTop,
				                          &cam->frustums._frustumBottom, &cam->frustums._frustumNear, &cam->frustums._frustumFar};
				for(int k = 0; k < 6; k++)
				{
					f.planes[k][0] = planes[k]->_normal.x; f.planes[k][1] = planes[k]->_normal.y; f.planes[k][2] = planes[k]->_normal.z;
					f.planes[k][3] = planes[k]->_d;
				}
				f.flags = static_cast<uint32>(cam->GetFlags());
				f.render_group = cam->GetRenderGroup();
				const Vector3 wp = cam->GetWorldPosition();
				positions[c * 3] = wp.x; positions[c * 3 + 1] = wp.y; positions[c * 3 + 2] = wp.z;
			}
			Check(rn_transform_cull(_scene, frusta, v, nullptr, nullptr));
			Check(rn_cull_sort(_scene, positions, v, nullptr));           // SortFrontToBack / SortTransparentBackToFront (:596-616)
			uint32 counts[RN_MAX_CAMERAS], capacities[RN_MAX_CAMERAS];
			uint32 *outs[RN_MAX_CAMERAS];
			for(uint32 c = 0; c < v; c++)
			{
				_listScratch[c].resize(n);
				outs[c] = _listScratch[c].data();
				capacities[c] = static_cast<uint32>(n);
			}
			Check(rn_cull_read_lists(_scene, v, outs, capacities, counts, nullptr));
			for(uint32 c = 0; c < v; c++)
			{
				Camera *camera = cams[c0 + c];
				const uint32 *list = outs[c];
				const uint32 count = counts[c];
				renderer->SubmitCamera(camera, [&] {
					for(uint32 e = 0; e < count; e++)
						_nodeOfSlot[list[e]]->Render(renderer, camera);
				});
			}
		}
		DidRender(renderer);
	}

	Matrix SceneCUDA::GetWorldTransformOf(SceneNode *node)
	{
		SceneCUDAInfo *info = static_cast<SceneCUDAInfo *>(node->GetSceneInfo());
		Matrix m;
		Check(rn_scene_download(_scene, RN_DL_WORLD_MATRIX, info->rank, 1, m.m));
		return m;
	}

	Vector3 SceneCUDA::GetWorldPositionOf(SceneNode *node)
	{
		SceneCUDAInfo *info = static_cast<SceneCUDAInfo *>(node->GetSceneInfo());
		float p[3];
		Check(rn_scene_download(_scene, RN_DL_WORLD_POSITION, info->rank, 1, p));
		return Vector3(p[0], p[1], p[2]);
	}
}

// module entry point, loaded by Module::Load (Source/Modules/RNModule.cpp:151-235; macro Source/Modules/RNModule.h:70-78)
RNModule(RayneCUDAScene, "net.uberpixel.rayne.cudascene")
