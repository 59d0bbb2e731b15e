/*
 * rn_cuda.h - C ABI of librn_cuda: the B200-native (sm_100a) implementation of Rayne's per-frame scene hot path.
 *
 * Drop-in boundary.  A Rayne `RN::Scene` subclass keeps RN::SceneNode /
 * RN::Camera / RN::ParticleEmitter as the front end and calls these entry points instead of walking its
 * node lists on the CPU.  No Rayne, torch or C++ types cross this boundary: plain pointers, sizes, PODs.
 *
 * Every function returns RN_OK (0) or a negative rn_status; the message is available per thread through
 * rn_last_error().  Nothing here throws, aborts or falls back to a CPU implementation: if no CUDA device is
 * usable, rn_scene_create fails with RN_ERR_CUDA.
 *
 * Index spaces
 *   rank : position of a node in the reference's render list (SceneBasic::_renderNodes head -> tail,
 *          Source/Scene/RNSceneBasic.cpp:652-676,683-736).  All host-facing node arrays are in rank order.
 *   slot : storage index inside the library (ranks are permuted inside small tiles so that each hierarchy
 *          level of a tile is contiguous, parents first).  Visibility lists hold *slots*, ordered by ascending *rank* - i.e.
 *          exactly the order in which SceneBasic::Render would have submitted the nodes.  rn_scene_slot_table
 *          converts between the two.
 */
#ifndef RN_CUDA_H
#define RN_CUDA_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RN_ABI_VERSION 4
#define RN_MAX_GROUP 16          /* GPUs of one rn_group */
#define RN_MAX_CAMERAS 8          /* views culled by one rn_cull call (stereo + 4 cascades = 6) */
#define RN_NO_PARENT 0xffffffffu

typedef enum rn_status {
	RN_OK = 0,
	RN_ERR_INVALID = -1,      /* bad argument */
	RN_ERR_CUDA = -2,         /* CUDA runtime error (message in rn_last_error) */
	RN_ERR_NOMEM = -3,
	RN_ERR_STATE = -4,        /* call order violated (e.g. cull before upload) */
	RN_ERR_CAPACITY = -5,     /* exceeds the capacities given at creation */
	RN_ERR_UNSUPPORTED = -7,  /* rn_scene_add_nodes: this edit needs a full rn_scene_upload (message says why) */
	RN_ERR_PEER = -6          /* rn_group: no usable transport between the GPUs (no peer access, no multicast object) */
} rn_status;

typedef struct rn_scene rn_scene_t;

/* RN::SceneNode::Flags bits the path looks at (Source/Scene/RNSceneNode.h:56-68). */
#define RN_NODE_STATIC     (1u << 0)
#define RN_NODE_HIDDEN     (1u << 1)
#define RN_NODE_NOCULLING  (1u << 2)
#define RN_NODE_OCCLUDER   (1u << 3)
/* RN::Camera::Flags bits the path looks at (Source/Scene/RNCamera.h:30-50). */
#define RN_CAMERA_RENDER_EARLY       (1u << 19)
#define RN_CAMERA_ORTHOGONAL         (1u << 21)
#define RN_CAMERA_USE_SIMPLE_CULLING (1u << 26)
#define RN_CAMERA_SORT_TRANSPARENT_BACK_TO_FRONT (1u << 5)   /* part of Camera::Flags::Defaults */
#define RN_CAMERA_SORT_FRONT_TO_BACK             (1u << 6)
/* Library extension (not a Rayne flag; bit 31 is unused by RN::Camera::Flags): additionally require the node's
 * transformed AABB to pass the plane test that Camera::InFrustum(const AABB&) carries as a comment
 * (Source/Scene/RNCamera.cpp:479-496; the live function returns true).  See rn_cull. */
#define RN_CAMERA_TEST_AABB          (1u << 31)
/* RN::SceneNode::RenderPriority values the sorts look at (Source/Scene/RNSceneNode.h:44-54). */
#define RN_RENDER_SKY         200000
#define RN_RENDER_TRANSPARENT 300000

typedef struct rn_scene_desc {
	int32_t  device;                     /* CUDA device ordinal */
	uint32_t max_nodes;                  /* capacity of the node columns */
	uint32_t tile_nodes;                 /* nodes per tile (0 = chosen per upload: 512 for large scenes, down to 64 for small ones; 64..1024) */
	uint32_t max_emitters;               /* 0 = no particle storage */
	uint32_t max_particles_per_emitter;  /* < 16383 (Source/Scene/RNParticleEmitter.cpp:115) */
	uint32_t tile_inner_nodes;           /* nodes WITH children per tile, kept in shared memory (0 = default 128; 16..1024) */
	uint32_t tile_unit_nodes;            /* largest subtree kept whole inside a tile (0 = default 128; 8..tile_nodes).  Smaller
	                                        units put more subtrees into a tile, which fills the 32-wide passes of
	                                        every hierarchy level; the nodes above the units go through the spine pre-pass */
	uint32_t flags;                      /* RN_SCENE_* bits, 0 = defaults */
} rn_scene_desc;
/* World MATRICES (rn_transform_propagate / rn_transform_cull, RN_DL_WORLD_MATRIX) are by default evaluated with fused
 * multiply-adds: one rounding less per term than the reference's strict-FP build, within the 1e-5 relative tolerance
 * the matrices carry (the reference's own release build, -funsafe-math-optimizations, differs from its strict build in
 * the same way).  With this flag they follow the reference operation by operation and match its strict-FP build bit
 * for bit.  World position / rotation / scale, bounding spheres and everything the cull reads are ALWAYS evaluated in
 * the reference's exact operation order, so visibility lists are bit-identical in both modes. */
#define RN_SCENE_STRICT_MATRICES (1u << 0)

/* Host node columns in RANK order.  Replaces the per-object state SceneNode keeps
 * (Source/Scene/RNSceneNode.h:204-246: _position/_scale/_rotation, _parent, _flags, _renderGroup,
 * _boundingSphere).  A NULL column is left untouched on partial updates. */
typedef struct rn_node_columns {
	const float    *position;         /* 3 floats per node                                   */
	const float    *scale;            /* 3 floats per node                                   */
	const float    *rotation;         /* 4 floats per node: quaternion x, y, z, w            */
	const uint32_t *parent;           /* rank of the parent or RN_NO_PARENT                  */
	const uint32_t *flags;            /* RN::SceneNode::Flags                                */
	const uint32_t *render_group;     /* low 8 bits = SceneNode::_renderGroup                */
	const float    *bounding_sphere;  /* 4 floats per node: local offset x, y, z and radius  */
	const float    *bounding_box;     /* optional, 6 floats per node: local AABB minExtend xyz, maxExtend xyz
	                                     (SceneNode::SetBoundingBox); NULL = the default AABB(-1, 1); full uploads only */
	const int32_t  *render_priority;  /* optional: SceneNode::_renderPriority; NULL = RenderNormal for all.
	                                     Only rn_cull_sort reads it (full uploads only)                        */
} rn_node_columns;

/* What Camera::PostUpdate / UpdateFrustum leave behind for the cull test
 * (Source/Scene/RNCamera.cpp:296-384,440-473): bounding sphere of the frustum, six planes in the order
 * left, right, top, bottom, near, far (normal xyz, d), camera flags and render group. */
typedef struct rn_camera_frustum {
	float    center[3];
	float    radius;
	float    planes[6][4];
	uint32_t flags;
	uint32_t render_group;
} rn_camera_frustum;

/* Camera parameters - the state RN::Camera holds (Source/Scene/RNCamera.cpp:60-95,136-228). */
typedef struct rn_camera_desc {
	uint32_t flags;
	uint32_t render_group;
	float    world_position[3];
	float    world_rotation[4];       /* x, y, z, w */
	float    scale[3];                /* the camera's LOCAL scale (RNCamera.cpp:305) */
	float    fov, aspect, clip_near, clip_far;
	float    ortho[4];                /* left, right, bottom, top */
	float    plane_offsets[4];        /* top, bottom, left, right (Camera::SetFrustumPlaneOffset) */
	uint32_t has_projection;          /* 1: `projection` was set through Camera::SetProjectionMatrix */
	float    projection[16];          /* column-major */
} rn_camera_desc;

/* Result of rn_cull for one camera.  `slots` is DEVICE memory owned by the scene, valid until the next
 * rn_cull / rn_transform_cull on the same scene; `count` is valid after the stream has been synchronised
 * (or use rn_cull_read_counts). */
typedef struct rn_vis_list {
	const uint32_t *slots;            /* device pointer, `capacity` entries available           */
	const uint32_t *count;            /* device pointer to the number of visible nodes           */
	uint32_t        capacity;
	uint32_t        reserved;
} rn_vis_list;

typedef enum rn_download_what {
	RN_DL_WORLD_POSITION = 0,   /* 3 floats per node  (SceneNode::GetWorldPosition)  */
	RN_DL_WORLD_ROTATION = 1,   /* 4 floats per node  (SceneNode::GetWorldRotation)  */
	RN_DL_WORLD_SCALE    = 2,   /* 3 floats per node  (SceneNode::GetWorldScale)     */
	RN_DL_WORLD_MATRIX   = 3,   /* 16 floats per node, column-major (GetWorldTransform) */
	RN_DL_SLOT_OF_RANK   = 4,   /* uint32 per node */
	RN_DL_RANK_OF_SLOT   = 5,   /* uint32 per node */
	RN_DL_INVERSE_WORLD_MATRIX = 6, /* 16 floats per node (GetInverseWorldTransform), after rn_transform_inverse */
	RN_DL_BOUNDING_SPHERE = 7,  /* 4 floats per node: position + offset (the tested centre) and radius, after rn_bounds_update */
	RN_DL_BOUNDING_BOX    = 8   /* 6 floats per node: world minExtend, maxExtend (relative to the world position) */
} rn_download_what;

/* ---- life cycle ------------------------------------------------------------------------------------------ */
const char *rn_last_error(void);
int rn_abi_version(void);
int rn_scene_create(rn_scene_t **out, const rn_scene_desc *desc);
void rn_scene_destroy(rn_scene_t *scene);

/* ---- nodes -------------------------------------------------------------------------------------------------
 * rn_scene_upload: replaces SceneBasic::AddNode's list building + the per-node state of SceneNode.
 *   cols->parent != NULL : full (re)definition of the scene; `first` must be 0 and `count` the node count.
 *                          Rebuilds the tiling.  All columns must be non-NULL.
 *   cols->parent == NULL : overwrite the given columns of ranks [first, first+count) (SetPosition & co).
 * rn_scene_update_nodes : the same partial update for a scattered set of ranks (the dirty set of a frame).
 * Host arrays are consumed before the call returns. */
int rn_scene_upload(rn_scene_t *scene, const rn_node_columns *cols, uint32_t first, uint32_t count);
int rn_scene_update_nodes(rn_scene_t *scene, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count);
/* The same on a caller-chosen stream: the staged copy and the scatter kernel are only ENQUEUED on `cuda_stream` (the
 * two calls above run on the default stream and return when the update is complete, so it is visible to later work on
 * any stream); work that must see the update has to follow on the same stream.  Host arrays are still consumed before
 * the call returns. */
int rn_scene_update_nodes_async(rn_scene_t *scene, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count,
                                void *cuda_stream);
int rn_scene_node_count(const rn_scene_t *scene, uint32_t *out_count);
/* Incremental edits - SceneBasic::AddNode / RemoveNode (Source/Scene/RNSceneBasic.cpp:683-746, :120-150) without a new
 * upload.  Nodes are named by their SLOT, which never changes while they exist (RN_DL_SLOT_OF_RANK gives the slots of an
 * upload; lists in slot mode hold them); requires slot lists (the default list mode).
 *   rn_scene_add_nodes   : `count` new nodes in the order SceneBasic::AddRenderNode would be CALLED for them (AddNode
 *                          recurses into the children first, Source/Scene/RNSceneNode.cpp:286-291, so a subtree arrives
 *                          children-before-parent).  cols->parent[i] = index of the parent inside this batch, or
 *                          RN_PARENT_IS_SLOT | slot of a node that is already in the scene (AddChild on an in-scene
 *                          parent), or RN_NO_PARENT.  One render priority per batch.  The batch lands where the
 *                          reference's insert rules put it (:652-676): in front of the list when its priority does not
 *                          exceed the list head's (in reverse call order - every PushFront lands before the previous
 *                          one), behind it when it is not below the tail's.  New tiles are appended to the node arrays;
 *                          nothing that exists moves.  out_slots[i] = slot of node i.
 *                          RN_ERR_UNSUPPORTED (do a full rn_scene_upload instead): a priority that belongs in the middle
 *                          of the list, a priority the upload did not contain, a subtree of more than tile_nodes nodes,
 *                          a parent that was itself added under a tile-computed node.  RN_ERR_CAPACITY: max_nodes (every
 *                          batch takes whole groups of 32 slots per tile) or the tile list's headroom is used up.
 *   rn_scene_remove_nodes: the nodes (a removed node's whole subtree, as RemoveNode -> UpdateSceneInfo(nullptr) takes it
 *                          out, Source/Scene/RNSceneNode.cpp:274-281) become holes: never visible, never listed.  Their
 *                          slots are not reused before the next full upload.
 *   rn_scene_update_nodes_by_slot : rn_scene_update_nodes addressed by slot. */
#define RN_PARENT_IS_SLOT 0x80000000u
int rn_scene_add_nodes(rn_scene_t *scene, const rn_node_columns *cols, uint32_t count, uint32_t *out_slots);
int rn_scene_remove_nodes(rn_scene_t *scene, const uint32_t *slots, uint32_t count);
int rn_scene_update_nodes_by_slot(rn_scene_t *scene, const rn_node_columns *cols, const uint32_t *slots, uint32_t count);
int rn_scene_edit_stats(const rn_scene_t *scene, uint32_t *live_nodes, uint32_t *slots_in_use, uint32_t *late_tiles, uint64_t *edits);

/* rn_transform_propagate: SceneNode::UpdateInternalData + UpdateInternalTransformData for every node
 * (Source/Scene/RNSceneNode.cpp:501-546): world position/rotation/scale and the world matrix.
 * rn_cull: SceneNode::CanRenderUtil + Camera::InFrustum(sphere) for every node and camera
 * (Source/Scene/RNSceneNode.cpp:405-417,570-588; Source/Scene/RNCamera.cpp:440-473), compacted in render-list
 * order.  Uses the world TRS left by the last rn_transform_propagate.
 * rn_transform_cull: both in one pass over the nodes (the reference evaluates transforms lazily inside the
 * cull loop, Source/Scene/RNSceneBasic.cpp:576-593, so fusing them is faithful to its frame).
 * All three are asynchronous on `cuda_stream` (a cudaStream_t, NULL = default stream). */
int rn_transform_propagate(rn_scene_t *scene, void *cuda_stream);
int rn_cull(rn_scene_t *scene, const rn_camera_frustum *cameras, uint32_t n_cameras, rn_vis_list *out_lists,
            void *cuda_stream);
int rn_transform_cull(rn_scene_t *scene, const rn_camera_frustum *cameras, uint32_t n_cameras,
                      rn_vis_list *out_lists, void *cuda_stream);
/* What the visibility lists hold.  ranks == 0 (default): slots (index the device outputs directly).
 * ranks != 0: render-list ranks plus `rank_offset` - the form in which the lists of the GPUs of one box are
 * exchanged (each GPU owns an ascending rank range starting at its rank_offset). */
int rn_scene_set_list_mode(rn_scene_t *scene, int ranks, uint32_t rank_offset);
/* Synchronises the stream and copies the per-camera visible counts of the last cull to the host. */
int rn_cull_read_counts(rn_scene_t *scene, uint32_t *out_counts, uint32_t n_cameras, void *cuda_stream);
/* Copies the visibility list of one camera to the host (synchronous). `out_slots` has `capacity` entries. */
int rn_cull_read_list(rn_scene_t *scene, uint32_t camera, uint32_t *out_slots, uint32_t capacity,
                      uint32_t *out_count, void *cuda_stream);

/* Copies the counts and the lists of cameras 0..n_cameras-1 of the last cull to the host in one go (one
 * synchronisation for the counts, then all list copies back to back).  out_slots[c] has capacities[c] entries. */
int rn_cull_read_lists(rn_scene_t *scene, uint32_t n_cameras, uint32_t *const *out_slots, const uint32_t *capacities,
                       uint32_t *out_counts, void *cuda_stream);

/* The same without the final wait: the counts are valid on return (one synchronisation), the list copies are only
 * enqueued on `cuda_stream` - the host buffers must be page-locked and stay untouched until that stream has been
 * synchronised (rn_stream_synchronize).  Lets the copy of one scene's lists overlap the kernels of another scene
 * on another stream (rayne_b200/pipeline.py). */
int rn_cull_read_lists_async(rn_scene_t *scene, uint32_t n_cameras, uint32_t *const *out_slots, const uint32_t *capacities,
                             uint32_t *out_counts, void *cuda_stream);
/* Streams for hosts that do not link the CUDA runtime themselves (on the current device).  urgency 0 = the lowest
 * stream priority, every step up is one priority level more urgent, clamped to what the device offers: when the
 * kernels of several scenes are in flight, the thread blocks of the more urgent stream are scheduled first. */
int rn_stream_create(void **out_stream, int urgency);
int rn_stream_destroy(void *cuda_stream);
int rn_stream_synchronize(void *cuda_stream);

/* Synchronous read-back in RANK order (parity tests, SetWorldPosition & co on the host side). */
int rn_scene_download(rn_scene_t *scene, rn_download_what what, uint32_t first, uint32_t count, void *dst);
/* Device pointers of the outputs in SLOT order (for GPU consumers: instance buffers etc.).
 * world_matrix: 16 floats per slot; world_pos_sx: float4 (x, y, z, scale.x); world_rot: float4;
 * world_scale_yz: float2. */
int rn_scene_device_outputs(rn_scene_t *scene, const float **world_matrix, const float **world_pos_sx,
                            const float **world_rot, const float **world_scale_yz);

/* SceneNode::UpdateInternalInverseTransformData for every node (Source/Scene/RNSceneNode.cpp:548-568): inverse local
 * matrix = WithScaling(1/scale or 0) . Rotate(conj(rot)) . Translate(-pos), inverse world = inverse local * parent's
 * inverse world.  Lazy in the reference (computed when GetInverseWorldTransform is called, e.g. by Drawable::Update);
 * here an explicit pass over the node columns, independent of rn_transform_propagate.  Storage is allocated on the
 * first call (64 bytes per node). */
int rn_transform_inverse(rn_scene_t *scene, void *cuda_stream);
/* SceneNode::UpdateInternalBoundsData for every node (Source/Scene/RNSceneNode.cpp:570-588): transformed bounding
 * sphere (Source/Math/RNSphere.h:62-73) and transformed AABB (Source/Math/RNAABB.h:108-114,152-177) from the world
 * TRS left by the last propagate.  The cull kernels evaluate the sphere on the fly and do not need this call; it
 * materialises what GetBoundingSphere / GetBoundingBox return (occlusion tests, picking, debugging). */
int rn_bounds_update(rn_scene_t *scene, void *cuda_stream);
/* Device pointers (slot order) of the optional outputs; NULL if the producing call has not run yet.
 * inverse_world_matrix: 16 floats; bounding_sphere: float4 (centre xyz, radius); bounding_box: 6 floats. */
int rn_scene_device_extras(rn_scene_t *scene, const float **inverse_world_matrix, const float **bounding_sphere,
                           const float **bounding_box);

/* The post-cull sorts of SceneBasic::Render (Source/Scene/RNSceneBasic.cpp:596-616) applied in place to the lists of
 * the last rn_cull / rn_transform_cull, per camera according to the flags it was culled with:
 *   RN_CAMERA_SORT_FRONT_TO_BACK: whole list by (render priority; squared distance of the node's world position to
 *       the camera ascending inside equal priorities below RenderSky);
 *   RN_CAMERA_SORT_TRANSPARENT_BACK_TO_FRONT: the range [first, last] of RenderTransparent entries by squared
 *       distance descending, where - as in the reference - an entry at list index 0 never counts as `first`.
 * camera_world_positions: 3 floats per camera of the last cull.  std::sort leaves the order of equal keys
 * unspecified; this implementation is stable (equal keys keep render-list order).  Needs the render_priority
 * column for anything but an all-RenderNormal scene.  Asynchronous on `cuda_stream`, no host synchronisation. */
int rn_cull_sort(rn_scene_t *scene, const float *camera_world_positions, uint32_t n_cameras, void *cuda_stream);

/* The visibility of the last cull as bitmasks in render-list order: bit (r & 31) of word (r >> 5) of camera c is set
 * iff the node of rank r is in camera c's list - the same information as the (unsorted) list of
 * SceneBasic::Render's cull loop (Source/Scene/RNSceneBasic.cpp:576-593) in 1/32 of the bytes per node.  Ranks are
 * local to the scene (rank_offset of rn_scene_set_list_mode is not applied).  This is what the multi-GPU exchange
 * ships; rn_masks_expand turns gathered masks back into lists.  Asynchronous on `cuda_stream`. */
int rn_cull_build_masks(rn_scene_t *scene, void *cuda_stream);
/* The same masks written into DEVICE memory of the caller (e.g. the send buffer of an all-gather): camera c at
 * device_out + c * row_stride_words, row_stride_words >= ceil(nodes / 32); words beyond the scene's are not touched. */
int rn_cull_build_masks_to(rn_scene_t *scene, uint32_t *device_out, uint32_t row_stride_words, void *cuda_stream);
/* The same, with `multicast_out` an NVLS MULTICAST address (CUDA multicast object / torch symmetric memory
 * `multicast_ptr`, pointing at this GPU's block of a buffer that every GPU of the group maps): the kernel stores with
 * multimem.st, the NVSwitch replicates every word into all GPUs' copies - building the masks IS the all-gather; what
 * remains is a barrier between the GPUs before the gathered buffer is read.  Only valid on such an address. */
int rn_cull_build_masks_multicast(rn_scene_t *scene, uint32_t *multicast_out, uint32_t row_stride_words, void *cuda_stream);
/* Device pointer of the masks (camera c at masks + c * words_per_camera), words per camera = ceil(nodes / 32), and
 * the number of cameras they hold; *masks = NULL before rn_cull_build_masks or after a newer cull. */
int rn_cull_device_masks(rn_scene_t *scene, const uint32_t **masks, uint32_t *words_per_camera, uint32_t *n_cameras);
/* Copies the masks of cameras 0..n_cameras-1 into out_words[n_cameras][words_per_camera] (host memory, pinned for a
 * truly asynchronous copy) behind the work of `cuda_stream`; the caller synchronises the stream. */
int rn_cull_read_masks_async(rn_scene_t *scene, uint32_t n_cameras, uint32_t *out_words, void *cuda_stream);
/* One rank range of a gathered mask: n_words words in DEVICE memory, bit b of word w = rank rank_base + 32 w + b. */
typedef struct rn_mask_segment {
	const uint32_t *device_words;
	uint32_t        n_words;
	uint32_t        rank_base;
} rn_mask_segment;
/* Appends the ranks of the set bits of the segments, in segment order and ascending inside a segment, to device_list
 * at position *device_count and adds their number to *device_count (a uint32 in DEVICE memory; no host round trip).
 * With the segments = the rank ranges of all GPUs in order (what the all-gather of rn_cull_build_masks' output
 * delivers) the result is the visibility list of the whole scene in global ranks, bit for bit.  Entries beyond
 * `capacity` are counted but not written.  `scene` only provides the device and scratch memory.  Asynchronous. */
int rn_masks_expand(rn_scene_t *scene, const rn_mask_segment *segments, uint32_t n_segments, uint32_t *device_list,
                    uint32_t capacity, uint32_t *device_count, void *cuda_stream);

/* rn_masks_expand for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same inputs,
 * outputs and result as rn_masks_expand, but it takes no scene: the per-CTA scratch of the three expand kernels is a
 * file-scope __device__ array of librn_cuda (64 KiB, held by every GPU that loads the library until the process exits),
 * so the call performs no cudaMalloc / cudaFree and no allocating library call.  `device` is the CUDA device ordinal
 * (-1 = the current device).  Limits: at most 32 segments per call, and the segments - each rounded up to a multiple of
 * 1024 words - may hold at most RN_MASKS_EXPAND_STATIC_MAX_WORDS mask words in total (16 Mi words = 536 870 912 ranks);
 * beyond that the call returns RN_ERR_CAPACITY and launches nothing (use rn_masks_expand).  All calls on one GPU share
 * the one scratch array: calls enqueued on the same stream are ordered and safe, calls that overlap on the GPU (different
 * streams without an event between them, or several host threads) race on it.  The array keeps the previous call's
 * contents; nothing in it has to start at zero, the first kernel of every call overwrites the entries the call reads.
 * The first call on a device looks the array's device address up (cudaGetSymbolAddress) - make that call before, not
 * inside, a stream capture.  Asynchronous. */
#define RN_MASKS_EXPAND_STATIC_MAX_WORDS (16u * 1024u * 1024u)
int rn_masks_expand_static(int device, const rn_mask_segment *segments, uint32_t n_segments, uint32_t *device_list,
                           uint32_t capacity, uint32_t *device_count, void *cuda_stream);

/* Per-drawable matrix records for the list of `camera` left by the last cull (and rn_cull_sort): what a renderer
 * writes per drawable before a draw - model and inverse model matrix (Drawable::Update,
 * Source/Rendering/RNRenderer.h:65-76), view * model and projection * view * model
 * (Modules/Vulkan/Sources/RNVulkanRenderer.cpp:1134-1168).  Entry e of the list becomes 64 floats at
 * device_out + e * 256 bytes: model[16], inverse model[16], view * model[16], proj_view * model[16] (column-major).
 * `device_out` is DEVICE memory of the caller (e.g. an instance / storage buffer mapped through graphics interop),
 * 16-byte aligned, room for `capacity` records; entries beyond `capacity` are not written.  The entry count stays on
 * the device (rn_cull_read_counts).  Needs rn_transform_propagate and rn_transform_inverse.  Asynchronous.
 * LOD selection, the other per-drawable step of Entity::Render, is rn_cull_select_lod below. */
int rn_cull_pack_instances(rn_scene_t *scene, uint32_t camera, const float *view, const float *proj_view,
                           void *device_out, uint32_t capacity, void *cuda_stream);

/* LOD selection per list entry - EXTENSION: the reference compiles it out (RN_MODEL_LOD_DISABLED 1,
 * Source/Rendering/RNModel.h:9; Entity::Render then always takes the single stage); this is its #else branch
 * (Source/Scene/RNEntity.cpp:98-106): lodDistance = |node.worldPosition - lodCamera.worldPosition| / lodCamera.clipFar,
 * and Model::GetLODStageForDistance (Source/Rendering/RNModel.cpp:260-281): stage 0 if the model has <= 1 stage or
 * lodDistance <= epsilon, else the first stage with lodDistance <= stage.distance, else the last stage.
 *   rn_lod_set_models      : model m owns stage_distances[stage_offsets[m] .. stage_offsets[m + 1]) (LODStage::GetDistance, ascending)
 *   rn_lod_set_node_models : Entity::_model of ranks [first, first + count) as a model index, 0xffff = not an Entity (stage 0);
 *                            a full rn_scene_upload re-tiles the scene and clears the table - set it again after one
 *   rn_cull_select_lod     : one byte per entry of `camera`'s list (after rn_cull_sort, in the sorted order) into DEVICE
 *                            memory of the caller; entries beyond `capacity` are not written.  Asynchronous. */
int rn_lod_set_models(rn_scene_t *scene, const uint32_t *stage_offsets, const float *stage_distances, uint32_t n_models);
int rn_lod_set_node_models(rn_scene_t *scene, const uint16_t *model_of_rank, uint32_t first, uint32_t count);
int rn_cull_select_lod(rn_scene_t *scene, uint32_t camera, const float *lod_camera_world_position, float lod_camera_clip_far,
                       uint8_t *device_out, uint32_t capacity, void *cuda_stream);

/* Software occlusion culling - the pass SceneBasic::Render runs for a camera that sees at least one Flags::Occluder node
 * (Source/Scene/RNSceneBasic.cpp:445-573; RasterizeClipSpaceTriangle :152-208, RasterizeMesh :210-318, TestBoundingBox
 * :320-405): the 30 occluders with the largest radius^2 / max(squared distance, 1), front to back, are box-tested and
 * rasterised (LOD 0 / mesh 0 of an Entity) into a 40 x 40 reverse-Z depth buffer; every node that passed CanRender is then
 * box-tested against it, except priorities >= RenderSky and the visible occluders; a node stays submitted for 50 frames
 * after it was last seen (SceneBasicInfo::occludedFrameCounter, kept per node on the device, shared by the cameras of a
 * frame - hence `camera_order`).  Runs AFTER rn_transform_cull / rn_cull and rn_bounds_update, BEFORE rn_cull_sort; it
 * clears visibility bits and re-emits the lists of ALL cameras.  A camera without a visible occluder keeps its list and
 * touches no counter, as in the reference.
 *   rn_occlusion_set_meshes      : triangle soups in model space, 9 floats per triangle (posA, posB, posC exactly as
 *                                  Mesh::GetTrianglesChunk()'s vertex iterator yields them in RasterizeMesh :213-224);
 *                                  mesh m owns triangles [triangle_offsets[m], triangle_offsets[m + 1])
 *   rn_occlusion_set_node_meshes : per rank the mesh of the Entity (its model's LOD 0 / mesh 0), -1 = not an Entity
 *                                  (an Occluder that is no Entity is box-tested but never rasterised, :516-525).  A full
 *                                  rn_scene_upload re-tiles the scene: set the table again (the counters restart at 0,
 *                                  like the SceneBasicInfo a node gets from AddNode)
 *   rn_cull_occlusion            : `cameras[c]` belongs to camera c of the last cull; `camera_order` (may be NULL = 0, 1, ...)
 *                                  is the order in which SceneBasic::Render visits them (:420-439: RenderEarly cameras, then
 *                                  the ordinary ones, then RenderLate; inside a class newest first)
 *   rn_occlusion_download        : number of occluders used by `camera` (0 = plain path), the jitter it took, the number of
 *                                  jitter draws the whole call consumed, the camera's depth buffer (1 600 floats as the
 *                                  reference indexes them), the counters of ranks [first, first + count); any may be NULL
 * The jitter of :488 comes from Rayne's shared random generator, and only a camera that found occluders draws one - which
 * the host cannot know before the cull has run.  The caller therefore passes the NEXT n_cameras draws of that generator
 * (`jitter`, x y z each, in draw order; GetRandomVector3Range((-0.15, -0.15, 0), (0.15, 0.15, 0))); the cameras take them in
 * order on the device, and rn_occlusion_download reports how many were consumed, so that a shim keeps the unconsumed ones
 * queued for the next frame and the sequence of jitters equals the reference's.  That also makes the pass reproducible. */
typedef struct rn_occlusion_camera
{
	float view[16];            /* Camera::GetViewMatrix() */
	float projection[16];      /* Camera::GetProjectionMatrix() */
	float pixel_size[2];       /* 1 / width, 1 / height of camera->GetRenderPass()->GetFrame() */
	float world_position[3];   /* camera->GetWorldPosition() */
	uint32_t enabled;          /* 0: leave this camera's list alone */
} rn_occlusion_camera;
int rn_occlusion_set_meshes(rn_scene_t *scene, const uint32_t *triangle_offsets, const float *triangles, uint32_t n_meshes);
int rn_occlusion_set_node_meshes(rn_scene_t *scene, const int32_t *mesh_of_rank, uint32_t first, uint32_t count);
int rn_occlusion_reset_counters(rn_scene_t *scene);
int rn_cull_occlusion(rn_scene_t *scene, const rn_occlusion_camera *cameras, const uint32_t *camera_order, uint32_t n_cameras,
                      const float *jitter, void *cuda_stream);
int rn_occlusion_download(rn_scene_t *scene, uint32_t camera, uint32_t *used, float *jitter_taken, uint32_t *draws_consumed,
                          float *depth_buffer, uint32_t *counters_by_rank, uint32_t first, uint32_t count);

/* ---- the pipelined frame: visibility in HOST memory, one GPU (rn_pipeline.cu) ------------------------------------------
 * The reference's Render hands its list to host code (Source/Scene/RNSceneBasic.cpp:576-640); when the consumer is the CPU
 * the frame is  dirty rows -> device, transform + cull, visibility -> host.  rn_pipeline cuts the render list into
 * `n_ranges` rank ranges where no parent / child edge crosses (rn_plan_ranges), gives every range its own rn_scene_t and
 * prioritised stream, and keeps `frames_in_flight` result buffers in mapped pinned memory, so that neither the kernels nor
 * the copy-out of a frame ever wait for the host:
 *     rn_pipeline_update_nodes(...);                    the frame's SetPosition & co (ranks 0-based in the given columns)
 *     rn_pipeline_submit(p, cameras, n, form);          enqueue; returns at once
 *     ... submit the next frame before looking at this one to overlap its copy-out with the next kernels ...
 *     rn_pipeline_acquire(p, &view);                    blocks until the OLDEST submitted frame is in host memory
 *     ... walk view.segments (or rn_frame_list) ...
 *     rn_pipeline_release(p);
 * view.segments[range * n_cameras + camera] describes one piece of a camera's visibility: `form` RN_VIS_LISTS = `count`
 * global ranks ascending; RN_VIS_MASKS = ceil(rank_count / 32) words, bit b of word w = rank rank_first + 32 w + b.  The
 * concatenation over the ranges in order is the reference's list.  RN_VIS_ADAPTIVE lets the device pick per segment the
 * smaller of the two encodings (a list wins below 1/32 visible).  Ranks are global: rank_base + position in the columns.
 * With n_ranges = 1 the scene keeps two sets of cull outputs and the copy-out runs on its own stream beside the next frame's
 * kernels; the masks then travel through the copy engine (a copy kernel would wait for the persistent tile kernel's CTAs),
 * so RN_VIS_ADAPTIVE delivers masks in that mode. */
typedef struct rn_pipeline rn_pipeline_t;
typedef enum rn_visibility_form { RN_VIS_LISTS = 0, RN_VIS_MASKS = 1, RN_VIS_ADAPTIVE = 2 } rn_visibility_form;
typedef struct rn_pipeline_desc
{
	int32_t  device;
	uint32_t n_ranges;            /* 0 = 1: one range, copy-out beside the next frame's kernels (measured on the 16.8 M-node scene:
	                                 1: 0.66 ms, 2: 0.85, 4: 0.80, 8: 0.73, 16: 0.75; device-only step 0.62) */
	uint32_t frames_in_flight;    /* 0 = 2 */
	uint32_t flags;               /* rn_scene_desc.flags of every range (RN_SCENE_STRICT_MATRICES) */
	uint32_t rank_base;           /* global rank of the first node (a GPU's own range in a multi-GPU split) */
	uint32_t tile_nodes, tile_inner_nodes, tile_unit_nodes;   /* rn_scene_desc, 0 = defaults */
} rn_pipeline_desc;
typedef struct rn_frame_segment
{
	uint32_t form;                /* RN_VIS_LISTS or RN_VIS_MASKS */
	uint32_t count;               /* visible nodes of this range for this camera */
	uint32_t rank_first, rank_count;
	const uint32_t *data;         /* host memory owned by the pipeline, valid until rn_pipeline_release */
} rn_frame_segment;
typedef struct rn_frame_view
{
	uint64_t frame;               /* 0, 1, 2 ... in submit order */
	uint32_t n_cameras, n_ranges;
	const rn_frame_segment *segments;
	uint64_t h2d_bytes, d2h_bytes;   /* what this frame moved over PCIe (dirty rows in, headers + payload out) */
} rn_frame_view;
int rn_plan_ranges(const uint32_t *parent, uint32_t n, uint32_t n_ranges, uint32_t *bounds /* n_ranges + 1 */);
int rn_pipeline_create(rn_pipeline_t **out, const rn_pipeline_desc *desc, const rn_node_columns *cols, uint32_t count);
void rn_pipeline_destroy(rn_pipeline_t *pipeline);
int rn_pipeline_ranges(const rn_pipeline_t *pipeline, uint32_t *bounds, uint32_t capacity, uint32_t *n_ranges);
int rn_pipeline_update_nodes(rn_pipeline_t *pipeline, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count);
int rn_pipeline_submit(rn_pipeline_t *pipeline, const rn_camera_frustum *cameras, uint32_t n_cameras, rn_visibility_form form);
int rn_pipeline_acquire(rn_pipeline_t *pipeline, rn_frame_view *view);
int rn_pipeline_release(rn_pipeline_t *pipeline);
/* Host decode of one camera of an acquired frame into a flat ascending list of global ranks (out_ranks may be NULL to get
 * the count only). */
int rn_frame_list(const rn_frame_view *view, uint32_t camera, uint32_t *out_ranks, uint32_t capacity, uint32_t *out_count);
/* The rn_scene_t and stream of one range (device-side consumers: rn_cull_pack_instances, rn_scene_device_outputs ...). */
int rn_pipeline_scene(rn_pipeline_t *pipeline, uint32_t range, rn_scene_t **scene, void **cuda_stream);

/* ---- several GPUs, ONE host process (rn_group.cu) -------------------------------------------------------------------
 * Rayne is a single-process engine: rn_group gives such a host the multi-GPU path of SURVEY.md §8(e).  The render list is
 * cut into one rank range per GPU (rn_plan_ranges), every GPU propagates and culls its range against all cameras, and the
 * one exchange of the path - every GPU learns the visibility of the whole scene - is fused into the kernel that builds the
 * rank-order bitmasks: its stores go through an NVLS multicast object (one multimem.st per word, replicated by the
 * NVSwitch; RN_GROUP_MULTICAST) or through peer-mapped pointers over NVLink (RN_GROUP_PEER).  No collective library; the
 * GPUs are ordered with CUDA events and the host never waits.
 *     rn_group_update_nodes(g, ...);                       the frame's setters, ranks 0-based in the given columns
 *     rn_group_frame(g, cameras, n);                       enqueue on every GPU, returns at once
 *     rn_group_expand(g, member, camera, ...);             that GPU's list of the WHOLE scene for one view (global ranks,
 *                                                          ascending = render-list order), enqueued on its stream
 *     rn_group_member(g, member, &scene, &stream, &masks)  the GPU's rn_scene_t (local lists, world matrices), its stream and
 *                                                          its copy of the gathered masks: [camera][row_words], GPU k's block
 *                                                          at word bounds-derived offset (rn_group_info)
 *     rn_group_synchronize(g);                             host wait for all GPUs */
typedef struct rn_group rn_group_t;
enum { RN_GROUP_AUTO = 0, RN_GROUP_PEER = 1, RN_GROUP_MULTICAST = 2 };
typedef struct rn_group_desc
{
	uint32_t n_devices;
	int32_t  devices[RN_MAX_GROUP];
	uint32_t flags;               /* rn_scene_desc.flags of every member */
	uint32_t transport;           /* RN_GROUP_AUTO: multicast if the driver and the GPUs offer it, else peer stores */
	uint32_t tile_nodes, tile_inner_nodes, tile_unit_nodes;
} rn_group_desc;
int rn_group_create(rn_group_t **out, const rn_group_desc *desc, const rn_node_columns *cols, uint32_t count);
void rn_group_destroy(rn_group_t *group);
int rn_group_info(const rn_group_t *group, uint32_t *transport, uint32_t *bounds, uint32_t capacity, uint32_t *row_words, const char **note);
int rn_group_update_nodes(rn_group_t *group, const rn_node_columns *cols, const uint32_t *ranks, uint32_t count);
int rn_group_frame(rn_group_t *group, const rn_camera_frustum *cameras, uint32_t n_cameras);
int rn_group_synchronize(rn_group_t *group);
int rn_group_member(rn_group_t *group, uint32_t member, rn_scene_t **scene, void **cuda_stream, const uint32_t **gathered_masks);
int rn_group_expand(rn_group_t *group, uint32_t member, uint32_t camera, uint32_t *device_list, uint32_t capacity, uint32_t *device_count,
                    const uint32_t **out_list, const uint32_t **out_count);
int rn_group_read_list(rn_group_t *group, uint32_t member, uint32_t camera, uint32_t *out_ranks, uint32_t capacity, uint32_t *out_count);

/* Camera::PostUpdate + UpdateProjection + UpdateFrustum on the host (Source/Scene/RNCamera.cpp:296-384).
 * Pure host function; V <= 8 cameras per frame make this negligible. */
int rn_camera_build_frustum(const rn_camera_desc *camera, rn_camera_frustum *out);

/* Camera::GetViewMatrix / GetProjectionMatrix as Camera::PostUpdate leaves them (Source/Scene/RNCamera.cpp:296-340)
 * and their product projection * view; 16 floats each, column-major, any of the outputs may be NULL.  Host function. */
int rn_camera_build_matrices(const rn_camera_desc *camera, float *view, float *projection, float *proj_view);

/* The shadow-cascade cameras of a directional light: Light::UpdateShadows (Source/Scene/RNLight.cpp:475-529: split
 * distances = blend of a linear and a logarithmic subdivision of the target camera's depth range, unless a split carries
 * its own maxDistance) and Camera::MakeShadowSplit (Source/Scene/RNCamera.cpp:240-287: bounding sphere of the split's
 * frustum slice, light camera placed directional_shadow_distance up the light direction, snapped to the shadow map's
 * texel grid in light space, square orthographic frustum).  Writes n_splits camera records ready for
 * rn_camera_build_frustum / rn_cull and, if out_proj_view != NULL, the 16 floats per split the reference keeps in
 * Light::_shadowCameraMatrices (projection * inverse world transform).  Pure host function. */
typedef struct rn_shadow_desc {
	uint32_t n_splits;                        /* ShadowParameter::splits.size(), 1..RN_MAX_CAMERAS */
	float    split_max_distance[RN_MAX_CAMERAS]; /* ShadowSplit::maxDistance (0 = derive from the blend) */
	float    resolution;                      /* ShadowParameter::resolution: width = height of a cascade's frame */
	float    clip_near, clip_far;             /* of the cascade cameras (ShadowParameter::clipNear / clipFar) */
	float    directional_shadow_distance, distance_blend_factor, max_shadow_dist;
	float    light_world_rotation[4];         /* the Light's GetWorldRotation(), x y z w */
} rn_shadow_desc;
int rn_light_build_cascades(const rn_camera_desc *target, const rn_shadow_desc *shadow, rn_camera_desc *out_cameras,
                            float *out_proj_view);

/* SceneBasic::AddNode / AddRenderNode list order (Source/Scene/RNSceneBasic.cpp:652-676,683-736 and
 * Source/Scene/RNSceneNode.cpp:270-294) for nodes given in CREATION order: parent[i] < i is the creation index of
 * the parent or -1; every tree is complete when its root is added; roots are added in creation order.
 * Writes order[rank] = creation index.  Pure host function. */
int rn_scenebasic_render_order(uint32_t n, const int32_t *parent, const int32_t *render_priority, uint32_t *order);

/* SceneWithVisibilityLists::AddNode order of the default volume (Source/Scene/RNSceneWithVisibilityLists.cpp:210-236):
 * children first (recursively, in child order), then the node; no priority sort.  Same conventions as
 * rn_scenebasic_render_order.  Upload the node columns in this order to get the lists RenderVolumeList would
 * submit for the default volume (:116-123). */
int rn_scenevis_default_volume_order(uint32_t n, const int32_t *parent, uint32_t *order);

/* SceneWithVisibilityLists volumes (Source/Scene/RNSceneWithVisibilityLists.cpp:29-37,157-207): AddVolume moves its
 * nodes out of the default volume, and a camera submits the default volume's nodes followed by those of the FIRST
 * volume that contains its world position.  Here: upload the node columns as [default volume in
 * rn_scenevis_default_volume_order][nodes of volume 1 in the order of its `nodes` vector][volume 2]..., tag every rank
 * with its volume (0 = default, k = volume k, 255 = in no volume: never rendered - the children of a node that was
 * added through AddVolume, :181-190), and give every camera of the next culls the volume it is in
 * (rn_scenevis_choose_volume; 0 = none).  The culls then drop the entries of the other volumes, and the lists are the
 * reference's submissions.  A node listed in several volumes cannot be expressed (it has one rank).
 * rn_scenevis_set_camera_volumes(scene, NULL, 0) switches the filter off; a full rn_scene_upload clears the tags. */
int rn_scenevis_set_node_volumes(rn_scene_t *scene, const uint8_t *volume_of_rank, uint32_t first, uint32_t count);
int rn_scenevis_set_camera_volumes(rn_scene_t *scene, const uint8_t *volume_per_camera, uint32_t n_cameras);
/* bounds: 6 floats per AxisAlignedBoxVolume (boundsMin xyz, boundsMax xyz), in AddVolume order.  Host function. */
int rn_scenevis_choose_volume(const float *bounds, uint32_t n_volumes, const float *camera_world_position, uint8_t *out_volume);

/* ---- particles ------------------------------------------------------------------------------------------- */
/* GenericParticleEmitter state (Source/Scene/RNParticleEmitter.cpp:24-36,379-395). */
typedef struct rn_emitter_desc {
	uint32_t seed;                    /* RandomNumberGenerator::Seed (Source/Math/RNRandom.cpp:313-323) */
	uint32_t max_particles;           /* SetMaxParticles: hard and soft maximum */
	float    spawn_rate;              /* seconds between spawns */
	float    life_span[2];
	float    start_color[4], end_color[4];
	float    start_size[2], end_size[2], start_rotation[2], end_rotation[2];
	float    gravity[3], velocity[3], velocity_min[3], velocity_max[3], position_min[3], position_max[3];
	uint32_t flags;                   /* RN_EMITTER_DESC_* */
	uint32_t max_particles_soft;      /* with RN_EMITTER_DESC_SOFT_MAX: SetMaxParticlesSoft (Source/Scene/RNParticleEmitter.cpp:128-132) */
	float    world_position[3];       /* with RN_EMITTER_DESC_NONLOCAL: the emitter node's GetWorldPosition() ... */
	float    world_scale[3];          /* ... and GetWorldScale() at spawn time (rn_particles_set_emitter_transforms updates them) */
} rn_emitter_desc;
/* !GetIsLocal() (Source/Scene/RNParticleEmitter.cpp:428-437): new particles are created in WORLD space - position scaled by
 * the emitter's world scale and moved to its world position, velocity and the particle's own gravity scaled, sizes
 * scaled by the largest scale component.  Costs 16 B more per particle and step (the per-particle gravity column). */
#define RN_EMITTER_DESC_NONLOCAL (1u << 0)
/* max_particles_soft is a soft maximum distinct from max_particles.  As in the reference (:204-205) the spawn count is
 * min(soft - live, spawn) in UNSIGNED arithmetic: with more live particles than the soft maximum the difference wraps
 * and the emitter spawns unbounded; here spawning additionally stops at max_particles_per_emitter, the storage a
 * segment has (rn_particles_counts reports such an emitter through RN_ERR_CAPACITY). 0 = no limit, as in the reference. */
#define RN_EMITTER_DESC_SOFT_MAX (1u << 1)

typedef enum rn_particle_field {
	RN_PF_POSITION = 0,   /* 3 floats */
	RN_PF_VELOCITY = 1,   /* 3 floats */
	RN_PF_TIMING   = 2,   /* 3 floats: time, lifespan, duration */
	RN_PF_RANGES   = 3,   /* 4 floats: size start, size end, rotation start, rotation end */
	RN_PF_DERIVED  = 4,   /* 7 floats: rotation, size x, size y, color r, g, b, a */
	RN_PF_SORTED   = 5,   /* 1 uint32: per-emitter depth-sorted particle indices (after rn_particles_sort) */
	RN_PF_GRAVITY  = 6    /* 3 floats: GenericParticle::gravity (the emitter's for local emitters, scaled at spawn otherwise) */
} rn_particle_field;

/* (Re)creates `count` emitters with no live particles (new GenericParticleEmitter + Seed). */
int rn_particles_configure(rn_scene_t *scene, const rn_emitter_desc *emitters, uint32_t count);
/* SetMaxParticlesSoft / the emitter node's world transform for emitters [first, first + count): host arrays of count (x3)
 * values, NULL = leave.  Takes effect with the next rn_particles_step (synchronous, tiny). */
int rn_particles_set_soft_max(rn_scene_t *scene, uint32_t first, uint32_t count, const uint32_t *max_particles_soft);
int rn_particles_set_emitter_transforms(rn_scene_t *scene, uint32_t first, uint32_t count, const float *world_position,
                                        const float *world_scale);
/* ParticleEmitter::UpdateParticles for every emitter (Source/Scene/RNParticleEmitter.cpp:180-211):
 * integrate/age/interpolate, order-preserving kill, spawn with the emitter's own MT19937 stream. */
int rn_particles_step(rn_scene_t *scene, float delta, void *cuda_stream);
/* Extension (no live reference behaviour, Source/Scene/RNParticleEmitter.cpp:364 is a comment): per-emitter
 * ascending sort of the live particles by distance to `camera_world_position`; ties keep vector order. */
int rn_particles_sort(rn_scene_t *scene, const float camera_world_position[3], void *cuda_stream);
int rn_particles_counts(rn_scene_t *scene, uint32_t *out_counts, float *out_times, uint32_t n_emitters);
/* Copies one field of the live particles of one emitter, in vector order (synchronous). */
int rn_particles_download(rn_scene_t *scene, uint32_t emitter, rn_particle_field field, void *dst,
                          uint32_t capacity_particles, uint32_t *out_count);

/* ParticleEmitter::UpdateMesh for every emitter (Source/Scene/RNParticleEmitter.cpp:213-336): per live particle four
 * vertices (position, colour, UV (0,0)(1,0)(0,1)(1,1), half-size corner vector in UV1 - rotated by the particle's
 * rotation when `can_roll`) and six uint16 indices; degenerate indices for unused slots.  Output is a headless
 * device buffer per attribute, emitter e at element offset e * max_particles_per_emitter * 4 (vertices) or * 6
 * (indices); an emitter's mesh capacity is its own max_particles.  world_scale_x: GetWorldScale().x per emitter
 * (NULL = 1).  flags per emitter (NULL = 0). */
#define RN_EMITTER_CAN_ROLL          (1u << 0)   /* ParticleEmitter::SetCanRollParticles   */
#define RN_EMITTER_RENDERED_INVERSED (1u << 1)   /* ParticleEmitter::SetIsRenderedInversed */
typedef struct rn_particle_mesh {
	const float    *position;   /* device, 3 floats per vertex */
	const float    *color;      /* device, 4 floats per vertex */
	const float    *uv0;        /* device, 2 floats per vertex */
	const float    *uv1;        /* device, 2 floats per vertex */
	const uint16_t *indices;    /* device, 6 per particle slot */
	uint32_t        vertex_stride_per_emitter;   /* = max_particles_per_emitter * 4 */
	uint32_t        index_stride_per_emitter;    /* = max_particles_per_emitter * 6 */
} rn_particle_mesh;
int rn_particles_build_mesh(rn_scene_t *scene, const float *world_scale_x, const uint32_t *flags, rn_particle_mesh *out,
                            void *cuda_stream);
/* Copies the mesh of one emitter to the host (synchronous): max_particles*4 vertices, max_particles*6 indices. */
int rn_particles_download_mesh(rn_scene_t *scene, uint32_t emitter, float *position, float *color, float *uv0, float *uv1,
                               uint16_t *indices);

/* The vertex stage of the reference's particle shader (Shaders/Particles.hlsl:52-71, `particles_vertex`) for every
 * vertex of the live quads of the mesh left by rn_particles_build_mesh:
 *   clip = modelViewProjection * (float4(position, 1) + inverseModelView * float4(texCoords2, 0, 0)),
 *   colour = colour * cameraAmbientColor.
 * inverse_model_view / model_view_projection: 16 floats PER EMITTER (column-major, as Rayne uploads them: each emitter
 * is a SceneNode with its own model matrix); ambient_color: 4 floats.  Output: headless device arrays, 4 + 4 floats per
 * vertex (vertices of dead slots stay zero).  The shader itself cannot run here (no graphics stack); parity is against
 * its C restatement in the oracle. */
typedef struct rn_particle_clip {
	const float *clip_position;   /* device, 4 floats per vertex */
	const float *color;           /* device, 4 floats per vertex */
	uint32_t     vertex_stride_per_emitter;
	uint32_t     reserved;
} rn_particle_clip;
int rn_particles_expand_billboards(rn_scene_t *scene, const float *inverse_model_view, const float *model_view_projection,
                                   const float *ambient_color, rn_particle_clip *out, void *cuda_stream);
int rn_particles_download_clip(rn_scene_t *scene, uint32_t emitter, float *clip_position, float *color);

/* ---- introspection for the bench / tests ---------------------------------------------------------------- */
typedef struct rn_scene_stats {
	uint32_t nodes, tiles, spine_nodes, spine_levels, max_depth, tile_nodes;
	uint64_t device_bytes;
	uint64_t kernel_launches;         /* kernels launched by this scene since creation */
} rn_scene_stats;
int rn_scene_get_stats(const rn_scene_t *scene, rn_scene_stats *out);
/* Optional in-stream timing of the dominant kernel (the tile kernel) with CUDA events, for the roofline line of
 * bench.py: enable, run frames, then read the summed duration (ms) and the number of timed launches (resets). */
int rn_scene_profile(rn_scene_t *scene, int enable);
int rn_scene_profile_read(rn_scene_t *scene, float *tile_ms_sum, uint32_t *tile_launches);

#ifdef __cplusplus
}
#endif
#endif /* RN_CUDA_H */
