"""ctypes binding of librn_cuda.so (include/rn_cuda.h) - the call path a Rayne `Scene` subclass would use.

There is deliberately no fallback: if the shared library is missing or a CUDA call fails, the error surfaces
as `RNError`.  numpy arrays are host buffers; everything on the device is owned by the library.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build
from .scenegen import CAMERA_DTYPE, EMITTER_DTYPE, Scene

RN_MAX_CAMERAS = 8
RN_NO_PARENT = 0xFFFFFFFF
PARENT_IS_SLOT = 0x80000000

DL_WORLD_POSITION, DL_WORLD_ROTATION, DL_WORLD_SCALE, DL_WORLD_MATRIX, DL_SLOT_OF_RANK, DL_RANK_OF_SLOT = range(6)
DL_INVERSE_WORLD_MATRIX, DL_BOUNDING_SPHERE, DL_BOUNDING_BOX = 6, 7, 8
CAM_SORT_TRANSPARENT_BACK_TO_FRONT, CAM_SORT_FRONT_TO_BACK, CAM_TEST_AABB = 1 << 5, 1 << 6, 1 << 31
EMITTER_CAN_ROLL, EMITTER_RENDERED_INVERSED = 1, 2
PF_POSITION, PF_VELOCITY, PF_TIMING, PF_RANGES, PF_DERIVED, PF_SORTED, PF_GRAVITY = range(7)

EXPORTS = [
    "rn_last_error", "rn_abi_version", "rn_scene_create", "rn_scene_destroy", "rn_scene_upload",
    "rn_scene_update_nodes", "rn_scene_update_nodes_async", "rn_scene_node_count", "rn_transform_propagate", "rn_cull", "rn_transform_cull",
    "rn_scene_set_list_mode", "rn_cull_read_counts", "rn_cull_read_list", "rn_cull_read_lists", "rn_cull_read_lists_async", "rn_stream_synchronize", "rn_stream_create", "rn_stream_destroy", "rn_scene_download", "rn_scene_device_outputs",
    "rn_camera_build_frustum", "rn_scenebasic_render_order", "rn_particles_configure", "rn_particles_step",
    "rn_particles_sort", "rn_particles_counts", "rn_particles_download", "rn_scene_get_stats",
    "rn_scene_profile", "rn_scene_profile_read",
    "rn_transform_inverse", "rn_bounds_update", "rn_scene_device_extras", "rn_cull_sort",
    "rn_scenevis_default_volume_order", "rn_particles_build_mesh", "rn_particles_download_mesh",
    "rn_cull_pack_instances", "rn_camera_build_matrices",
    "rn_scenevis_set_node_volumes", "rn_scenevis_set_camera_volumes", "rn_scenevis_choose_volume",
    "rn_particles_expand_billboards", "rn_particles_download_clip", "rn_particles_set_soft_max",
    "rn_particles_set_emitter_transforms", "rn_light_build_cascades", "rn_lod_set_models", "rn_lod_set_node_models", "rn_cull_select_lod",
    "rn_occlusion_set_meshes", "rn_occlusion_set_node_meshes", "rn_occlusion_reset_counters", "rn_cull_occlusion", "rn_occlusion_download",
    "rn_plan_ranges", "rn_pipeline_create", "rn_pipeline_destroy", "rn_pipeline_ranges", "rn_pipeline_update_nodes", "rn_pipeline_submit",
    "rn_pipeline_acquire", "rn_pipeline_release", "rn_frame_list", "rn_pipeline_scene",
    "rn_scene_add_nodes", "rn_scene_remove_nodes", "rn_scene_update_nodes_by_slot", "rn_scene_edit_stats",
    "rn_group_create", "rn_group_destroy", "rn_group_info", "rn_group_update_nodes", "rn_group_frame", "rn_group_synchronize",
    "rn_group_member", "rn_group_expand", "rn_group_read_list",
    "rn_cull_build_masks", "rn_cull_build_masks_to", "rn_cull_build_masks_multicast", "rn_cull_device_masks", "rn_cull_read_masks_async", "rn_masks_expand",
    "rn_masks_expand_static",
]


class RNError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"rn_status {status}: {message}")
        self.status = status


class SceneDesc(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_nodes", C.c_uint32), ("tile_nodes", C.c_uint32),
                ("max_emitters", C.c_uint32), ("max_particles_per_emitter", C.c_uint32), ("tile_inner_nodes", C.c_uint32),
                ("tile_unit_nodes", C.c_uint32), ("flags", C.c_uint32)]


class NodeColumns(C.Structure):
    _fields_ = [("position", C.c_void_p), ("scale", C.c_void_p), ("rotation", C.c_void_p), ("parent", C.c_void_p),
                ("flags", C.c_void_p), ("render_group", C.c_void_p), ("bounding_sphere", C.c_void_p),
                ("bounding_box", C.c_void_p), ("render_priority", C.c_void_p)]


class ParticleMesh(C.Structure):
    _fields_ = [("position", C.c_void_p), ("color", C.c_void_p), ("uv0", C.c_void_p), ("uv1", C.c_void_p),
                ("indices", C.c_void_p), ("vertex_stride_per_emitter", C.c_uint32), ("index_stride_per_emitter", C.c_uint32)]


class ParticleClip(C.Structure):
    _fields_ = [("clip_position", C.c_void_p), ("color", C.c_void_p), ("vertex_stride_per_emitter", C.c_uint32),
                ("reserved", C.c_uint32)]


class MaskSegment(C.Structure):
    _fields_ = [("device_words", C.c_void_p), ("n_words", C.c_uint32), ("rank_base", C.c_uint32)]


class VisList(C.Structure):
    _fields_ = [("slots", C.c_void_p), ("count", C.c_void_p), ("capacity", C.c_uint32), ("reserved", C.c_uint32)]


class SceneStats(C.Structure):
    _fields_ = [("nodes", C.c_uint32), ("tiles", C.c_uint32), ("spine_nodes", C.c_uint32), ("spine_levels", C.c_uint32),
                ("max_depth", C.c_uint32), ("tile_nodes", C.c_uint32), ("device_bytes", C.c_uint64),
                ("kernel_launches", C.c_uint64)]


OCCLUSION_CAMERA_DTYPE = np.dtype([("view", "<f4", 16), ("projection", "<f4", 16), ("pixel_size", "<f4", 2),
                                   ("world_position", "<f4", 3), ("enabled", "<u4")])
assert OCCLUSION_CAMERA_DTYPE.itemsize == 152
FRUSTUM_DTYPE = np.dtype([("center", "<f4", 3), ("radius", "<f4"), ("planes", "<f4", (6, 4)),
                          ("flags", "<u4"), ("render_group", "<u4")])
assert FRUSTUM_DTYPE.itemsize == 120
# rn_camera_desc has the same layout as scenegen.CAMERA_DTYPE (pos/rot are the WORLD position/rotation)
assert CAMERA_DTYPE.itemsize == 164
assert EMITTER_DTYPE.itemsize == 188

_lib = None


def lib_path() -> str:
    return _build.LIB


def load(build: bool = True) -> C.CDLL:
    """Loads (building first if the sources are newer) librn_cuda.so.  Raises if it cannot be had."""
    global _lib
    if _lib is not None:
        return _lib
    if build and os.path.isdir(_build.CSRC):
        try:
            _build.build()
        except Exception:
            if not os.path.exists(_build.LIB):
                raise
    if not os.path.exists(_build.LIB):
        raise RNError(-2, f"{_build.LIB} is missing: build it with `python -m rayne_b200.build` (no CPU fallback exists)")
    lib = C.CDLL(_build.LIB)
    lib.rn_last_error.restype = C.c_char_p
    for name in EXPORTS:
        getattr(lib, name)  # raises AttributeError if the ABI is incomplete
    _lib = lib
    return lib


def _check(rc: int) -> None:
    if rc != 0:
        raise RNError(rc, load().rn_last_error().decode("utf-8", "replace"))


def stream_create(urgency: int = 0) -> int:
    out = C.c_void_p()
    _check(load().rn_stream_create(C.byref(out), int(urgency)))
    return int(out.value or 0)


def stream_destroy(stream: int) -> None:
    _check(load().rn_stream_destroy(C.c_void_p(stream)))


def stream_synchronize(stream: int) -> None:
    _check(load().rn_stream_synchronize(C.c_void_p(stream)))


def _f32(a, shape_last=None) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape_last is not None:
        assert a.shape[-1] == shape_last, a.shape
    return a


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def frustum_from_harness(fr28: np.ndarray, flags: int, group: int) -> np.ndarray:
    """Packs the 28 frustum floats dumped by the reference harness into an rn_camera_frustum record."""
    f = np.zeros((), dtype=FRUSTUM_DTYPE)
    f["center"] = fr28[0:3]
    f["radius"] = fr28[3]
    f["planes"] = fr28[4:28].reshape(6, 4)
    f["flags"] = flags
    f["render_group"] = group
    return f


def build_frustums(cameras: np.ndarray) -> np.ndarray:
    """rn_camera_build_frustum for an array of camera records (scenegen.CAMERA_DTYPE)."""
    lib = load()
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    out = np.zeros(cameras.shape[0], dtype=FRUSTUM_DTYPE)
    for i in range(cameras.shape[0]):
        _check(lib.rn_camera_build_frustum(C.c_void_p(cameras[i:i + 1].ctypes.data), C.c_void_p(out[i:i + 1].ctypes.data)))
    return out


def build_camera_matrices(cameras: np.ndarray):
    """rn_camera_build_matrices for an array of camera records: (view, projection, proj_view), each [V, 16]."""
    lib = load()
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    v = np.zeros((cameras.shape[0], 16), np.float32); p = np.zeros_like(v); pv = np.zeros_like(v)
    for i in range(cameras.shape[0]):
        _check(lib.rn_camera_build_matrices(C.c_void_p(cameras[i:i + 1].ctypes.data), C.c_void_p(v[i].ctypes.data),
                                            C.c_void_p(p[i].ctypes.data), C.c_void_p(pv[i].ctypes.data)))
    return v, p, pv


def build_cascades(target: np.ndarray, shadow: np.ndarray):
    """rn_light_build_cascades: (camera records [n_splits] (scenegen.CAMERA_DTYPE), proj_view [n_splits, 16])."""
    from .scenegen import SHADOW_DTYPE
    lib = load()
    t = np.ascontiguousarray(np.atleast_1d(target), dtype=CAMERA_DTYPE)[:1]
    sh = np.ascontiguousarray(np.atleast_1d(shadow), dtype=SHADOW_DTYPE)[:1]
    n = int(sh["n_splits"][0])
    out = np.zeros(max(n, 1), dtype=CAMERA_DTYPE)
    pv = np.zeros((max(n, 1), 16), np.float32)
    _check(lib.rn_light_build_cascades(C.c_void_p(t.ctypes.data), C.c_void_p(sh.ctypes.data), C.c_void_p(out.ctypes.data),
                                       C.c_void_p(pv.ctypes.data)))
    return out[:n], pv[:n]


def render_order(parent: np.ndarray, priority: np.ndarray | None = None) -> np.ndarray:
    """rn_scenebasic_render_order: order[rank] = creation index."""
    lib = load()
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    prio = None if priority is None else np.ascontiguousarray(priority, dtype=np.int32)
    order = np.zeros(parent.shape[0], dtype=np.uint32)
    _check(lib.rn_scenebasic_render_order(C.c_uint32(parent.shape[0]), _ptr(parent), _ptr(prio), _ptr(order)))
    return order


def choose_volume(bounds: np.ndarray, camera_position) -> int:
    """rn_scenevis_choose_volume: index + 1 of the first AxisAlignedBoxVolume containing the position, 0 = none."""
    b = _f32(np.asarray(bounds, np.float32).reshape(-1, 6))
    p = _f32(camera_position).reshape(3)
    out = C.c_uint8()
    _check(load().rn_scenevis_choose_volume(_ptr(b), C.c_uint32(b.shape[0]), _ptr(p), C.byref(out)))
    return int(out.value)


def vislist_order(parent: np.ndarray) -> np.ndarray:
    """rn_scenevis_default_volume_order: order[rank] = creation index (SceneWithVisibilityLists default volume)."""
    lib = load()
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    order = np.zeros(parent.shape[0], dtype=np.uint32)
    _check(lib.rn_scenevis_default_volume_order(C.c_uint32(parent.shape[0]), _ptr(parent), _ptr(order)))
    return order


class RankedScene:
    """A scenegen.Scene re-ordered into the reference's render-list order (what a Scene subclass hands over).
    vislists=True: the default-volume order of SceneWithVisibilityLists instead of SceneBasic's render list."""

    def __init__(self, scene: Scene, vislists: bool = False, volume_of_node: np.ndarray | None = None):
        self.order = vislist_order(scene.parent) if vislists else render_order(scene.parent, scene.priority)
        self.volume = None
        if volume_of_node is not None:
            # SceneWithVisibilityLists with user volumes: the default volume's nodes in its order, then every volume's
            # nodes in the order of its `nodes` vector (creation order here); self.volume = volume of every rank
            assert vislists
            vol = np.asarray(volume_of_node, dtype=np.uint8)
            parts = [self.order[vol[self.order] == 0]]
            for k in sorted(set(int(v) for v in np.unique(vol)) - {0}):
                parts.append(np.nonzero(vol == k)[0].astype(np.uint32))
            self.order = np.concatenate(parts).astype(np.uint32)
            self.volume = np.ascontiguousarray(vol[self.order])
        self.rank_of = np.empty(scene.n, dtype=np.uint32)                # creation id -> rank
        self.rank_of[self.order] = np.arange(scene.n, dtype=np.uint32)
        o = self.order
        self.position = np.ascontiguousarray(scene.pos[o])
        self.scale = np.ascontiguousarray(scene.scale[o])
        self.rotation = np.ascontiguousarray(scene.rot[o])
        p = scene.parent[o]
        self.parent = np.where(p >= 0, self.rank_of[np.maximum(p, 0)], RN_NO_PARENT).astype(np.uint32)
        self.flags = np.ascontiguousarray(scene.flags[o])
        self.group = np.ascontiguousarray(scene.group[o])
        self.sphere = np.ascontiguousarray(scene.sphere[o])
        self.box = None if scene.box is None else np.ascontiguousarray(scene.box[o])
        self.priority = np.ascontiguousarray(scene.priority[o]) if np.any(scene.priority != 0) else None
        self.n = scene.n
        self.roots = np.nonzero(self.parent == RN_NO_PARENT)[0].astype(np.uint32)


SCENE_STRICT_MATRICES = 1 << 0
VIS_LISTS, VIS_MASKS, VIS_ADAPTIVE = 0, 1, 2


class PipelineDesc(C.Structure):
    _fields_ = [("device", C.c_int32), ("n_ranges", C.c_uint32), ("frames_in_flight", C.c_uint32), ("flags", C.c_uint32),
                ("rank_base", C.c_uint32), ("tile_nodes", C.c_uint32), ("tile_inner_nodes", C.c_uint32), ("tile_unit_nodes", C.c_uint32)]


class FrameSegment(C.Structure):
    _fields_ = [("form", C.c_uint32), ("count", C.c_uint32), ("rank_first", C.c_uint32), ("rank_count", C.c_uint32), ("data", C.c_void_p)]


class FrameView(C.Structure):
    _fields_ = [("frame", C.c_uint64), ("n_cameras", C.c_uint32), ("n_ranges", C.c_uint32), ("segments", C.POINTER(FrameSegment)),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64)]


RN_MAX_GROUP = 16
GROUP_AUTO, GROUP_PEER, GROUP_MULTICAST = 0, 1, 2


class GroupDesc(C.Structure):
    _fields_ = [("n_devices", C.c_uint32), ("devices", C.c_int32 * RN_MAX_GROUP), ("flags", C.c_uint32), ("transport", C.c_uint32),
                ("tile_nodes", C.c_uint32), ("tile_inner_nodes", C.c_uint32), ("tile_unit_nodes", C.c_uint32)]


class Group:
    """Owns one rn_group_t: several GPUs driven by this one process (rn_group.cu)."""

    def __init__(self, rs, devices, transport: int = GROUP_AUTO, strict_matrices: bool = False):
        self.lib = load()
        self._h = C.c_void_p()
        desc = GroupDesc()
        desc.n_devices = len(devices)
        for k, d in enumerate(devices):
            desc.devices[k] = int(d)
        desc.flags = SCENE_STRICT_MATRICES if strict_matrices else 0
        desc.transport = transport
        keep = [_f32(rs.position), _f32(rs.scale), _f32(rs.rotation), np.ascontiguousarray(rs.parent, np.uint32),
                np.ascontiguousarray(rs.flags, np.uint32), np.ascontiguousarray(rs.group, np.uint32), _f32(rs.sphere)]
        cols = NodeColumns(*[_ptr(a) for a in keep], None, None)
        _check(self.lib.rn_group_create(C.byref(self._h), C.byref(desc), C.byref(cols), C.c_uint32(rs.n)))
        self.n, self.n_dev = rs.n, len(devices)
        t = C.c_uint32(0); rw = C.c_uint32(0); note = C.c_char_p()
        b = np.zeros(self.n_dev + 1, np.uint32)
        _check(self.lib.rn_group_info(self._h, C.byref(t), _ptr(b), C.c_uint32(b.shape[0]), C.byref(rw), C.byref(note)))
        self.transport, self.bounds, self.row_words, self.note = int(t.value), b, int(rw.value), (note.value or b"").decode()

    def close(self) -> None:
        if self._h:
            self.lib.rn_group_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def update_nodes(self, ranks, position=None, scale=None, rotation=None, flags=None) -> None:
        ranks = np.ascontiguousarray(ranks, dtype=np.uint32)
        keep = [None if a is None else _f32(a) for a in (position, scale, rotation)]
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint32)
        cols = NodeColumns(_ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), None, _ptr(fl), None, None)
        _check(self.lib.rn_group_update_nodes(self._h, C.byref(cols), _ptr(ranks), C.c_uint32(ranks.shape[0])))

    def frame(self, frustums: np.ndarray) -> None:
        fr = np.ascontiguousarray(np.atleast_1d(frustums), dtype=FRUSTUM_DTYPE)
        _check(self.lib.rn_group_frame(self._h, _ptr(fr), C.c_uint32(fr.shape[0])))

    def synchronize(self) -> None:
        _check(self.lib.rn_group_synchronize(self._h))

    def read_list(self, member: int, camera: int) -> np.ndarray:
        """The whole scene's list of one view as GPU `member` built it from its gathered masks (global ranks), on the host."""
        out = np.zeros(max(self.n, 1), np.uint32)
        cnt = C.c_uint32(0)
        _check(self.lib.rn_group_read_list(self._h, C.c_uint32(member), C.c_uint32(camera), _ptr(out), C.c_uint32(out.shape[0]), C.byref(cnt)))
        return out[:cnt.value].copy()


def plan_ranges(parent: np.ndarray, n_ranges: int) -> np.ndarray:
    """rn_plan_ranges: n_ranges + 1 bounds of contiguous rank ranges that no parent / child edge crosses."""
    parent = np.ascontiguousarray(parent, np.uint32)
    bounds = np.zeros(n_ranges + 1, np.uint32)
    _check(load().rn_plan_ranges(_ptr(parent), C.c_uint32(parent.shape[0]), C.c_uint32(n_ranges), _ptr(bounds)))
    return bounds


MASKS_EXPAND_STATIC_MAX_WORDS = 16 * 1024 * 1024


def masks_expand_static(segments, device_list: int, capacity: int, device_count: int, stream: int = 0, device: int = -1) -> None:
    """rn_masks_expand_static: rn_masks_expand with its scratch in a __device__ array of the library - no scene, no
    allocation.  segments: iterable of (device pointer of the words, n_words, rank_base), at most 32."""
    segments = list(segments)
    segs = (MaskSegment * max(len(segments), 1))()
    for k, (p, nw, base) in enumerate(segments):
        segs[k].device_words, segs[k].n_words, segs[k].rank_base = int(p), int(nw), int(base)
    _check(load().rn_masks_expand_static(C.c_int(device), segs, C.c_uint32(len(segments)), C.c_void_p(device_list),
                                         C.c_uint32(capacity), C.c_void_p(device_count), C.c_void_p(stream)))


class Pipeline:
    """Owns one rn_pipeline_t: the frame with its visibility delivered to host memory (rn_pipeline.cu)."""

    def __init__(self, rs, n_ranges: int = 4, frames_in_flight: int = 2, device: int = 0, rank_base: int = 0, strict_matrices: bool = False,
                 tile_nodes: int = 0, tile_inner_nodes: int = 0, tile_unit_nodes: int = 0):
        self.lib = load()
        self._h = C.c_void_p()
        desc = PipelineDesc(device, n_ranges, frames_in_flight, SCENE_STRICT_MATRICES if strict_matrices else 0, rank_base, tile_nodes,
                            tile_inner_nodes, tile_unit_nodes)
        keep = [_f32(rs.position), _f32(rs.scale), _f32(rs.rotation), np.ascontiguousarray(rs.parent, np.uint32),
                np.ascontiguousarray(rs.flags, np.uint32), np.ascontiguousarray(rs.group, np.uint32), _f32(rs.sphere)]
        box = getattr(rs, "box", None)
        box = None if box is None else _f32(box)
        prio = getattr(rs, "priority", None)
        prio = None if prio is None else np.ascontiguousarray(prio, np.int32)
        cols = NodeColumns(*[_ptr(a) for a in keep], _ptr(box), _ptr(prio))
        _check(self.lib.rn_pipeline_create(C.byref(self._h), C.byref(desc), C.byref(cols), C.c_uint32(rs.n)))
        self.n = rs.n
        self.rank_base = rank_base
        nr = C.c_uint32(0)
        b = np.zeros(n_ranges + 1, np.uint32)
        _check(self.lib.rn_pipeline_ranges(self._h, _ptr(b), C.c_uint32(b.shape[0]), C.byref(nr)))
        self.bounds = b

    def close(self) -> None:
        if self._h:
            self.lib.rn_pipeline_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def update_nodes(self, ranks, position=None, scale=None, rotation=None, flags=None, group=None, sphere=None) -> None:
        ranks = np.ascontiguousarray(ranks, dtype=np.uint32)
        keep = [None if a is None else _f32(a) for a in (position, scale, rotation, sphere)]
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint32)
        gr = None if group is None else np.ascontiguousarray(group, dtype=np.uint32)
        cols = NodeColumns(_ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), None, _ptr(fl), _ptr(gr), _ptr(keep[3]))
        _check(self.lib.rn_pipeline_update_nodes(self._h, C.byref(cols), _ptr(ranks), C.c_uint32(ranks.shape[0])))

    def submit(self, frustums: np.ndarray, form: int = VIS_ADAPTIVE) -> None:
        fr = np.ascontiguousarray(np.atleast_1d(frustums), dtype=FRUSTUM_DTYPE)
        _check(self.lib.rn_pipeline_submit(self._h, _ptr(fr), C.c_uint32(fr.shape[0]), C.c_int(form)))

    def acquire(self) -> FrameView:
        view = FrameView()
        _check(self.lib.rn_pipeline_acquire(self._h, C.byref(view)))
        return view

    def release(self) -> None:
        _check(self.lib.rn_pipeline_release(self._h))

    def frame_lists(self, view: FrameView) -> list:
        """Every camera of an acquired frame as a flat array of global ranks (host decode, rn_frame_list)."""
        out = []
        for c in range(view.n_cameras):
            cnt = C.c_uint32(0)
            _check(self.lib.rn_frame_list(C.byref(view), C.c_uint32(c), None, C.c_uint32(0), C.byref(cnt)))
            a = np.zeros(max(cnt.value, 1), np.uint32)
            _check(self.lib.rn_frame_list(C.byref(view), C.c_uint32(c), _ptr(a), C.c_uint32(a.shape[0]), C.byref(cnt)))
            out.append(a[:cnt.value])
        return out

    def forms(self, view: FrameView) -> np.ndarray:
        """[n_ranges, n_cameras] of the encodings the device chose."""
        return np.array([[view.segments[k * view.n_cameras + c].form for c in range(view.n_cameras)] for k in range(view.n_ranges)])


class SceneHandle:
    """Owns one rn_scene_t."""

    def __init__(self, max_nodes: int, device: int = 0, tile_nodes: int = 0, max_emitters: int = 0,
                 max_particles_per_emitter: int = 0, tile_inner_nodes: int = 0, tile_unit_nodes: int = 0,
                 strict_matrices: bool = False):
        """strict_matrices: RN_SCENE_STRICT_MATRICES - world matrices operation by operation like the reference's
        strict-FP build (bit-exact) instead of the default fused multiply-adds (within 1e-5)."""
        self.lib = load()
        desc = SceneDesc(device, max_nodes, tile_nodes, max_emitters, max_particles_per_emitter, tile_inner_nodes,
                         tile_unit_nodes, SCENE_STRICT_MATRICES if strict_matrices else 0)
        self._h = C.c_void_p()
        _check(self.lib.rn_scene_create(C.byref(self._h), C.byref(desc)))
        self.max_nodes = max_nodes
        self.n = 0

    def close(self) -> None:
        if getattr(self, "_h", None):
            self.lib.rn_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- nodes ----
    def upload(self, rs) -> None:
        """rs: RankedScene or shard.ShardColumns (node columns in rank order)."""
        cols = NodeColumns(_ptr(_f32(rs.position, 3)), _ptr(_f32(rs.scale, 3)), _ptr(_f32(rs.rotation, 4)),
                           _ptr(np.ascontiguousarray(rs.parent, dtype=np.uint32)),
                           _ptr(np.ascontiguousarray(rs.flags, dtype=np.uint32)),
                           _ptr(np.ascontiguousarray(rs.group, dtype=np.uint32)), _ptr(_f32(rs.sphere, 4)))
        box = getattr(rs, "box", None)
        prio = getattr(rs, "priority", None)
        keep_box = None if box is None else _f32(box, 6)
        keep_prio = None if prio is None else np.ascontiguousarray(prio, dtype=np.int32)
        cols.bounding_box = _ptr(keep_box)
        cols.render_priority = _ptr(keep_prio)
        _check(self.lib.rn_scene_upload(self._h, C.byref(cols), 0, rs.n))
        self.n = rs.n

    def update_nodes(self, ranks: np.ndarray, position=None, scale=None, rotation=None, flags=None, group=None,
                     sphere=None, stream: int | None = None) -> None:
        ranks = np.ascontiguousarray(ranks, dtype=np.uint32)
        keep = [None if a is None else _f32(a) for a in (position, scale, rotation, sphere)]
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint32)
        gr = None if group is None else np.ascontiguousarray(group, dtype=np.uint32)
        cols = NodeColumns(_ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), None, _ptr(fl), _ptr(gr), _ptr(keep[3]))
        if stream is None:
            _check(self.lib.rn_scene_update_nodes(self._h, C.byref(cols), _ptr(ranks), ranks.shape[0]))
        else:
            _check(self.lib.rn_scene_update_nodes_async(self._h, C.byref(cols), _ptr(ranks), ranks.shape[0], C.c_void_p(stream)))

    def add_nodes(self, position, scale, rotation, parent, flags, group, sphere, box=None, priority=None) -> np.ndarray:
        """rn_scene_add_nodes: the batch in AddRenderNode call order; parent = index into the batch, PARENT_IS_SLOT | slot, or
        RN_NO_PARENT.  Returns the slots of the new nodes."""
        keep = [_f32(position), _f32(scale), _f32(rotation), np.ascontiguousarray(parent, np.uint32), np.ascontiguousarray(flags, np.uint32),
                np.ascontiguousarray(group, np.uint32), _f32(sphere)]
        bx = None if box is None else _f32(box)
        pr = None if priority is None else np.ascontiguousarray(priority, np.int32)
        cols = NodeColumns(*[_ptr(a) for a in keep], _ptr(bx), _ptr(pr))
        out = np.zeros(keep[3].shape[0], np.uint32)
        _check(self.lib.rn_scene_add_nodes(self._h, C.byref(cols), C.c_uint32(out.shape[0]), _ptr(out)))
        return out

    def remove_nodes(self, slots) -> None:
        sl = np.ascontiguousarray(slots, np.uint32)
        _check(self.lib.rn_scene_remove_nodes(self._h, _ptr(sl), C.c_uint32(sl.shape[0])))

    def update_nodes_by_slot(self, slots, position=None, scale=None, rotation=None, flags=None) -> None:
        sl = np.ascontiguousarray(slots, np.uint32)
        keep = [None if a is None else _f32(a) for a in (position, scale, rotation)]
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint32)
        cols = NodeColumns(_ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), None, _ptr(fl), None, None)
        _check(self.lib.rn_scene_update_nodes_by_slot(self._h, C.byref(cols), _ptr(sl), C.c_uint32(sl.shape[0])))

    def edit_stats(self) -> dict:
        live = C.c_uint32(0); used = C.c_uint32(0); late = C.c_uint32(0); edits = C.c_uint64(0)
        _check(self.lib.rn_scene_edit_stats(self._h, C.byref(live), C.byref(used), C.byref(late), C.byref(edits)))
        return {"live_nodes": live.value, "slots_in_use": used.value, "late_tiles": late.value, "edits": edits.value}

    def update_range(self, first: int, count: int, position=None, scale=None, rotation=None) -> None:
        keep = [None if a is None else _f32(a) for a in (position, scale, rotation)]
        cols = NodeColumns(_ptr(keep[0]), _ptr(keep[1]), _ptr(keep[2]), None, None, None, None)
        _check(self.lib.rn_scene_upload(self._h, C.byref(cols), first, count))

    def transform_propagate(self, stream: int = 0) -> None:
        _check(self.lib.rn_transform_propagate(self._h, C.c_void_p(stream)))

    def _cull(self, fn, frustums: np.ndarray, stream: int):
        fr = np.ascontiguousarray(np.atleast_1d(frustums), dtype=FRUSTUM_DTYPE)
        lists = (VisList * RN_MAX_CAMERAS)()
        _check(fn(self._h, _ptr(fr), C.c_uint32(fr.shape[0]), lists, C.c_void_p(stream)))
        self.ncams = fr.shape[0]
        return lists

    def cull(self, frustums: np.ndarray, stream: int = 0):
        return self._cull(self.lib.rn_cull, frustums, stream)

    def transform_cull(self, frustums: np.ndarray, stream: int = 0):
        return self._cull(self.lib.rn_transform_cull, frustums, stream)

    def transform_inverse(self, stream: int = 0) -> None:
        _check(self.lib.rn_transform_inverse(self._h, C.c_void_p(stream)))

    def bounds_update(self, stream: int = 0) -> None:
        _check(self.lib.rn_bounds_update(self._h, C.c_void_p(stream)))

    def cull_sort(self, camera_positions: np.ndarray, stream: int = 0) -> None:
        cp = _f32(np.atleast_2d(camera_positions), 3)
        _check(self.lib.rn_cull_sort(self._h, _ptr(cp), C.c_uint32(cp.shape[0]), C.c_void_p(stream)))

    def pack_instances(self, camera: int, view: np.ndarray, proj_view: np.ndarray, device_out: int, capacity: int,
                       stream: int = 0) -> None:
        """rn_cull_pack_instances: 256-byte matrix records of camera's list into DEVICE memory at `device_out`."""
        v = _f32(view).reshape(16); pv = _f32(proj_view).reshape(16)
        _check(self.lib.rn_cull_pack_instances(self._h, C.c_uint32(camera), _ptr(v), _ptr(pv), C.c_void_p(device_out),
                                               C.c_uint32(capacity), C.c_void_p(stream)))

    def lod_set_models(self, stage_offsets, stage_distances) -> None:
        so = np.ascontiguousarray(stage_offsets, dtype=np.uint32); sd = _f32(stage_distances)
        _check(self.lib.rn_lod_set_models(self._h, _ptr(so), _ptr(sd), C.c_uint32(so.shape[0] - 1)))

    def lod_set_node_models(self, model_of_rank, first: int = 0) -> None:
        m = np.ascontiguousarray(model_of_rank, dtype=np.uint16)
        _check(self.lib.rn_lod_set_node_models(self._h, _ptr(m), C.c_uint32(first), C.c_uint32(m.shape[0])))

    def select_lod(self, camera: int, lod_camera_position, clip_far: float, device_out: int, capacity: int, stream: int = 0) -> None:
        cp = _f32(lod_camera_position).reshape(3)
        _check(self.lib.rn_cull_select_lod(self._h, C.c_uint32(camera), _ptr(cp), C.c_float(clip_far), C.c_void_p(device_out),
                                           C.c_uint32(capacity), C.c_void_p(stream)))

    def occlusion_set_meshes(self, tri_ofs: np.ndarray, tris: np.ndarray) -> None:
        to = np.ascontiguousarray(tri_ofs, np.uint32); tr = np.ascontiguousarray(tris, np.float32)
        _check(self.lib.rn_occlusion_set_meshes(self._h, _ptr(to), _ptr(tr), C.c_uint32(to.shape[0] - 1)))

    def occlusion_set_node_meshes(self, mesh_of_rank: np.ndarray, first: int = 0) -> None:
        m = np.ascontiguousarray(mesh_of_rank, np.int32)
        _check(self.lib.rn_occlusion_set_node_meshes(self._h, _ptr(m), C.c_uint32(first), C.c_uint32(m.shape[0])))

    def occlusion_reset_counters(self) -> None:
        _check(self.lib.rn_occlusion_reset_counters(self._h))

    def cull_occlusion(self, cameras: np.ndarray, jitter: np.ndarray, order=None, stream: int = 0) -> None:
        """cameras: OCCLUSION_CAMERA_DTYPE records, one per camera of the last cull; jitter: the next len(cameras) draws
        [V, 3]; order: SceneBasic::Render's camera order."""
        cams = np.ascontiguousarray(np.atleast_1d(cameras), dtype=OCCLUSION_CAMERA_DTYPE)
        j = np.ascontiguousarray(jitter, np.float32).reshape(-1, 3)
        assert j.shape[0] >= cams.shape[0]
        o = None if order is None else np.ascontiguousarray(order, np.uint32)
        _check(self.lib.rn_cull_occlusion(self._h, C.c_void_p(cams.ctypes.data), None if o is None else _ptr(o),
                                          C.c_uint32(cams.shape[0]), _ptr(j), C.c_void_p(stream)))

    def occlusion_download(self, camera: int, n: int):
        """(occluders used, jitter taken [3], draws consumed by the call, depth buffer [40, 40], counters of ranks [0, n))"""
        used = C.c_uint32(0); draws = C.c_uint32(0)
        taken = np.zeros(3, np.float32)
        depth = np.zeros(1600, np.float32)
        counters = np.zeros(max(n, 1), np.uint32)
        _check(self.lib.rn_occlusion_download(self._h, C.c_uint32(camera), C.byref(used), _ptr(taken), C.byref(draws), _ptr(depth),
                                              _ptr(counters), C.c_uint32(0), C.c_uint32(n)))
        return int(used.value), taken, int(draws.value), depth.reshape(40, 40), counters[:n]

    def set_node_volumes(self, volume_of_rank: np.ndarray, first: int = 0) -> None:
        v = np.ascontiguousarray(volume_of_rank, dtype=np.uint8)
        _check(self.lib.rn_scenevis_set_node_volumes(self._h, _ptr(v), C.c_uint32(first), C.c_uint32(v.shape[0])))

    def set_camera_volumes(self, volume_per_camera) -> None:
        if volume_per_camera is None:
            _check(self.lib.rn_scenevis_set_camera_volumes(self._h, None, 0))
            return
        v = np.ascontiguousarray(volume_per_camera, dtype=np.uint8)
        _check(self.lib.rn_scenevis_set_camera_volumes(self._h, _ptr(v), C.c_uint32(v.shape[0])))

    def build_masks(self, stream: int = 0) -> None:
        _check(self.lib.rn_cull_build_masks(self._h, C.c_void_p(stream)))

    def build_masks_to(self, device_out: int, row_stride_words: int, stream: int = 0) -> None:
        _check(self.lib.rn_cull_build_masks_to(self._h, C.c_void_p(device_out), C.c_uint32(row_stride_words), C.c_void_p(stream)))

    def build_masks_multicast(self, multicast_out: int, row_stride_words: int, stream: int = 0) -> None:
        _check(self.lib.rn_cull_build_masks_multicast(self._h, C.c_void_p(multicast_out), C.c_uint32(row_stride_words), C.c_void_p(stream)))

    def device_masks(self):
        """(device pointer, words per camera, cameras) of the rank-order visibility bitmasks."""
        p = C.c_void_p(); w = C.c_uint32(); k = C.c_uint32()
        _check(self.lib.rn_cull_device_masks(self._h, C.byref(p), C.byref(w), C.byref(k)))
        return int(p.value or 0), int(w.value), int(k.value)

    def read_masks(self, stream: int = 0, out: np.ndarray | None = None, wait: bool = True) -> np.ndarray:
        """[ncams, words] uint32; `out` may be a pinned host array for an asynchronous copy (wait=False)."""
        _, words, k = self.device_masks()
        if out is None:
            out = np.zeros((k, words), dtype=np.uint32)
        _check(self.lib.rn_cull_read_masks_async(self._h, C.c_uint32(out.shape[0]), C.c_void_p(out.ctypes.data), C.c_void_p(stream)))
        if wait:
            stream_synchronize(stream)
        return out

    def read_masks_into(self, host_ptr: int, n_cameras: int, stream: int = 0) -> None:
        """rn_cull_read_masks_async into raw (pinned) host memory; the caller synchronises the stream."""
        _check(self.lib.rn_cull_read_masks_async(self._h, C.c_uint32(n_cameras), C.c_void_p(host_ptr), C.c_void_p(stream)))

    def masks_expand(self, segments, device_list: int, capacity: int, device_count: int, stream: int = 0) -> None:
        """segments: iterable of (device pointer of the words, n_words, rank_base), expanded in order."""
        segs = (MaskSegment * max(len(segments), 1))()
        for k, (p, nw, base) in enumerate(segments):
            segs[k].device_words, segs[k].n_words, segs[k].rank_base = int(p), int(nw), int(base)
        _check(self.lib.rn_masks_expand(self._h, segs, C.c_uint32(len(segments)), C.c_void_p(device_list),
                                        C.c_uint32(capacity), C.c_void_p(device_count), C.c_void_p(stream)))

    def set_list_mode(self, ranks: bool, rank_offset: int = 0) -> None:
        _check(self.lib.rn_scene_set_list_mode(self._h, int(ranks), C.c_uint32(rank_offset)))

    def read_counts(self, stream: int = 0) -> np.ndarray:
        out = np.zeros(self.ncams, dtype=np.uint32)
        _check(self.lib.rn_cull_read_counts(self._h, _ptr(out), self.ncams, C.c_void_p(stream)))
        return out

    def read_list(self, camera: int, stream: int = 0) -> np.ndarray:
        cnt = C.c_uint32()
        _check(self.lib.rn_cull_read_list(self._h, camera, None, 0, C.byref(cnt), C.c_void_p(stream)))
        out = np.zeros(cnt.value, dtype=np.uint32)
        if cnt.value:
            _check(self.lib.rn_cull_read_list(self._h, camera, _ptr(out), cnt.value, C.byref(cnt), C.c_void_p(stream)))
        return out

    def read_list_into(self, camera: int, out_ptr: int, capacity: int, stream: int = 0) -> int:
        cnt = C.c_uint32()
        _check(self.lib.rn_cull_read_list(self._h, camera, C.c_void_p(out_ptr), capacity, C.byref(cnt), C.c_void_p(stream)))
        return cnt.value

    def read_lists_into(self, out_ptrs: list[int], capacities: list[int], stream: int = 0, wait: bool = True) -> np.ndarray:
        """rn_cull_read_lists[_async] into caller-owned (pinned) host buffers; returns the counts."""
        n = len(out_ptrs)
        ptrs = (C.c_void_p * n)(*out_ptrs)
        caps = (C.c_uint32 * n)(*capacities)
        counts = np.zeros(n, dtype=np.uint32)
        fn = self.lib.rn_cull_read_lists if wait else self.lib.rn_cull_read_lists_async
        _check(fn(self._h, n, ptrs, caps, _ptr(counts), C.c_void_p(stream)))
        return counts

    def download(self, what: int, first: int = 0, count: int | None = None) -> np.ndarray:
        if count is None:
            total = C.c_uint32(0)
            _check(self.lib.rn_scene_node_count(self._h, C.byref(total)))      # grows with rn_scene_add_nodes
            count = total.value - first
        width, dt = {DL_WORLD_POSITION: (3, np.float32), DL_WORLD_ROTATION: (4, np.float32), DL_WORLD_SCALE: (3, np.float32),
                     DL_WORLD_MATRIX: (16, np.float32), DL_SLOT_OF_RANK: (1, np.uint32), DL_RANK_OF_SLOT: (1, np.uint32),
                     DL_INVERSE_WORLD_MATRIX: (16, np.float32), DL_BOUNDING_SPHERE: (4, np.float32),
                     DL_BOUNDING_BOX: (6, np.float32)}[what]
        out = np.zeros((count, width), dtype=dt)
        _check(self.lib.rn_scene_download(self._h, what, first, count, _ptr(out)))
        return out[:, 0] if width == 1 else out

    def stats(self) -> SceneStats:
        st = SceneStats()
        _check(self.lib.rn_scene_get_stats(self._h, C.byref(st)))
        return st

    def profile(self, enable: bool) -> None:
        _check(self.lib.rn_scene_profile(self._h, int(enable)))

    def profile_read(self) -> tuple[float, int]:
        ms = C.c_float(); cnt = C.c_uint32()
        _check(self.lib.rn_scene_profile_read(self._h, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    # ---- particles ----
    def particles_configure(self, emitters: np.ndarray) -> None:
        from .scenegen import upgrade_emitters
        em = np.ascontiguousarray(upgrade_emitters(emitters), dtype=EMITTER_DTYPE)
        _check(self.lib.rn_particles_configure(self._h, _ptr(em), em.shape[0]))
        self.n_emitters = em.shape[0]

    def particles_set_soft_max(self, soft, first: int = 0) -> None:
        v = np.ascontiguousarray(soft, dtype=np.uint32)
        _check(self.lib.rn_particles_set_soft_max(self._h, C.c_uint32(first), C.c_uint32(v.shape[0]), _ptr(v)))

    def particles_set_emitter_transforms(self, world_position=None, world_scale=None, first: int = 0) -> None:
        wp = None if world_position is None else _f32(world_position, 3)
        ws = None if world_scale is None else _f32(world_scale, 3)
        cnt = (wp if wp is not None else ws).shape[0]
        _check(self.lib.rn_particles_set_emitter_transforms(self._h, C.c_uint32(first), C.c_uint32(cnt), _ptr(wp), _ptr(ws)))

    def particles_step(self, delta: float, stream: int = 0) -> None:
        _check(self.lib.rn_particles_step(self._h, C.c_float(delta), C.c_void_p(stream)))

    def particles_sort(self, camera_pos, stream: int = 0) -> None:
        cp = (C.c_float * 3)(*[float(x) for x in camera_pos])
        _check(self.lib.rn_particles_sort(self._h, cp, C.c_void_p(stream)))

    def particles_build_mesh(self, world_scale_x=None, flags=None, stream: int = 0) -> ParticleMesh:
        sc = None if world_scale_x is None else _f32(world_scale_x)
        fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint32)
        out = ParticleMesh()
        _check(self.lib.rn_particles_build_mesh(self._h, _ptr(sc), _ptr(fl), C.byref(out), C.c_void_p(stream)))
        return out

    def particles_download_mesh(self, emitter: int, max_particles: int) -> dict:
        out = {"mesh_pos": np.zeros((max_particles * 4, 3), np.float32), "mesh_color": np.zeros((max_particles * 4, 4), np.float32),
               "mesh_uv0": np.zeros((max_particles * 4, 2), np.float32), "mesh_uv1": np.zeros((max_particles * 4, 2), np.float32),
               "mesh_idx": np.zeros(max_particles * 6, np.uint16)}
        _check(self.lib.rn_particles_download_mesh(self._h, emitter, _ptr(out["mesh_pos"]), _ptr(out["mesh_color"]),
                                                   _ptr(out["mesh_uv0"]), _ptr(out["mesh_uv1"]), _ptr(out["mesh_idx"])))
        return out

    def particles_expand_billboards(self, inverse_model_view: np.ndarray, model_view_projection: np.ndarray,
                                    ambient=(1.0, 1.0, 1.0, 1.0), stream: int = 0) -> "ParticleClip":
        """rn_particles_expand_billboards; matrices [n_emitters, 16] column-major."""
        im = _f32(inverse_model_view).reshape(self.n_emitters, 16)
        mvp = _f32(model_view_projection).reshape(self.n_emitters, 16)
        amb = _f32(ambient).reshape(4)
        out = ParticleClip()
        _check(self.lib.rn_particles_expand_billboards(self._h, _ptr(im), _ptr(mvp), _ptr(amb), C.byref(out), C.c_void_p(stream)))
        return out

    def particles_download_clip(self, emitter: int, max_particles: int) -> dict:
        out = {"clip_pos": np.zeros((max_particles * 4, 4), np.float32), "clip_color": np.zeros((max_particles * 4, 4), np.float32)}
        _check(self.lib.rn_particles_download_clip(self._h, emitter, _ptr(out["clip_pos"]), _ptr(out["clip_color"])))
        return out

    def particles_counts(self) -> tuple[np.ndarray, np.ndarray]:
        counts = np.zeros(self.n_emitters, dtype=np.uint32)
        times = np.zeros(self.n_emitters, dtype=np.float32)
        _check(self.lib.rn_particles_counts(self._h, _ptr(counts), _ptr(times), self.n_emitters))
        return counts, times

    def particles_download(self, emitter: int, field: int) -> np.ndarray:
        cnt = C.c_uint32()
        _check(self.lib.rn_particles_download(self._h, emitter, field, None, 0, C.byref(cnt)))
        width, dt = {PF_POSITION: (3, np.float32), PF_VELOCITY: (3, np.float32), PF_TIMING: (3, np.float32),
                     PF_RANGES: (4, np.float32), PF_DERIVED: (7, np.float32), PF_SORTED: (1, np.uint32),
                     PF_GRAVITY: (3, np.float32)}[field]
        out = np.zeros((cnt.value, width), dtype=dt)
        if cnt.value:
            _check(self.lib.rn_particles_download(self._h, emitter, field, _ptr(out), cnt.value, C.byref(cnt)))
        return out[:, 0] if width == 1 else out
