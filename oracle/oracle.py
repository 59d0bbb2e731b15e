"""TEST INFRASTRUCTURE - ctypes wrapper of oracle/rn_oracle.c (the CPU restatement of the reference path).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from rayne_b200.scenegen import CAMERA_DTYPE, EMITTER_DTYPE, Scene

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "rn_oracle.c")
OUT_DIR = os.path.join(HERE, "_build")
LIB = os.path.join(OUT_DIR, "librn_oracle.so")

FRUSTUM_DTYPE = np.dtype([("center", "<f4", 3), ("radius", "<f4"), ("planes", "<f4", (6, 4)),
                          ("flags", "<u4"), ("render_group", "<u4")])
RN_NO_PARENT = 0xFFFFFFFF
REC = 24

_lib = None


def build(force: bool = False) -> str:
    """gcc -O2 -ffp-contract=off: strict IEEE evaluation, one rounding per operation."""
    os.makedirs(OUT_DIR, exist_ok=True)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-std=c11", "-Wall", "-o", LIB, SRC, "-lm"],
                       check=True)
    return LIB


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if os.path.exists(SRC):
            try:
                build()
            except Exception:
                if not os.path.exists(LIB):
                    raise
        _lib = C.CDLL(LIB)
        _lib.rn_oracle_cull.restype = C.c_uint32
        _lib.rn_oracle_emitters_create.restype = C.c_void_p
        _lib.rn_oracle_emitter_read.restype = C.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


RENDER_ORDER_WALK_LIMIT = 100_000


def render_order(parent: np.ndarray, priority: np.ndarray | None, method: str = "auto") -> np.ndarray:
    """SceneBasic's render-list order.  method "walk": the literal restatement of AddRenderNode's three insert rules
    (quadratic when priorities are scattered); "fast": the same order derived from those rules with one sort
    (rn_oracle_render_order_fast; equal to the walk, tests/test_cpu_oracle_golden.py); "auto": the walk up to
    RENDER_ORDER_WALK_LIMIT nodes, the sort above (the GPU suite's 1 M-node scenes)."""
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    prio = None if priority is None else np.ascontiguousarray(priority, dtype=np.int32)
    order = np.zeros(parent.shape[0], dtype=np.uint32)
    if method == "auto":
        method = "walk" if parent.shape[0] <= RENDER_ORDER_WALK_LIMIT else "fast"
    fn = load().rn_oracle_render_order if method == "walk" else load().rn_oracle_render_order_fast
    rc = fn(C.c_uint32(parent.shape[0]), _p(parent), _p(prio), _p(order))
    assert rc == 0
    return order


def world(parent_u32: np.ndarray, pos: np.ndarray, scale: np.ndarray, rot: np.ndarray):
    """World TRS + matrices for arrays in any common order (parent = index into the same arrays / RN_NO_PARENT)."""
    n = parent_u32.shape[0]
    parent_u32 = np.ascontiguousarray(parent_u32, dtype=np.uint32)
    pos = np.ascontiguousarray(pos, dtype=np.float32); scale = np.ascontiguousarray(scale, dtype=np.float32)
    rot = np.ascontiguousarray(rot, dtype=np.float32)
    wpos = np.zeros((n, 3), np.float32); wrot = np.zeros((n, 4), np.float32); wscale = np.zeros((n, 3), np.float32)
    wmat = np.zeros((n, 16), np.float32)
    load().rn_oracle_world(C.c_uint32(n), _p(parent_u32), _p(pos), _p(scale), _p(rot), _p(wpos), _p(wrot), _p(wscale), _p(wmat))
    return wpos, wrot, wscale, wmat


def inverse(parent_u32: np.ndarray, pos: np.ndarray, scale: np.ndarray, rot: np.ndarray) -> np.ndarray:
    """A3: inverse world matrices (same index conventions as world())."""
    n = parent_u32.shape[0]
    parent_u32 = np.ascontiguousarray(parent_u32, dtype=np.uint32)
    pos = np.ascontiguousarray(pos, dtype=np.float32); scale = np.ascontiguousarray(scale, dtype=np.float32)
    rot = np.ascontiguousarray(rot, dtype=np.float32)
    winv = np.zeros((n, 16), np.float32)
    load().rn_oracle_inverse(C.c_uint32(n), _p(parent_u32), _p(pos), _p(scale), _p(rot), _p(winv))
    return winv


def bounds(wrot: np.ndarray, wscale: np.ndarray, box: np.ndarray | None) -> np.ndarray:
    """A4: transformed AABB extents (minExtend, maxExtend relative to the world position) per node."""
    n = wrot.shape[0]
    wrot = np.ascontiguousarray(wrot, dtype=np.float32); wscale = np.ascontiguousarray(wscale, dtype=np.float32)
    box = None if box is None else np.ascontiguousarray(box, dtype=np.float32)
    out = np.zeros((n, 6), np.float32)
    load().rn_oracle_bounds(C.c_uint32(n), _p(wrot), _p(wscale), _p(box), _p(out))
    return out


def cull_aabb(wpos, aabb, flags, frustum: np.ndarray, in_list: np.ndarray) -> np.ndarray:
    """Extension C5a: filters `in_list` (indices into the arrays) with the AABB-vs-planes test."""
    wpos = np.ascontiguousarray(wpos, dtype=np.float32); aabb = np.ascontiguousarray(aabb, dtype=np.float32)
    flags = np.ascontiguousarray(flags, dtype=np.uint32)
    in_list = np.ascontiguousarray(in_list, dtype=np.uint32)
    fr = np.ascontiguousarray(frustum, dtype=FRUSTUM_DTYPE).reshape(1)
    out = np.zeros(in_list.shape[0], dtype=np.uint32)
    cnt = load().rn_oracle_cull_aabb(C.c_uint32(wpos.shape[0]), _p(wpos), _p(aabb), _p(flags), C.c_void_p(fr.ctypes.data),
                                     _p(in_list), C.c_uint32(in_list.shape[0]), _p(out))
    return out[:cnt].copy()


def vislist_order(parent: np.ndarray) -> np.ndarray:
    """A10: SceneWithVisibilityLists default-volume order (creation ids)."""
    parent = np.ascontiguousarray(parent, dtype=np.int32)
    order = np.zeros(parent.shape[0], dtype=np.uint32)
    rc = load().rn_oracle_vislist_order(C.c_uint32(parent.shape[0]), _p(parent), _p(order))
    assert rc == 0
    return order


def sort_list(lst: np.ndarray, prio: np.ndarray, wpos: np.ndarray, cam_pos, front_to_back: bool,
              transparent_back_to_front: bool) -> np.ndarray:
    """A11: the post-cull sorts of SceneBasic::Render on one list of indices into prio / wpos."""
    lst = np.ascontiguousarray(lst, dtype=np.uint32).copy()
    prio = np.ascontiguousarray(prio, dtype=np.int32); wpos = np.ascontiguousarray(wpos, dtype=np.float32)
    cam = (C.c_float * 3)(*[float(x) for x in cam_pos])
    load().rn_oracle_sort_list(_p(lst), C.c_uint32(lst.shape[0]), _p(prio), _p(wpos), cam, C.c_int(int(front_to_back)),
                               C.c_int(int(transparent_back_to_front)))
    return lst


def instances(ids, wmat, winv, view, projview) -> np.ndarray:
    """Per list entry 64 floats: model, inverse model, view * model, projView * model."""
    ids = np.ascontiguousarray(ids, dtype=np.uint32)
    wmat = np.ascontiguousarray(wmat, dtype=np.float32); winv = np.ascontiguousarray(winv, dtype=np.float32)
    view = np.ascontiguousarray(view, dtype=np.float32); projview = np.ascontiguousarray(projview, dtype=np.float32)
    out = np.zeros((ids.shape[0], 64), np.float32)
    load().rn_oracle_instances(C.c_uint32(ids.shape[0]), _p(ids), _p(wmat), _p(winv), _p(view), _p(projview), _p(out))
    return out


def billboards(pos, color, uv1, inv_mv, mvp, ambient):
    """Shaders/Particles.hlsl particles_vertex for n vertices: (clip [n, 4], colour [n, 4])."""
    pos = np.ascontiguousarray(pos, np.float32); color = np.ascontiguousarray(color, np.float32)
    uv1 = np.ascontiguousarray(uv1, np.float32)
    inv_mv = np.ascontiguousarray(inv_mv, np.float32); mvp = np.ascontiguousarray(mvp, np.float32)
    ambient = np.ascontiguousarray(ambient, np.float32)
    n = pos.shape[0]
    clip = np.zeros((n, 4), np.float32); col = np.zeros((n, 4), np.float32)
    load().rn_oracle_billboards(C.c_uint32(n), _p(pos), _p(color), _p(uv1), _p(inv_mv), _p(mvp), _p(ambient), _p(clip), _p(col))
    return clip, col


def choose_volume(bounds, camera_position) -> int:
    """SceneWithVisibilityLists::Render's volume choice (Source/Scene/RNSceneWithVisibilityLists.cpp:157-166) with
    AxisAlignedBoxVolume::ContainsPosition (:34-37): index + 1 of the FIRST box that contains the position, 0 = none."""
    x, y, z = (np.float32(v) for v in camera_position)
    for k, b in enumerate(np.asarray(bounds, np.float32).reshape(-1, 6)):
        if z >= b[2] and y >= b[1] and x >= b[0] and x <= b[3] and y <= b[4] and z <= b[5]:
            return k + 1
    return 0


def volume_lists(order, volume_of_node, chosen, visible_ids):
    """RenderVolumeList(default) then RenderVolumeList(chosen volume) (:116-123,168-175) for one camera: `order` = the
    default volume's nodes followed by every volume's nodes (creation ids), visible_ids = the ids CanRender accepts."""
    vol = np.asarray(volume_of_node)
    vis = np.zeros(vol.shape[0], bool)
    vis[np.asarray(visible_ids, dtype=np.int64)] = True
    o = np.asarray(order, dtype=np.int64)
    keep = vis[o] & ((vol[o] == 0) | (vol[o] == chosen))
    return o[keep].astype(np.uint32)


def camera_matrices(cameras: np.ndarray):
    """(view, projection, proj_view) per camera, each [V, 16]."""
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    v = np.zeros((cameras.shape[0], 16), np.float32); p = np.zeros_like(v); pv = np.zeros_like(v)
    for i in range(cameras.shape[0]):
        load().rn_oracle_camera_matrices(C.c_void_p(cameras[i:i + 1].ctypes.data), C.c_void_p(v[i].ctypes.data),
                                         C.c_void_p(p[i].ctypes.data), C.c_void_p(pv[i].ctypes.data))
    return v, p, pv


def frustums(cameras: np.ndarray) -> np.ndarray:
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    out = np.zeros(cameras.shape[0], dtype=FRUSTUM_DTYPE)
    for i in range(cameras.shape[0]):
        load().rn_oracle_frustum(C.c_void_p(cameras[i:i + 1].ctypes.data), C.c_void_p(out[i:i + 1].ctypes.data))
    return out


def cull(wpos, wrot, wscale, sphere, flags, group, frustum: np.ndarray, want_sphere: bool = False):
    """Visible indices (into the given arrays, ascending) for ONE frustum record."""
    n = wpos.shape[0]
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (wpos, wrot, wscale, sphere)]
    flags = np.ascontiguousarray(flags, dtype=np.uint32); group = np.ascontiguousarray(group, dtype=np.uint32)
    fr = np.ascontiguousarray(frustum, dtype=FRUSTUM_DTYPE).reshape(1)
    out = np.zeros(n, dtype=np.uint32)
    sph = np.zeros((n, 4), np.float32) if want_sphere else None
    cnt = load().rn_oracle_cull(C.c_uint32(n), _p(arrs[0]), _p(arrs[1]), _p(arrs[2]), _p(arrs[3]), _p(flags), _p(group),
                                C.c_void_p(fr.ctypes.data), _p(out), _p(sph))
    return (out[:cnt].copy(), sph) if want_sphere else out[:cnt].copy()


def scene_reference(scene: Scene, frustum_records: np.ndarray, cameras: np.ndarray | None = None,
                    vislists: bool = False, aabb_test: bool = False) -> dict:
    """Full oracle frame for a scenegen.Scene: order, world TRS/matrices/inverse/AABB (creation order), lists of
    creation ids.  `cameras` (scenegen.CAMERA_DTYPE records): also apply the post-cull sorts of SceneBasic::Render
    with each camera's flags | Camera::Flags::Defaults (A11).  vislists: SceneWithVisibilityLists default-volume
    order, no sorts (A10).  aabb_test: the RN_CAMERA_TEST_AABB extension on every camera."""
    from rayne_b200.scenegen import CAM_DEFAULTS
    order = vislist_order(scene.parent) if vislists else render_order(scene.parent, scene.priority)
    parent = np.where(scene.parent >= 0, scene.parent, RN_NO_PARENT).astype(np.uint32)
    wpos, wrot, wscale, wmat = world(parent, scene.pos, scene.scale, scene.rot)
    winv = inverse(parent, scene.pos, scene.scale, scene.rot)
    aabb = bounds(wrot, wscale, scene.box)
    lists, unsorted = [], []
    for c, fr in enumerate(np.atleast_1d(frustum_records)):
        vis = cull(wpos[order], wrot[order], wscale[order], scene.sphere[order], scene.flags[order], scene.group[order], fr)
        lst = order[vis]
        if aabb_test:
            lst = cull_aabb(wpos, aabb, scene.flags, fr, lst)
        unsorted.append(lst)
        if cameras is not None and not vislists:
            flags = int(np.atleast_1d(cameras)[c]["flags"]) | CAM_DEFAULTS
            lst = sort_list(lst, scene.priority, wpos, np.atleast_1d(cameras)[c]["pos"], bool(flags & (1 << 6)), bool(flags & (1 << 5)))
        lists.append(lst)
    return {"order": order, "wpos": wpos, "wrot": wrot, "wscale": wscale, "wmat": wmat, "winv": winv, "aabb": aabb,
            "lists": lists, "lists_unsorted": unsorted}


class Emitters:
    def __init__(self, emitters: np.ndarray):
        from rayne_b200.scenegen import upgrade_emitters
        self.em = np.ascontiguousarray(upgrade_emitters(emitters), dtype=EMITTER_DTYPE)
        self.n = self.em.shape[0]
        self.h = C.c_void_p(load().rn_oracle_emitters_create(_p(self.em), C.c_uint32(self.n)))

    def step(self, delta: float, steps: int = 1) -> None:
        for _ in range(steps):
            load().rn_oracle_emitters_step(self.h, C.c_uint32(self.n), C.c_float(delta))

    def set_soft_max(self, k: int, soft: int) -> None:
        load().rn_oracle_emitter_set_soft_max(self.h, C.c_uint32(k), C.c_uint32(soft))

    def set_transform(self, k: int, world_position=None, world_scale=None) -> None:
        wp = None if world_position is None else np.ascontiguousarray(world_position, np.float32)
        ws = None if world_scale is None else np.ascontiguousarray(world_scale, np.float32)
        load().rn_oracle_emitter_set_transform(self.h, C.c_uint32(k), _p(wp) if wp is not None else None,
                                               _p(ws) if ws is not None else None)

    def read(self, k: int):
        t = C.c_float()
        cap = int(load().rn_oracle_emitter_read(self.h, C.c_uint32(k), None, C.c_uint32(0), C.byref(t))) + 1
        rec = np.zeros((cap, REC), np.float32)
        cnt = load().rn_oracle_emitter_read(self.h, C.c_uint32(k), _p(rec), C.c_uint32(cap), C.byref(t))
        return rec[:cnt].copy(), np.float32(t.value)

    def sort(self, k: int, cam):
        rec, _ = self.read(k)
        idx = np.zeros(rec.shape[0], np.uint32); key = np.zeros(rec.shape[0], np.float32)
        camv = (C.c_float * 3)(*[float(x) for x in cam])
        load().rn_oracle_emitter_sort(self.h, C.c_uint32(k), camv, _p(idx), _p(key))
        return idx, key

    def mesh(self, k: int, world_scale_x: float = 1.0, roll: bool = False, inversed: bool = False) -> dict:
        """P6: the render mesh ParticleEmitter::UpdateMesh would build for emitter k."""
        mx = int(self.em["max_particles"][k])
        out = {"mesh_pos": np.zeros((mx * 4, 3), np.float32), "mesh_color": np.zeros((mx * 4, 4), np.float32),
               "mesh_uv0": np.zeros((mx * 4, 2), np.float32), "mesh_uv1": np.zeros((mx * 4, 2), np.float32),
               "mesh_idx": np.zeros(mx * 6, np.uint16)}
        load().rn_oracle_emitter_mesh(self.h, C.c_uint32(k), C.c_float(world_scale_x), C.c_int(int(roll)), C.c_int(int(inversed)),
                                      _p(out["mesh_pos"]), _p(out["mesh_color"]), _p(out["mesh_uv0"]), _p(out["mesh_uv1"]),
                                      _p(out["mesh_idx"]))
        return out

    def close(self):
        if self.h:
            load().rn_oracle_emitters_destroy(self.h, C.c_uint32(self.n))
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def select_lod(ids: np.ndarray, wpos: np.ndarray, model_of_node: np.ndarray, stage_offsets: np.ndarray, stage_distances: np.ndarray,
               cam_pos, clip_far: float) -> np.ndarray:
    """LOD stage per list entry (extension, see rn_oracle_select_lod); ids index wpos / model_of_node."""
    ids = np.ascontiguousarray(ids, np.uint32); wpos = np.ascontiguousarray(wpos, np.float32)
    mo = np.ascontiguousarray(model_of_node, np.uint16)
    so = np.ascontiguousarray(stage_offsets, np.uint32); sd = np.ascontiguousarray(stage_distances, np.float32)
    cam = np.ascontiguousarray(cam_pos, np.float32)
    out = np.zeros(ids.shape[0], np.uint8)
    load().rn_oracle_select_lod(_p(ids), C.c_uint32(ids.shape[0]), _p(wpos), _p(mo), _p(so), _p(sd), C.c_uint32(so.shape[0] - 1),
                                _p(cam), C.c_float(clip_far), _p(out))
    return out


OCC_PARAMS_DTYPE = np.dtype([("view", "<f4", 16), ("projection", "<f4", 16), ("offset", "<f4", 3), ("pixel", "<f4", 2),
                             ("cam_pos", "<f4", 3)])
FLAG_OCCLUDER = 8


def jitter(seed: int, n_draws: int) -> np.ndarray:
    """The occlusion pass's camera offsets as the reference's shared generator yields them after Seed(seed)."""
    out = np.zeros((n_draws, 3), np.float32)
    load().rn_oracle_jitter(C.c_uint32(seed), C.c_uint32(n_draws), _p(out))
    return out


def occlusion(order, can_render, flags, priority, wpos, radius, aabb, wmat, mesh_of, tri_ofs, tris, counter, view, projection,
              offset, pixel, cam_pos):
    """C8 for ONE camera and frame (rn_oracle_occlusion).  Arrays in creation order; `counter` is updated in place.
    Returns (used, list of creation ids, depth buffer [40, 40]); used False = no occluder passed CanRender (plain cull)."""
    lib = load()
    lib.rn_oracle_occlusion.restype = C.c_uint32
    n = len(order)
    order = np.ascontiguousarray(order, np.uint32)
    can_render = np.ascontiguousarray(can_render, np.uint8)
    flags = np.ascontiguousarray(flags, np.uint32); priority = np.ascontiguousarray(priority, np.int32)
    wpos = np.ascontiguousarray(wpos, np.float32); radius = np.ascontiguousarray(radius, np.float32)
    aabb = np.ascontiguousarray(aabb, np.float32); wmat = np.ascontiguousarray(wmat, np.float32)
    mesh_of = np.ascontiguousarray(mesh_of, np.int32)
    tri_ofs = np.ascontiguousarray(tri_ofs, np.uint32); tris = np.ascontiguousarray(tris, np.float32)
    assert counter.dtype == np.uint32 and counter.flags.c_contiguous
    p = np.zeros((), OCC_PARAMS_DTYPE)
    p["view"] = view; p["projection"] = projection; p["offset"] = offset; p["pixel"] = pixel; p["cam_pos"] = cam_pos
    out = np.zeros(max(n, 1), np.uint32)
    depth = np.zeros(1600, np.float32)
    used = C.c_int(0)
    cnt = lib.rn_oracle_occlusion(C.c_uint32(n), _p(order), _p(can_render), _p(flags), _p(priority), _p(wpos), _p(radius), _p(aabb),
                                  _p(wmat), _p(mesh_of), _p(tri_ofs), _p(tris), _p(counter), C.c_void_p(p.ctypes.data), _p(out),
                                  _p(depth), C.byref(used))
    return bool(used.value), out[:cnt].copy(), depth.reshape(40, 40)


def occlusion_frames(scene: Scene, cameras: np.ndarray, tri_ofs, tris, mesh_of, frames: int, seed: int,
                     frame_size=(1280.0, 720.0), root_translate=(0.0, 0.0, 0.0), camera_translate=(0.0, 0.0, 0.0)) -> list:
    """SceneBasic::Update + Render over `frames` frames with the software occlusion pass, restated: per frame a list of
    per-camera dicts (used, offset, list of creation ids after the post-cull sorts, view, projection, cam_pos, frustum)
    plus the counters after the frame.  Cameras are visited in SceneBasic::Render's order; every camera that takes the
    occlusion path consumes one jitter draw of the generator seeded with `seed`."""
    from rayne_b200.scenegen import CAM_DEFAULTS, camera_render_order
    cams = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE).copy()
    order = render_order(scene.parent, scene.priority)
    parent = np.where(scene.parent >= 0, scene.parent, RN_NO_PARENT).astype(np.uint32)
    pos = scene.pos.copy()
    roots = np.nonzero(scene.parent < 0)[0]
    counter = np.zeros(scene.n, np.uint32)
    draws = jitter(seed, frames * len(cams))
    drawn = 0
    pixel = np.float32(1.0) / np.float32(frame_size)
    visit = camera_render_order(cams["flags"])
    out = []
    for f in range(frames):
        if f > 0:
            pos[roots] += np.float32(root_translate)
            cams["pos"] += np.float32(camera_translate)
        wpos, wrot, wscale, wmat = world(parent, pos, scene.scale, scene.rot)
        aabb = bounds(wrot, wscale, scene.box)
        fr = frustums(cams)
        view, proj, _ = camera_matrices(cams)
        frame = {"cams": [None] * len(cams)}
        for c in visit:
            vis, sph = cull(wpos[order], wrot[order], wscale[order], scene.sphere[order], scene.flags[order], scene.group[order], fr[c],
                            want_sphere=True)
            radius = np.zeros(scene.n, np.float32); radius[order] = sph[:, 3]
            can = np.zeros(scene.n, np.uint8); can[order[vis]] = 1
            offset = draws[drawn]
            used, lst, depth = occlusion(order, can, scene.flags, scene.priority, wpos, radius, aabb, wmat, mesh_of, tri_ofs, tris,
                                         counter, view[c], proj[c], offset, pixel, cams["pos"][c])
            if used:
                drawn += 1
            else:
                lst, offset = order[vis], np.zeros(3, np.float32)
            flags = int(cams[c]["flags"]) | CAM_DEFAULTS
            unsorted = lst
            lst = sort_list(lst, scene.priority, wpos, cams["pos"][c], bool(flags & (1 << 6)), bool(flags & (1 << 5)))
            frame["cams"][c] = {"used": int(used), "offset": offset.copy(), "list": lst, "list_unsorted": unsorted, "view": view[c].copy(),
                                "projection": proj[c].copy(), "cam_pos": cams["pos"][c].copy(), "frustum": fr[c].copy(),
                                "depth": depth if used else None}
        frame["counters"] = counter.copy()
        frame["wpos"] = wpos; frame["aabb"] = aabb; frame["wmat"] = wmat
        out.append(frame)
    return out
