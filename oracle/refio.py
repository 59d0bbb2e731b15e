"""TEST INFRASTRUCTURE - file IO for oracle/_ref/ref_harness_* (the reference itself, compiled by build_ref.sh).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.

Scene input  ("RNSI" v1; + f32 box[n,6] after sphere when options & OPT_BOX): header u32 magic, version, n_nodes, n_cameras, n_frames, options; f32 root_translate[3];
    i32 parent[n]; f32 pos[n,3]; f32 scale[n,3]; f32 rot[n,4]; u32 flags[n]; u32 group[n]; i32 priority[n];
    f32 sphere[n,4]; CAMERA_DTYPE cameras[n_cameras]           (all creation order, little endian)
Scene output ("RNSO" v2: + f32 aabb[n,6] (world minExtend, maxExtend) after winv): u32 magic, version, n_nodes, n_cameras; u32 n_order; u32 order[n_order] (render list,
    head->tail, creation ids); f32 wpos[n,3]; wrot[n,4]; wscale[n,3]; wmat[n,16]; sphere[n,4] (centre, radius);
    u32 has_inv; f32 winv[n,16] if has_inv; per camera: f32 frustum[28] (centre3, radius, 6x(normal3, d) in the
    order left,right,top,bottom,near,far), u32 flags, u32 group, u32 count, u32 list[count] (creation ids)
Particle input ("RNPI" v3): u32 magic, version, n_emitters, n_steps; f32 delta; u32 options, soft_after_steps; EMITTER_DTYPE emitters[n]
Particle output ("RNPO" v2): u32 magic, version, n_emitters; per emitter u32 count, f32 time, f32 rec[count,24],
    u32 mesh_max (0 = no mesh); if mesh_max: f32 pos[4m,3], color[4m,4], uv0[4m,2], uv1[4m,2], u16 idx[6m]
"""
from __future__ import annotations

import json
import os
import subprocess
import tempfile

import numpy as np

from rayne_b200.scenegen import CAMERA_DTYPE, EMITTER_DTYPE, Scene

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

OPT_INVERSE = 1
OPT_TOUCH_MATRICES = 2
OPT_VISLISTS = 4      # SceneWithVisibilityLists (default volume) instead of SceneBasic
OPT_BOX = 8           # the input carries scene.box (local AABB column)
OPT_VOLUMES = 32      # with OPT_VISLISTS: volumes = (bounds [V, 6], volume_of_node [n]) -> SceneWithVisibilityLists::AddVolume
OPT_INSTANCES = 16    # dump view / projView per camera and view*world, projView*world of every node for camera 0
POPT_ROLL, POPT_INVERSED, POPT_MESH = 1, 2, 4

# columns of the 24-float particle record written by the harness
P_POS, P_VEL, P_TIME, P_LIFE, P_DUR = slice(0, 3), slice(3, 6), 6, 7, 8
P_SIZE0, P_SIZE1, P_ROT0, P_ROT1, P_ROT = 9, 10, 11, 12, 13
P_SIZE, P_COLOR, P_GRAVITY = slice(14, 16), slice(16, 20), slice(20, 23)


def harness_path(kind: str = "strict") -> str:
    return os.path.join(REF_DIR, f"ref_harness_{kind}")


def have_harness(kind: str = "strict") -> bool:
    return os.access(harness_path(kind), os.X_OK)


def write_scene_input(path: str, scene: Scene, cameras: np.ndarray, frames: int = 1, options: int = 0,
                      root_translate=(0.0, 0.0, 0.0), volumes=None) -> None:
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    with open(path, "wb") as f:
        f.write(np.array([0x49534E52, 1, scene.n, cameras.shape[0], frames, options], dtype="<u4").tobytes())
        f.write(np.asarray(root_translate, dtype="<f4").tobytes())
        f.write(np.ascontiguousarray(scene.parent, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(scene.pos, dtype="<f4").tobytes())
        f.write(np.ascontiguousarray(scene.scale, dtype="<f4").tobytes())
        f.write(np.ascontiguousarray(scene.rot, dtype="<f4").tobytes())
        f.write(np.ascontiguousarray(scene.flags, dtype="<u4").tobytes())
        f.write(np.ascontiguousarray(scene.group, dtype="<u4").tobytes())
        f.write(np.ascontiguousarray(scene.priority, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(scene.sphere, dtype="<f4").tobytes())
        if options & OPT_BOX:
            f.write(np.ascontiguousarray(scene.box, dtype="<f4").tobytes())
        if options & OPT_VOLUMES:
            bounds, volume_of_node = volumes
            bounds = np.ascontiguousarray(bounds, dtype="<f4").reshape(-1, 6)
            f.write(np.array([bounds.shape[0]], dtype="<u4").tobytes())
            f.write(bounds.tobytes())
            f.write(np.ascontiguousarray(volume_of_node, dtype="<u4").tobytes())
        f.write(cameras.tobytes())


def read_scene_output(path: str, with_instances: bool = False) -> dict:
    buf = np.fromfile(path, dtype=np.uint8)
    off = 0

    def take(dtype, count):
        nonlocal off
        dt = np.dtype(dtype)
        a = np.frombuffer(buf, dtype=dt, count=count, offset=off)
        off += dt.itemsize * count
        return a

    magic, version, n, ncam = take("<u4", 4)
    assert magic == 0x4F534E52 and version == 2
    n = int(n); ncam = int(ncam)
    norder = int(take("<u4", 1)[0])
    out = {"order": take("<u4", norder).copy()}
    out["wpos"] = take("<f4", n * 3).reshape(n, 3).copy()
    out["wrot"] = take("<f4", n * 4).reshape(n, 4).copy()
    out["wscale"] = take("<f4", n * 3).reshape(n, 3).copy()
    out["wmat"] = take("<f4", n * 16).reshape(n, 16).copy()
    out["sphere"] = take("<f4", n * 4).reshape(n, 4).copy()
    has_inv = int(take("<u4", 1)[0])
    out["winv"] = take("<f4", n * 16).reshape(n, 16).copy() if has_inv else None
    out["aabb"] = take("<f4", n * 6).reshape(n, 6).copy()
    out["frustum"] = np.zeros((ncam, 28), dtype=np.float32)
    out["cam_flags"] = np.zeros(ncam, dtype=np.uint32)
    out["cam_group"] = np.zeros(ncam, dtype=np.uint32)
    out["lists"] = []
    for c in range(ncam):
        out["frustum"][c] = take("<f4", 28)
        out["cam_flags"][c], out["cam_group"][c], cnt = take("<u4", 3)
        out["lists"].append(take("<u4", int(cnt)).copy())
    if with_instances:
        out["view"] = np.zeros((ncam, 16), np.float32); out["projview"] = np.zeros((ncam, 16), np.float32)
        for c in range(ncam):
            out["view"][c] = take("<f4", 16); out["projview"][c] = take("<f4", 16)
            if c == 0:
                out["mv"] = take("<f4", n * 16).reshape(n, 16).copy()
                out["mvp"] = take("<f4", n * 16).reshape(n, 16).copy()
    assert off == buf.shape[0], (off, buf.shape[0])
    return out


def run_scene(scene: Scene, cameras: np.ndarray, frames: int = 1, options: int = 0,
              root_translate=(0.0, 0.0, 0.0), kind: str = "strict", volumes=None) -> dict:
    """Run the real reference (SceneBasic::Update + Render) on the scene; returns the parsed dump."""
    with tempfile.TemporaryDirectory() as tmp:
        fin, fout = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out.bin")
        write_scene_input(fin, scene, cameras, frames, options, root_translate, volumes)
        subprocess.run([harness_path(kind), "scene", fin, fout], check=True, stdout=subprocess.DEVNULL)
        return read_scene_output(fout, with_instances=bool(options & OPT_INSTANCES))


def bench_scene(scene: Scene, cameras: np.ndarray, frames: int, warmup: int,
                root_translate=(0.01, 0.0, 0.0), kind: str = "release") -> dict:
    with tempfile.TemporaryDirectory() as tmp:
        fin = os.path.join(tmp, "in.bin")
        write_scene_input(fin, scene, cameras, 1, OPT_TOUCH_MATRICES, root_translate)
        r = subprocess.run([harness_path(kind), "bench-scene", fin, str(frames), str(warmup)], check=True,
                           capture_output=True, text=True)
        return json.loads(r.stdout.strip().splitlines()[-1])


def run_cascades(target: np.ndarray, shadow: np.ndarray, kind: str = "strict") -> dict:
    """The reference's Light::UpdateShadows + Camera::MakeShadowSplit (harness command `cascades`).  target: one
    CAMERA_DTYPE record; shadow: one scenegen.SHADOW_DTYPE record.  Returns per split: pos [n,3], rot [n,4], ortho
    [n,4] (left, right, bottom, top), clip [n,2], flags [n], frustum [n,28], projview [n,16]."""
    from rayne_b200.scenegen import SHADOW_DTYPE
    t = np.ascontiguousarray(np.atleast_1d(target), dtype=CAMERA_DTYPE)[:1]
    sh = np.ascontiguousarray(np.atleast_1d(shadow), dtype=SHADOW_DTYPE)[0]
    n = int(sh["n_splits"])
    with tempfile.TemporaryDirectory() as tmp:
        fin, fout = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.array([0x49434E52, 1], dtype="<u4").tobytes())
            f.write(t.tobytes())
            f.write(np.asarray(sh["light_rot"], "<f4").tobytes())
            f.write(np.array([n], dtype="<u4").tobytes())
            f.write(np.asarray(sh["split_max_distance"][:n], "<f4").tobytes())
            f.write(np.array([sh["resolution"], sh["clip_near"], sh["clip_far"], sh["directional_shadow_distance"],
                              sh["distance_blend_factor"], sh["max_shadow_dist"]], dtype="<f4").tobytes())
        subprocess.run([harness_path(kind), "cascades", fin, fout], check=True, stdout=subprocess.DEVNULL)
        buf = np.fromfile(fout, dtype=np.uint8)
    magic, version, ns = np.frombuffer(buf, dtype="<u4", count=3)
    assert magic == 0x4F434E52 and version == 1 and ns == n
    rec = np.dtype([("pos", "<f4", 3), ("rot", "<f4", 4), ("ortho", "<f4", 4), ("clip", "<f4", 2), ("flags", "<u4"),
                    ("frustum", "<f4", 28), ("projview", "<f4", 16)])
    arr = np.frombuffer(buf, dtype=rec, count=n, offset=12)
    return {k: arr[k].copy() for k in rec.names}


def write_particle_input(path: str, emitters: np.ndarray, steps: int, delta: float, options: int = 0,
                         soft_after_steps: int = 0) -> None:
    from rayne_b200.scenegen import upgrade_emitters
    emitters = np.ascontiguousarray(upgrade_emitters(emitters), dtype=EMITTER_DTYPE)
    with open(path, "wb") as f:
        f.write(np.array([0x49504E52, 3, emitters.shape[0], steps], dtype="<u4").tobytes())
        f.write(np.array([delta], dtype="<f4").tobytes())
        f.write(np.array([options, soft_after_steps], dtype="<u4").tobytes())
        f.write(emitters.tobytes())


def read_particle_output(path: str) -> list[dict]:
    buf = np.fromfile(path, dtype=np.uint8)
    magic, version, ne = np.frombuffer(buf, dtype="<u4", count=3)
    assert magic == 0x4F504E52 and version == 2
    off = 12
    out = []
    for _ in range(int(ne)):
        cnt = int(np.frombuffer(buf, dtype="<u4", count=1, offset=off)[0])
        time = float(np.frombuffer(buf, dtype="<f4", count=1, offset=off + 4)[0])
        off += 8
        rec = np.frombuffer(buf, dtype="<f4", count=cnt * 24, offset=off).reshape(cnt, 24).copy()
        off += cnt * 96
        item = {"count": cnt, "time": np.float32(time), "rec": rec}
        mx = int(np.frombuffer(buf, dtype="<u4", count=1, offset=off)[0])
        off += 4
        if mx:
            for name, width in (("mesh_pos", 3), ("mesh_color", 4), ("mesh_uv0", 2), ("mesh_uv1", 2)):
                item[name] = np.frombuffer(buf, dtype="<f4", count=mx * 4 * width, offset=off).reshape(mx * 4, width).copy()
                off += mx * 4 * width * 4
            item["mesh_idx"] = np.frombuffer(buf, dtype="<u2", count=mx * 6, offset=off).copy()
            off += mx * 12
        out.append(item)
    assert off == buf.shape[0]
    return out


def run_particles(emitters: np.ndarray, steps: int, delta: float = 1.0 / 60.0, kind: str = "strict",
                  options: int = 0, soft_after_steps: int = 0) -> list[dict]:
    """soft_after_steps: emitters with EMITTER_SOFT_MAX get SetMaxParticlesSoft only after that many steps (0 = at
    creation) - lowering the soft maximum below the live count exercises the reference's unsigned wrap."""
    with tempfile.TemporaryDirectory() as tmp:
        fin, fout = os.path.join(tmp, "in.bin"), os.path.join(tmp, "out.bin")
        write_particle_input(fin, emitters, steps, delta, options, soft_after_steps)
        subprocess.run([harness_path(kind), "particles", fin, fout], check=True, stdout=subprocess.DEVNULL)
        return read_particle_output(fout)


def bench_particles(emitters: np.ndarray, steps: int, warmup: int, delta: float = 1.0 / 60.0,
                    kind: str = "release") -> dict:
    with tempfile.TemporaryDirectory() as tmp:
        fin = os.path.join(tmp, "in.bin")
        write_particle_input(fin, emitters, 0, delta)
        r = subprocess.run([harness_path(kind), "bench-particles", fin, str(steps), str(warmup)], check=True,
                           capture_output=True, text=True)
        return json.loads(r.stdout.strip().splitlines()[-1])


MESH_DESC_DTYPE = np.dtype([("kind", "<u4"), ("p", "<f4", 4)])   # kind 0: Mesh::WithColoredCube(size xyz)


def run_occlusion(scene: Scene, cameras: np.ndarray, meshes: np.ndarray, mesh_of_node: np.ndarray, frames: int, seed: int,
                  frame_size=(1280.0, 720.0), root_translate=(0.0, 0.0, 0.0), camera_translate=(0.0, 0.0, 0.0),
                  kind: str = "strict", custom_meshes=()) -> dict:
    """SceneBasic::Render with Flags::Occluder entities (software occlusion culling, RNSceneBasic.cpp:445-573) for
    `frames` frames; harness command `occlusion`.  meshes: MESH_DESC_DTYPE records (cubes); custom_meshes: (vertices [V, 3],
    uint16 indices) pairs built through the reference's Mesh class, ids continuing after the cubes; mesh_of_node: i32 [n],
    -1 = plain node.
    Returns order, mesh triangle soups (tri_ofs, tris [T, 9]) and per frame and camera: used, offset, view, projection,
    cam_pos, frustum, flags, group, list; per frame the counters and the depth buffer the last camera left."""
    cameras = np.ascontiguousarray(np.atleast_1d(cameras), dtype=CAMERA_DTYPE)
    meshes = np.ascontiguousarray(np.atleast_1d(meshes), dtype=MESH_DESC_DTYPE)
    with tempfile.TemporaryDirectory() as tmp:
        fin, focc, fout = (os.path.join(tmp, x) for x in ("in.bin", "occ.bin", "out.bin"))
        write_scene_input(fin, scene, cameras, frames, OPT_BOX if scene.box is not None else 0, root_translate)
        with open(focc, "wb") as f:
            f.write(np.array([0x494F4E52, 1, seed], dtype="<u4").tobytes())
            f.write(np.asarray(list(frame_size) + list(camera_translate), dtype="<f4").tobytes())
            f.write(np.array([meshes.shape[0]], dtype="<u4").tobytes())
            f.write(meshes.tobytes())
            f.write(np.ascontiguousarray(mesh_of_node, dtype="<i4").tobytes())
            f.write(np.array([len(custom_meshes)], dtype="<u4").tobytes())
            for verts, idx in custom_meshes:
                verts = np.ascontiguousarray(verts, dtype="<f4").reshape(-1, 3)
                idx = np.ascontiguousarray(idx, dtype="<u2")
                f.write(np.array([verts.shape[0], idx.shape[0]], dtype="<u4").tobytes())
                f.write(verts.tobytes())
                f.write(idx.tobytes())
        subprocess.run([harness_path(kind), "occlusion", fin, focc, fout], check=True, stdout=subprocess.DEVNULL)
        buf = np.fromfile(fout, dtype=np.uint8)
    off = 0

    def take(dtype, count):
        nonlocal off
        dt = np.dtype(dtype)
        a = np.frombuffer(buf, dtype=dt, count=count, offset=off)
        off += dt.itemsize * count
        return a

    magic, version, n, ncam, nframes = (int(x) for x in take("<u4", 5))
    assert magic == 0x4F4F4E52 and version == 1
    out = {"order": take("<u4", int(take("<u4", 1)[0])).copy()}
    nm = int(take("<u4", 1)[0])
    tri_ofs, tris = [0], []
    for _ in range(nm):
        nt = int(take("<u4", 1)[0])
        tris.append(take("<f4", nt * 9).reshape(nt, 9).copy())
        tri_ofs.append(tri_ofs[-1] + nt)
    out["tri_ofs"] = np.asarray(tri_ofs, np.uint32)
    out["tris"] = np.concatenate(tris) if tris else np.zeros((0, 9), np.float32)
    out["frames"] = []
    for _ in range(nframes):
        fr = {"cams": []}
        for _c in range(ncam):
            cam = {"used": int(take("<u4", 1)[0]), "offset": take("<f4", 3).copy(), "view": take("<f4", 16).copy(),
                   "projection": take("<f4", 16).copy(), "cam_pos": take("<f4", 3).copy(), "frustum": take("<f4", 28).copy()}
            cam["flags"], cam["group"], cnt = (int(x) for x in take("<u4", 3))
            cam["list"] = take("<u4", cnt).copy()
            fr["cams"].append(cam)
        fr["counters"] = take("<u4", n).copy()
        fr["depth"] = take("<f4", 1600).reshape(40, 40).copy()
        out["frames"].append(fr)
    assert off == buf.shape[0], (off, buf.shape[0])
    return out
