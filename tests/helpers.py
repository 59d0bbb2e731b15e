"""Shared helpers for the parity tests: run the CUDA path through the C ABI and compare it with the oracle."""
from __future__ import annotations

import os

import numpy as np

from oracle import oracle
from rayne_b200 import capi, scenegen as sg

REL_TOL = 1e-5  # north star: world matrices / particle state within 1e-5 relative fp32
GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden_scene(name: str):
    """(scene, cameras, npz) of a fixture written by tests/golden/make_golden.py (outputs of the reference)."""
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    scene = sg.Scene(**{k: g["in_" + k] for k in ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")})
    if "in_box" in g.files:
        scene.box = g["in_box"]
    return scene, g["in_cameras"], g


def same_floats(a, b) -> bool:
    """Bit-exact up to the sign of zero (the CUDA path drops the reference's `* 0` terms, see rn_math.cuh)."""
    a = np.ascontiguousarray(a, np.float32); b = np.ascontiguousarray(b, np.float32)
    return bool(((a == b) | (a.view(np.uint32) == b.view(np.uint32))).all())


def rel_close(a, b, tol=REL_TOL) -> bool:
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    scale = max(1.0, float(np.abs(b).max()) if b.size else 1.0)
    return bool((np.abs(a - b) <= tol * np.maximum(np.abs(b), 1e-3 * scale)).all())


def matrices_close(a, b, tol=REL_TOL) -> bool:
    """World matrices [n, 16] (column-major) within `tol` RELATIVE: every rotation-scale column against the largest
    entry of that column in the reference, every translation against max(its largest entry, the scene extent) - the
    natural scales of the two kinds of entries (an entry that is small by cancellation carries the rounding error of
    the terms it was summed from, in the reference's own release build just as here)."""
    a = np.asarray(a, np.float64).reshape(-1, 4, 4); b = np.asarray(b, np.float64).reshape(-1, 4, 4)
    if a.shape != b.shape:
        return False
    if a.size == 0:
        return True
    d = np.abs(a - b)
    col_scale = np.abs(b[:, :3, :3]).max(axis=2, keepdims=True)                   # [n, 3 columns, 1]
    ok_rot = (d[:, :3, :3] <= tol * np.maximum(col_scale, 1e-30)).all()
    extent = max(float(np.abs(b[:, 3, :3]).max()), 1.0)
    ok_tr = (d[:, 3, :3] <= tol * np.maximum(np.abs(b[:, 3, :3]).max(axis=1, keepdims=True), extent)).all()
    ok_last = (a[:, :, 3] == b[:, :, 3]).all()
    return bool(ok_rot and ok_tr and ok_last)


def standard_cameras() -> np.ndarray:
    return np.stack([
        sg.make_camera(),
        sg.make_camera(pos=(10, 5, -3), euler=(30, -20, 10), fov=90, aspect=1.3, near=0.5, far=300,
                       plane_offsets=(0.1, 0.2, -0.3, 0.05), render_group=3),
        sg.make_camera(flags=sg.CAM_USE_SIMPLE_CULLING),
    ] + list(sg.six_view_rig())[:5])


def same_list_mod_ties(got: np.ndarray, ref: np.ndarray, priority: np.ndarray, front_to_back: bool) -> bool:
    """Equality of two post-sort lists (creation ids) up to what std::sort leaves unspecified: under SortFrontToBack
    all entries of one priority >= RenderSky compare equal (RNSceneBasic.cpp:601-605), so the reference's order
    inside such a group is an accident of introsort.  RenderTransparent entries are re-sorted by distance afterwards
    and stay exact."""
    if got.shape != ref.shape:
        return False
    if not front_to_back:
        return bool(np.array_equal(got, ref))
    pr = priority[ref]
    loose = (pr >= sg.RENDER_SKY) & (pr != sg.RENDER_TRANSPARENT)
    if not np.array_equal(got[~loose], ref[~loose]):
        return False
    if not np.array_equal(priority[got[loose]], pr[loose]):
        return False
    return all(np.array_equal(np.sort(got[loose][pr[loose] == p]), np.sort(ref[loose][pr[loose] == p])) for p in np.unique(pr[loose]))


def sort_cameras() -> np.ndarray:
    """Cameras that exercise the post-cull sorts (A11): SortFrontToBack on top of the default
    SortTransparentBackToFront, one with simple culling (large list), one ordinary frustum."""
    ftb = 1 << 6
    return np.stack([
        sg.make_camera(pos=(3, 2, 1), flags=sg.CAM_USE_SIMPLE_CULLING | ftb),
        sg.make_camera(pos=(-30, 2, 10), euler=(40, 0, 0), flags=ftb, far=900),
    ])


def run_cuda(scene: sg.Scene, frustums: np.ndarray, tile_nodes: int = 0, fused: bool = True, strict: bool = True):
    """Returns dict with order, world TRS/matrix in CREATION order, lists of creation ids, stats.
    strict: RN_SCENE_STRICT_MATRICES (world matrices bit-exact with the reference's strict-FP build); False = the
    library default (fused multiply-adds on the matrix chain, within 1e-5)."""
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=max(scene.n, 1), tile_nodes=tile_nodes, strict_matrices=strict) as h:
        h.upload(rs)
        if fused:
            h.transform_cull(frustums)
        else:
            h.transform_propagate()
            h.cull(frustums)
        out = {"order": rs.order}
        rank_of_slot = h.download(capi.DL_RANK_OF_SLOT)
        for key, what in (("wpos", capi.DL_WORLD_POSITION), ("wrot", capi.DL_WORLD_ROTATION),
                          ("wscale", capi.DL_WORLD_SCALE), ("wmat", capi.DL_WORLD_MATRIX)):
            ranked = h.download(what)
            full = np.zeros_like(ranked)
            full[rs.order] = ranked
            out[key] = full
        out["counts"] = h.read_counts()
        out["lists"] = [rs.order[rank_of_slot[h.read_list(c)]] for c in range(len(np.atleast_1d(frustums)))]
        out["slot_lists_rank"] = [rank_of_slot[h.read_list(c)] for c in range(len(np.atleast_1d(frustums)))]
        st = h.stats()
        out["stats"] = {k: getattr(st, k) for k, _ in st._fields_}
    return out


def check_against_oracle(scene: sg.Scene, cameras: np.ndarray, strict: bool = True, ref=None, **kw):
    """Lists, world position / rotation / scale: bit-exact in both modes.  World matrices: bit-exact with
    strict=True, within 1e-5 relative (north star) with the library default (strict=False)."""
    fr = capi.build_frustums(cameras)
    assert fr.tobytes() == oracle.frustums(cameras).tobytes()
    got = run_cuda(scene, fr, strict=strict, **kw)
    if ref is None:
        ref = oracle.scene_reference(scene, fr)
    assert np.array_equal(got["order"], ref["order"])
    for key in ("wpos", "wrot", "wscale"):
        assert same_floats(got[key], ref[key]), key
    if strict:
        assert same_floats(got["wmat"], ref["wmat"]), "wmat (strict)"
    else:
        assert matrices_close(got["wmat"], ref["wmat"]), "wmat (fma)"
    for c, (a, b) in enumerate(zip(got["lists"], ref["lists"])):
        assert a.shape == b.shape and np.array_equal(a, b), f"visibility list of camera {c} differs"
        assert got["counts"][c] == len(b)
        r = got["slot_lists_rank"][c]
        assert (np.diff(r.astype(np.int64)) > 0).all(), "list is not in render-list order"
    return got, ref


def load_golden_occlusion():
    """(scene, cameras, mesh_of, tri_ofs, tris, case dict, npz) of tests/golden/occlusion_basic.npz."""
    from tests.golden.make_golden import OCCLUSION_CASE
    g = np.load(os.path.join(GOLDEN_DIR, "occlusion_basic.npz"))
    scene = sg.Scene(**{k: g["in_" + k] for k in ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")})
    scene.box = g["in_box"]
    return scene, g["in_cameras"], g["in_mesh_of"], g["ref_tri_ofs"], g["ref_tris"], dict(OCCLUSION_CASE), g


def run_cuda_occlusion(scene: sg.Scene, cameras: np.ndarray, tri_ofs, tris, mesh_of, frames: int, seed: int,
                       frame_size=(1280.0, 720.0), root_translate=(0.0, 0.0, 0.0), camera_translate=(0.0, 0.0, 0.0),
                       rank_mode: bool = False):
    """The frames of oracle.occlusion_frames through the C ABI: dirty roots -> rn_scene_update_nodes, rn_transform_cull,
    rn_bounds_update, rn_cull_occlusion (jitter queue as a shim would keep it), rn_cull_sort.  Returns per frame a dict
    with cams[c] = {used, offset, list (creation ids), depth}, counters (creation order)."""
    cams = np.ascontiguousarray(np.atleast_1d(cameras), dtype=sg.CAMERA_DTYPE).copy()
    rs = capi.RankedScene(scene)
    roots = np.nonzero(scene.parent < 0)[0]
    pos = scene.pos.copy()
    draws = oracle.jitter(seed, frames * len(cams) + len(cams))      # what Rayne's shared generator would yield
    drawn = 0
    visit = np.asarray(sg.camera_render_order(cams["flags"]), np.uint32)
    pixel = np.float32(1.0) / np.float32(frame_size)
    out = []
    with capi.SceneHandle(max_nodes=scene.n, strict_matrices=True) as h:
        h.upload(rs)
        if rank_mode:
            h.set_list_mode(True, 500)
        h.occlusion_set_meshes(tri_ofs, tris)
        h.occlusion_set_node_meshes(np.asarray(mesh_of, np.int32)[rs.order])
        rank_of_slot = h.download(capi.DL_RANK_OF_SLOT)
        for f in range(frames):
            if f > 0:
                pos[roots] += np.float32(root_translate)
                cams["pos"] += np.float32(camera_translate)
                h.update_nodes(rs.rank_of[roots], position=pos[roots])
            fr = capi.build_frustums(cams)
            fr["flags"] |= sg.CAM_DEFAULTS
            view, proj, _ = capi.build_camera_matrices(cams)
            oc = np.zeros(len(cams), capi.OCCLUSION_CAMERA_DTYPE)
            oc["view"] = view; oc["projection"] = proj; oc["pixel_size"] = pixel; oc["world_position"] = cams["pos"]; oc["enabled"] = 1
            h.transform_cull(fr)
            h.bounds_update()
            h.cull_occlusion(oc, draws[drawn:drawn + len(cams)], order=visit)
            h.cull_sort(np.ascontiguousarray(cams["pos"], np.float32))
            frame = {"cams": []}
            consumed = 0
            for c in range(len(cams)):
                used, taken, consumed, depth, counters = h.occlusion_download(c, scene.n)
                ids = h.read_list(c)
                ranks = (ids - 500) if rank_mode else rank_of_slot[ids]
                frame["cams"].append({"used": used, "offset": taken if used else np.zeros(3, np.float32), "list": rs.order[ranks], "depth": depth})
            full = np.zeros(scene.n, np.uint32); full[rs.order] = counters
            frame["counters"] = full
            drawn += consumed
            out.append(frame)
    return out
