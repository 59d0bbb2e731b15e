"""GPU parity of the rows next to transform + cull (§8(f) particle mesh):
inverse world matrices, transformed bounds, SceneWithVisibilityLists order, post-cull sorts, the AABB extension and
the particle render mesh - against the golden fixtures written by the unmodified reference and against the oracle."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle
from rayne_b200 import capi, scenegen as sg
from tests.helpers import (GOLDEN_DIR, load_golden_scene, same_floats, same_list_mod_ties, sort_cameras,
                           standard_cameras)

pytestmark = pytest.mark.gpu

SCENE_FIXTURES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "scene_*.npz")))


def by_creation(rs, ranked):
    full = np.zeros_like(ranked)
    full[rs.order] = ranked
    return full


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_inverse_and_bounds_match_golden(name):
    """A3 + A4 bit-exact against the reference's GetInverseWorldTransform / GetBoundingSphere / GetBoundingBox."""
    scene, cams, g = load_golden_scene(name)
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n, strict_matrices=True) as h:
        h.upload(rs)
        with pytest.raises(capi.RNError):
            h.download(capi.DL_INVERSE_WORLD_MATRIX)     # not computed yet: loud, not stale
        h.transform_inverse()
        h.transform_propagate()
        with pytest.raises(capi.RNError):
            h.download(capi.DL_BOUNDING_BOX)
        h.bounds_update()
        assert same_floats(by_creation(rs, h.download(capi.DL_INVERSE_WORLD_MATRIX)), g["ref_winv"])
        assert same_floats(by_creation(rs, h.download(capi.DL_BOUNDING_BOX)), g["ref_aabb"])
        assert same_floats(by_creation(rs, h.download(capi.DL_BOUNDING_SPHERE)), g["ref_sphere"])
        assert same_floats(by_creation(rs, h.download(capi.DL_WORLD_MATRIX)), g["ref_wmat"])


@pytest.mark.parametrize("kw", [
    dict(n_trees=3, branching=4, depth=8, seed=2, nonuniform_scale=True, random_boxes=True),
    dict(ragged_nodes=60000, depth=8, seed=5, random_bounds=True),
    dict(n_trees=2, branching=700, depth=3, seed=8),
    dict(n_trees=2, branching=1, depth=40, seed=7, nonuniform_scale=True),
])
@pytest.mark.parametrize("tile", [0, 64])
def test_inverse_and_bounds_match_oracle(kw, tile):
    scene = sg.make_scene(**kw)
    # `scale != 0` false -> zero inverse scale (RNSceneNode.cpp:552).  A single zero component would give 1/0 = inf
    # and then NaN through the reference's `inf * 0` terms, which this library drops (rn_math.cuh): not compared.
    scene.scale[::97] = 0.0
    ref = oracle.scene_reference(scene, oracle.frustums(standard_cameras()[:1]))
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n, tile_nodes=tile) as h:
        h.upload(rs)
        h.transform_inverse()
        h.transform_propagate()
        h.bounds_update()
        winv = by_creation(rs, h.download(capi.DL_INVERSE_WORLD_MATRIX))
        aabb = by_creation(rs, h.download(capi.DL_BOUNDING_BOX))
    for name, got, want in (("winv", winv, ref["winv"]), ("aabb", aabb, ref["aabb"])):
        bad = ~((got == want) | (got.view(np.uint32) == want.view(np.uint32)))
        rows = np.nonzero(bad.any(axis=1))[0]
        assert rows.size == 0, (name, rows.size, rows[:5].tolist(), got[rows[:2]].tolist(), want[rows[:2]].tolist())


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_vislist_order_and_lists_match_golden(name):
    """A10: nodes uploaded in SceneWithVisibilityLists' default-volume order give RenderVolumeList's lists."""
    scene, cams, g = load_golden_scene(name)
    rs = capi.RankedScene(scene, vislists=True)
    assert np.array_equal(rs.order, g["ref_vis_order"])
    fr = capi.build_frustums(cams)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        ros = h.download(capi.DL_RANK_OF_SLOT)
        for c0 in range(0, len(cams), capi.RN_MAX_CAMERAS):       # at most 8 views per call
            h.transform_cull(fr[c0:c0 + capi.RN_MAX_CAMERAS])
            for c in range(c0, min(c0 + capi.RN_MAX_CAMERAS, len(cams))):
                assert np.array_equal(rs.order[ros[h.read_list(c - c0)]], g[f"ref_vis_list{c}"]), f"camera {c}"


def cull_sorted(scene, cams, rank_mode=False, fused=True):
    if len(cams) > capi.RN_MAX_CAMERAS:                           # at most 8 views per call
        a1, b1 = cull_sorted(scene, cams[:capi.RN_MAX_CAMERAS], rank_mode, fused)
        a2, b2 = cull_sorted(scene, cams[capi.RN_MAX_CAMERAS:], rank_mode, fused)
        return a1 + a2, b1 + b2
    fr = capi.build_frustums(cams)
    fr["flags"] |= sg.CAM_DEFAULTS          # Camera::GetFlags() of a Rayne camera carries Flags::Defaults
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        if rank_mode:
            h.set_list_mode(True, 1000)
        if fused:
            h.transform_cull(fr)
        else:
            h.transform_propagate()
            h.cull(fr)
        before = [h.read_list(c) for c in range(len(cams))]
        h.cull_sort(cams["pos"])
        after = [h.read_list(c) for c in range(len(cams))]
        ros = h.download(capi.DL_RANK_OF_SLOT)
    conv = (lambda l: rs.order[l - 1000]) if rank_mode else (lambda l: rs.order[ros[l]])
    return [conv(l) for l in before], [conv(l) for l in after]


def test_sorted_lists_match_golden():
    """A11 against the reference: its recorded lists are post-sort (SceneBasic::Render :596-616)."""
    scene, cams, g = load_golden_scene("scene_boxes_transparent")
    for rank_mode in (False, True):
        before, after = cull_sorted(scene, cams, rank_mode=rank_mode)
        for c in range(len(cams)):
            ftb = bool(int(cams[c]["flags"]) & (1 << 6))
            assert same_list_mod_ties(after[c], g[f"ref_list{c}"], scene.priority, ftb), f"camera {c}"
            assert np.array_equal(np.sort(before[c]), np.sort(after[c]))


@pytest.mark.parametrize("kw", [
    dict(n_trees=40, branching=4, depth=6, seed=31, random_bounds=True, priorities=True, transparent_frac=0.2),
    dict(n_trees=3000, branching=1, depth=1, seed=32, transparent_frac=1.0),        # everything transparent
    dict(n_trees=300, branching=3, depth=3, seed=33, transparent_frac=0.001),       # 0-2 transparent nodes
    dict(n_trees=64, branching=4, depth=5, seed=34),                                # no priorities at all
])
def test_sorted_lists_match_oracle(kw):
    scene = sg.make_scene(**kw)
    cams = np.concatenate([standard_cameras()[:3], sort_cameras()])
    ref = oracle.scene_reference(scene, oracle.frustums(cams), cameras=cams)
    _, after = cull_sorted(scene, cams, fused=False)
    for c in range(len(cams)):
        # the oracle sorts stably, and so does rn_cull_sort: exact equality, ties included
        assert np.array_equal(after[c], ref["lists"][c]), f"camera {c}"


def test_sort_full_size_properties():
    """1 M nodes, front-to-back over ~all of them: permutation of the cull result, keys monotone per priority."""
    scene = sg.make_scene(n_trees=48, branching=4, depth=8, seed=12345, priorities=True, transparent_frac=0.05)
    cams = sort_cameras()[:1]
    before, after = cull_sorted(scene, cams)
    a, b = before[0], after[0]
    assert a.shape[0] > 900000 and np.array_equal(np.sort(a), np.sort(b))
    pr = scene.priority[b]
    assert (np.diff(pr.astype(np.int64)) >= 0).all()
    parent = np.where(scene.parent >= 0, scene.parent, 0xFFFFFFFF).astype(np.uint32)
    wpos = oracle.world(parent, scene.pos, scene.scale, scene.rot)[0]
    d = wpos[b] - cams[0]["pos"]
    key = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    for p in np.unique(pr):
        k = key[pr == p]
        if p < sg.RENDER_SKY:
            assert (np.diff(k) >= 0).all()
        elif p == sg.RENDER_TRANSPARENT:
            assert (np.diff(k[1:]) <= 0).all()


@pytest.mark.parametrize("fused", [True, False])
def test_aabb_extension_matches_oracle(fused):
    """C5a extension (no live reference behaviour): the AABB-vs-planes filter on top of the sphere predicate."""
    scene = sg.make_scene(n_trees=30, branching=4, depth=6, seed=41, random_bounds=True, random_boxes=True,
                          nonuniform_scale=True, random_groups=True)
    scene.sphere[:, 3] *= 4.0       # fat spheres, so that the boxes decide
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    ref = oracle.scene_reference(scene, fr, aabb_test=True)
    plain = oracle.scene_reference(scene, fr)
    assert sum(len(a) for a in ref["lists"]) < sum(len(a) for a in plain["lists"])
    fr2 = fr.copy()
    fr2["flags"][::2] |= capi.CAM_TEST_AABB          # every other camera carries the flag
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        if fused:
            h.transform_cull(fr2)
        else:
            h.transform_propagate()
            h.cull(fr2)
        ros = h.download(capi.DL_RANK_OF_SLOT)
        counts = h.read_counts()
        for c in range(len(cams)):
            want = (ref if c % 2 == 0 else plain)["lists"][c]
            got = rs.order[ros[h.read_list(c)]]
            assert np.array_equal(got, want), f"camera {c}"
            assert counts[c] == len(want)


def test_particle_mesh_matches_golden_and_oracle():
    """P6: plain / rendered-inversed meshes bit-exact against the reference; rolled within 2 ulp-scale tolerance
    (cosf / sinf of the device are not libm's)."""
    g = np.load(os.path.join(GOLDEN_DIR, "particles_varied.npz"))
    em = g["in_emitters"]
    with capi.SceneHandle(max_nodes=1, max_emitters=len(em), max_particles_per_emitter=int(em["max_particles"].max())) as h:
        h.particles_configure(em)
        for _ in range(int(g["in_steps"])):
            h.particles_step(float(np.float32(g["in_delta"])))
        counts, _ = h.particles_counts()
        for tag, flags in (("", 0), ("_inv", capi.EMITTER_RENDERED_INVERSED), ("_roll", capi.EMITTER_CAN_ROLL)):
            h.particles_build_mesh(flags=np.full(len(em), flags, np.uint32))
            for k in range(len(em)):
                mx = int(em["max_particles"][k])
                m = h.particles_download_mesh(k, mx)
                live = 4 * min(int(counts[k]), mx)
                for key in ("mesh_pos", "mesh_color", "mesh_uv0"):
                    assert same_floats(m[key][:live], g[f"ref_{key}{tag}{k}"][:live]), (tag, k, key)
                if tag == "_roll":
                    assert np.allclose(m["mesh_uv1"][:live], g[f"ref_mesh_uv1{tag}{k}"][:live], rtol=0, atol=2e-6)
                else:
                    assert same_floats(m["mesh_uv1"][:live], g[f"ref_mesh_uv1{tag}{k}"][:live]), (tag, k)
                assert np.array_equal(m["mesh_idx"], g[f"ref_mesh_idx{tag}{k}"]), (tag, k)
    # world scale and a bigger set against the oracle
    em = sg.make_emitters(9, max_particles=900, spawn_rate=0.0015, vary=True)
    o = oracle.Emitters(em)
    o.step(np.float32(1 / 60), 80)
    scale = np.linspace(0.5, 3.0, len(em)).astype(np.float32)
    with capi.SceneHandle(max_nodes=1, max_emitters=9, max_particles_per_emitter=900) as h:
        h.particles_configure(em)
        for _ in range(80):
            h.particles_step(float(np.float32(1 / 60)))
        h.particles_build_mesh(world_scale_x=scale)
        for k in range(len(em)):
            ref = o.mesh(k, world_scale_x=float(scale[k]))
            m = h.particles_download_mesh(k, int(em["max_particles"][k]))
            live = 4 * o.read(k)[0].shape[0]
            for key in ("mesh_pos", "mesh_color", "mesh_uv0", "mesh_uv1"):
                assert same_floats(m[key][:live], ref[key][:live]), (k, key)
            assert np.array_equal(m["mesh_idx"], ref["mesh_idx"])


def _pack(h, camera, view, projview, count, capacity):
    import torch
    buf = torch.full((capacity + 2, 64), float("nan"), dtype=torch.float32, device="cuda")
    h.pack_instances(camera, view, projview, buf.data_ptr(), capacity)
    torch.cuda.synchronize()
    got = buf.cpu().numpy()
    assert np.isnan(got[min(count, capacity):]).all(), "wrote past min(count, capacity)"
    return got[:min(count, capacity)]


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_instance_records_match_golden(name):
    """§8(f) drawable uniform packing: per list entry model, inverse model, view * model, projView * model - bit-exact
    against the products the reference's Matrix class gives for camera 0 of the fixture."""
    scene, cams, g = load_golden_scene(name)
    rs = capi.RankedScene(scene)
    fr = capi.build_frustums(cams[:capi.RN_MAX_CAMERAS])
    v, _, pv = capi.build_camera_matrices(cams)
    assert same_floats(v, g["ref_view"]) and same_floats(pv, g["ref_projview"])
    with capi.SceneHandle(max_nodes=scene.n, strict_matrices=True) as h:
        h.upload(rs)
        h.transform_cull(fr)
        with pytest.raises(capi.RNError):
            _pack(h, 0, v[0], pv[0], 0, 4)           # inverse matrices not computed yet: loud
        h.transform_inverse()
        lst = h.read_list(0)
        ids = rs.order[h.download(capi.DL_RANK_OF_SLOT)[lst]]
        got = _pack(h, 0, v[0], pv[0], len(lst), scene.n)
        assert len(got) == len(lst) > 0
        assert same_floats(got[:, 0:16], g["ref_wmat"][ids]) and same_floats(got[:, 16:32], g["ref_winv"][ids])
        assert same_floats(got[:, 32:48], g["ref_mv"][ids]) and same_floats(got[:, 48:64], g["ref_mvp"][ids])
        # a buffer smaller than the list: the first `capacity` entries, nothing behind them
        half = _pack(h, 0, v[0], pv[0], len(lst), len(lst) // 2)
        assert same_floats(half, got[:len(lst) // 2])
        with pytest.raises(capi.RNError):
            h.pack_instances(len(fr), v[0], pv[0], 256, 1)


@pytest.mark.parametrize("rank_mode", [False, True])
def test_instance_records_match_oracle_after_sort(rank_mode):
    """Records follow the list order left by rn_cull_sort, in slot and in rank list mode, for every camera."""
    scene = sg.make_scene(n_trees=4, branching=5, depth=7, seed=12, nonuniform_scale=True, random_bounds=True,
                          priorities=True, transparent_frac=0.2)
    cams = np.concatenate([standard_cameras()[:2], sort_cameras()])[:capi.RN_MAX_CAMERAS]
    fr = capi.build_frustums(cams)
    fr["flags"] |= sg.CAM_DEFAULTS
    v, _, pv = capi.build_camera_matrices(cams)
    ov, _, opv = oracle.camera_matrices(cams)
    assert same_floats(v, ov) and same_floats(pv, opv)
    ref = oracle.scene_reference(scene, oracle.frustums(cams), cameras=cams)
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n, strict_matrices=True) as h:
        h.upload(rs)
        if rank_mode:
            h.set_list_mode(True, 77)
        h.transform_cull(fr)
        h.transform_inverse()
        h.cull_sort(cams["pos"])
        ros = h.download(capi.DL_RANK_OF_SLOT)
        for c in range(len(cams)):
            lst = h.read_list(c)
            ids = rs.order[lst - 77] if rank_mode else rs.order[ros[lst]]
            want = oracle.instances(ids, ref["wmat"], ref["winv"], ov[c], opv[c])
            got = _pack(h, c, v[c], pv[c], len(lst), len(lst) + 5)
            assert same_floats(got, want), f"camera {c}"


def test_particle_billboards_match_oracle():
    """§8(f) rank 2, second half: the vertex stage of Shaders/Particles.hlsl (camera-facing quad expansion) over the
    mesh of rn_particles_build_mesh - bit-exact against the C restatement (the shader itself cannot run here), and
    the four corners of a quad really are centre +- the camera's right / up vectors in clip space."""
    em = sg.make_emitters(6, max_particles=700, spawn_rate=0.003, vary=True)
    cams = standard_cameras()
    view, _, projview = capi.build_camera_matrices(cams[:1])
    rng = np.random.default_rng(4)
    n = len(em)
    # every emitter its own model matrix (translation + uniform scale): inverse model-view and MVP per emitter
    inv_mv = np.zeros((n, 16), np.float32); mvp = np.zeros((n, 16), np.float32)
    for k in range(n):
        model = np.eye(4); model[:3, 3] = rng.uniform(-5, 5, 3); model[:3, :3] *= rng.uniform(0.5, 2.0)
        mv = view[0].reshape(4, 4).T.astype(np.float64) @ model
        inv_mv[k] = np.linalg.inv(mv).T.reshape(16).astype(np.float32)
        mvp[k] = (projview[0].reshape(4, 4).T.astype(np.float64) @ model).T.reshape(16).astype(np.float32)
    ambient = np.float32([0.9, 0.8, 0.7, 1.0])
    with capi.SceneHandle(max_nodes=1, max_emitters=n, max_particles_per_emitter=700) as h:
        h.particles_configure(em)
        with pytest.raises(capi.RNError):
            h.particles_expand_billboards(inv_mv, mvp, ambient)       # no mesh yet
        for _ in range(50):
            h.particles_step(float(np.float32(1 / 60)))
        counts, _ = h.particles_counts()
        h.particles_build_mesh(flags=np.full(n, capi.EMITTER_CAN_ROLL, np.uint32))
        h.particles_expand_billboards(inv_mv, mvp, ambient)
        for k in range(n):
            mx = int(em["max_particles"][k])
            live = 4 * min(int(counts[k]), mx)
            assert live > 0
            m = h.particles_download_mesh(k, mx)
            got = h.particles_download_clip(k, mx)
            clip, col = oracle.billboards(m["mesh_pos"][:live], m["mesh_color"][:live], m["mesh_uv1"][:live], inv_mv[k], mvp[k], ambient)
            assert same_floats(got["clip_pos"][:live], clip), k
            assert same_floats(got["clip_color"][:live], col), k
            assert not got["clip_pos"][live:].any() and not got["clip_color"][live:].any()
            # float64 check of the formula itself
            p4 = np.concatenate([m["mesh_pos"][:live].astype(np.float64), np.ones((live, 1))], axis=1)
            t4 = np.concatenate([m["mesh_uv1"][:live].astype(np.float64), np.zeros((live, 2))], axis=1)
            want = (p4 + t4 @ inv_mv[k].reshape(4, 4).astype(np.float64)) @ mvp[k].reshape(4, 4).astype(np.float64)
            assert np.allclose(got["clip_pos"][:live], want, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("rank_mode", [False, True])
@pytest.mark.parametrize("fused", [True, False])
def test_volumes_match_golden(rank_mode, fused):
    """A10 with user volumes: nodes uploaded as [default volume][volume 1][volume 2].., tagged with their volume; every
    camera gets the first volume containing it; lists == what SceneWithVisibilityLists::Render submitted."""
    g = np.load(os.path.join(GOLDEN_DIR, "volumes_basic.npz"))
    scene = sg.Scene(**{k: g["in_" + k] for k in ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")})
    cams, vol, bounds = g["in_cameras"], g["in_volume"], g["in_bounds"]
    rs = capi.RankedScene(scene, vislists=True, volume_of_node=vol)
    fr = capi.build_frustums(cams)
    chosen = np.array([capi.choose_volume(bounds, c["pos"]) for c in cams], np.uint8)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        with pytest.raises(capi.RNError):
            h.set_camera_volumes(chosen[:1])                  # no node volumes yet
        h.set_node_volumes(rs.volume)
        if rank_mode:
            h.set_list_mode(True, 500)
        ros = h.download(capi.DL_RANK_OF_SLOT)
        conv = (lambda l: rs.order[l - 500]) if rank_mode else (lambda l: rs.order[ros[l]])
        plain = None
        for c0 in range(0, len(cams), capi.RN_MAX_CAMERAS):
            sub = slice(c0, min(c0 + capi.RN_MAX_CAMERAS, len(cams)))
            h.set_camera_volumes(chosen[sub])
            if fused:
                h.transform_cull(fr[sub])
            else:
                h.transform_propagate()
                h.cull(fr[sub])
            counts = h.read_counts()
            for c in range(sub.start, sub.stop):
                got = conv(h.read_list(c - c0))
                assert np.array_equal(got, g[f"ref_list{c}"]), f"camera {c}"
                assert counts[c - c0] == len(got)
            # masks follow the filtered visibility too
            h.build_masks()
            m = h.read_masks()
            for c in range(sub.start, sub.stop):
                bits = np.nonzero(np.unpackbits(m[c - c0].view(np.uint8), bitorder="little")[:scene.n])[0]
                assert np.array_equal(rs.order[bits], g[f"ref_list{c}"])
            if c0 == 0:
                h.set_camera_volumes(None)                    # filter off: every volume's nodes are back
                h.transform_cull(fr[sub])
                plain = [conv(h.read_list(c)) for c in range(sub.stop - sub.start)]
        ref = oracle.scene_reference(scene, oracle.frustums(cams), vislists=True)
        for c, l in enumerate(plain):
            vis = np.zeros(scene.n, bool); vis[ref["lists"][c]] = True
            assert np.array_equal(l, rs.order[vis[rs.order]]), f"camera {c} (filter off)"


@pytest.mark.parametrize("rank_mode", [False, True])
def test_lod_selection_extension_matches_oracle(rank_mode):
    """rn_cull_select_lod (extension: Entity::Render's #else branch + Model::GetLODStageForDistance, compiled out in the
    reference) against the C restatement, per entry of an unsorted and of a front-to-back sorted list; nodes without a
    model, single-stage models, a camera sitting exactly on a node (distance <= epsilon) and distances beyond the last
    stage included."""
    import torch
    from oracle import oracle
    scene = sg.make_scene(n_trees=20, branching=4, depth=5, seed=31, random_bounds=True)
    cams = np.stack([sg.make_camera(flags=sg.CAM_USE_SIMPLE_CULLING), sg.make_camera(pos=(20, 3, 4), far=800, flags=capi.CAM_SORT_FRONT_TO_BACK)])
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    rng = np.random.default_rng(9)
    stage_counts = np.array([1, 2, 3, 5, 4], np.uint32)
    so = np.concatenate([[0], np.cumsum(stage_counts)]).astype(np.uint32)
    sd = np.concatenate([np.sort(rng.random(int(c)).astype(np.float32) * np.float32(0.8)) for c in stage_counts]).astype(np.float32)
    model = rng.integers(0, 6, size=scene.n).astype(np.uint16)
    model[model == 5] = 0xffff                                       # not an Entity
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        if rank_mode:
            h.set_list_mode(True, 1000)
        h.lod_set_models(so, sd)
        h.lod_set_node_models(model[rs.order])
        ref = oracle.scene_reference(scene, fr)
        # the LOD camera of view 0 sits exactly on a node of its list: distance 0 <= epsilon -> stage 0
        probe = int(ref["lists"][0][len(ref["lists"][0]) // 2])
        lod_pos = [ref["wpos"][probe], cams["pos"][1]]
        h.transform_cull(fr)
        h.cull_sort(np.ascontiguousarray(cams["pos"], np.float32))
        rank_of_slot = h.download(capi.DL_RANK_OF_SLOT)
        for c in range(2):
            ids = h.read_list(c)
            ranks = (ids - 1000) if rank_mode else rank_of_slot[ids]
            creation = rs.order[ranks]
            out = torch.zeros(len(ids) + 8, dtype=torch.uint8, device="cuda")
            h.select_lod(c, lod_pos[c], float(cams["clip_far"][c]), out.data_ptr(), len(ids))
            torch.cuda.synchronize()
            got = out.cpu().numpy()
            want = oracle.select_lod(creation, ref["wpos"], model, so, sd, lod_pos[c], float(cams["clip_far"][c]))
            assert np.array_equal(got[:len(ids)], want), f"camera {c}"
            assert (got[len(ids):] == 0).all()                        # nothing written beyond the list
            assert len(np.unique(want)) >= 3
        assert oracle.select_lod(np.array([probe], np.uint32), ref["wpos"], np.full(scene.n, 3, np.uint16), so, sd, lod_pos[0], 500.0)[0] == 0
        with pytest.raises(capi.RNError):
            h.select_lod(5, lod_pos[0], 500.0, 256, 1)
