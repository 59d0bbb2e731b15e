"""CPU: the oracle (oracle/rn_oracle.c) against the golden fixtures produced by the unmodified reference
(tests/golden/make_golden.py), and - where oracle/_ref exists - against the reference harness run live.
Everything is bit-exact: the oracle and the strict-FP reference build evaluate the same fp32 operation sequence."""
import glob
import os

import numpy as np
import pytest

from oracle import oracle, refio
from rayne_b200 import capi, scenegen as sg
from tests.helpers import same_floats, same_list_mod_ties, standard_cameras, load_golden_scene, GOLDEN_DIR

SCENE_FIXTURES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "scene_*.npz")))
PARTICLE_FIXTURES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "particles_*.npz")))


def test_fixtures_present():
    assert len(SCENE_FIXTURES) >= 4 and len(PARTICLE_FIXTURES) >= 3


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_oracle_scene_matches_golden(name):
    scene, cams, g = load_golden_scene(name)
    fr = oracle.frustums(cams)
    assert np.array_equal(fr["center"], g["ref_frustum"][:, 0:3])
    assert np.array_equal(fr["radius"], g["ref_frustum"][:, 3])
    assert np.array_equal(fr["planes"].reshape(-1, 24), g["ref_frustum"][:, 4:28])
    # the reference camera also carries its default flags (RNCamera.h:50); only these two reach the cull test
    path_flags = sg.CAM_ORTHOGONAL | sg.CAM_USE_SIMPLE_CULLING
    assert np.array_equal(fr["flags"] & path_flags, g["ref_cam_flags"] & path_flags)
    assert np.array_equal(fr["render_group"], g["ref_cam_group"])
    # the reference's lists are what node->Render saw: AFTER the post-cull sorts of SceneBasic::Render (A11)
    o = oracle.scene_reference(scene, fr, cameras=cams)
    assert np.array_equal(o["order"], g["ref_order"])
    for k in ("wpos", "wrot", "wscale", "wmat", "winv", "aabb"):   # A1, A2, A3, A4
        assert same_floats(o[k], g["ref_" + k]), k
    for c in range(len(cams)):
        ftb = bool(int(cams[c]["flags"]) & (1 << 6))
        assert same_list_mod_ties(o["lists"][c], g[f"ref_list{c}"], scene.priority, ftb), f"camera {c}"


def test_oracle_render_order_sort_equals_the_literal_walk():
    """oracle.render_order has two forms: the literal walk of AddRenderNode's three insert rules (pinned by the golden
    `ref_order` arrays above) and, for the 1 M-node scenes of the GPU suite, the same order derived from those rules with
    one sort (the walk is quadratic when priorities are scattered: 11 minutes for 1 M nodes).  They must agree on every
    input: random forests with no, few, many distinct and negative priorities, the golden scenes, a 60 k-node scene."""
    rng = np.random.default_rng(1)
    for trial in range(240):
        n = int(rng.integers(1, 300))
        parent = np.full(n, -1, np.int32)
        for i in range(1, n):
            if rng.random() < 0.85:
                parent[i] = rng.integers(0, i)
        prio = (None, rng.integers(-2, 3, n).astype(np.int32), rng.integers(-10 ** 6, 10 ** 6, n).astype(np.int32),
                rng.choice(np.int32([0, 0, 0, 0, 1000000, 1]), n).astype(np.int32))[trial % 4]
        assert np.array_equal(oracle.render_order(parent, prio, "walk"), oracle.render_order(parent, prio, "fast")), trial
    for name in SCENE_FIXTURES:
        scene, _, g = load_golden_scene(name)
        assert np.array_equal(oracle.render_order(scene.parent, scene.priority, "fast"), g["ref_order"]), name
    big = sg.make_scene(n_trees=3, branching=4, depth=8, seed=12345, priorities=True, transparent_frac=0.05)
    assert np.array_equal(oracle.render_order(big.parent, big.priority, "walk"), oracle.render_order(big.parent, big.priority, "fast"))
    assert oracle.RENDER_ORDER_WALK_LIMIT >= 65535      # the golden-size and test-size scenes below it keep the literal walk


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_oracle_instance_matrices_match_golden(name):
    """Camera view / projection * view and the per-drawable products view * world, projView * world, bit-exact
    against the reference's Camera and Matrix classes; the host side of the library (rn_camera_build_matrices) too."""
    scene, cams, g = load_golden_scene(name)
    v, _, pv = oracle.camera_matrices(cams)
    assert same_floats(v, g["ref_view"]) and same_floats(pv, g["ref_projview"])
    lv, _, lpv = capi.build_camera_matrices(cams)
    assert same_floats(lv, g["ref_view"]) and same_floats(lpv, g["ref_projview"])
    ids = np.arange(scene.n, dtype=np.uint32)
    inst = oracle.instances(ids, g["ref_wmat"], g["ref_winv"], v[0], pv[0])
    assert same_floats(inst[:, 0:16], g["ref_wmat"]) and same_floats(inst[:, 16:32], g["ref_winv"])
    assert same_floats(inst[:, 32:48], g["ref_mv"]) and same_floats(inst[:, 48:64], g["ref_mvp"])


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_oracle_vislist_order_and_lists_match_golden(name):
    """A10: SceneWithVisibilityLists default volume - post-order, no priority sort, no post-cull sorts."""
    scene, cams, g = load_golden_scene(name)
    o = oracle.scene_reference(scene, oracle.frustums(cams), vislists=True)
    assert np.array_equal(o["order"], g["ref_vis_order"])
    for c in range(len(cams)):
        assert np.array_equal(o["lists"][c], g[f"ref_vis_list{c}"]), f"camera {c}"


def test_sorted_fixture_really_sorts():
    """The A11 fixture must exercise both sorts: lists differ from render-list order, and the transparent range of
    the plain cameras is back to front."""
    scene, cams, g = load_golden_scene("scene_boxes_transparent")
    o = oracle.scene_reference(scene, oracle.frustums(cams), cameras=cams)
    changed = [not np.array_equal(a, b) for a, b in zip(o["lists"], o["lists_unsorted"])]
    assert changed[2] and changed[8] and changed[9]
    lst = o["lists"][2]
    tr = np.nonzero(scene.priority[lst] == sg.RENDER_TRANSPARENT)[0]
    d = ((o["wpos"][lst[tr]] - cams[2]["pos"]) ** 2).sum(axis=1)
    assert (np.diff(d[1:]) <= 1e-3).all()


def test_oracle_particle_mesh_matches_golden():
    """P6: ParticleEmitter::UpdateMesh - plain and rendered-inversed bit-exact (no transcendental), rolled bit-exact
    too on this libm (same cosf/sinf on both sides)."""
    g = np.load(os.path.join(GOLDEN_DIR, "particles_varied.npz"))
    em = g["in_emitters"]
    o = oracle.Emitters(em)
    o.step(np.float32(g["in_delta"]), int(g["in_steps"]))
    for tag, kw in (("", {}), ("_inv", dict(inversed=True)), ("_roll", dict(roll=True))):
        for k in range(len(em)):
            m = o.mesh(k, **kw)
            live = 4 * min(int(g[f"ref_rec{k}"].shape[0]), int(em["max_particles"][k]))
            for key in ("mesh_pos", "mesh_color", "mesh_uv0", "mesh_uv1"):
                assert same_floats(m[key][:live], g[f"ref_{key}{tag}{k}"][:live]), (tag, k, key)
            assert np.array_equal(m["mesh_idx"], g[f"ref_mesh_idx{tag}{k}"]), (tag, k)


@pytest.mark.parametrize("name", SCENE_FIXTURES)
def test_oracle_bounding_sphere_matches_golden(name):
    """SceneNode::GetBoundingSphere (RNSceneNode.cpp:570-588): centre = position + rotated offset, scaled radius."""
    scene, cams, g = load_golden_scene(name)
    fr = oracle.frustums(cams)[:1]
    _, sph = oracle.cull(g["ref_wpos"], g["ref_wrot"], g["ref_wscale"], scene.sphere, scene.flags, scene.group, fr[0],
                         want_sphere=True)
    assert same_floats(sph, g["ref_sphere"])


def run_emitter_fixture(o, g) -> None:
    """Steps an oracle.Emitters through a particle fixture; `in_soft_after` (particles_nonlocal_soft): the soft maxima
    are only applied after that many steps (SetMaxParticlesSoft on a running emitter)."""
    em = sg.upgrade_emitters(g["in_emitters"])
    steps, delta = int(g["in_steps"]), np.float32(g["in_delta"])
    soft_after = int(g["in_soft_after"]) if "in_soft_after" in g.files else 0
    if soft_after:
        # the emitters start WITHOUT their soft maximum: the oracle is created from records whose flag is cleared
        o.step(delta, soft_after)
        for k in np.nonzero(em["flags"] & sg.EMITTER_SOFT_MAX)[0]:
            o.set_soft_max(int(k), int(em["max_particles_soft"][k]))
        o.step(delta, steps - soft_after)
    else:
        o.step(delta, steps)


def emitters_without_late_soft_max(g) -> np.ndarray:
    """The records an implementation is configured with at step 0 of a fixture: a soft maximum that the fixture applies
    later is not set yet."""
    em = sg.upgrade_emitters(g["in_emitters"]).copy()
    if "in_soft_after" in g.files and int(g["in_soft_after"]):
        em["flags"] &= ~np.uint32(sg.EMITTER_SOFT_MAX)
    return em


@pytest.mark.parametrize("name", PARTICLE_FIXTURES)
def test_oracle_particles_match_golden(name):
    g = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    em = g["in_emitters"]
    o = oracle.Emitters(emitters_without_late_soft_max(g))
    run_emitter_fixture(o, g)
    for k in range(len(em)):
        rec, t = o.read(k)
        ref = g[f"ref_rec{k}"]
        assert rec.shape == ref.shape, (k, rec.shape, ref.shape)
        assert same_floats(rec[:, :23], ref[:, :23]), f"emitter {k}"
        assert same_floats(t, g[f"ref_time{k}"])


def test_mt_stream_known_answer():
    """MT19937 with Rayne's Seed (RNRandom.cpp:313-323): _offset = 0, so the first outputs are the tempered
    seed-initialised words (NOT the standard first outputs of mt19937).  Known answer computed by hand here."""
    seed = 5489
    mt = [seed]
    for k in range(1, 624):
        mt.append((1812433253 * (mt[-1] ^ (mt[-1] >> 30)) + k) & 0xFFFFFFFF)

    def temper(y):
        y ^= y >> 11
        y ^= (y << 7) & 0x9D2C5680
        y ^= (y << 15) & 0xEFC60000
        y ^= y >> 18
        return y & 0x7FFFFFFF

    em = sg.make_emitters(1, max_particles=10, spawn_rate=1.0 / 60.0)
    em["seed"] = seed
    em["pos_min"] = (0, 0, 0); em["pos_max"] = (1, 1, 1)
    o = oracle.Emitters(em)
    o.step(np.float32(1.0 / 60.0 + 1e-4), 1)
    rec, _ = o.read(0)
    assert rec.shape[0] == 1
    want = [np.float32(float(temper(mt[i])) * (1.0 / 2147483648.0) * 1.0 + 0.0) for i in range(3)]
    # the particle was advanced by Update(0): position unchanged by a zero step
    assert same_floats(rec[0, 0:3], np.array(want, np.float32))


@pytest.mark.ref
@pytest.mark.skipif(not refio.have_harness("strict"), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("kw", [
    dict(n_trees=10, branching=9, depth=4, seed=12345),
    dict(n_trees=4, branching=4, depth=7, seed=9, nonuniform_scale=True, random_bounds=True, random_groups=True,
         priorities=True),
    dict(ragged_nodes=20000, depth=8, seed=13, random_bounds=True),
])
def test_oracle_matches_live_reference(kw):
    scene = sg.make_scene(**kw)
    cams = standard_cameras()
    ref = refio.run_scene(scene, cams)
    fr = oracle.frustums(cams)
    assert np.array_equal(fr["planes"].reshape(-1, 24), ref["frustum"][:, 4:28])
    o = oracle.scene_reference(scene, fr)
    assert np.array_equal(o["order"], ref["order"])
    for k in ("wpos", "wrot", "wscale", "wmat"):
        assert same_floats(o[k], ref[k]), k
    for a, b in zip(o["lists"], ref["lists"]):
        assert np.array_equal(a, b)


@pytest.mark.ref
@pytest.mark.skipif(not refio.have_harness("strict"), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_particles_match_live_reference():
    em = sg.make_emitters(6, max_particles=600, spawn_rate=0.0017, vary=True)
    ref = refio.run_particles(em, 150)
    o = oracle.Emitters(em)
    o.step(np.float32(1.0 / 60.0), 150)
    for k in range(len(em)):
        rec, t = o.read(k)
        assert rec.shape == ref[k]["rec"].shape
        assert same_floats(rec[:, :23], ref[k]["rec"][:, :23])
        assert same_floats(t, ref[k]["time"])


@pytest.mark.ref
@pytest.mark.skipif(not refio.have_harness("strict"), reason="oracle/_ref not built (needs /root/reference)")
def test_oracle_emitter_modes_match_live_reference():
    """!GetIsLocal() and a soft maximum set at creation (RNParticleEmitter.cpp:128-132,204-205,428-437) against the reference."""
    em = sg.make_emitters(5, max_particles=400, spawn_rate=0.0013, vary=True, seed=3)
    em["flags"][:3] = sg.EMITTER_NONLOCAL
    em["world_position"][:3] = [(1, 2, 3), (-50, 0.5, 9), (0, 0, 0)]
    em["world_scale"][:3] = [(1, 1, 1), (0.25, 4, 2), (3, 3, -3)]
    em["flags"][3:] |= sg.EMITTER_SOFT_MAX
    em["max_particles_soft"][3:] = (37, 0)          # 0: no limit at all (the reference skips the min)
    em["max_particles"][4] = 90
    em["spawn_rate"][4] = 0.001
    em["life_span"][4] = (2.0, 3.0)
    ref = refio.run_particles(em, 130)
    o = oracle.Emitters(em)
    o.step(np.float32(1.0 / 60.0), 130)
    for k in range(len(em)):
        rec, t = o.read(k)
        assert rec.shape == ref[k]["rec"].shape, k
        assert same_floats(rec[:, :23], ref[k]["rec"][:, :23]), k
        assert same_floats(t, ref[k]["time"])
    assert ref[3]["count"] == 37 and ref[4]["count"] > 90


def test_oracle_billboard_vertex_stage():
    """Shaders/Particles.hlsl particles_vertex restated (unpinned: the shader cannot run here) against the formula in
    float64: clip = MVP * (float4(pos, 1) + invMV * float4(uv1, 0, 0)), colour * ambient."""
    rng = np.random.default_rng(1)
    n = 64
    pos = rng.normal(size=(n, 3)).astype(np.float32); col = rng.random((n, 4)).astype(np.float32)
    uv = rng.normal(size=(n, 2)).astype(np.float32)
    im = rng.normal(size=16).astype(np.float32); mvp = rng.normal(size=16).astype(np.float32)
    amb = np.float32([0.5, 0.6, 0.7, 1.0])
    clip, c = oracle.billboards(pos, col, uv, im, mvp, amb)
    p4 = np.concatenate([pos.astype(np.float64), np.ones((n, 1))], axis=1)
    t4 = np.concatenate([uv.astype(np.float64), np.zeros((n, 2))], axis=1)
    want = (p4 + t4 @ im.reshape(4, 4).astype(np.float64)) @ mvp.reshape(4, 4).astype(np.float64)
    assert np.allclose(clip, want, rtol=1e-5, atol=1e-5)
    assert np.array_equal(c, col * amb)
    # a quad's corners: centre + invMV * (+-hx, +-hy): opposite corners are symmetric about the centre before MVP
    centre = np.float32([[1, 2, 3]] * 4); corners = np.float32([[-1, -2], [1, -2], [-1, 2], [1, 2]])
    ident = np.eye(4, dtype=np.float32).reshape(16)
    q, _ = oracle.billboards(centre, np.ones((4, 4), np.float32), corners, ident, ident, np.ones(4, np.float32))
    assert np.array_equal(q[:, :3], centre + np.concatenate([corners, np.zeros((4, 1), np.float32)], axis=1))


def _volumes_fixture():
    g = np.load(os.path.join(GOLDEN_DIR, "volumes_basic.npz"))
    scene = sg.Scene(**{k: g["in_" + k] for k in ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")})
    return scene, g["in_cameras"], g["in_volume"], g["in_bounds"], g


def test_volumes_match_golden():
    """SceneWithVisibilityLists with AddVolume: default-volume order after the move, the first-containing-volume choice
    (closed box test, library and restatement), and per camera default volume ++ chosen volume - against the lists
    the reference's Render submitted."""
    scene, cams, vol, bounds, g = _volumes_fixture()
    rs = capi.RankedScene(scene, vislists=True, volume_of_node=vol)
    nd = int((vol == 0).sum())
    assert np.array_equal(rs.order[:nd], g["ref_default_order"])
    chosen = [oracle.choose_volume(bounds, c["pos"]) for c in cams]
    assert chosen == [capi.choose_volume(bounds, c["pos"]) for c in cams]
    assert set(chosen) == {0, 1, 2, 3}                  # volume 3 contains everything but is only the first match far out
    ref = oracle.scene_reference(scene, oracle.frustums(cams), vislists=True)
    for c in range(len(cams)):
        want = oracle.volume_lists(rs.order, vol, chosen[c], ref["lists"][c])
        assert np.array_equal(want, g[f"ref_list{c}"]), f"camera {c}"
    # closed bounds, first match wins
    b = np.float32([[0, 0, 0, 1, 1, 1], [0, 0, 0, 2, 2, 2]])
    for p, want in (((1, 1, 1), 1), ((0, 0, 0), 1), ((1.5, 1, 1), 2), ((2.0001, 1, 1), 0), ((-0.0001, 0.5, 0.5), 0)):
        assert capi.choose_volume(b, p) == want == oracle.choose_volume(b, p)
    assert capi.choose_volume(np.zeros((0, 6), np.float32), (0, 0, 0)) == 0


def test_oracle_lod_selection_restatement():
    """The LOD extension's C restatement against a literal numpy evaluation of RNEntity.cpp:98-106 + RNModel.cpp:260-281."""
    rng = np.random.default_rng(4)
    n = 500
    wpos = (rng.random((n, 3)).astype(np.float32) - np.float32(0.5)) * np.float32(400)
    so = np.array([0, 1, 4, 6], np.uint32)
    sd = np.array([0.5, 0.05, 0.2, 0.6, 0.1, 0.3], np.float32)
    model = rng.integers(0, 4, size=n).astype(np.uint16)
    model[model == 3] = 0xffff
    cam = wpos[7].copy()
    ids = rng.permutation(n).astype(np.uint32)
    got = oracle.select_lod(ids, wpos, model, so, sd, cam, 300.0)
    for e, i in enumerate(ids):
        m = int(model[i])
        if m == 0xffff:
            want = 0
        else:
            dist = sg_dist(wpos[i], cam) / np.float32(300.0)
            stages = sd[so[m]:so[m + 1]]
            if len(stages) <= 1 or dist <= np.finfo(np.float32).eps:
                want = 0
            else:
                hit = np.nonzero(dist <= stages)[0]
                want = int(hit[0]) if len(hit) else len(stages) - 1
        assert got[e] == want, (e, i, m)
    assert got[list(ids).index(7)] == 0


def sg_dist(a, b) -> np.float32:
    d = (a - b).astype(np.float32)
    return np.sqrt(np.float32(np.float32(np.float32(d[0] * d[0]) + np.float32(d[1] * d[1])) + np.float32(d[2] * d[2])))


def test_occlusion_oracle_matches_golden():
    """C8: the C restatement of the software occlusion pass against the reference's own run (60 frames, three cameras):
    the jitter sequence of the seeded shared generator, which cameras take the pass, every list after the transparent sort,
    every node's occludedFrameCounter after every frame and the depth buffer the last camera left."""
    from rayne_b200 import scenegen as sg
    from tests.helpers import load_golden_occlusion
    scene, cams, mesh_of, tri_ofs, tris, kw, g = load_golden_occlusion()
    frames = oracle.occlusion_frames(scene, cams, tri_ofs, tris, mesh_of, kw["frames"], kw["rng_seed"], kw["frame_size"],
                                     kw["root_translate"], kw["camera_translate"])
    assert np.array_equal(frames[0]["cams"][0]["projection"], g["ref_projection"][0])
    nc = len(cams)
    ofs = g["ref_list_ofs"]
    visit = sg.camera_render_order(cams["flags"])
    assert visit == [2, 1, 0]
    for f, fr in enumerate(frames):
        for c in range(nc):
            k = f * nc + c
            cam = fr["cams"][c]
            assert cam["used"] == int(g["ref_used"][f, c]), (f, c)
            assert np.array_equal(cam["offset"], g["ref_offset"][f, c]), (f, c)
            assert np.array_equal(cam["view"], g["ref_view"][f, c]), (f, c)
            assert np.array_equal(cam["list"], g["ref_lists"][ofs[k]:ofs[k + 1]]), (f, c)
        assert np.array_equal(fr["counters"], g["ref_counters"][f]), f
        last = [c for c in visit if fr["cams"][c]["used"]][-1]
        assert np.array_equal(fr["cams"][last]["depth"], g["ref_depth"][f]), f
    assert g["ref_counters"].max() >= 50 and g["ref_used"][:, 2].sum() == 0


def test_occlusion_oracle_matches_the_live_reference():
    """The same comparison on another scene where oracle/_ref exists (skipped on machines without it)."""
    from oracle import refio
    from rayne_b200 import scenegen as sg
    if not refio.have_harness("strict"):
        pytest.skip("oracle/_ref not built here")
    scene, meshes, custom, mesh_of = sg.make_occlusion_scene(seed=21, n_small=1500, n_occluders=64, n_groups=6)
    cams = np.stack([sg.make_camera(pos=(0, 1, 3)), sg.make_camera(pos=(-3, 0, 8), euler=(-6, 2, 0), flags=sg.CAM_RENDER_EARLY)])
    ref = refio.run_occlusion(scene, cams, meshes, mesh_of, frames=5, seed=9, frame_size=(800.0, 600.0), root_translate=(0.1, 0, 0),
                              camera_translate=(0, 0, -2.0), custom_meshes=custom)
    got = oracle.occlusion_frames(scene, cams, ref["tri_ofs"], ref["tris"], mesh_of, 5, 9, (800.0, 600.0), (0.1, 0, 0), (0, 0, -2.0))
    for f in range(5):
        for c in range(2):
            r = ref["frames"][f]["cams"][c]
            assert got[f]["cams"][c]["used"] == r["used"]
            assert np.array_equal(got[f]["cams"][c]["offset"], r["offset"])
            assert np.array_equal(got[f]["cams"][c]["list"], r["list"]), (f, c)
        assert np.array_equal(got[f]["counters"], ref["frames"][f]["counters"])
