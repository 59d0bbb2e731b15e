"""GPU parity: librn_cuda (through the C ABI) against the oracle on seeded synthetic scenes.
Visibility lists bit-exact, world TRS bit-exact, world matrices within 1e-5 relative (north star)."""
import numpy as np
import pytest

from rayne_b200 import capi, scenegen as sg
from tests.helpers import check_against_oracle, standard_cameras, run_cuda, same_floats

pytestmark = pytest.mark.gpu

CASES = {
    "c1_10k_depth4": dict(n_trees=10, branching=(9, 10, 10), depth=4, seed=12345),   # SURVEY's C1: 10 x (1+9+90+900) = 10 000
    "uniform_4ary_depth8": dict(n_trees=3, branching=4, depth=8, seed=2),
    "nonuniform_groups_prio": dict(n_trees=12, branching=4, depth=6, seed=4, nonuniform_scale=True,
                                   random_bounds=True, random_groups=True, priorities=True),
    "ragged": dict(ragged_nodes=60000, depth=8, seed=5, random_bounds=True),
    "flat": dict(n_trees=3000, branching=1, depth=1, seed=6),
    "chain": dict(n_trees=2, branching=1, depth=40, seed=7),
    "wide": dict(n_trees=2, branching=700, depth=3, seed=8),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("strict", [True, False])
def test_scene_parity(name, fused, strict):
    """strict: RN_SCENE_STRICT_MATRICES, matrices bit-exact; otherwise the library default (FMA matrix chain, 1e-5).
    Lists and world TRS are bit-exact in both."""
    scene = sg.make_scene(**CASES[name])
    got, ref = check_against_oracle(scene, standard_cameras(), fused=fused, strict=strict)
    assert got["stats"]["nodes"] == scene.n
    if name == "c1_10k_depth4":
        assert scene.n == 10000


def test_c2_full_size_against_the_oracle():
    """Config C2 exactly (48 complete 4-ary depth-8 trees = 1 048 560 nodes): lists, world TRS and matrices against
    the C oracle for the single default camera BASELINE.json names AND for the 6-view rig, in both matrix modes,
    fused and unfused.  (And against the reference itself where oracle/_ref travels with the snapshot.)"""
    from oracle import oracle, refio
    scene = sg.config_scene("C2")
    assert scene.n == 1048560
    for cams in (np.atleast_1d(sg.make_camera()), sg.six_view_rig()):
        fr = capi.build_frustums(cams)
        ref = oracle.scene_reference(scene, fr)
        check_against_oracle(scene, cams, strict=True, ref=ref)
        check_against_oracle(scene, cams, strict=False, ref=ref)
        check_against_oracle(scene, cams, strict=True, ref=ref, fused=False)
        if refio.have_harness("strict"):
            real = refio.run_scene(scene, cams, kind="strict")
            assert np.array_equal(real["order"], ref["order"])
            assert same_floats(real["wmat"], ref["wmat"])
            for a, b in zip(real["lists"], ref["lists"]):
                assert np.array_equal(a, b)


def test_scattered_priorities_1m_repeated():
    """1 M nodes with random render priorities: the render list separates children from their parents, so ~20 % of the
    nodes are spine nodes (7 spine depths) and the tiles are mostly single-leaf units whose parents sit in global memory -
    the shape on which the tile kernel once read a stage the next TMA copy was already overwriting (generic-proxy read
    vs async-proxy write without a proxy fence; profiles/r02_tile_proxy_fence.txt).  Fresh handle per repetition, fused
    and unfused, everything bit-exact against the oracle."""
    from oracle import oracle
    scene = sg.make_scene(n_trees=48, branching=4, depth=8, seed=12345, priorities=True, transparent_frac=0.05)
    cams = standard_cameras()[:3]
    ref = oracle.scene_reference(scene, capi.build_frustums(cams))
    for rep in range(2):
        for fused in (True, False):
            got, _ = check_against_oracle(scene, cams, strict=True, ref=ref, fused=fused)
    assert got["stats"]["spine_nodes"] > 100000


@pytest.mark.parametrize("tile", [64, 256, 1024])
def test_tile_size_independent(tile):
    scene = sg.make_scene(n_trees=5, branching=4, depth=7, seed=11, random_bounds=True)
    check_against_oracle(scene, standard_cameras(), tile_nodes=tile)


def test_single_node_and_empty_lists():
    scene = sg.make_scene(n_trees=1, branching=1, depth=1, seed=1)
    cams = np.stack([sg.make_camera(pos=(1e6, 0, 0)), sg.make_camera()])
    got, ref = check_against_oracle(scene, cams)
    assert got["counts"][0] == 0


def test_partial_update_matches_full_upload():
    scene = sg.make_scene(n_trees=40, branching=4, depth=5, seed=21)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    moved = sg.Scene(**{k: getattr(scene, k).copy() for k in ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")})
    roots = np.nonzero(scene.parent < 0)[0]
    moved.pos[roots] += np.float32([0.01, 0, 0])
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        h.transform_cull(fr)
        h.update_nodes(rs.rank_of[roots], position=moved.pos[roots])
        h.transform_cull(fr)
        a = h.download(capi.DL_WORLD_MATRIX)
        la = [h.read_list(c) for c in range(len(fr))]
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(capi.RankedScene(moved))
        h.transform_cull(fr)
        b = h.download(capi.DL_WORLD_MATRIX)
        lb = [h.read_list(c) for c in range(len(fr))]
    assert same_floats(a, b)
    assert all(np.array_equal(x, y) for x, y in zip(la, lb))


def test_errors_are_loud():
    with capi.SceneHandle(max_nodes=16) as h:
        with pytest.raises(capi.RNError):
            h.transform_propagate()          # nothing uploaded
        scene = sg.make_scene(n_trees=1, branching=2, depth=3, seed=1)
        h.upload(capi.RankedScene(scene))
        with pytest.raises(capi.RNError):
            h.cull(capi.build_frustums(sg.make_camera()))   # cull before propagate
        big = sg.make_scene(n_trees=1, branching=2, depth=6, seed=1)
        with pytest.raises(capi.RNError):
            h.upload(capi.RankedScene(big))  # exceeds capacity


def test_full_size_properties():
    """1 M nodes (config C2): order-independent properties (the oracle comparison is the test above)."""
    scene = sg.make_scene(n_trees=48, branching=4, depth=8, seed=12345)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    got = run_cuda(scene, fr)
    for c, r in enumerate(got["slot_lists_rank"]):
        assert (np.diff(r.astype(np.int64)) > 0).all()
        assert got["counts"][c] == r.shape[0]
    # hidden nodes never appear, NoCulling nodes of a matching group always do
    hidden = np.nonzero(scene.flags & sg.FLAG_HIDDEN)[0]
    nocull = np.nonzero((scene.flags & sg.FLAG_NOCULLING) != 0)[0]
    for c in (0, 3):
        s = set(got["lists"][c].tolist())
        assert not s.intersection(hidden.tolist())
        assert set(nocull.tolist()).issubset(s)
    # the simple-culling camera sees a superset of nothing smaller than the plane-tested camera at the same pose
    # unfused path gives the same lists
    got2 = run_cuda(scene, fr, fused=False)
    assert all(np.array_equal(a, b) for a, b in zip(got["lists"], got2["lists"]))
    assert same_floats(got["wmat"], got2["wmat"])


def test_sharded_matches_single():
    """rank ranges cut at subtree boundaries, each culled on its own handle (one per GPU in
    production), lists in global-rank form; their concatenation is the single-GPU / reference list."""
    from rayne_b200 import shard
    from oracle import oracle
    scene = sg.make_scene(n_trees=40, branching=4, depth=6, seed=17, random_bounds=True, random_groups=True)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    ref = oracle.scene_reference(scene, fr)
    want = [rs.rank_of[l] for l in ref["lists"]]
    for world in (1, 3, 8):
        parts = [[] for _ in fr]
        for a, b in shard.plan_shards(rs.parent, world):
            with capi.SceneHandle(max_nodes=b - a) as h:
                h.upload(shard.extract_shard(rs, a, b))
                h.set_list_mode(True, a)
                h.transform_cull(fr)
                for c in range(len(fr)):
                    parts[c].append(h.read_list(c))
        for c in range(len(fr)):
            assert np.array_equal(np.concatenate(parts[c]), want[c]), (world, c)


def test_classifier_edge_cases():
    """Negative and huge scales, zero radii, nodes exactly on planes, clusters straddling every plane: the warp-level
    classifier must never change a decision of the exact test."""
    scene = sg.make_scene(n_trees=30, branching=4, depth=5, seed=77, random_bounds=True, nonuniform_scale=True)
    rng = np.random.Generator(np.random.PCG64(5))
    n = scene.n
    neg = rng.random(n) < 0.05
    scene.scale[neg] *= np.float32(-1.0)                     # negative world scale -> negative radius (signed max)
    big = rng.random(n) < 0.01
    scene.scale[big] *= np.float32(1e4)
    scene.sphere[rng.random(n) < 0.05, 3] = 0.0
    # shrink the scene so that many subtrees straddle the frustum planes of the cameras at the origin
    scene.pos *= np.float32(0.05)
    check_against_oracle(scene, standard_cameras())
    # positions that make plane distances exactly equal to the radius for axis-aligned planes
    flat = sg.make_scene(n_trees=2000, branching=1, depth=1, seed=3)
    flat.pos[:, 0] = np.float32(0.0)
    flat.pos[:, 1] = np.float32(0.0)
    flat.pos[:, 2] = -np.linspace(0.0, 510.0, flat.n, dtype=np.float32)
    flat.sphere[:, 3] = np.float32(1.0)
    flat.scale[:] = np.float32(1.0)
    check_against_oracle(flat, standard_cameras())
