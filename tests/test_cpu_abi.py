"""CPU: the C-ABI library loads, exports every symbol include/rn_cuda.h declares, its PODs have the layout the
ctypes binding assumes, the host-side entry points (camera frustum, render-list order) match the reference, and the
device entry points fail loudly - not silently on a CPU path - when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from oracle import oracle
from rayne_b200 import capi, scenegen as sg
from tests.conftest import have_gpu
from tests.helpers import GOLDEN_DIR, load_golden_scene, standard_cameras

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rn_cuda.h")


def declared_functions() -> list[str]:
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rn_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_north_star_entry_points():
    names = declared_functions()
    for must in ("rn_scene_upload", "rn_transform_propagate", "rn_cull", "rn_particles_step"):
        assert must in names


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    names = declared_functions()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} is declared in include/rn_cuda.h but not exported by librn_cuda.so"
    assert sorted(capi.EXPORTS) == names, "capi.EXPORTS and the header disagree"
    lib.rn_abi_version.restype = C.c_int
    assert lib.rn_abi_version() == 4


def test_pod_layouts_match_the_header():
    # sizes computed from the header's field lists (all 4-byte members, pointers 8 bytes)
    assert C.sizeof(capi.SceneDesc) == 8 * 4
    assert C.sizeof(capi.NodeColumns) == 9 * 8
    assert C.sizeof(capi.ParticleMesh) == 5 * 8 + 8
    assert C.sizeof(capi.VisList) == 2 * 8 + 8
    assert capi.FRUSTUM_DTYPE.itemsize == (3 + 1 + 24) * 4 + 8
    assert sg.CAMERA_DTYPE.itemsize == (2 + 3 + 4 + 3 + 4 + 4 + 4 + 1 + 16) * 4
    assert sg.EMITTER_DTYPE.itemsize == (2 + 1 + 2 + 8 + 8 + 18 + 2 + 6) * 4
    assert C.sizeof(capi.SceneStats) == 6 * 4 + 2 * 8
    assert capi.OCCLUSION_CAMERA_DTYPE.itemsize == (16 + 16 + 2 + 3 + 1) * 4


def test_build_frustum_matches_oracle_and_golden():
    cams = standard_cameras()
    got = capi.build_frustums(cams)
    assert got.tobytes() == oracle.frustums(cams).tobytes()
    _, gcams, g = load_golden_scene("scene_ragged")
    fr = capi.build_frustums(gcams)
    assert np.array_equal(fr["center"], g["ref_frustum"][:, 0:3])
    assert np.array_equal(fr["radius"], g["ref_frustum"][:, 3])
    assert np.array_equal(fr["planes"].reshape(-1, 24), g["ref_frustum"][:, 4:28])


@pytest.mark.parametrize("name", ["scene_uniform", "scene_nonuniform_groups_prio", "scene_ragged", "scene_deep_chain"])
def test_render_order_matches_golden(name):
    scene, _, g = load_golden_scene(name)
    assert np.array_equal(capi.render_order(scene.parent, scene.priority), g["ref_order"])


def test_render_order_probe_from_the_survey():
    """tree 0{1{3,4},2{5,6}}, then 7, 9 (Transparent), 8 (Early), 10 (Normal)."""
    parent = np.array([-1, 0, 0, 1, 1, 2, 2, -1, -1, -1, -1], np.int32)
    prio = np.zeros(11, np.int32)
    # creation order of the last three roots: 7, then 9 (transparent), 8 (early), 10 (normal)
    # AddNode is called in creation order of the ROOTS, so renumber: ids 8 and 9 swap roles via priorities
    prio[8] = sg.RENDER_TRANSPARENT   # added third  (the survey's node 9)
    prio[9] = sg.RENDER_EARLY         # added fourth (the survey's node 8)
    order = capi.render_order(parent, prio)
    assert order.tolist() == [9, 7, 0, 2, 6, 5, 1, 4, 3, 10, 8]
    assert np.array_equal(order, oracle.render_order(parent, prio))


def test_render_order_matches_the_linked_list_walk_on_random_forests():
    """rn_scenebasic_render_order keeps one deque per priority instead of the reference's walk from the tail; the
    oracle restates the walk literally (and is pinned by the golden orders).  Heavy ties, descending minima, many
    roots - the cases where PushFront / PushBack / insert-after differ."""
    rng = np.random.default_rng(0)
    for trial in range(40):
        n = int(rng.integers(1, 2500))
        parent = np.array([-1 if (i == 0 or rng.random() < 0.05) else int(rng.integers(0, i)) for i in range(n)], np.int32)
        prio = rng.choice([-5, 0, 0, 0, 3, 7, 100000, 200000, 300000], size=n).astype(np.int32)
        assert np.array_equal(capi.render_order(parent, prio), oracle.render_order(parent, prio)), trial


def test_render_order_rejects_bad_hierarchy():
    with pytest.raises(capi.RNError):
        capi.render_order(np.array([1, -1], np.int32), None)   # parent after child


def test_argument_errors_do_not_need_a_device():
    lib = capi.load()
    assert lib.rn_scene_create(None, None) == -1
    assert b"NULL" in lib.rn_last_error()
    out = C.c_void_p()
    desc = capi.SceneDesc(0, 0, 0, 0, 0, 0)
    assert lib.rn_scene_create(C.byref(out), C.byref(desc)) == -1      # max_nodes = 0
    desc = capi.SceneDesc(0, 16, 100, 0, 0, 0)
    assert lib.rn_scene_create(C.byref(out), C.byref(desc)) == -1      # tile_nodes not a multiple of 16


def test_static_expand_checks_its_limits_before_it_needs_a_device():
    """rn_masks_expand_static: the segment count and the size of its __device__ scratch array are argument errors,
    reported before any CUDA call (no compute here: every call below is rejected or has nothing to do)."""
    limit = capi.MASKS_EXPAND_STATIC_MAX_WORDS
    assert limit % 1024 == 0 and f"({limit // (1024 * 1024)}u * 1024u * 1024u)" in open(HEADER).read()
    fake = 0x1000      # never dereferenced: the calls fail before any launch
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(fake, 1, 0)] * 33, fake, 4, fake)
    assert e.value.status == -1 and "32" in str(e.value)
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(fake, limit + 1, 0)], fake, 4, fake)
    assert e.value.status == -5 and "RN_MASKS_EXPAND_STATIC_MAX_WORDS" in str(e.value)
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(fake, limit - 1024, 0), (fake, 1, 0), (fake, 1, 0)], fake, 4, fake)   # 1 + 1 word round up to 2 CTAs
    assert e.value.status == -5
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(0, 5, 0)], fake, 4, fake)              # NULL words
    assert e.value.status == -1
    with pytest.raises(capi.RNError) as e:
        capi.masks_expand_static([(fake, 5, 0)], 0, 4, fake)              # NULL list
    assert e.value.status == -1
    capi.masks_expand_static([], fake, 4, fake)                           # nothing to do: no device needed
    capi.masks_expand_static([(fake, 0, 0)], fake, 4, fake)


@pytest.mark.skipif(have_gpu(), reason="only meaningful on a machine without a CUDA device")
def test_no_cpu_fallback_without_a_device():
    """On a CPU-only box the product path must fail loudly (RN_ERR_CUDA), never compute on the host."""
    with pytest.raises(capi.RNError) as e:
        capi.SceneHandle(max_nodes=16)
    assert e.value.status == -2
    assert "CUDA" in str(e.value)


def test_product_sources_never_touch_the_oracle():
    """oracle/ is test infrastructure: nothing under rayne_b200/ may import, link or execute it."""
    pkg = os.path.join(ROOT, "rayne_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "rn_oracle" not in text and "oracle/_ref" not in text, f


def test_light_cascades_match_golden():
    """rn_light_build_cascades (Light::UpdateShadows + Camera::MakeShadowSplit) against the reference's own cascade
    cameras (tests/golden/cascades_basic.npz): world position after the texel snap, rotation, orthographic extents,
    clip planes, flags, the 28 frustum floats rn_camera_build_frustum derives from them, and _shadowCameraMatrices -
    all bit-exact."""
    import os
    from tests.helpers import GOLDEN_DIR
    g = np.load(os.path.join(GOLDEN_DIR, "cascades_basic.npz"))
    k = 0
    while f"in_target{k}" in g.files:
        cams, pv = capi.build_cascades(g[f"in_target{k}"], g[f"in_shadow{k}"])
        assert cams["pos"].tobytes() == g[f"ref_pos{k}"].tobytes()
        assert cams["rot"].tobytes() == g[f"ref_rot{k}"].tobytes()
        assert cams["ortho"].tobytes() == g[f"ref_ortho{k}"].tobytes()
        assert np.stack([cams["clip_near"], cams["clip_far"]], 1).astype(np.float32).tobytes() == g[f"ref_clip{k}"].tobytes()
        assert np.array_equal(cams["flags"], g[f"ref_flags{k}"])
        assert pv.tobytes() == g[f"ref_projview{k}"].tobytes()
        fr = capi.build_frustums(cams)
        got28 = np.concatenate([fr["center"], fr["radius"][:, None], fr["planes"].reshape(-1, 24)], 1).astype(np.float32)
        assert got28.tobytes() == g[f"ref_frustum{k}"].tobytes()
        k += 1
    assert k == 2
    # the C4 rig built from the reference-derived cascades: 2 eye cameras + 4 orthogonal cascade cameras
    rig = sg.vr_cascade_rig()
    assert rig.shape[0] == 6 and (rig["flags"][2:] & sg.CAM_ORTHOGONAL).all()


def test_plan_ranges_matches_the_python_planner():
    """rn_plan_ranges (C) against shard.plan_shards on random forests: same cuts, all edge-free."""
    from rayne_b200 import shard
    for seed in range(6):
        scene = sg.make_scene(ragged_nodes=3000 + 500 * seed, depth=6, seed=seed) if seed % 2 else sg.make_scene(n_trees=17 + seed, branching=3, depth=5, seed=seed)
        rs = capi.RankedScene(scene)
        for world in (1, 2, 3, 8):
            try:
                want = shard.plan_shards(rs.parent, world)
            except ValueError:
                with pytest.raises(capi.RNError):
                    capi.plan_ranges(rs.parent, world)
                continue
            got = capi.plan_ranges(rs.parent, world)
            assert [int(a) for a, _ in want] + [rs.n] == got.tolist(), (seed, world)
    with pytest.raises(capi.RNError):
        capi.plan_ranges(capi.RankedScene(sg.make_scene(n_trees=1, branching=2, depth=4, seed=1)).parent, 2)
