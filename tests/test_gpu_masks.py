"""GPU parity of the compact visibility form (rank-order bitmasks, rn_masks.cu): masks == the lists of the cull, and
masks expanded rank range after rank range == the single-scene / reference list, which is what the multi-GPU exchange
relies on (shard.PipelinedMaskExchange)."""
import numpy as np
import pytest

from oracle import oracle
from rayne_b200 import capi, scenegen as sg, shard
from tests.helpers import load_golden_scene, standard_cameras

pytestmark = pytest.mark.gpu


def bits_of(words: np.ndarray, n: int) -> np.ndarray:
    return np.nonzero(np.unpackbits(np.ascontiguousarray(words).view(np.uint8), bitorder="little")[:n])[0].astype(np.uint32)


@pytest.mark.parametrize("name", ["scene_ragged", "scene_nonuniform_groups_prio", "scene_deep_chain"])
def test_masks_match_golden_lists(name):
    """Set bits (ranks) of camera c's mask == the reference's visibility list of camera c."""
    scene, cams, g = load_golden_scene(name)
    cams = cams[:capi.RN_MAX_CAMERAS]
    rs = capi.RankedScene(scene)
    fr = capi.build_frustums(cams)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        with pytest.raises(capi.RNError):
            h.build_masks()                      # no cull yet
        h.transform_cull(fr)
        assert h.device_masks()[0] == 0          # not built yet: NULL, not stale
        with pytest.raises(capi.RNError):
            h.read_masks()
        h.build_masks()
        ptr, words, k = h.device_masks()
        assert ptr != 0 and words == (scene.n + 31) // 32 and k == len(cams)
        m = h.read_masks()
        for c in range(len(cams)):
            assert np.array_equal(rs.order[bits_of(m[c], scene.n)], g[f"ref_list{c}"]), f"camera {c}"
        h.transform_cull(fr[:2])                 # a newer cull invalidates the masks
        assert h.device_masks()[0] == 0


@pytest.mark.parametrize("kw", [
    dict(n_trees=40, branching=4, depth=6, seed=17, random_bounds=True, random_groups=True),
    dict(ragged_nodes=70001, depth=8, seed=5, random_bounds=True),
    dict(n_trees=33, branching=1, depth=1, seed=3),          # 33 single nodes: the last word is partial
])
@pytest.mark.parametrize("fused", [True, False])
def test_masks_and_expansion_match_lists(kw, fused):
    import torch
    scene = sg.make_scene(**kw)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        h.set_list_mode(True, 0)
        if fused:
            h.transform_cull(fr)
        else:
            h.transform_propagate()
            h.cull(fr)
        h.build_masks()
        ptr, words, k = h.device_masks()
        m = h.read_masks()
        # the same masks into a caller-owned buffer with a wider row (the send buffer of the all-gather)
        stride = words + 5
        ext = torch.full((len(cams), stride), 0x5a5a5a5a, dtype=torch.int32, device="cuda")
        h.build_masks_to(ext.data_ptr(), stride)
        torch.cuda.synchronize()
        ext = ext.cpu().numpy().view(np.uint32)
        assert np.array_equal(ext[:, :words], m) and (ext[:, words:] == 0x5a5a5a5a).all()
        with pytest.raises(capi.RNError):
            h.build_masks_to(1024, words - 1)
        out = torch.full((scene.n + 8,), -1, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        for c in range(len(cams)):
            want = h.read_list(c)
            assert np.array_equal(bits_of(m[c], scene.n), want), f"camera {c}"
            out.fill_(-1); cnt.zero_()
            torch.cuda.synchronize()
            h.masks_expand([(ptr + 4 * c * words, words, 0)], out.data_ptr(), scene.n, cnt.data_ptr())
            torch.cuda.synchronize()
            got = out.cpu().numpy()
            assert int(cnt.item()) == len(want)
            assert np.array_equal(got[:len(want)].astype(np.uint32), want) and (got[len(want):] == -1).all(), f"camera {c}"
        # capacity smaller than the list: counted in full, written up to the capacity only
        c = 2                                    # simple-culling camera: (nearly) every node
        want = h.read_list(c)
        out.fill_(-1); cnt.zero_()
        torch.cuda.synchronize()
        h.masks_expand([(ptr + 4 * c * words, words, 0)], out.data_ptr(), len(want) // 2, cnt.data_ptr())
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert int(cnt.item()) == len(want)
        assert np.array_equal(got[:len(want) // 2].astype(np.uint32), want[:len(want) // 2]) and (got[len(want) // 2:] == -1).all()


@pytest.mark.parametrize("world", [2, 5])
def test_sharded_masks_expand_to_the_single_scene_list(world):
    """SURVEY.md §8(e) with the compact exchange: every rank range builds its masks; expanding the ranges' masks one
    behind the other (device-side running count, rank_base = first rank of the range) gives the oracle's list."""
    import torch
    scene = sg.make_scene(n_trees=40, branching=4, depth=6, seed=23, random_bounds=True, random_groups=True)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    ref = oracle.scene_reference(scene, fr)
    want = [rs.rank_of[l] for l in ref["lists"]]
    ranges = shard.plan_shards(rs.parent, world)
    handles, masks = [], []
    try:
        for a, b in ranges:
            h = capi.SceneHandle(max_nodes=b - a)
            h.upload(shard.extract_shard(rs, a, b))
            h.transform_cull(fr)
            h.build_masks()
            handles.append(h)
            # what the all-gather would deliver: a device copy of the range's masks
            masks.append(torch.from_numpy(h.read_masks().view(np.int32)).cuda())
        out = torch.full((scene.n,), -1, dtype=torch.int32, device="cuda")
        cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
        for c in range(len(cams)):
            out.fill_(-1); cnt.zero_()
            torch.cuda.synchronize()
            segs = [(mk[c].data_ptr(), mk.shape[1], a) for (a, b), mk in zip(ranges, masks)]
            if c % 2:       # one call with all ranges, or one call per range on the running device-side count
                handles[0].masks_expand(segs, out.data_ptr(), scene.n, cnt.data_ptr())
            else:
                for sg_ in segs:
                    handles[0].masks_expand([sg_], out.data_ptr(), scene.n, cnt.data_ptr())
            torch.cuda.synchronize()
            k = int(cnt.item())
            assert k == len(want[c])
            assert np.array_equal(out.cpu().numpy()[:k].astype(np.uint32), want[c]), f"camera {c}"
    finally:
        for h in handles:
            h.close()
