"""rn_group: several GPUs driven by ONE process through the C ABI (rn_group.cu).  Runs where at least two GPUs are visible
(`gpurun --gpus 2`); with one GPU only the single-member group is exercised."""
import numpy as np
import pytest

from rayne_b200 import capi, scenegen as sg
from tests.helpers import standard_cameras

pytestmark = pytest.mark.gpu


def gpu_count() -> int:
    import torch
    return torch.cuda.device_count()


def reference_lists(rs, fr, roots=None, moved=None):
    with capi.SceneHandle(max_nodes=rs.n) as h:
        h.upload(rs)
        h.set_list_mode(True, 0)
        if roots is not None:
            h.update_nodes(roots, position=moved)
        h.transform_cull(fr)
        return [h.read_list(c) for c in range(len(fr))]


@pytest.mark.parametrize("transport", [capi.GROUP_AUTO, capi.GROUP_PEER])
def test_group_frames_match_the_single_scene(transport):
    n_dev = min(gpu_count(), 8)
    scene = sg.make_scene(n_trees=64, branching=4, depth=6, seed=23, random_bounds=True, random_groups=True)
    cams = standard_cameras()[:6]
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    roots = rs.roots
    moves = [None] + [rs.position[roots] + np.float32([0.75 * k, 0, -0.5 * k]) for k in range(1, 4)]
    with capi.Group(rs, list(range(n_dev)), transport=transport) as g:
        assert g.bounds[0] == 0 and g.bounds[-1] == rs.n
        if n_dev > 1:
            assert g.transport in (capi.GROUP_PEER, capi.GROUP_MULTICAST), g.note
            if transport == capi.GROUP_PEER:
                assert g.transport == capi.GROUP_PEER
        for k, mv in enumerate(moves):
            if mv is not None:
                g.update_nodes(roots, position=mv)
            g.frame(fr)          # differing frames back to back: a stale or torn exchange would show
            want = reference_lists(rs, fr, roots if mv is not None else None, mv)
            for member in range(n_dev):
                for c in range(len(fr)):
                    assert np.array_equal(g.read_list(member, c), want[c]), (k, member, c, g.note)
        print("transport:", g.note)
