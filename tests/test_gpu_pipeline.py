"""The pipelined frame of the C ABI (rn_pipeline.cu): rank ranges on prioritised streams, visibility written into mapped
pinned host memory by the device in list / mask / adaptive form, a ring of frames in flight - against the single-scene lists."""
import numpy as np
import pytest

from rayne_b200 import capi, scenegen as sg
from tests.helpers import standard_cameras

pytestmark = pytest.mark.gpu


def single_scene_lists(rs, fr, roots=None, moved=None, rank_base=0):
    with capi.SceneHandle(max_nodes=rs.n) as h:
        h.upload(rs)
        h.set_list_mode(True, rank_base)
        if roots is not None:
            h.update_nodes(roots, position=moved)
        h.transform_cull(fr)
        return [h.read_list(c) for c in range(len(fr))]


@pytest.mark.parametrize("form", [capi.VIS_LISTS, capi.VIS_MASKS, capi.VIS_ADAPTIVE])
@pytest.mark.parametrize("n_ranges", [1, 3, 5])
def test_pipeline_frames_match_the_single_scene(form, n_ranges):
    scene = sg.make_scene(n_trees=40, branching=4, depth=6, seed=19, random_bounds=True, random_groups=True)
    cams = standard_cameras()
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(scene)
    roots = rs.roots
    want0 = single_scene_lists(rs, fr, rank_base=700)
    moves = [rs.position[roots] + np.float32([0.5 * k, 0, -0.25 * k]) for k in range(1, 5)]
    wants = [want0] + [single_scene_lists(rs, fr, roots, m, rank_base=700) for m in moves]
    with capi.Pipeline(rs, n_ranges=n_ranges, frames_in_flight=2, rank_base=700) as p:
        assert p.bounds[0] == 0 and p.bounds[-1] == rs.n and (np.diff(p.bounds.astype(np.int64)) > 0).all()
        # two frames in flight: frame k + 1 is submitted before frame k is looked at
        p.submit(fr, form)
        for k in range(5):
            if k < 4:
                p.update_nodes(roots, position=moves[k])
                p.submit(fr, form)
            view = p.acquire()
            assert view.frame == k and view.n_cameras == len(fr) and view.n_ranges == n_ranges
            got = p.frame_lists(view)
            forms = p.forms(view)
            if form != capi.VIS_ADAPTIVE:
                assert (forms == form).all()
            elif n_ranges == 1:
                assert (forms == capi.VIS_MASKS).all()    # single range: the copy engine moves the (fixed-size) masks
            else:
                for kk in range(n_ranges):
                    words = (int(p.bounds[kk + 1]) - int(p.bounds[kk]) + 31) // 32
                    for c in range(len(fr)):
                        seg = view.segments[kk * view.n_cameras + c]
                        assert seg.form == (capi.VIS_LISTS if seg.count <= words else capi.VIS_MASKS)
            assert view.d2h_bytes > 0 and (k == 0 or view.h2d_bytes > 0)
            for c in range(len(fr)):
                assert np.array_equal(got[c], wants[k][c]), (k, c)
            p.release()
        with pytest.raises(capi.RNError):
            p.acquire()                      # nothing in flight


def test_pipeline_ring_is_bounded_and_errors_are_loud():
    scene = sg.make_scene(n_trees=6, branching=3, depth=4, seed=2)
    rs = capi.RankedScene(scene)
    fr = capi.build_frustums(standard_cameras()[:2])
    with capi.Pipeline(rs, n_ranges=2, frames_in_flight=2) as p:
        p.submit(fr)
        p.submit(fr)
        with pytest.raises(capi.RNError):
            p.submit(fr)                     # ring full
        p.acquire()
        with pytest.raises(capi.RNError):
            p.acquire()                      # the acquired frame must be released first
        p.release()
        p.acquire(); p.release()
        with pytest.raises(capi.RNError):
            p.update_nodes(np.array([rs.n], np.uint32), position=np.zeros((1, 3), np.float32))
    with pytest.raises(capi.RNError):
        capi.Pipeline(capi.RankedScene(sg.make_scene(n_trees=2, branching=1, depth=6, seed=1)), n_ranges=5)   # only two independent pieces
