"""Incremental edits through the C ABI (rn_scene_add_nodes / rn_scene_remove_nodes / rn_scene_update_nodes_by_slot): after
every edit the lists and world transforms must equal what the oracle computes for the resulting scene from scratch, in the
order SceneBasic::AddRenderNode gives (front insertion in reverse call order, back insertion in call order), and - where the
final scene can be expressed as one - what a full upload of it gives."""
import numpy as np
import pytest

from oracle import oracle
from rayne_b200 import capi, scenegen as sg
from tests.helpers import standard_cameras, same_floats

pytestmark = pytest.mark.gpu

COLS = ("parent", "pos", "scale", "rot", "flags", "group", "priority", "sphere")


def call_sequence(parent):
    """AddRenderNode call order of a forest given in creation order with complete trees: post-order, roots in order."""
    n = len(parent)
    kids = [[] for _ in range(n)]
    roots = []
    for i, p in enumerate(parent):
        (kids[p] if p >= 0 else roots).append(i)
    seq = []
    for r in roots:
        stack = [(r, 0)]
        while stack:
            node, k = stack.pop()
            if k < len(kids[node]):
                stack.append((node, k + 1))
                stack.append((kids[node][k], 0))
            else:
                seq.append(node)
    return np.asarray(seq, np.int64)


class World:
    """Host-side truth: all nodes ever created (creation ids), their current list order, which are alive."""

    def __init__(self, base: sg.Scene):
        self.cols = {k: getattr(base, k).copy() for k in COLS}
        self.order = list(oracle.render_order(base.parent, base.priority))     # list order of live ids
        self.alive = np.ones(base.n, bool)

    @property
    def n(self):
        return self.cols["parent"].shape[0]

    def append(self, extra: sg.Scene, parent_ids):
        """extra: a forest in creation order whose roots hang under parent_ids[k] (creation id or -1); returns new ids."""
        first = self.n
        p = extra.parent.copy()
        roots = np.nonzero(p < 0)[0]
        p[p >= 0] += first
        for r, pid in zip(roots, parent_ids):
            p[r] = pid
        for k in COLS:
            v = p if k == "parent" else getattr(extra, k)
            self.cols[k] = np.concatenate([self.cols[k], v])
        self.alive = np.concatenate([self.alive, np.ones(extra.n, bool)])
        return first + np.arange(extra.n)

    def scene(self):
        return sg.Scene(**self.cols)

    def expected(self, fr):
        sc = self.scene()
        par = np.where(sc.parent >= 0, sc.parent, 0xFFFFFFFF).astype(np.uint32)
        wpos, wrot, wscale, wmat = oracle.world(par, sc.pos, sc.scale, sc.rot)
        order = np.asarray(self.order, np.int64)
        lists = []
        for f in np.atleast_1d(fr):
            vis = oracle.cull(wpos[order], wrot[order], wscale[order], sc.sphere[order], sc.flags[order], sc.group[order], f)
            lists.append(order[vis])
        return lists, wmat


def check(h, world, fr, slot_of_id, id_of_slot):
    want, wmat = world.expected(fr)
    h.transform_cull(fr)
    for c in range(len(fr)):
        got = id_of_slot[h.read_list(c)]
        assert np.array_equal(got, want[c]), (c, len(got), len(want[c]))
    # world matrices of every live node, read back by (padded) rank
    ros = h.download(capi.DL_RANK_OF_SLOT)
    mats = h.download(capi.DL_WORLD_MATRIX)
    live = np.nonzero(world.alive)[0]
    assert same_floats(mats[ros[slot_of_id[live]]], wmat[live])


def test_add_and_remove_match_the_from_scratch_scene():
    base = sg.make_scene(n_trees=30, branching=4, depth=5, seed=51, random_bounds=True, random_groups=True)
    cams = standard_cameras()[:4]
    fr = capi.build_frustums(cams)
    rs = capi.RankedScene(base)
    world = World(base)
    with capi.SceneHandle(max_nodes=base.n + 6000, strict_matrices=True) as h:
        h.upload(rs)
        sor = h.download(capi.DL_SLOT_OF_RANK)[:base.n]
        slot_of_id = np.full(base.n + 6000, -1, np.int64)
        slot_of_id[rs.order] = sor
        id_of_slot = np.full(base.n + 6000, -1, np.int64)
        id_of_slot[sor] = rs.order
        check(h, world, fr, slot_of_id, id_of_slot)

        def add(extra, parent_ids):
            ids = world.append(extra, parent_ids)
            seq = call_sequence(extra.parent)                 # batch-local indices in AddRenderNode call order
            pos_in_seq = np.empty(extra.n, np.int64); pos_in_seq[seq] = np.arange(extra.n)
            par = np.full(extra.n, capi.RN_NO_PARENT, np.uint32)
            roots = list(np.nonzero(extra.parent < 0)[0])
            for k, i in enumerate(seq):
                if extra.parent[i] >= 0:
                    par[k] = pos_in_seq[extra.parent[i]]
                elif parent_ids[roots.index(i)] >= 0:
                    par[k] = capi.PARENT_IS_SLOT | int(slot_of_id[parent_ids[roots.index(i)]])
            slots = h.add_nodes(extra.pos[seq], extra.scale[seq], extra.rot[seq], par, extra.flags[seq], extra.group[seq], extra.sphere[seq])
            slot_of_id[ids[seq]] = slots
            id_of_slot[slots] = ids[seq]
            # all RenderNormal: every node is pushed to the FRONT in call order -> the batch appears reversed before the rest
            world.order = list(ids[seq][::-1]) + world.order
            return ids

        # 1. new trees (roots without a parent), several per batch and several batches
        for seed in (1, 2):
            add(sg.make_scene(n_trees=3, branching=3, depth=4, seed=100 + seed, random_bounds=True, hidden_frac=0.05), [-1, -1, -1])
            check(h, world, fr, slot_of_id, id_of_slot)
        # 2. children under nodes of the upload: under a root (large trees: the root is a spine node or the unit root) and under
        #    leaves / inner nodes that a tile computes (late tiles)
        leaves = [i for i in range(base.n) if not (base.parent == i).any()][:40]
        inner = [int(base.parent[l]) for l in leaves[:10]]
        kids = sg.make_scene(n_trees=25, branching=2, depth=2, seed=7, random_bounds=True)
        kid_roots = np.nonzero(kids.parent < 0)[0]
        targets = (leaves + inner)[:len(kid_roots)]
        new_ids = add(kids, targets)
        st = h.edit_stats()
        assert st["late_tiles"] >= 1 and st["live_nodes"] == int(world.alive.sum())
        check(h, world, fr, slot_of_id, id_of_slot)
        # 3. setters on old and new nodes by slot
        movers = np.concatenate([rs.order[rs.roots][:5], new_ids[:7]])
        world.cols["pos"][movers] += np.float32([3.0, -1.0, 0.5])
        h.update_nodes_by_slot(slot_of_id[movers].astype(np.uint32), position=world.cols["pos"][movers])
        check(h, world, fr, slot_of_id, id_of_slot)
        # 4. removal of whole subtrees: an uploaded tree and an added one
        def subtree(root):
            par = world.cols["parent"]
            out, todo = [], [root]
            while todo:
                x = todo.pop()
                out.append(x)
                todo.extend(np.nonzero(par == x)[0].tolist())
            return np.asarray(out, np.int64)
        for root in (int(rs.order[rs.roots][2]), int(new_ids[0])):
            gone = subtree(root)
            gone = gone[world.alive[gone]]
            h.remove_nodes(slot_of_id[gone].astype(np.uint32))
            world.alive[gone] = False
            dead = set(gone.tolist())
            world.order = [i for i in world.order if i not in dead]
            check(h, world, fr, slot_of_id, id_of_slot)
        # 5. the same final scene from a full upload gives the same lists (as creation ids)
        live = np.nonzero(world.alive)[0]
        want, _ = world.expected(fr)
    remap = np.full(world.n, -1, np.int64); remap[live] = np.arange(len(live))
    sc = world.scene()
    final = sg.Scene(parent=np.where(sc.parent[live] >= 0, remap[np.maximum(sc.parent[live], 0)], -1).astype(np.int32),
                     **{k: world.cols[k][live] for k in COLS if k != "parent"})
    with capi.SceneHandle(max_nodes=final.n, strict_matrices=True) as h2:
        # rank order = the order the edits produced (what SceneBasic's list would be), not render_order of a rebuilt forest
        rs2 = capi.RankedScene(final)
        order2 = remap[np.asarray(world.order, np.int64)]
        rs2.order = order2.astype(np.uint32)
        rs2.rank_of = np.empty(final.n, np.uint32); rs2.rank_of[rs2.order] = np.arange(final.n, dtype=np.uint32)
        o = rs2.order
        rs2.position, rs2.scale, rs2.rotation = final.pos[o], final.scale[o], final.rot[o]
        p = final.parent[o]
        rs2.parent = np.where(p >= 0, rs2.rank_of[np.maximum(p, 0)], capi.RN_NO_PARENT).astype(np.uint32)
        rs2.flags, rs2.group, rs2.sphere = final.flags[o], final.group[o], final.sphere[o]
        h2.upload(rs2)
        h2.transform_cull(fr)
        ros2 = h2.download(capi.DL_RANK_OF_SLOT)
        for c in range(len(fr)):
            assert np.array_equal(live[rs2.order[ros2[h2.read_list(c)]]], want[c]), c


def test_edit_errors_and_fallbacks_are_loud():
    base = sg.make_scene(n_trees=4, branching=3, depth=4, seed=3, priorities=True)
    rs = capi.RankedScene(base)
    one = sg.make_scene(n_trees=1, branching=1, depth=1, seed=1)
    args = lambda sc, par: (sc.pos, sc.scale, sc.rot, par, sc.flags, sc.group, sc.sphere)
    with capi.SceneHandle(max_nodes=base.n + 64) as h:
        h.upload(rs)
        with pytest.raises(capi.RNError) as e:      # a priority in the middle of the list: full upload needed
            h.add_nodes(*args(one, np.array([capi.RN_NO_PARENT], np.uint32)), priority=np.array([sg.RENDER_LATE - 1], np.int32))
        assert e.value.status == -7
        big = sg.make_scene(n_trees=1, branching=30, depth=3, seed=2)          # 931 nodes: does not fit a tile
        with pytest.raises(capi.RNError):
            h.add_nodes(*args(big, np.where(big.parent >= 0, big.parent, capi.RN_NO_PARENT).astype(np.uint32)))
        with pytest.raises(capi.RNError):
            h.remove_nodes(np.array([base.n + 5], np.uint32))                  # not a node
        slots = h.add_nodes(*args(one, np.array([capi.RN_NO_PARENT], np.uint32)), priority=np.array([sg.RENDER_EARLY], np.int32))
        h.remove_nodes(slots)
        with pytest.raises(capi.RNError):
            h.remove_nodes(slots)                                              # already a hole
        h.set_list_mode(True, 0)
        with pytest.raises(capi.RNError):
            h.add_nodes(*args(one, np.array([capi.RN_NO_PARENT], np.uint32)), priority=np.array([sg.RENDER_EARLY], np.int32))
