"""Diagnostic for rn_cull_sort at 1 M nodes: where do the sorted keys stop being monotone, and is it the sort or the keys?"""
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle import oracle  # noqa: E402
from rayne_b200 import capi, scenegen as sg  # noqa: E402
from tests.helpers import sort_cameras  # noqa: E402

scene = sg.make_scene(n_trees=48, branching=4, depth=8, seed=12345, priorities=True, transparent_frac=0.05)
cams = sort_cameras()[:1]
fr = capi.build_frustums(cams)
fr["flags"] |= sg.CAM_DEFAULTS
rs = capi.RankedScene(scene)
for rep in range(3):
    with capi.SceneHandle(max_nodes=scene.n) as h:
        h.upload(rs)
        h.transform_cull(fr)
        before = h.read_list(0)
        h.cull_sort(cams["pos"])
        after = h.read_list(0)
        ros = h.download(capi.DL_RANK_OF_SLOT)
        wpos_ranked = h.download(capi.DL_WORLD_POSITION)
    a, b = rs.order[ros[before]], rs.order[ros[after]]
    parent = np.where(scene.parent >= 0, scene.parent, 0xFFFFFFFF).astype(np.uint32)
    wpos = oracle.world(parent, scene.pos, scene.scale, scene.rot)[0]
    gw = np.zeros_like(wpos_ranked); gw[rs.order] = wpos_ranked
    print("rep", rep, "entries", len(a), "perm", np.array_equal(np.sort(a), np.sort(b)), "wpos equal", np.array_equal(gw.view(np.uint32), wpos.view(np.uint32)))
    pr = scene.priority[b]
    print("  priorities monotone", bool((np.diff(pr.astype(np.int64)) >= 0).all()), "distinct", np.unique(pr).tolist())
    d = wpos[b] - cams[0]["pos"]
    key = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2]
    for p in np.unique(pr):
        idx = np.nonzero(pr == p)[0]
        k = key[idx]
        if p < sg.RENDER_SKY:
            bad = np.nonzero(np.diff(k) < 0)[0]
        elif p == sg.RENDER_TRANSPARENT:
            bad = np.nonzero(np.diff(k[1:]) > 0)[0] + 1
        else:
            continue
        print(f"  priority {p}: {len(k)} entries, {len(bad)} violations")
        for j in bad[:8]:
            print("    at", int(idx[j]), "keys", k[j - 1:j + 3].tolist(), "bits", [hex(x) for x in k[j - 1:j + 3].view(np.uint32)])

# ---- which nodes carry a wrong world position, and is it their own computation or inherited? ----
print("---- transform check (3 fresh handles, strict and default)")
n = scene.n
par = scene.parent
depth = np.zeros(n, np.int32)
for i in range(n):                      # parents precede children in creation order
    if par[i] >= 0:
        depth[i] = depth[par[i]] + 1
rank_of = np.empty(n, np.int64); rank_of[rs.order] = np.arange(n)
for rep in range(3):
    for mode in ("cull", "propagate"):
        with capi.SceneHandle(max_nodes=scene.n) as h:
            h.upload(rs)
            if mode == "cull":
                h.transform_cull(fr)
            else:
                h.transform_propagate()
            ros = h.download(capi.DL_RANK_OF_SLOT)
            sor = h.download(capi.DL_SLOT_OF_RANK)
            wr = h.download(capi.DL_WORLD_POSITION)
            st = h.stats()
        gw = np.zeros_like(wr); gw[rs.order] = wr
        bad = np.nonzero((gw.view(np.uint32) != wpos.view(np.uint32)).any(axis=1))[0]
        own = bad[(par[bad] < 0) | ~np.isin(par[bad], bad)]
        print(f"rep {rep} {mode}: tiles {st.tiles} spine {st.spine_nodes} levels {st.spine_levels}; wrong {len(bad)}, wrong with a correct parent {len(own)}")
        for i in own[:12]:
            r = int(rank_of[i]); s = int(sor[r])
            pr = int(par[i])
            prk = int(rank_of[pr]) if pr >= 0 else -1
            print(f"    id {int(i)} depth {int(depth[i])} prio {int(scene.priority[i])} rank {r} slot {s} | parent id {pr} rank {prk} slot {int(sor[prk]) if pr >= 0 else -1} prio {int(scene.priority[pr]) if pr >= 0 else 0}"
                  f" | got {gw[i].tolist()} want {wpos[i].tolist()}")
