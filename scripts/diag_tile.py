"""Diagnostic: world positions of the 1 M-node priorities scene after rn_transform_propagate / rn_transform_cull, fresh handle per
repetition; prints how many nodes differ from the oracle and the slot runs they form."""
import os
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle import oracle  # noqa: E402
from rayne_b200 import capi, scenegen as sg  # noqa: E402
from tests.helpers import sort_cameras  # noqa: E402

scene = sg.make_scene(n_trees=48, branching=4, depth=8, seed=12345, priorities=True, transparent_frac=0.05)
fr = capi.build_frustums(sort_cameras()[:1])
rs = capi.RankedScene(scene)
parent = np.where(scene.parent >= 0, scene.parent, 0xFFFFFFFF).astype(np.uint32)
wpos = oracle.world(parent, scene.pos, scene.scale, scene.rot)[0]
tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("RN_TILE"))
for mode in ("propagate", "cull"):
    counts = []
    runs_txt = ""
    for rep in range(4):
        with capi.SceneHandle(max_nodes=scene.n) as h:
            h.upload(rs)
            if mode == "cull":
                h.transform_cull(fr)
            else:
                h.transform_propagate()
            sor = h.download(capi.DL_SLOT_OF_RANK)
            wr = h.download(capi.DL_WORLD_POSITION)
        want = wpos[rs.order]
        bad_rank = np.nonzero((wr.view(np.uint32) != want.view(np.uint32)).any(axis=1))[0]
        counts.append(len(bad_rank))
        if rep == 0 and len(bad_rank):
            slots = np.sort(sor[bad_rank])
            brk = np.nonzero(np.diff(slots) != 1)[0]
            starts = np.concatenate([[0], brk + 1]); ends = np.concatenate([brk, [len(slots) - 1]])
            runs_txt = " runs(slot,len): " + " ".join(f"({int(slots[a])},{int(b - a + 1)})" for a, b in list(zip(starts, ends))[:10])
    print(f"[{tag}] {mode}: wrong per rep {counts}{runs_txt}", flush=True)
