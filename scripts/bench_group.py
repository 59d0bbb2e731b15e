"""rn_group (several GPUs, ONE process, C ABI only) on the benchmark scene: python scripts/bench_group.py <n_gpus> [trees] [steps]
Strong scaling: the fixed scene is cut into one rank range per GPU; a frame = rn_group_update_nodes (all roots) +
rn_group_frame (transform + cull + mask exchange fused into the mask kernel's stores); host wall clock over `steps` frames
with one rn_group_synchronize at the end.  Prints one JSON line per transport."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from rayne_b200 import capi, scenegen as sg  # noqa: E402

n_gpus = int(sys.argv[1]) if len(sys.argv) > 1 else 2
trees = int(sys.argv[2]) if len(sys.argv) > 2 else 768
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
scene = sg.config_scene("C4", trees=trees)
rs = capi.RankedScene(scene)
fr = capi.build_frustums(sg.vr_cascade_rig())
roots = rs.roots
pos = rs.position[roots].copy()
for transport, name in ((capi.GROUP_AUTO, "auto"), (capi.GROUP_PEER, "peer")):
    if n_gpus == 1 and transport == capi.GROUP_PEER:
        continue
    with capi.Group(rs, list(range(n_gpus)), transport=transport) as g:
        def frame():
            pos[:, 0] += np.float32(0.01)
            g.update_nodes(roots, position=pos)
            g.frame(fr)
        for _ in range(5):
            frame()
        g.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            frame()
        g.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        # the whole scene's list of the largest view on the last GPU equals the sum of what the single GPUs report
        n_last = len(g.read_list(n_gpus - 1, len(fr) - 1))
        print(json.dumps({"what": "rn_group frame (one process)", "gpus": n_gpus, "nodes": rs.n, "views": len(fr), "transport": g.note,
                          "requested": name, "ms_per_frame": ms, "Mnodes_per_s": rs.n / ms / 1e3, "largest_view_entries_on_last_gpu": n_last,
                          "bounds": g.bounds.tolist()}), flush=True)
