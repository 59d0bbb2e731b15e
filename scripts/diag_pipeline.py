"""Diagnostic: host-side timeline of the pipelined end-to-end frame (not part of the bench)."""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from rayne_b200 import capi, pipeline, scenegen as sg

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4
scene = sg.make_scene(n_trees=768, branching=4, depth=8, seed=12345)
cams = sg.six_view_rig(); fr = capi.build_frustums(cams)
rs = capi.RankedScene(scene); n = scene.n; del scene
pipe = pipeline.PipelinedScene(rs, S)
pinned = [torch.empty(n // 2, dtype=torch.int32).pin_memory() for _ in fr]
ptrs = [p.data_ptr() for p in pinned]; caps = [p.numel() for p in pinned]
roots = rs.roots; rootpos = rs.position[roots].copy()
# raw D2H bandwidth into the same pinned buffers
dev = torch.empty(n // 2, dtype=torch.int32, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for p in pinned: p.copy_(dev, non_blocking=True)
torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"raw D2H {6 * dev.numel() * 4 / 1e6:.1f} MB in {dt * 1e3:.3f} ms = {6 * dev.numel() * 4 / dt / 1e9:.1f} GB/s")
for it in range(6):
    torch.cuda.synchronize()
    t = [time.perf_counter()]
    pipe.update_nodes(roots, position=rootpos); t.append(time.perf_counter())
    pipe.transform_cull(fr); t.append(time.perf_counter())
    total = np.zeros(len(fr), dtype=np.int64)
    for h, st in zip(pipe.handles, pipe.streams):
        pp = [int(p) + 4 * int(x) for p, x in zip(ptrs, total)]
        cc = [int(c) - int(x) for c, x in zip(caps, total)]
        total += h.read_lists_into(pp, cc, st, wait=False); t.append(time.perf_counter())
    for st in pipe.streams:
        capi.stream_synchronize(st)
    t.append(time.perf_counter())
    print("iter", it, " ".join(f"{(b - t[0]) * 1e3:.3f}" for b in t[1:]), "entries", int(total.sum()))
# device-only time of the S shards enqueued back to back
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): pipe.transform_cull(fr)
torch.cuda.synchronize(); print(f"device only, {S} shards: {(time.perf_counter() - t0) * 100:.3f} ms per frame")
