"""Diagnostic (N>1, torchrun): raw all_gather time and the effect of the pipelined exchange on the frame loop."""
import datetime, os, sys, time
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, ".")
from rayne_b200 import capi, scenegen as sg, shard
import bench

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
opts = dist.ProcessGroupNCCL.Options(); opts.is_high_priority_stream = True
dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120), pg_options=opts)
def log(*a):
    if rank == 0: print(*a, flush=True)
# raw collective
for mb in (1, 16, 77):
    src = torch.empty(mb * 250000, dtype=torch.int32, device="cuda"); dst = torch.empty(world * src.numel(), dtype=torch.int32, device="cuda")
    for _ in range(3): dist.all_gather_into_tensor(dst, src)
    torch.cuda.synchronize(); a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): dist.all_gather_into_tensor(dst, src)
    b.record(); torch.cuda.synchronize()
    log(f"all_gather {mb} MB/rank: {a.elapsed_time(b) / 10:.3f} ms")
scene = sg.make_scene(n_trees=768, branching=4, depth=8, seed=12345 + rank)
fr = capi.build_frustums(sg.six_view_rig()); ncam = len(fr)
rs = capi.RankedScene(scene); n = scene.n; del scene
h = capi.SceneHandle(max_nodes=n, device=local); h.upload(rs); h.set_list_mode(True, rank * n)
stream = torch.cuda.current_stream().cuda_stream
lists = h.transform_cull(fr, stream); h.read_counts(stream)
count_t = torch.as_tensor(bench.DevArray(lists[0].count, ncam), device="cuda")
list_t = [torch.as_tensor(bench.DevArray(lists[c].slots, n), device="cuda") for c in range(ncam)]
def loop(name, fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(reps): fn()
    b.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    log(f"{name}: {a.elapsed_time(b) / reps:.3f} ms/frame (host enqueue {(t1 - t0) * 1e3 / reps:.3f} ms)")
loop("compute only", lambda: h.transform_cull(fr, stream))
count_t = torch.as_tensor(bench.DevArray(lists[0].count, ncam), device="cuda")
for depth in (1, 2, 3):
    pex = shard.PipelinedListExchange(depth=depth)
    loop(f"compute + pipelined exchange depth {depth}", lambda: (h.transform_cull(fr, stream), pex.submit(count_t, list_t)))
    pex.flush()
    log("overflows", pex.overflows, "sizes", pex.sizes)
torch.cuda.synchronize(); dist.barrier(); dist.destroy_process_group()
