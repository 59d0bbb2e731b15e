#!/usr/bin/env python
"""bench.py - transform + cull throughput of librn_cuda on the north-star workload (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--trees T]

Workload (config C4): a synthetic forest of complete 4-ary depth-8 trees (21 845 nodes per tree), 768 trees =
16 776 960 SceneNodes PER GPU (weak scaling), 6 views (2 eye cameras + 4 shadow-cascade cameras).  A step is one
frame: every root moved by (0.01, 0, 0) -> world-transform propagation of every node -> 6-view frustum cull ->
ordered visibility lists.  `value` times the frame with all inputs resident in HBM; `e2e` times the same frame
through the C ABI with host buffers (dirty roots uploaded from pinned memory, counts + lists downloaded to pinned
memory).  One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from rayne_b200 import scenegen as sg  # noqa: E402

NODES_PER_TREE = 21845
BYTES_PER_NODE = 168.0          # 64 B read + 104 B written per node
METRIC = "Mnodes/s transform+cull"
CPU_SAMPLE_TREES = 12           # bounded CPU sample: 12 trees = 262 140 nodes x 6 views


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak() -> tuple[float, str]:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.index = index
        self.rows: list[list[str]] = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def wait_first(self, timeout: float = 5.0) -> None:
        t0 = time.time()
        while self.proc and not self.rows and time.time() - t0 < timeout:
            time.sleep(0.01)

    def mark(self) -> int:
        return len(self.rows)

    def stop(self, lo: int = 0, hi: int | None = None) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows[lo:hi]:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for k, nm in enumerate(names):
                    if r[3 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def ncu_traffic(nodes: int, views: int):
    """dram__bytes_read + dram__bytes_write of the tile kernel from the committed ncu capture (same workload only)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            t = json.load(f)["tile_kernel"]
        if t["nodes"] == nodes and t["views"] == views:
            return t["dram_bytes_read"] + t["dram_bytes_write"], t["source"]
    except Exception:
        pass
    return None, None


def build_scene(trees: int, seed: int):
    t0 = time.time()
    scene = sg.make_scene(n_trees=trees, branching=4, depth=8, seed=seed)
    log(f"[bench] scene: {scene.n} nodes generated in {time.time() - t0:.1f}s")
    return scene


def cpu_reference(trees: int, frames: int, warmup: int) -> dict:
    """Times the reference's own CPU path (oracle/_ref/ref_harness_release = unmodified Rayne sources) on a bounded
    sample of the workload; falls back to the C restatement (kind "port") if the harness binary is absent."""
    from oracle import refio
    scene = sg.make_scene(n_trees=trees, branching=4, depth=8, seed=12345)
    cams = sg.six_view_rig()
    sample = f"{trees} of 768 trees ({scene.n} nodes) x 6 views, {frames} frames after {warmup} warm-up"
    cores = os.cpu_count() or 1
    if refio.have_harness("release"):
        r = refio.bench_scene(scene, cams, frames, warmup)
        ms = r["ms_mean"]
        return {"value": scene.n / ms / 1e3, "unit": "Mnodes/s", "cores": cores, "kind": "reference",
                "threads_used": "update batches of 8192 nodes on the WorkQueue pool, cull loop single-threaded (RNSceneBasic.cpp:62-105,576-593)",
                "sample": sample, "ms_per_frame": ms, "ms_best": r["ms_best"], "visible": r["visible"]}
    from oracle import oracle
    fr = oracle.frustums(cams)
    oracle.scene_reference(scene, fr)
    t0 = time.perf_counter()
    for _ in range(max(1, frames // 4)):
        oracle.scene_reference(scene, fr)
    ms = (time.perf_counter() - t0) * 1e3 / max(1, frames // 4)
    return {"value": scene.n / ms / 1e3, "unit": "Mnodes/s", "cores": 1, "kind": "port", "sample": sample, "ms_per_frame": ms}


class DevArray:
    """Wraps a raw device pointer for torch.as_tensor (plumbing for the NCCL allgather only)."""

    def __init__(self, ptr: int, n: int, typestr: str = "<i4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist
    from rayne_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: librn_cuda has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        import datetime
        # a collective that does not complete within two minutes is a bug here, not a slow link: fail fast
        # NCCL's stream gets a high priority: the gather of frame i must be scheduled AHEAD of the thousands of pending
        # thread blocks of frame i+1's transform kernel, or it only starts in that kernel's tail
        opts = dist.ProcessGroupNCCL.Options()
        opts.is_high_priority_stream = True
        if args.nccl_ctas > 0:
            # the gather is small and hides under the next frame: a few CTAs are enough, and every CTA NCCL takes is an
            # SM slot the transform kernel of the next frame does not get
            opts.config.max_ctas = args.nccl_ctas
            opts.config.min_ctas = min(args.nccl_ctas, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120),
                                pg_options=opts)

    trees = args.trees
    scene = build_scene(trees, seed=12345 + rank)
    cams = sg.six_view_rig()
    fr = capi.build_frustums(cams)
    ncam = len(fr)
    t0 = time.time()
    rs = capi.RankedScene(scene)
    h = capi.SceneHandle(max_nodes=scene.n, device=local, max_emitters=args.emitters, max_particles_per_emitter=4000,
                         tile_nodes=args.tile_nodes, tile_inner_nodes=args.inner_nodes, tile_unit_nodes=args.unit_nodes)
    h.upload(rs)
    st = h.stats()
    log(f"[bench] rank {rank}: upload+tiling {time.time() - t0:.1f}s tiles={st.tiles} spine={st.spine_nodes} levels={st.spine_levels}")
    stream = torch.cuda.current_stream().cuda_stream
    n = scene.n
    del scene

    if world > 1:
        h.set_list_mode(True, rank * n)   # lists travel as global render-list ranks; rank r owns [r*n, (r+1)*n)
    lists = h.transform_cull(fr, stream)
    counts = h.read_counts(stream)
    vis_frac = [float(c) / n for c in counts]
    exch = None
    use_masks = False
    use_multicast = False
    if world > 1:
        from rayne_b200 import shard
        use_masks = args.exchange == "masks"
        count_t = torch.as_tensor(DevArray(lists[0].count, ncam), device="cuda")
        if use_masks:
            # compact form: rank-order visibility bitmasks (rn_cull_build_masks), 1 bit per node and view
            h.build_masks(stream)
            mptr, mwords, _ = h.device_masks()
            mask_t = torch.as_tensor(DevArray(mptr, ncam * mwords), device="cuda")
            pexch = None
            if args.exchange_transport == "multicast":
                try:
                    pexch = shard.MulticastMaskExchange([mwords] * world, ncam)
                except Exception as e:                      # no NVLS / no symmetric memory: say so and use NCCL
                    log(f"[bench] rank {rank}: multicast exchange unavailable ({type(e).__name__}: {e}); using NCCL")
                # every rank must take the same path: agree on the minimum
                okt = torch.tensor([1 if pexch is not None else 0], dtype=torch.int32, device="cuda")
                dist.all_reduce(okt, op=dist.ReduceOp.MIN)
                if int(okt.item()) == 0:
                    pexch = None
            use_multicast = pexch is not None
            if pexch is None:
                pexch = shard.PipelinedMaskExchange([mwords] * world, ncam)
        else:
            exch = shard.ListExchange()
            pexch = shard.PipelinedListExchange()
            list_t = [torch.as_tensor(DevArray(lists[c].slots, n), device="cuda") for c in range(ncam)]

    def submit_masks():
        if use_multicast:
            # the mask kernel stores through the NVLS multicast address: building the masks IS the gather
            h.build_masks_multicast(pexch.begin(), pexch.row, stream)
        else:
            # the masks go straight into the send buffer of this frame's NCCL gather (no staging copy)
            h.build_masks_to(pexch.begin().data_ptr(), pexch.row, stream)
        pexch.submit()

    def exchange():
        # the path's only collective, synchronous form: all-gather of the visibility of every view
        if use_masks:
            submit_masks()
            return pexch.flush()
        return exch.exchange(count_t, list_t)

    def step():
        # every frame's visibility IS gathered on every GPU; the gather of frame i runs on a side stream under the
        # kernels of frame i+1 (two frames in flight) and the last ones are flushed inside the timed region
        h.transform_cull(fr, stream)
        if world > 1:
            if use_masks:
                submit_masks()
            else:
                pexch.submit(count_t, list_t)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = h.stats().kernel_launches
    h.profile(True)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    barrier()
    m0 = sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    if world > 1:
        pexch.flush()
    e1.record()
    barrier()
    m1 = sampler.mark()
    ms_total = e0.elapsed_time(e1)
    tile_ms, tile_n = h.profile_read()
    h.profile(False)
    launches = h.stats().kernel_launches - launches0
    clocks = None
    if rank == 0 and m1 - m0 >= 3:
        clocks = sampler.stop(m0, m1)
        clocks["sampled"] = "during the timed region"
    # a timed region shorter than a few nvidia-smi periods: keep running the same step (untimed) under the sampler
    # rank 0 decides - and how many steps - for everybody: every step holds collectives, so the count must agree
    reps0 = max(1, int(400.0 / max(ms_total / args.steps, 1e-3))) if (rank == 0 and clocks is None) else 0
    need_ext = torch.tensor([reps0], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.broadcast(need_ext, 0)
    reps = int(need_ext.item())
    if reps > 0:
        m0 = sampler.mark()
        for _ in range(reps):
            step()
        barrier()
        if rank == 0:
            clocks = sampler.stop(m0, sampler.mark())
            clocks["sampled"] = f"{reps} more untimed steps of the same loop right after the timed region (it lasted only {ms_total:.1f} ms)"
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps

    # the gathered masks expand to the global lists: count check against the all-reduced counts, and the cost of
    # materialising all views' lists on one GPU (outside the timed region; a consumer expands what it draws)
    exchange_info = None
    if world > 1 and use_masks:
        gathered = pexch.flush()
        tot = count_t.clone().to(torch.int64)
        dist.all_reduce(tot)
        tot = tot.cpu().tolist()
        starts = [r * n for r in range(world)]
        out_l = torch.empty(max(tot) + 1, dtype=torch.int32, device="cuda")
        out_c = torch.zeros(1, dtype=torch.int32, device="cuda")
        ok = True
        for c in range(ncam):
            pexch.expand(h, gathered, c, starts, out_l, out_c, stream)
            k = int(out_c.item())
            asc = bool((out_l[1:k] > out_l[:k - 1]).all().item()) if k > 1 else True
            ok = ok and k == tot[c] and asc
        x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        x0.record()
        for _ in range(5):
            for c in range(ncam):
                pexch.expand(h, gathered, c, starts, out_l, out_c, stream)
        x1.record()
        torch.cuda.synchronize()
        exchange_info = {"kind": ("rank-order visibility bitmasks stored through the NVLS multicast address by the mask kernel itself "
                                  "(rn_cull_build_masks_multicast) + a symmetric-memory barrier, two frames in flight"
                                  if use_multicast else
                                  "NCCL all-gather of rank-order visibility bitmasks (rn_cull_build_masks), fixed size, "
                                  "two frames in flight on a side stream"),
                         "transport": "nvls-multicast" if use_multicast else "nccl",
                         "bytes_per_gpu_and_frame": int(ncam * mwords * 4), "global_list_entries_per_view": tot,
                         "expanded_lists_match_counts_and_ascend": ok,
                         "expand_all_views_ms": x0.elapsed_time(x1) / 5,
                         "expand_note": "rn_masks_expand of all views' global lists on one GPU; not part of the step"}
        del out_l
    elif world > 1:
        exchange_info = {"kind": "NCCL all-gather of per-camera counts + index lists, two frames in flight on a side stream",
                         "overflows": int(pexch.overflows), "entries_per_rank_and_camera": list(pexch.sizes or [])}

    # ---- end to end through the C ABI with host buffers (pinned) ----
    roots = rs.roots
    rootpos = rs.position[roots].copy()
    pinned = [torch.empty(max(int(c * 1.05) + 1024, 1024), dtype=torch.int32).pin_memory() for c in counts]
    pin_ptrs = [p.data_ptr() for p in pinned]
    pin_caps = [p.numel() for p in pinned]
    d2h = 0

    # One GPU: the frame is pipelined over S rank ranges on S streams (rayne_b200/pipeline.py) so that the PCIe copy of
    # range k's lists overlaps the kernels of range k+1.  With several GPUs the lists are gathered on the devices
    # (NCCL) from the single per-GPU scene, so the unpipelined frame is used there.
    shards = args.e2e_shards if world == 1 else 1
    pipe = None
    if shards > 1:
        from rayne_b200 import pipeline
        torch.cuda.synchronize()
        pipe = pipeline.PipelinedScene(rs, shards, device=local, tile_nodes=args.tile_nodes,
                                       tile_inner_nodes=args.inner_nodes, tile_unit_nodes=args.unit_nodes)

    def e2e_step():
        nonlocal d2h
        rootpos[:, 0] += np.float32(0.01)
        if pipe is not None:
            got = pipe.frame(fr, pin_ptrs, pin_caps, dirty_ranks=roots, position=rootpos)
            d2h = int(got.sum()) * 4 + ncam * 4 * shards
            return
        h.update_nodes(roots, position=rootpos)
        h.transform_cull(fr, stream)
        got = h.read_lists_into(pin_ptrs, pin_caps, stream)
        d2h = int(got.sum()) * 4 + ncam * 4
        if world > 1:
            exchange()

    for _ in range(min(3, args.warmup)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / args.steps
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    # the same frame with the visibility read back as rank-order bitmasks (1 bit per node and view instead of 4 B per
    # visible node): the compact form of the same result, 12.6 MB instead of 61 MB over PCIe
    e2e_masks = None
    if world > 1 and use_masks:
        # several GPUs: the same frame with the compact read-back (one scene per GPU, unpipelined) + the mask gather
        mw = (n + 31) // 32
        pinned_masks = torch.empty(ncam * mw, dtype=torch.int32).pin_memory()

        def e2e_mask_step_n():
            rootpos[:, 0] += np.float32(0.01)
            h.update_nodes(roots, position=rootpos)
            h.transform_cull(fr, stream)
            submit_masks()
            h.build_masks(stream)
            h.read_masks_into(pinned_masks.data_ptr(), ncam, stream)
            pexch.flush()
            torch.cuda.current_stream().synchronize()

        for _ in range(min(3, args.warmup)):
            e2e_mask_step_n()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_mask_step_n()
        barrier()
        mms = (time.perf_counter() - t0) * 1e3 / args.steps
        t = torch.tensor([mms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        mms = float(t.item())
        e2e_masks = {"value": n * world / mms / 1e3, "unit": "Mnodes/s", "ms_per_step": mms,
                     "h2d_bytes_per_step": int(roots.shape[0] * 16), "d2h_bytes_per_step": int(ncam * mw * 4),
                     "what": "per GPU: rn_scene_update_nodes + rn_transform_cull + rn_cull_build_masks + "
                             "rn_cull_read_masks_async into pinned host memory, and the all-gather of the masks"}
    if pipe is not None:
        offs, total_words = pipe.mask_layout(ncam)
        pinned_masks = torch.empty(total_words, dtype=torch.int32).pin_memory()

        def e2e_mask_step():
            rootpos[:, 0] += np.float32(0.01)
            pipe.frame_masks(fr, pinned_masks.data_ptr(), dirty_ranks=roots, position=rootpos)

        for _ in range(min(3, args.warmup)):
            e2e_mask_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_mask_step()
        barrier()
        mms = (time.perf_counter() - t0) * 1e3 / args.steps
        # same visibility as the lists of the last list frame?  (the roots moved 0.01 per frame in between, so compare
        # the population counts of a fresh pair of frames on the same positions)
        pipe.frame(fr, pin_ptrs, pin_caps, dirty_ranks=roots, position=rootpos)
        got_lists = pipe.frame(fr, pin_ptrs, pin_caps, dirty_ranks=roots, position=rootpos)
        pipe.frame_masks(fr, pinned_masks.data_ptr(), dirty_ranks=roots, position=rootpos)
        mk = pinned_masks.numpy().view(np.uint32)
        pop = np.zeros(ncam, dtype=np.int64)
        for k, (a, b) in enumerate(pipe.ranges):
            w = (b - a + 31) // 32
            blk = mk[offs[k]: offs[k] + ncam * w].reshape(ncam, w)
            pop += np.unpackbits(blk.view(np.uint8), axis=1).sum(axis=1, dtype=np.int64)
        e2e_masks = {"value": n / mms / 1e3, "unit": "Mnodes/s", "ms_per_step": mms,
                     "h2d_bytes_per_step": int(roots.shape[0] * 16), "d2h_bytes_per_step": int(total_words * 4),
                     "what": "as e2e, but rn_cull_build_masks + rn_cull_read_masks_async: the views' visibility as "
                             "rank-order bitmasks in pinned host memory",
                     "set_bits_equal_list_counts": bool(np.array_equal(pop, np.asarray(got_lists, dtype=np.int64)))}
        pipe.close()

    # ---- particles (config C3 on rank 0's GPU share; reported beside the headline) ----
    particles = None
    if args.emitters > 0:
        em = sg.make_emitters(args.emitters, seed_base=1000 + rank * args.emitters)
        h.particles_configure(em)
        for _ in range(240):
            h.particles_step(1.0 / 60.0, stream)
        torch.cuda.synchronize()
        live = int(h.particles_counts()[0].sum())
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        for _ in range(30):
            h.particles_step(1.0 / 60.0, stream)
        p1.record()
        torch.cuda.synchronize()
        pms = p0.elapsed_time(p1) / 30
        peak, _ = measured_peak()
        particles = {"metric": "Mparticles/s", "value": live * world / pms / 1e3, "live_per_gpu": live, "ms_per_step": pms,
                     "roofline_frac": live * 140.0 / (pms * 1e-3) / 1e9 / peak, "bytes_per_particle": 140}

    if rank != 0:
        if world > 1:
            dist.barrier()          # rank 0 is still timing its side kernels on its GPU
            dist.destroy_process_group()
        return {}
    peak, peak_kind = measured_peak()

    # ---- the rows next to the headline (§8(f) P6): kernel times on rank 0 ----
    def timed(fn, reps=10):
        reps = max(1, min(reps, args.extras_reps))
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    extras = None
    if not args.no_extras:
        extras = {}
        if world > 1:
            h.set_list_mode(False, 0)
        h.transform_cull(fr, stream)
        ms = timed(lambda: h.transform_inverse(stream))
        extras["transform_inverse"] = {"ms": ms, "bytes_per_node": 112, "roofline_frac": n * 112.0 / (ms * 1e-3) / 1e9 / peak}
        ms = timed(lambda: h.bounds_update(stream))
        extras["bounds_update"] = {"ms": ms, "bytes_per_node": 96, "roofline_frac": n * 96.0 / (ms * 1e-3) / 1e9 / peak}
        fr_sort = fr.copy()
        fr_sort["flags"][:] = 0
        fr_sort["flags"][ncam - 1] = capi.CAM_SORT_FRONT_TO_BACK
        h.transform_cull(fr_sort, stream)
        entries = int(h.read_counts(stream)[ncam - 1])
        campos = np.ascontiguousarray(cams["pos"], dtype=np.float32)
        ms = timed(lambda: h.cull_sort(campos, stream), reps=5)
        # 4 radix passes x (key + entry read by the histogram... counted once: read 8 + written 8; the last pass writes
        # the entry only) + key build (entry + world position in, key + entry out)
        sort_bytes = entries * (4 * 16.0 - 4.0 + 28.0)
        extras["cull_sort_front_to_back"] = {"ms": ms, "entries": entries, "Mentries_per_s": entries / ms / 1e3,
                                             "roofline_frac": sort_bytes / (ms * 1e-3) / 1e9 / peak}
        # visibility as rank-order bitmasks (the multi-GPU exchange format) and back: 5 B + V/8 B per node; expansion of
        # the largest view reads its words twice and writes 4 B per entry
        ms = timed(lambda: h.build_masks(stream))
        extras["cull_build_masks"] = {"ms": ms, "bytes_per_node": 5 + ncam / 8.0,
                                      "roofline_frac": n * (5 + ncam / 8.0) / (ms * 1e-3) / 1e9 / peak}
        mptr_x, mwords_x, _ = h.device_masks()
        xl = torch.empty(max(entries, 1), dtype=torch.int32, device="cuda")
        xc = torch.zeros(1, dtype=torch.int32, device="cuda")

        def expand_largest():
            xc.zero_()
            h.masks_expand([(mptr_x + 4 * (ncam - 1) * mwords_x, mwords_x, 0)], xl.data_ptr(), xl.numel(), xc.data_ptr(), stream)

        ms = timed(expand_largest)
        extras["masks_expand_largest_view"] = {"ms": ms, "entries": int(xc.item()),
                                               "roofline_frac": (mwords_x * 8.0 + entries * 4.0) / (ms * 1e-3) / 1e9 / peak}
        del xl
        # §8(f) drawable uniform packing: 256-byte matrix records of the largest list
        view, _, projview = capi.build_camera_matrices(cams)
        inst = torch.empty((max(entries, 1), 64), dtype=torch.float32, device="cuda")
        ms = timed(lambda: h.pack_instances(ncam - 1, view[ncam - 1], projview[ncam - 1], inst.data_ptr(), entries, stream))
        extras["cull_pack_instances"] = {"ms": ms, "entries": entries, "bytes_per_entry": 388,
                                         "roofline_frac": entries * 388.0 / (ms * 1e-3) / 1e9 / peak}
        del inst
        if particles:
            ms = timed(lambda: h.particles_build_mesh(stream=stream))
            extras["particles_build_mesh"] = {"ms": ms, "bytes_per_particle": 236,
                                              "roofline_frac": particles["live_per_gpu"] * 236.0 / (ms * 1e-3) / 1e9 / peak}
            # vertex stage of Shaders/Particles.hlsl over that mesh: 36 B read + 32 B written per vertex
            ident = np.tile(np.eye(4, dtype=np.float32).reshape(16), (args.emitters, 1))
            ms = timed(lambda: h.particles_expand_billboards(ident, ident, stream=stream))
            extras["particles_expand_billboards"] = {"ms": ms, "bytes_per_particle": 272,
                                                     "roofline_frac": particles["live_per_gpu"] * 272.0 / (ms * 1e-3) / 1e9 / peak}
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
    traffic, traffic_src = ncu_traffic(n, ncam)
    alg_bytes = n * (BYTES_PER_NODE + 4.0 * sum(vis_frac))
    tile_avg_ms = tile_ms / max(tile_n, 1)
    achieved = alg_bytes / (tile_avg_ms * 1e-3) / 1e9
    out = {
        "metric": METRIC, "value": n * world / ms_step / 1e3, "unit": "Mnodes/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"C4: {n} SceneNodes per GPU ({trees} complete 4-ary depth-8 trees), 6-view cull "
                               "(2 eye + 4 shadow-cascade cameras), every node re-propagated each frame",
                   "nodes_per_gpu": n, "views": ncam, "visible_fraction_per_view": [round(v, 5) for v in vis_frac],
                   "l2": "inputs (1.07 GB of node columns) exceed the 126 MB L2; no flush needed",
                   "parallelism": (f"node-range sharding x{world}, " + ("visibility bitmasks multicast through NVLS by the mask kernel" if use_multicast else
                                                                          "NCCL all-gather of " + ("visibility bitmasks" if args.exchange == "masks" else "counts + lists"))) if world > 1 else "single GPU"},
        "e2e": {"value": n * world / e2e_ms / 1e3, "unit": "Mnodes/s", "ms_per_step": e2e_ms,
                "h2d_bytes_per_step": int(roots.shape[0] * 16), "d2h_bytes_per_step": int(d2h),
                "what": ("rn_scene_update_nodes(dirty roots, pinned staging) + rn_transform_cull + rn_cull_read_lists (counts + 6 lists) "
                         "into pinned host buffers" + (f"; frame pipelined over {shards} rank ranges on {shards} streams "
                                                       "(rayne_b200/pipeline.py), every range's copy inside the timed region" if shards > 1 else "")),
                "shards": shards},
        "e2e_masks": e2e_masks,
        "gpu_launches": int(launches),
        "exchange": exchange_info,
        "roofline": {"bound": "hbm", "kernel": "rn::tile_kernel<true,true>", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "peak_kind": peak_kind, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": alg_bytes, "kernel_ms": tile_avg_ms,
                     "kernel_share_of_step": tile_avg_ms / ms_step,
                     "step_frac": alg_bytes / (ms_step * 1e-3) / 1e9 / peak},
        "clocks": clocks,
    }
    if particles:
        out["particles"] = particles
    if extras:
        out["extras"] = extras
    if not args.no_cpu:
        t0 = time.time()
        out["cpu_baseline"] = cpu_reference(CPU_SAMPLE_TREES, 6, 2)
        log(f"[bench] cpu baseline took {time.time() - t0:.1f}s")
    return out


def run_reference(args) -> dict:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # bounded sample: keep the whole --steps/--warmup run near two minutes (the reference does ~0.9 Mnodes/s here)
    frames = max(1, args.steps) + max(0, args.warmup)
    trees = int(min(CPU_SAMPLE_TREES, max(2, 4500 // frames)))
    r = cpu_reference(trees, max(1, args.steps), max(0, args.warmup))
    v = r["value"]
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mnodes/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_frame"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C4 sample: " + r["sample"] + " on the host CPU (SceneBasic::Update + Render, unmodified Rayne sources)"},
            "cpu_baseline": r, "gpu_launches": 0,
            "e2e": {"value": v, "unit": "Mnodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--trees", type=int, default=768, help="trees per GPU (768 = 16.8 M nodes)")
    ap.add_argument("--emitters", type=int, default=1000, help="particle emitters per GPU for the side metric (0 = skip)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--e2e-shards", type=int, default=4, help="rank ranges / streams of the pipelined end-to-end frame (1 = off)")
    ap.add_argument("--exchange-transport", choices=("nccl", "multicast"), default="multicast",
                    help="N>1, masks: multimem stores of the mask kernel through NVLS multicast (falls back to NCCL where "
                         "symmetric memory / NVLS is unavailable), or the NCCL all-gather")
    ap.add_argument("--nccl-ctas", type=int, default=0, help="N>1: cap on NCCL's CTAs per collective (0 = NCCL's default)")
    ap.add_argument("--exchange", choices=("masks", "lists"), default="masks",
                    help="N>1: what the per-frame all-gather ships (rank-order bitmasks, or counts + index lists)")
    ap.add_argument("--extras-reps", type=int, default=10, help="timed repetitions of each side kernel (1 for an ncu capture)")
    ap.add_argument("--no-extras", action="store_true", help="skip the side timings of inverse / bounds / sort / mesh")
    ap.add_argument("--tile-nodes", type=int, default=0, help="rn_scene_desc.tile_nodes (0 = library default)")
    ap.add_argument("--inner-nodes", type=int, default=0, help="rn_scene_desc.tile_inner_nodes (0 = library default)")
    ap.add_argument("--unit-nodes", type=int, default=0, help="rn_scene_desc.tile_unit_nodes (0 = library default)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    # stdout carries exactly ONE line, the JSON: libraries that print to fd 1 (NCCL's version banner) go to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    os.close(real_stdout)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
