#!/usr/bin/env python
"""bench.py - transform + cull (and particle) throughput of librn_cuda on the north-star workloads (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--config C1|C2|C3|C4|C4-strong|C5] [--verify-full]

Configs (BASELINE.json `configs`); the default - and what the driver runs - is C4:
  C1         10 x (1+9+90+900) = 10 000 SceneNodes, depth 4, 1 camera           (the reference's own CPU-sized case)
  C2         48 complete 4-ary depth-8 trees = 1 048 560 nodes, 1 camera, one GPU
  C3         1 000 ParticleEmitters x 4 000 particles: step + depth sort per frame  (metric Mparticles/s)
  C4         768 trees = 16 776 960 nodes PER GPU (weak scaling), 6 views (2 eye + 4 shadow-cascade cameras)
  C4-strong  the same fixed 16 776 960-node scene cut over the N GPUs by shard.plan_shards (strong scaling)
  C5         3 072 trees = 67 107 840 nodes + 4 000 emitters x 4 000 particles, one shared scene cut over the N GPUs

A step is one frame: (every root moved by (0.01, 0, 0) in the end-to-end loop) -> world-transform propagation of every
node -> frustum cull for all views -> ordered visibility lists [-> at N > 1 the exchange of every view's visibility
between the GPUs].  `value` times the frame with all inputs resident in HBM; `e2e` times the same frame through the
C ABI with host buffers (dirty roots uploaded from pinned memory, counts + lists downloaded to pinned memory).
Every run carries a parity gate (`parity`): the GPU result restricted to the first trees of the scene is compared
with the reference / the C oracle run on exactly those trees.  One JSON line is printed by rank 0.
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
TILE_BYTES_PER_NODE = 168.0     # the tile kernel reads 64 B and writes 104 B per node
METRIC = "Mnodes/s transform+cull"
CPU_SAMPLE_TREES = 12           # bounded CPU sample of the depth-8 configs: 12 trees = 262 140 nodes
PARTICLE_BYTES = 140.0          # particles_step_kernel: algorithmic bytes per live particle and step
CPU_SAMPLE_EMITTERS = 100       # bounded CPU sample of the particle configs: 100 emitters x 4 000

CONFIGS = {
    "C1": dict(trees=10, branching=(9, 10, 10), depth=4, views=1, scaling="weak"),
    "C2": dict(trees=48, branching=4, depth=8, views=1, scaling="weak"),
    "C4": dict(trees=768, branching=4, depth=8, views=6, scaling="weak"),
    "C4-strong": dict(trees=768, branching=4, depth=8, views=6, scaling="strong"),
    "C5": dict(trees=3072, branching=4, depth=8, views=6, scaling="strong", emitters=4000),
}


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


RIG = "reference"


def config_cameras(cfg: dict) -> np.ndarray:
    """6 views: the two VR eye cameras + the four cascade cameras the reference's Light::UpdateShadows derives for a
    directional light with the default ShadowParameter and the head camera as target (rn_light_build_cascades, pinned by
    tests/golden/cascades_basic.npz).  --rig legacy: round 1's hand-parameterised cascade cameras (scenegen.six_view_rig)."""
    if cfg["views"] != 6:
        return np.atleast_1d(sg.make_camera())
    return sg.vr_cascade_rig() if RIG == "reference" else sg.six_view_rig()


def build_scene(cfg: dict, trees: int, seed: int):
    t0 = time.time()
    scene = sg.make_scene(n_trees=trees, branching=cfg["branching"], depth=cfg["depth"], seed=seed)
    log(f"[bench] scene: {scene.n} nodes generated in {time.time() - t0:.1f}s")
    return scene


def first_trees(scene, trees: int, cfg: dict):
    """The first `trees` trees of a forest scene as a scene of their own (trees are stored back to back in creation
    order, so this is a prefix of every column)."""
    per_tree = scene.n // max(1, int((scene.parent < 0).sum()))
    m = trees * per_tree
    return sg.Scene(parent=scene.parent[:m].copy(), pos=scene.pos[:m], scale=scene.scale[:m], rot=scene.rot[:m],
                    flags=scene.flags[:m], group=scene.group[:m], priority=scene.priority[:m], sphere=scene.sphere[:m])


# ------------------------------------------------------------------------------------------------------------------
# CPU arms: the reference's own code (oracle/_ref/ref_harness_release = unmodified Rayne sources) or the C restatement
# ------------------------------------------------------------------------------------------------------------------
def cpu_reference_scene(scene, cams, frames: int, warmup: int, what: str) -> dict:
    """Times the reference's CPU path on a bounded sample scene; falls back to the C restatement (kind "port")."""
    from oracle import refio
    sample = f"{what} ({scene.n} nodes) x {len(cams)} views, {frames} frames after {warmup} warm-up"
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
    reps = max(1, frames // 4)
    t0 = time.perf_counter()
    for _ in range(reps):
        oracle.scene_reference(scene, fr)
    ms = (time.perf_counter() - t0) * 1e3 / reps
    return {"value": scene.n / ms / 1e3, "unit": "Mnodes/s", "cores": 1, "kind": "port", "sample": sample, "ms_per_frame": ms}


def cpu_reference_particles(emitters: int, steps: int, warmup: int) -> dict:
    """The reference's GenericParticleEmitter::Update on a bounded sample of the emitters (same parameters, same seeds)."""
    from oracle import refio
    em = sg.make_emitters(emitters)
    sample = f"{emitters} emitters x 4000 particles, {steps} steps after {warmup} warm-up steps"
    if refio.have_harness("release"):
        r = refio.bench_particles(em, steps, warmup)
        return {"value": r["live"] / r["ms_mean"] / 1e3, "unit": "Mparticles/s", "cores": os.cpu_count() or 1, "kind": "reference",
                "threads_used": "one thread (ParticleEmitter::Update runs on the main thread, RNParticleEmitter.cpp:339-347)",
                "sample": sample, "ms_per_step": r["ms_mean"], "ms_best": r["ms_best"], "live": r["live"]}
    from oracle import oracle
    o = oracle.Emitters(em)
    o.step(np.float32(1 / 60), warmup)
    t0 = time.perf_counter()
    o.step(np.float32(1 / 60), steps)
    ms = (time.perf_counter() - t0) * 1e3 / steps
    live = int(sum(o.read(k)[0].shape[0] for k in range(emitters)))
    return {"value": live / ms / 1e3, "unit": "Mparticles/s", "cores": 1, "kind": "port", "sample": sample, "ms_per_step": ms, "live": live}


class DevArray:
    """Wraps a raw device pointer for torch.as_tensor (plumbing for the exchange and the parity gate only)."""

    def __init__(self, ptr: int, n: int, typestr: str = "<i4"):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}


# ------------------------------------------------------------------------------------------------------------------
# parity gate
# ------------------------------------------------------------------------------------------------------------------
def list_hash(a: np.ndarray) -> str:
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint32).tobytes()).hexdigest()[:16]


def parity_gate(h, rs, scene, cams, fr, cfg, sample_trees, stream, strict_flag: bool) -> dict:
    """Compares what the GPU holds for the first `sample_trees` trees of the scene (their visibility-list entries in
    list order, world position / rotation / scale and world matrices) with the reference run on exactly those trees
    (oracle/_ref/ref_harness_strict where present, the C oracle otherwise).  Trees are independent and keep their
    relative order in the render list, so the sub-lists must be identical.  sample_trees = None: the whole scene
    through the C oracle (--verify-full)."""
    from rayne_b200 import capi
    from oracle import oracle, refio
    from tests.helpers import matrices_close, same_floats
    t0 = time.time()
    total_trees = int((scene.parent < 0).sum())
    full = sample_trees is None or sample_trees >= total_trees
    sub = scene if full else first_trees(scene, sample_trees, cfg)
    m = sub.n
    use_harness = refio.have_harness("strict") and not full
    if use_harness:
        ref = refio.run_scene(sub, cams, kind="strict")
        kind = "reference (oracle/_ref/ref_harness_strict, unmodified Rayne sources)"
    else:
        ref = oracle.scene_reference(sub, oracle.frustums(cams))
        kind = "port (oracle/rn_oracle.c)"
    h.transform_cull(fr, stream)
    rank_of_slot = h.download(capi.DL_RANK_OF_SLOT)
    lists_equal, hashes = True, []
    for c in range(len(fr)):
        ids = rs.order[rank_of_slot[h.read_list(c, stream)]]         # creation ids in list order
        mine = ids[ids < m] if not full else ids
        ok = mine.shape == ref["lists"][c].shape and bool(np.array_equal(mine, ref["lists"][c]))
        lists_equal = lists_equal and ok
        hashes.append({"gpu": list_hash(mine), "ref": list_hash(ref["lists"][c]), "entries": int(ref["lists"][c].shape[0])})
    # world transforms of the sample: the sample's creation ids are ranks rs.rank_of[0:m]; download in rank order
    ranks = rs.rank_of[:m].astype(np.int64)
    lo, hi = int(ranks.min()), int(ranks.max()) + 1
    out = {}
    for key, what in (("wpos", capi.DL_WORLD_POSITION), ("wrot", capi.DL_WORLD_ROTATION), ("wscale", capi.DL_WORLD_SCALE),
                      ("wmat", capi.DL_WORLD_MATRIX)):
        out[key] = h.download(what, lo, hi - lo)[ranks - lo]
    trs_equal = all(same_floats(out[k], ref[k]) for k in ("wpos", "wrot", "wscale"))
    d = np.abs(out["wmat"].astype(np.float64) - ref["wmat"].astype(np.float64))
    extent = max(float(np.abs(ref["wmat"]).max()), 1.0)
    res = {"sample_trees": total_trees if full else sample_trees, "sample_nodes": m, "against": kind,
           "lists_equal": bool(lists_equal), "world_trs_bit_exact": bool(trs_equal),
           "matrices_bit_exact": bool(same_floats(out["wmat"], ref["wmat"])),
           "matrices_within_1e-5": bool(matrices_close(out["wmat"], ref["wmat"])),
           "matrices_max_rel": float(d.max() / extent), "matrix_mode": "strict" if strict_flag else "fma (library default)",
           "lists": hashes, "seconds": round(time.time() - t0, 2)}
    return res


# ------------------------------------------------------------------------------------------------------------------
# particles
# ------------------------------------------------------------------------------------------------------------------
def time_particles(h, torch, dist, world: int, emitters: int, seed_base: int, stream, steps: int = 30, with_sort: bool = True) -> dict:
    """config C3 on this rank's share of the emitters: 240 warm-up steps, then `steps` timed frames of
    rn_particles_step (+ rn_particles_sort: C3 names a depth sort per frame).  Times are max over ranks."""
    em = sg.make_emitters(emitters, seed_base=seed_base)
    h.particles_configure(em)
    cam = (0.0, 0.0, 0.0)
    for _ in range(240):
        h.particles_step(1.0 / 60.0, stream)
    torch.cuda.synchronize()
    live = int(h.particles_counts()[0].sum())

    def timed(fn):
        fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            fn()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / steps
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    step_ms = timed(lambda: h.particles_step(1.0 / 60.0, stream))
    out = {"metric": "Mparticles/s", "live_per_gpu": live, "emitters_per_gpu": emitters, "step_ms": step_ms,
           "bytes_per_particle": PARTICLE_BYTES}
    total_live = live
    if world > 1:
        t = torch.tensor([live], dtype=torch.int64, device="cuda")
        dist.all_reduce(t)
        total_live = int(t.item())
    peak, _ = measured_peak()
    out["step_value"] = total_live / step_ms / 1e3
    out["step_roofline_frac"] = live * PARTICLE_BYTES / (step_ms * 1e-3) / 1e9 / peak
    if with_sort:
        def frame():
            h.particles_step(1.0 / 60.0, stream)
            h.particles_sort(cam, stream)
        sort_ms = timed(lambda: h.particles_sort(cam, stream))
        frame_ms = timed(frame)
        out.update({"sort_ms": sort_ms, "ms_per_step": frame_ms, "value": total_live / frame_ms / 1e3,
                    "what": "rn_particles_step + rn_particles_sort (per-emitter depth sort) per frame; max over ranks"})
    else:
        out.update({"ms_per_step": step_ms, "value": out["step_value"], "what": "rn_particles_step per frame; max over ranks"})
    out["live_total"] = total_live
    return out


# ------------------------------------------------------------------------------------------------------------------
# ours
# ------------------------------------------------------------------------------------------------------------------
def init_dist(args):
    import torch
    import torch.distributed as dist
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
            opts.config.max_ctas = args.nccl_ctas
            opts.config.min_ctas = min(args.nccl_ctas, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=300),
                                pg_options=opts)
    return torch, dist, world, rank, local


def run_particles_config(args) -> dict:
    """--config C3: 1 000 emitters x 4 000 particles per GPU."""
    from rayne_b200 import capi
    torch, dist, world, rank, local = init_dist(args)
    stream = torch.cuda.current_stream().cuda_stream
    emitters = args.emitters or 1000
    h = capi.SceneHandle(max_nodes=1, device=local, max_emitters=emitters, max_particles_per_emitter=4000)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    m0 = sampler.mark()
    launches0 = h.stats().kernel_launches
    p = time_particles(h, torch, dist, world, emitters, 1000 + rank * emitters, stream, steps=max(args.steps, 30))
    launches = h.stats().kernel_launches - launches0
    # end to end: the frame + the read-back of every emitter's live count and sorted index list heads (what a CPU renderer needs
    # to draw: counts) - the particle state itself stays on the GPU (the mesh is built there, rn_particles_build_mesh)
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        h.particles_step(1.0 / 60.0, stream)
        h.particles_sort((0.0, 0.0, 0.0), stream)
        h.particles_counts()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / reps
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    clocks = sampler.stop(m0, sampler.mark()) if rank == 0 else None
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return {}
    peak, peak_kind = measured_peak()
    out = {"metric": "Mparticles/s", "value": p["value"], "unit": "Mparticles/s", "n_gpus": world, "steps": max(args.steps, 30),
           "warmup": 240, "ms_per_step": p["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic",
           "config": {"workload": f"C3: {emitters} RN::ParticleEmitters x 4000 particles per GPU ({p['live_per_gpu']} live), "
                                  "integrate/age/kill/spawn + per-emitter depth sort per frame",
                      "l2": "particle state (0.56 GB per generation, double-buffered) exceeds the 126 MB L2; no flush needed",
                      "parallelism": f"emitters split over {world} GPUs, no exchange" if world > 1 else "single GPU"},
           "e2e": {"value": p["live_total"] / e2e_ms / 1e3, "unit": "Mparticles/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 4,
                   "d2h_bytes_per_step": int(emitters * 8),
                   "what": "rn_particles_step(delta) + rn_particles_sort + rn_particles_counts (live count and emitter time of every "
                           "emitter to the host); the particle state and its render mesh stay in HBM"},
           "gpu_launches": int(launches),
           "roofline": {"bound": "hbm", "kernel": "rn::particles_step_kernel", "achieved": p["live_per_gpu"] * PARTICLE_BYTES / (p["step_ms"] * 1e-3) / 1e9,
                        "peak": peak, "unit": "GB/s", "frac": p["step_roofline_frac"], "peak_kind": peak_kind, "traffic": None,
                        "kernel_ms": p["step_ms"], "algorithmic_bytes_per_launch": p["live_per_gpu"] * PARTICLE_BYTES},
           "particles": p, "clocks": clocks}
    if not args.no_cpu:
        out["cpu_baseline"] = cpu_reference_particles(CPU_SAMPLE_EMITTERS, 30, 240)
    return out


def run_ours(args) -> dict:
    if args.config == "C3":
        return run_particles_config(args)
    from rayne_b200 import capi
    torch, dist, world, rank, local = init_dist(args)
    cfg = dict(CONFIGS[args.config])
    strong = cfg["scaling"] == "strong"
    trees = args.trees or cfg["trees"]
    cams = config_cameras(cfg)
    fr = capi.build_frustums(cams)
    ncam = len(fr)
    n_emitters = args.emitters if args.emitters is not None else cfg.get("emitters", 1000 if args.config == "C4" else 0)

    # ---- the scene of this rank ----
    t0 = time.time()
    if strong and world > 1:
        # ONE shared scene, cut into contiguous rank ranges where no parent/child edge crosses (shard.plan_shards);
        # every rank builds the same scene from the same seed and keeps its range
        from rayne_b200 import shard
        scene_all = build_scene(cfg, trees, seed=12345)
        rs_all = capi.RankedScene(scene_all)
        ranges = shard.plan_shards(rs_all.parent, world)
        a, b = ranges[rank]
        rs = shard.extract_shard(rs_all, a, b)
        rank_offset = a
        range_sizes = [int(y - x) for x, y in ranges]
        # the parity sample (first trees in creation order) sits at the END of the render list: the last rank holds it
        parity_rank = world - 1
        scene = scene_all
        order_all, rank_of_all = rs_all.order, rs_all.rank_of
        local_roots = np.nonzero(rs.parent == capi.RN_NO_PARENT)[0].astype(np.uint32)
        n = rs.n
        n_total = scene_all.n
        del rs_all
    else:
        # weak scaling: rank r generates and owns its own `trees` trees (ranks [r*n, (r+1)*n) of the N x larger scene)
        scene = build_scene(cfg, trees, seed=12345 + rank)
        rs = capi.RankedScene(scene)
        n = scene.n
        rank_offset = rank * n
        range_sizes = [n] * world
        parity_rank = 0
        local_roots = rs.roots
        n_total = n * world
    h = capi.SceneHandle(max_nodes=n + 65536, device=local, max_emitters=n_emitters, max_particles_per_emitter=4000 if n_emitters else 0,   # + room for incremental edits
                         tile_nodes=args.tile_nodes, tile_inner_nodes=args.inner_nodes, tile_unit_nodes=args.unit_nodes,
                         strict_matrices=args.strict_matrices)
    h.upload(rs)
    st = h.stats()
    log(f"[bench] rank {rank}: {n} nodes, upload+tiling {time.time() - t0:.1f}s tiles={st.tiles} spine={st.spine_nodes} levels={st.spine_levels}")
    stream = torch.cuda.current_stream().cuda_stream

    if world > 1:
        h.set_list_mode(True, rank_offset)   # lists travel as global render-list ranks
    lists = h.transform_cull(fr, stream)
    counts = h.read_counts(stream)
    vis_frac = [float(c) / n for c in counts]
    use_multicast = False
    pexch = None
    if world > 1:
        from rayne_b200 import shard
        words = [(s + 31) // 32 for s in range_sizes]
        count_t = torch.as_tensor(DevArray(lists[0].count, ncam), device="cuda")
        if args.exchange_transport == "multicast":
            try:
                pexch = shard.MulticastMaskExchange(words, ncam)
            except Exception as e:                      # no NVLS / no symmetric memory: say so and use NCCL
                log(f"[bench] rank {rank}: multicast exchange unavailable ({type(e).__name__}: {e}); using NCCL")
            # every rank must take the same path: agree on the minimum
            okt = torch.tensor([1 if pexch is not None else 0], dtype=torch.int32, device="cuda")
            dist.all_reduce(okt, op=dist.ReduceOp.MIN)
            if int(okt.item()) == 0:
                pexch = None
        use_multicast = pexch is not None
        if pexch is None:
            pexch = shard.PipelinedMaskExchange(words, ncam)

    def submit_masks():
        if use_multicast:
            # the mask kernel stores through the NVLS multicast address: building the masks IS the gather
            h.build_masks_multicast(pexch.begin(), pexch.row, stream)
        else:
            # the masks go straight into the send buffer of this frame's NCCL gather (no staging copy)
            h.build_masks_to(pexch.begin().data_ptr(), pexch.row, stream)
        pexch.submit()

    do_particles_in_step = args.config == "C5" and n_emitters > 0
    emitters_local = 0
    if n_emitters:
        emitters_local = n_emitters // world + (1 if rank < n_emitters % world else 0) if strong else n_emitters
    if do_particles_in_step:
        em = sg.make_emitters(emitters_local, seed_base=1000 + sum(n_emitters // world + (1 if r < n_emitters % world else 0) for r in range(rank)))
        h.particles_configure(em)
        for _ in range(240):
            h.particles_step(1.0 / 60.0, stream)
        torch.cuda.synchronize()

    def step():
        # every frame's visibility IS gathered on every GPU; the gather of frame i runs on a side stream under the
        # kernels of frame i+1 (two frames in flight) and the last ones are flushed inside the timed region
        h.transform_cull(fr, stream)
        if world > 1:
            submit_masks()
        if do_particles_in_step:
            h.particles_step(1.0 / 60.0, stream)

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

    # ---- the exchange, verified by EQUALITY (outside the timed region) ----
    # For three frames with the roots moved in between (so that consecutive frames differ and a stale or misplaced ring
    # slot cannot pass): every rank's local masks are all-gathered over NCCL and compared word for word with the blocks
    # the transport under test delivered; the expanded global lists are checked against the all-reduced counts.
    roots = local_roots
    rootpos = rs.position[roots].copy()
    exchange_info = None
    if world > 1:
        blocks_equal, frames_differ, prev = True, False, None
        for _ in range(3):
            rootpos[:, 0] += np.float32(7.5)
            h.update_nodes(roots, position=rootpos)
            h.transform_cull(fr, stream)
            submit_masks()
            gathered = pexch.flush()                                   # [world, ncams, row]
            h.build_masks(stream)
            mptr, mwords, _ = h.device_masks()
            local_m = torch.as_tensor(DevArray(mptr, ncam * mwords), device="cuda").view(ncam, mwords)
            send = torch.zeros(ncam, pexch.row, dtype=torch.int32, device="cuda")
            send[:, :mwords] = local_m
            ref_g = torch.empty(world, ncam, pexch.row, dtype=torch.int32, device="cuda")
            dist.all_gather_into_tensor(ref_g.view(-1), send.view(-1))
            for r in range(world):
                w = (range_sizes[r] + 31) // 32
                blocks_equal = blocks_equal and bool(torch.equal(gathered[r, :, :w], ref_g[r, :, :w]))
            if prev is not None and not torch.equal(prev, ref_g):
                frames_differ = True
            prev = ref_g.clone()
        tot = count_t.clone().to(torch.int64)
        dist.all_reduce(tot)
        tot = tot.cpu().tolist()
        starts = np.concatenate([[0], np.cumsum(range_sizes)[:-1]]).tolist() if strong else [r * n for r in range(world)]
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
        okt = torch.tensor([1 if (blocks_equal and ok) else 0], dtype=torch.int32, device="cuda")
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
        exchange_info = {"kind": ("rank-order visibility bitmasks stored through the NVLS multicast address by the mask kernel itself "
                                  "(rn_cull_build_masks_multicast) + a symmetric-memory barrier, two frames in flight"
                                  if use_multicast else
                                  "NCCL all-gather of rank-order visibility bitmasks (rn_cull_build_masks_to), fixed size, "
                                  "two frames in flight on a side stream"),
                         "transport": "nvls-multicast" if use_multicast else "nccl",
                         "bytes_per_gpu_and_frame": int(ncam * ((n + 31) // 32) * 4), "global_list_entries_per_view": tot,
                         "blocks_equal": bool(blocks_equal), "blocks_equal_on_every_rank": bool(int(okt.item()) == 1),
                         "verified_frames": 3, "verified_frames_differ": bool(frames_differ),
                         "blocks_equal_how": "3 frames with moved roots: every GPU's local masks all-gathered over NCCL and compared "
                                             "word for word with the blocks the transport delivered",
                         "expanded_lists_match_counts_and_ascend": ok,
                         "expand_all_views_ms": x0.elapsed_time(x1) / 5,
                         "expand_note": "rn_masks_expand of all views' global lists on one GPU; not part of the step"}
        del out_l

    # ---- parity gate on the timed workload (outside the timed region) ----
    parity = None
    if rank == parity_rank and not args.no_parity:
        if strong and world > 1:
            # this rank holds the tail of the render list = the first trees in creation order; translate ids through the
            # global order of the shared scene
            class _View:           # what parity_gate needs of a RankedScene, in GLOBAL terms restricted to this shard
                pass
            v = _View()
            v.order = order_all[rank_offset: rank_offset + n]          # local rank -> creation id
            full_rank_of = rank_of_all
            v.rank_of = (full_rank_of.astype(np.int64) - rank_offset).astype(np.int64)   # creation id -> local rank (valid for ids in the shard)
            h.set_list_mode(False, 0)
            # restore the positions the scene was generated with (the exchange check moved the roots)
            h.update_nodes(local_roots, position=rs.position[local_roots])
            parity = parity_gate(h, v, scene, cams, fr, cfg, min(CPU_SAMPLE_TREES, max(1, trees // world // 2)), stream, args.strict_matrices)
            h.set_list_mode(True, rank_offset)
        else:
            if world > 1:
                h.set_list_mode(False, 0)
            # restore the positions the scene was generated with (the exchange check moved the roots)
            h.update_nodes(roots, position=rs.position[roots])
            parity = parity_gate(h, rs, scene, cams, fr, cfg, None if args.verify_full else min(CPU_SAMPLE_TREES, trees), stream,
                                 args.strict_matrices)
            if world > 1:
                h.set_list_mode(True, rank_offset)
        log(f"[bench] parity gate: {json.dumps({k: v for k, v in parity.items() if k != 'lists'})}")
    if world > 1:
        # every rank waits for the parity rank; the result travels to rank 0
        obj = [parity]
        dist.broadcast_object_list(obj, src=parity_rank)
        parity = obj[0]
    h.update_nodes(roots, position=rs.position[roots])
    rootpos = rs.position[roots].copy()

    # ---- end to end through the C ABI with host buffers: rn_pipeline (rayne_b200/csrc/rn_pipeline.cu) ----
    # dirty rows -> device, transform + cull over `shards` rank ranges on prioritised streams, visibility written into mapped
    # pinned HOST memory by the device, two frames in flight (frame i's copy-out runs under frame i + 1's kernels).  Three
    # encodings of the same visibility are timed; `e2e` is the adaptive one (per range and view the smaller of list and
    # rank-order bitmask, chosen on the device), the one a host consumer would ask for.
    shards = max(1, min(args.e2e_shards, len(roots)))
    torch.cuda.synchronize()
    pipe = capi.Pipeline(rs, n_ranges=shards, frames_in_flight=2, device=local, rank_base=rank_offset if world > 1 else 0,
                         strict_matrices=args.strict_matrices, tile_nodes=args.tile_nodes, tile_inner_nodes=args.inner_nodes,
                         tile_unit_nodes=args.unit_nodes)

    def time_form(form):
        """steady state with two frames in flight: submit frame i + 1, then wait for frame i"""
        bytes_io = [0, 0]

        def produce():
            rootpos[:, 0] += np.float32(0.01)
            pipe.update_nodes(roots, position=rootpos)
            pipe.submit(fr, form)

        def consume():
            view = pipe.acquire()
            bytes_io[0], bytes_io[1] = int(view.h2d_bytes), int(view.d2h_bytes)
            pipe.release()

        produce()
        for _ in range(min(3, args.warmup)):
            produce(); consume()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            produce(); consume()
        barrier()
        ms = (time.perf_counter() - t0) * 1e3 / args.steps
        consume()
        io = list(bytes_io)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            t = torch.tensor(io, dtype=torch.int64, device="cuda")
            dist.all_reduce(t)
            io = [int(x) for x in t.tolist()]
        return ms, io

    e2e_forms = {}
    for name, form in (("adaptive", capi.VIS_ADAPTIVE), ("lists", capi.VIS_LISTS), ("masks", capi.VIS_MASKS)):
        e2e_forms[name] = time_form(form)
    # the three encodings carry the same lists, and those are the lists of the plain handle on the same positions
    h.update_nodes(roots, position=rootpos)
    h.set_list_mode(True, rank_offset if world > 1 else 0)        # global ranks, like the pipeline's segments
    h.transform_cull(fr, stream)
    want = [h.read_list(c) for c in range(ncam)]
    if world == 1:
        h.set_list_mode(False, 0)
    forms_equal = True
    for form in (capi.VIS_ADAPTIVE, capi.VIS_LISTS, capi.VIS_MASKS):
        pipe.update_nodes(roots, position=rootpos)
        pipe.submit(fr, form)
        view = pipe.acquire()
        got = pipe.frame_lists(view)
        forms_equal = forms_equal and all(np.array_equal(a, b) for a, b in zip(got, want))
        if form == capi.VIS_ADAPTIVE:
            taken = pipe.forms(view)
            forms_taken = ["".join("L" if taken[k, c] == capi.VIS_LISTS else "M" for k in range(taken.shape[0])) for c in range(ncam)]
        pipe.release()
    pipe.close()
    e2e_ms, (h2d_total, d2h_total) = e2e_forms["adaptive"]

    def e2e_entry(name, what):
        ms, (hb, db) = e2e_forms[name]
        return {"value": n_total / ms / 1e3, "unit": "Mnodes/s", "ms_per_step": ms, "h2d_bytes_per_step": hb, "d2h_bytes_per_step": db,
                "form": name, "what": what}

    e2e_common = (f"rn_pipeline_update_nodes (dirty roots) + rn_pipeline_submit + rn_pipeline_acquire, 2 frames in flight, visibility into pinned host "
                  "memory; " + (f"{shards} rank ranges on prioritised streams, stored by the device (pack kernel); " if shards > 1 else
                                "one range, two sets of cull outputs, the copy-out of frame i runs on its own stream beside the kernels of frame i + 1 "
                                "(masks through the copy engine, lists through the pack kernel); "))
    e2e_lists = e2e_entry("lists", e2e_common + "every view as its list of render-list ranks (4 B per visible node) - the container the reference's Render fills")
    e2e_masks = e2e_entry("masks", e2e_common + "every view as a rank-order bitmask (1 bit per node)")
    e2e_adaptive = e2e_entry("adaptive", e2e_common + ("per range and view the smaller of list and bitmask, chosen on the device from the compaction's count; "
                                                       if shards > 1 else "RN_VIS_ADAPTIVE, which delivers rank-order bitmasks in the one-range mode; ") +
                             "rn_frame_list decodes either into the reference's list")
    e2e_adaptive["encodings_equal_and_equal_to_rn_cull_read_list"] = bool(forms_equal)
    e2e_adaptive["encoding_per_view_and_range"] = forms_taken

    # ---- particles beside the headline (config C3 on every GPU; C5 times them inside the step) ----
    particles = None
    if n_emitters and not do_particles_in_step:
        particles = time_particles(h, torch, dist, world, n_emitters, 1000 + rank * n_emitters, stream)
    elif do_particles_in_step:
        live = int(h.particles_counts()[0].sum())
        t = torch.tensor([live], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(t)
        particles = {"metric": "Mparticles/s", "live_total": int(t.item()), "live_per_gpu": live, "emitters_per_gpu": emitters_local,
                     "value": int(t.item()) / ms_step / 1e3,
                     "what": "rn_particles_step runs inside every timed step of this config; value = live particles / step time"}

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
        ms = timed(lambda: h.transform_propagate(stream))
        extras["transform_propagate"] = {"ms": ms, "bytes_per_node": 152, "roofline_frac": n * 152.0 / (ms * 1e-3) / 1e9 / peak}
        ms = timed(lambda: h.cull(fr, stream))
        extras["cull_stored"] = {"ms": ms, "bytes_per_node": 66, "roofline_frac": n * (65.0 + 1.0) / (ms * 1e-3) / 1e9 / peak,
                                 "what": "rn_cull on stored transforms: cull_kernel + emit_kernel"}
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

        def cull_and_sort():
            h.transform_cull(fr_sort, stream)
            h.cull_sort(campos, stream)

        ms_both = timed(cull_and_sort, reps=5)
        ms_cull = timed(lambda: h.transform_cull(fr_sort, stream), reps=5)
        ms = max(ms_both - ms_cull, 1e-6)
        # 4 radix passes (key + entry read and written) + key build (entry + world position in, key + entry out)
        sort_bytes = entries * (4 * 16.0 - 4.0 + 28.0)
        extras["cull_sort_front_to_back"] = {"ms": ms, "entries": entries, "Mentries_per_s": entries / ms / 1e3,
                                             "roofline_frac": sort_bytes / (ms * 1e-3) / 1e9 / peak,
                                             "how": "time of (rn_transform_cull + rn_cull_sort) minus rn_transform_cull: every repetition sorts a fresh, unsorted list"}
        # the 1 %-dirty frame - one node in a hundred (spread over the whole render list) gets a new local
        # position through rn_scene_update_nodes, then the usual frame.  The host wall clock covers the staging copy, the
        # H2D copy, the scatter kernel and rn_transform_cull (which re-propagates every node: there is no dirty shortcut).
        dirty = np.arange(0, n, 100, dtype=np.uint32)
        dpos = np.ascontiguousarray(rs.position[dirty])

        def dirty_frame():
            h.update_nodes(dirty, position=dpos)
            h.transform_cull(fr, stream)

        for _ in range(3):
            dirty_frame()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = max(1, min(20, args.extras_reps * 2))
        for _ in range(reps):
            dirty_frame()
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / reps
        extras["frame_1pct_dirty"] = {"ms": ms, "dirty_nodes": int(dirty.shape[0]), "h2d_bytes": int(dirty.shape[0]) * 16,
                                      "Mnodes_per_s": n / ms / 1e3,
                                      "what": "rn_scene_update_nodes (1 % of the nodes, host arrays) + rn_transform_cull, host wall clock"}
        # visibility as rank-order bitmasks (the multi-GPU exchange format) and back: 1 B + V/8 B per node; expansion of
        # the largest view reads its words twice and writes 4 B per entry
        ms = timed(lambda: h.build_masks(stream))
        extras["cull_build_masks"] = {"ms": ms, "bytes_per_node": 1 + ncam / 8.0,
                                      "roofline_frac": n * (1 + ncam / 8.0) / (ms * 1e-3) / 1e9 / peak}
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
        h.transform_inverse(stream)
        inst = torch.empty((max(entries, 1), 64), dtype=torch.float32, device="cuda")
        ms = timed(lambda: h.pack_instances(ncam - 1, view[ncam - 1], projview[ncam - 1], inst.data_ptr(), entries, stream))
        extras["cull_pack_instances"] = {"ms": ms, "entries": entries, "bytes_per_entry": 388,
                                         "roofline_frac": entries * 388.0 / (ms * 1e-3) / 1e9 / peak}
        del inst
        if particles and n_emitters:
            ms = timed(lambda: h.particles_build_mesh(stream=stream))
            extras["particles_build_mesh"] = {"ms": ms, "bytes_per_particle": 236,
                                              "roofline_frac": particles["live_per_gpu"] * 236.0 / (ms * 1e-3) / 1e9 / peak}
            # vertex stage of Shaders/Particles.hlsl over that mesh: 36 B read + 32 B written per vertex
            ne = emitters_local if do_particles_in_step else n_emitters
            ident = np.tile(np.eye(4, dtype=np.float32).reshape(16), (ne, 1))
            ms = timed(lambda: h.particles_expand_billboards(ident, ident, stream=stream))
            extras["particles_expand_billboards"] = {"ms": ms, "bytes_per_particle": 272,
                                                     "roofline_frac": particles["live_per_gpu"] * 272.0 / (ms * 1e-3) / 1e9 / peak}
        # incremental edits (SceneBasic::AddNode / RemoveNode without a new upload): 1 000 new nodes - 10 trees of 100 - into the
        # uploaded scene, a frame, then out again.  Host wall clock around each call (they synchronise internally).
        if world == 1 and not args.strict_matrices:
            h.set_list_mode(False, 0)
            ex = sg.make_scene(n_trees=10, branching=99, depth=2, seed=777)
            exp = np.where(ex.parent >= 0, ex.parent, capi.RN_NO_PARENT).astype(np.uint32)
            # call order = children before their parent: reverse creation order of every (root, 99 children) tree
            seq = np.arange(ex.n)[::-1].copy()
            pos_in = np.empty(ex.n, np.int64); pos_in[seq] = np.arange(ex.n)
            par = np.where(ex.parent[seq] >= 0, pos_in[np.maximum(ex.parent[seq], 0)], capi.RN_NO_PARENT).astype(np.uint32)
            h.transform_cull(fr, stream)
            counts_before = h.read_counts()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            try:
                slots = h.add_nodes(ex.pos[seq], ex.scale[seq], ex.rot[seq], par, ex.flags[seq], ex.group[seq], ex.sphere[seq])
            except capi.RNError as err:
                # small scenes get small tiles (rn_scene_desc.tile_nodes = 0): a 100-node tree then needs a full upload
                slots = None
                extras["incremental_edits"] = {"nodes": int(ex.n), "fast_path": False, "why": str(err)}
            t_add = (time.perf_counter() - t0) * 1e3
            if slots is not None:
                ms_after = timed(lambda: h.transform_cull(fr, stream))
                counts_after = h.read_counts()
                t0 = time.perf_counter()
                h.remove_nodes(slots)
                t_rem = (time.perf_counter() - t0) * 1e3
                h.transform_cull(fr, stream)
                counts_removed = h.read_counts()
                extras["incremental_edits"] = {"nodes": int(ex.n), "add_ms": t_add, "remove_ms": t_rem, "frame_ms_with_the_new_tiles": ms_after,
                                               "visible_after_add": [int(c) for c in counts_after], "visible_after_remove": [int(c) for c in counts_removed],
                                               "counts_back_to_before_the_edit": bool(np.array_equal(counts_removed, counts_before)),
                                               "what": "rn_scene_add_nodes / rn_scene_remove_nodes on the uploaded scene, host wall clock; "
                                                       "a full rn_scene_upload of this scene takes seconds (host re-tiling)"}
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
    traffic, traffic_src = ncu_traffic(n, ncam)
    tile_bytes = n * TILE_BYTES_PER_NODE                                     # the dominant kernel's algorithmic bytes per launch
    step_bytes = n * (TILE_BYTES_PER_NODE + 1.0 + 5.0 + 4.0 * sum(vis_frac))   # + mask byte, compaction reads, list entries
    tile_avg_ms = tile_ms / max(tile_n, 1)
    achieved = tile_bytes / (tile_avg_ms * 1e-3) / 1e9
    desc = {"C1": "C1: 10 x (1+9+90+900) = 10 000 SceneNodes, depth 4, 1 camera",
            "C2": f"C2: {n} SceneNodes ({trees} complete 4-ary depth-8 trees), 1 camera frustum cull with visibility list",
            "C4": f"C4: {n} SceneNodes per GPU ({trees} complete 4-ary depth-8 trees), 6-view cull (2 eye + 4 shadow-cascade cameras"
                  + (" as Light::UpdateShadows derives them)" if RIG == "reference" else ", round-1 hand-set rig)"),
            "C4-strong": f"C4-strong: one fixed scene of {n_total} SceneNodes ({trees} trees) cut over {world} GPU(s) by plan_shards, 6-view cull",
            "C5": f"C5: one scene of {n_total} SceneNodes ({trees} trees) + {n_emitters} emitters x 4000 particles cut over {world} GPU(s), "
                  "6-view cull + particle step per frame"}[args.config]
    if world > 1:
        par = ("one shared scene cut into contiguous rank ranges by shard.plan_shards, one range per GPU" if strong else
               f"independent per-GPU rank ranges (weak scaling: rank r generates and owns its own {trees} trees)")
        par += ("; visibility bitmasks of every view multicast to all GPUs through NVLS by the mask kernel" if use_multicast else
                "; NCCL all-gather of every view's visibility bitmasks")
    else:
        par = "single GPU"
    out = {
        "metric": METRIC, "value": n_total / ms_step / 1e3, "unit": "Mnodes/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc + "; every node re-propagated each frame (no dirty shortcut exists); the e2e loop also moves every root",
                   "name": args.config, "nodes_per_gpu": n, "nodes_total": n_total, "views": ncam,
                   "visible_fraction_per_view": [round(v, 5) for v in vis_frac],
                   "matrices": "strict (bit-exact)" if args.strict_matrices else "fma (library default, within 1e-5; lists and world TRS bit-exact)",
                   "l2": (f"inputs ({n * 64 / 1e9:.2f} GB of node columns) exceed the 126 MB L2; no flush needed" if n * 64 > 126e6 else
                          f"inputs + outputs ({n * 169 / 1e6:.1f} MB) FIT in the 126 MB L2 at this config's size: the number is an L2-resident one, "
                          "as the reference's own frame at this size is cache-resident on the CPU"),
                   "parallelism": par},
        "e2e": dict(e2e_adaptive, shards=shards, gpus="every GPU reads back its own rank range; byte counts are totals over all GPUs" if world > 1 else "one GPU"),
        "e2e_lists": e2e_lists,
        "e2e_masks": e2e_masks,
        "gpu_launches": int(launches),
        "parity": parity,
        "exchange": exchange_info,
        "roofline": {"bound": "hbm", "kernel": "rn::tile_kernel<true,%s>" % ("false" if args.strict_matrices else "true"),
                     "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "peak_kind": peak_kind, "traffic": traffic, "traffic_source": traffic_src,
                     "algorithmic_bytes_per_launch": tile_bytes, "bytes_per_node": TILE_BYTES_PER_NODE, "kernel_ms": tile_avg_ms,
                     "kernel_share_of_step": tile_avg_ms / ms_step,
                     "step_frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak, "step_algorithmic_bytes": step_bytes,
                     "step_frac_note": "whole step (spine + tile + compaction [+ mask kernel]) against the same peak: the headline fraction"},
        "clocks": clocks,
    }
    if particles:
        out["particles"] = particles
    if extras:
        out["extras"] = extras
    if not args.no_cpu:
        t0 = time.time()
        if args.config == "C1":
            sample_scene, what = scene, "the whole C1 scene"
        else:
            k = min(CPU_SAMPLE_TREES, trees)
            sample_scene, what = first_trees(scene, k, cfg), f"the first {k} of the GPU scene's {trees} trees"
        out["cpu_baseline"] = cpu_reference_scene(sample_scene, cams, 6, 2, what)
        if particles:
            particles["cpu_baseline"] = cpu_reference_particles(CPU_SAMPLE_EMITTERS, 30, 240)
        log(f"[bench] cpu baseline took {time.time() - t0:.1f}s")
    return out


def run_reference(args) -> dict:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.config == "C3":
        r = cpu_reference_particles(CPU_SAMPLE_EMITTERS, max(1, args.steps), 240)
        v = r["value"]
        return {"impl": "reference", "metric": "Mparticles/s", "value": v, "unit": "Mparticles/s", "n_gpus": world, "steps": args.steps,
                "warmup": 240, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "C3 sample: " + r["sample"] + " on the host CPU (GenericParticleEmitter::Update, unmodified Rayne sources)"},
                "cpu_baseline": r, "gpu_launches": 0,
                "e2e": {"value": v, "unit": "Mparticles/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    cfg = CONFIGS[args.config]
    cams = config_cameras(cfg)
    # bounded sample: keep the whole --steps/--warmup run near two minutes (the reference does ~0.9 Mnodes/s here)
    frames = max(1, args.steps) + max(0, args.warmup)
    if args.config == "C1":
        scene = sg.make_scene(n_trees=10, branching=cfg["branching"], depth=cfg["depth"], seed=12345)
        what = "the whole C1 scene"
    else:
        trees = int(min(CPU_SAMPLE_TREES, max(2, 4500 // frames), cfg["trees"]))
        # the same trees the GPU arm's parity gate and cpu_baseline use: the first trees of the rank-0 scene
        scene = sg.make_scene(n_trees=trees, branching=cfg["branching"], depth=cfg["depth"], seed=12345)
        what = f"{trees} of {cfg['trees']} trees"
    r = cpu_reference_scene(scene, cams, max(1, args.steps), max(0, args.warmup), what)
    v = r["value"]
    return {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mnodes/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_frame"], "higher_is_better": True, "scaling": cfg["scaling"],
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.config} sample: " + r["sample"] + " on the host CPU (SceneBasic::Update + Render, unmodified Rayne sources)",
                       "name": args.config},
            "cpu_baseline": r, "gpu_launches": 0,
            "e2e": {"value": v, "unit": "Mnodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="C4", choices=["C1", "C2", "C3", "C4", "C4-strong", "C5"],
                    help="BASELINE.json config (default C4: 16.8 M nodes per GPU, 6 views, weak scaling)")
    ap.add_argument("--trees", type=int, default=0, help="override the number of trees of the config (0 = the config's)")
    ap.add_argument("--emitters", type=int, default=None, help="particle emitters (per GPU for C3/C4 side metric; total for C5); 0 = skip")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity gate")
    ap.add_argument("--verify-full", action="store_true", help="parity gate over the WHOLE scene through the C oracle (~25 s at 16.8 M nodes; N = 1)")
    ap.add_argument("--strict-matrices", action="store_true", help="RN_SCENE_STRICT_MATRICES: world matrices bit-exact (no FMA on the matrix chain)")
    ap.add_argument("--rig", choices=("reference", "legacy"), default="reference",
                    help="6-view configs: cascade cameras derived like the reference's Light::UpdateShadows (default) or round 1's hand-set ones")
    ap.add_argument("--e2e-shards", type=int, default=1, help="rank ranges / streams of the pipelined end-to-end frame (1 = off)")
    ap.add_argument("--exchange-transport", choices=("nccl", "multicast"), default="multicast",
                    help="N>1: multimem stores of the mask kernel through NVLS multicast (falls back to NCCL where "
                         "symmetric memory / NVLS is unavailable), or the NCCL all-gather")
    ap.add_argument("--nccl-ctas", type=int, default=0, help="N>1: cap on NCCL's CTAs per collective (0 = NCCL's default)")
    ap.add_argument("--extras-reps", type=int, default=10, help="timed repetitions of each side kernel (1 for an ncu capture)")
    ap.add_argument("--no-extras", action="store_true", help="skip the side timings of inverse / bounds / sort / mesh")
    ap.add_argument("--tile-nodes", type=int, default=0, help="rn_scene_desc.tile_nodes (0 = library default)")
    ap.add_argument("--inner-nodes", type=int, default=0, help="rn_scene_desc.tile_inner_nodes (0 = library default)")
    ap.add_argument("--unit-nodes", type=int, default=0, help="rn_scene_desc.tile_unit_nodes (0 = library default)")
    args = ap.parse_args()
    global RIG
    RIG = args.rig
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
