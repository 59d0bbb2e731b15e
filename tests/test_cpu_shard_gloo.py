"""CPU, world_size 2 over gloo: the multi-GPU host logic (rank-range planning at subtree boundaries, shard extraction,
count + list all-gather) reproduces the single-process visibility lists.  The per-rank cull itself is stood in by the
oracle here (there is no GPU); on the GPU box tests/test_gpu_scene.py::test_sharded_matches_single runs the same
planner against librn_cuda."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle
from rayne_b200 import capi, scenegen as sg, shard
from tests.helpers import standard_cameras

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_valid_cuts_and_plan():
    #      0        3      5
    #     / \       |
    #    1   2      4          ranks given directly (parent by rank)
    parent = np.array([shard.RN_NO_PARENT, 0, 0, shard.RN_NO_PARENT, 3, shard.RN_NO_PARENT], np.uint32)
    cuts = shard.valid_cuts(parent)
    assert cuts.tolist() == [True, False, False, True, False, True, True]
    assert shard.plan_shards(parent, 2) == [(0, 3), (3, 6)]
    assert shard.plan_shards(parent, 3) == [(0, 3), (3, 5), (5, 6)]
    with pytest.raises(ValueError):
        shard.plan_shards(parent, 5)
    # a child ranked before its parent (render-list order is reverse post-order in the reference)
    parent = np.array([1, shard.RN_NO_PARENT, 3, shard.RN_NO_PARENT], np.uint32)
    assert shard.valid_cuts(parent).tolist() == [True, False, True, False, True]


def test_plan_balances_a_forest():
    scene = sg.make_scene(n_trees=64, branching=4, depth=5, seed=5)
    rs = capi.RankedScene(scene)
    for world in (2, 4, 8):
        ranges = shard.plan_shards(rs.parent, world)
        sizes = [b - a for a, b in ranges]
        assert sum(sizes) == scene.n and ranges[0][0] == 0 and ranges[-1][1] == scene.n
        assert max(sizes) - min(sizes) <= 341  # one tree
        for a, b in ranges:
            sh = shard.extract_shard(rs, a, b)
            assert sh.n == b - a
            p = sh.parent[sh.parent != shard.RN_NO_PARENT]
            assert (p < sh.n).all()


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, q):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = sg.make_scene(n_trees=24, branching=4, depth=5, seed=31, random_bounds=True, random_groups=True)
        cams = standard_cameras()
        fr = oracle.frustums(cams)
        rs = capi.RankedScene(scene)
        ranges = shard.plan_shards(rs.parent, world)
        a, b = ranges[rank]
        sh = shard.extract_shard(rs, a, b)
        # stand-in for librn_cuda on this rank's GPU: world transforms + cull of the local range only
        wpos, wrot, wscale, _ = oracle.world(sh.parent, sh.position, sh.scale, sh.rotation)
        local = [oracle.cull(wpos, wrot, wscale, sh.sphere, sh.flags, sh.group, f) for f in fr]
        cap = max(sh.n, 1)
        lists = []
        for l in local:
            t = torch.zeros(scene.n, dtype=torch.int32)      # capacity >= any rank's count
            t[: len(l)] = torch.from_numpy((l.astype(np.int64) + a).astype(np.int32))
            lists.append(t)
        counts = torch.tensor([len(l) for l in local], dtype=torch.int32)
        ex = shard.ListExchange()
        all_counts, gathered = ex.exchange(counts, lists)
        merged = [torch.cat(parts).numpy().astype(np.uint32) for parts in gathered]
        # the one-frame-deep variant: frame 0 full lists, frame 1 half of every list, frame 2 full lists again with a
        # tight margin (-> reported as an overflow, because the sizes came from frame 1's halved counts)
        pex = shard.PipelinedListExchange(cuda=False, margin=1.0, slack=0, depth=1)
        frames = []
        for f, keep in enumerate((1.0, 0.5, 1.0)):
            cnt = torch.tensor([int(len(l) * keep) for l in local], dtype=torch.int32)
            pex.submit(cnt, lists)
            if f > 0:
                frames.append((pex.result[0].numpy().copy(), pex.overflows))
        table, outs, sizes = pex.flush()
        frames.append((table.numpy().copy(), pex.overflows))
        pmerged = [torch.cat(v).numpy().astype(np.uint32) for v in shard.PipelinedListExchange.views((table, outs, sizes), world)]
        # the compact exchange: rank-order bitmasks of the local range (what rn_cull_build_masks leaves), static sizes,
        # two frames in flight; frame 0 = empty masks, frames 1 and 2 = the real ones
        words = [(e - s + 31) // 32 for s, e in ranges]
        bits = np.zeros((len(fr), words[rank] * 32), np.uint8)
        for c, l in enumerate(local):
            bits[c, l] = 1
        mask = np.packbits(bits, axis=1, bitorder="little").view(np.uint32).reshape(-1)
        mex = shard.PipelinedMaskExchange(words, len(fr), cuda=False, depth=2)
        mex.submit(torch.zeros(mask.shape[0], dtype=torch.int32))
        first = mex.flush().clone()
        mex.submit(torch.from_numpy(mask.view(np.int32)))
        mex.submit(torch.from_numpy(mask.view(np.int32)))
        g = mex.flush().numpy().view(np.uint32)
        mmerged = []
        for c in range(len(fr)):
            parts = []
            for r, (s, e) in enumerate(ranges):
                seg = np.unpackbits(np.ascontiguousarray(g[r, c, :words[r]]).view(np.uint8), bitorder="little")
                parts.append(np.nonzero(seg)[0].astype(np.uint32) + np.uint32(s))
            mmerged.append(np.concatenate(parts))
        if rank == 0:
            q.put((all_counts.numpy(), merged, ranges, frames, pmerged, sizes, mmerged, int(first.abs().sum())))
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_matches_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    all_counts, merged, ranges, frames, pmerged, sizes, mmerged, first_sum = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    scene = sg.make_scene(n_trees=24, branching=4, depth=5, seed=31, random_bounds=True, random_groups=True)
    cams = standard_cameras()
    ref = oracle.scene_reference(scene, oracle.frustums(cams))
    rs = capi.RankedScene(scene)
    assert len(ranges) == 2 and all_counts.shape == (2, len(cams))
    for c in range(len(cams)):
        want_ranks = rs.rank_of[ref["lists"][c]]            # creation ids -> global ranks, already ascending
        assert np.array_equal(merged[c], want_ranks), f"camera {c}"
        assert int(all_counts[:, c].sum()) == len(want_ranks)
    # bitmask exchange: gathered masks expand to exactly the single-process lists
    assert first_sum == 0
    for c in range(len(cams)):
        assert np.array_equal(mmerged[c], rs.rank_of[ref["lists"][c]]), f"camera {c} (masks)"
    # pipelined exchange: tables arrive one frame late and equal the submitted counts; the third frame was sized from
    # the halved counts of the second one with no margin -> flagged as an overflow, its lists cut at `sizes`
    assert np.array_equal(frames[0][0], all_counts) and frames[0][1] == 0
    assert np.array_equal(frames[1][0], all_counts // 2 + 0 * all_counts) or np.array_equal(frames[1][0], (all_counts * 0.5).astype(all_counts.dtype))
    assert frames[1][1] == 0
    assert np.array_equal(frames[2][0], all_counts) and frames[2][1] == 1
    for c in range(len(cams)):
        want_ranks = rs.rank_of[ref["lists"][c]]
        got = pmerged[c]
        # per rank only the first sizes[c] entries travelled: the merged list is a prefix-per-rank subset
        assert set(got.tolist()) <= set(want_ranks.tolist())
