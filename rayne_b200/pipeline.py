"""Frame pipelining on ONE GPU: the render list is cut into S rank ranges (the same edge-free cuts the multi-GPU
path uses, shard.plan_shards), each range lives in its own rn_scene_t on its own CUDA stream, and a frame is

    for every range:  rn_scene_update_nodes_async (dirty rows, pinned staging)  ->  rn_transform_cull
    for every range:  rn_cull_read_lists into the caller's host buffers, behind the lists of the earlier ranges

Everything is enqueued before the first read-back blocks, so the device-to-host copy of range k's lists (the PCIe-bound
part of a frame that hands its visibility lists to a CPU renderer) runs while the SMs work on range k+1.  Lists travel
as global render-list ranks (rn_scene_set_list_mode), and because the ranges ascend and every range's compaction is
order-stable, the concatenation per camera IS the single-scene list (tests/test_gpu_scene.py::test_pipelined_*).

Host-side plumbing only: transforms, culling and compaction are librn_cuda's kernels.
"""
from __future__ import annotations

import numpy as np

from . import capi, shard


class PipelinedScene:
    def __init__(self, rs, n_ranges: int = 4, device: int = 0, rank_base: int = 0, **scene_kw):
        """rs: capi.RankedScene (node columns in render-list order).  Range k gets its own stream, earlier ranges
        more urgent ones: without priorities the compaction of range k queues behind the (already launched) transform
        kernels of ranges k+1.., and the first copy would only start when nearly all kernels are done."""
        self.n = rs.n
        self.ranges = shard.plan_shards(rs.parent, n_ranges)
        self.handles = []
        self.streams = []
        self.bounds = np.array([a for a, _ in self.ranges] + [rs.n], dtype=np.int64)
        for (a, b) in self.ranges:
            h = capi.SceneHandle(max_nodes=b - a, device=device, **scene_kw)    # makes `device` current
            self.streams.append(capi.stream_create(n_ranges - 1 - len(self.streams)))
            h.upload(shard.extract_shard(rs, a, b))
            h.set_list_mode(True, rank_base + a)    # rank_base: first global rank of `rs` when it is itself one GPU's range
            self.handles.append(h)
        self._split_cache = None

    def close(self) -> None:
        for h in self.handles:
            h.close()
        for st in self.streams:
            capi.stream_destroy(st)
        self.handles = []
        self.streams = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _split(self, ranks: np.ndarray):
        """Per range: (selector into `ranks`, local ranks).  Cached for a repeated dirty set (same array object)."""
        if self._split_cache is not None and self._split_cache[0] is ranks:
            return self._split_cache[1]
        r = np.asarray(ranks, dtype=np.int64)
        which = np.searchsorted(self.bounds, r, side="right") - 1
        parts = []
        for k, (a, _) in enumerate(self.ranges):
            sel = np.nonzero(which == k)[0]
            parts.append((sel, (r[sel] - a).astype(np.uint32)))
        self._split_cache = (ranks, parts)
        return parts

    def update_nodes(self, ranks: np.ndarray, position=None, scale=None, rotation=None) -> None:
        """SetPosition / SetScale / SetRotation on a dirty set of GLOBAL ranks, enqueued per range on its stream."""
        for (sel, local), h, st in zip(self._split(ranks), self.handles, self.streams):
            if sel.shape[0] == 0:
                continue
            pick = lambda a: None if a is None else np.ascontiguousarray(a[sel])
            h.update_nodes(local, position=pick(position), scale=pick(scale), rotation=pick(rotation), stream=st)

    def transform_cull(self, frustums: np.ndarray) -> None:
        for h, st in zip(self.handles, self.streams):
            h.transform_cull(frustums, st)
        self.ncams = len(np.atleast_1d(frustums))

    def frame(self, frustums: np.ndarray, out_ptrs: list[int], capacities: list[int], dirty_ranks=None, position=None,
              scale=None, rotation=None) -> np.ndarray:
        """One whole frame: per range (update of its dirty rows ->) transform + cull, enqueued range after range so
        that range 0 is already running while the host still prepares range 1.., then the read-back."""
        parts = self._split(dirty_ranks) if dirty_ranks is not None else [None] * len(self.handles)
        fr = np.ascontiguousarray(np.atleast_1d(frustums), dtype=capi.FRUSTUM_DTYPE)
        for part, h, st in zip(parts, self.handles, self.streams):
            if part is not None and part[0].shape[0]:
                sel, local = part
                pick = lambda a: None if a is None else np.ascontiguousarray(a[sel])
                h.update_nodes(local, position=pick(position), scale=pick(scale), rotation=pick(rotation), stream=st)
            h.transform_cull(fr, st)
        self.ncams = fr.shape[0]
        return self.read_lists_into(out_ptrs, capacities)

    def frame_masks(self, frustums: np.ndarray, out_ptr: int, dirty_ranks=None, position=None, scale=None,
                    rotation=None) -> None:
        """The same frame with the visibility read back in its compact form: per range the rank-order bitmasks of
        every view (rn_cull_build_masks), copied to host memory at out_ptr + 4 * mask_offsets[k] as [views, words_k]
        (mask_layout()).  Bit b of word w of range k = global rank ranges[k][0] + 32 w + b.  Blocks until the last
        copy has finished."""
        parts = self._split(dirty_ranks) if dirty_ranks is not None else [None] * len(self.handles)
        fr = np.ascontiguousarray(np.atleast_1d(frustums), dtype=capi.FRUSTUM_DTYPE)
        offs, _ = self.mask_layout(fr.shape[0])
        for k, (part, h, st) in enumerate(zip(parts, self.handles, self.streams)):
            if part is not None and part[0].shape[0]:
                sel, local = part
                pick = lambda a: None if a is None else np.ascontiguousarray(a[sel])
                h.update_nodes(local, position=pick(position), scale=pick(scale), rotation=pick(rotation), stream=st)
            h.transform_cull(fr, st)
            h.build_masks(st)
            h.read_masks_into(int(out_ptr) + 4 * offs[k], fr.shape[0], st)
        self.ncams = fr.shape[0]
        for st in self.streams:
            capi.stream_synchronize(st)

    def mask_layout(self, n_views: int):
        """(word offset of every range's block in the host buffer, total words) for frame_masks."""
        offs, total = [], 0
        for a, b in self.ranges:
            offs.append(total)
            total += n_views * ((b - a + 31) // 32)
        return offs, total

    def read_lists_into(self, out_ptrs: list[int], capacities: list[int]) -> np.ndarray:
        """Per camera one host buffer (pinned for an asynchronous copy): the ranges' lists land one behind the other.
        Returns the total count per camera.  Blocks until the last copy has finished."""
        total = np.zeros(self.ncams, dtype=np.int64)
        for h, st in zip(self.handles, self.streams):
            ptrs = [int(p) + 4 * int(t) for p, t in zip(out_ptrs, total)]
            caps = [max(int(c) - int(t), 0) for c, t in zip(capacities, total)]
            total += h.read_lists_into(ptrs, caps, st, wait=False)
        for st in self.streams:
            capi.stream_synchronize(st)
        return total
