"""Multi-GPU sharding of the scene path: one process per GPU, node ranges cut at subtree
boundaries, no data-path collective until the per-camera visibility lists are gathered.

* ``plan_shards``    - contiguous ranges of render-list RANKS, cut only where no parent/child edge crosses, so every
                       rank's world transforms depend on local nodes alone (no halo, no transform exchange).
* ``extract_shard``  - the node columns of one range with the parent column rebased to local ranks.
* ``ListExchange``   - the path's only collective: all-gather of the per-camera visible counts, then of the index
                       lists (padded to the largest count), over ``torch.distributed`` (NCCL on the GPUs, gloo in the
                       CPU tests).  Lists travel as GLOBAL ranks; concatenating the ranks' lists in rank order IS the
                       reference's render-list order because every rank owns an ascending rank range and its local
                       compaction is order-stable.

Nothing here computes transforms or visibility: that is librn_cuda's job on each GPU.
"""
from __future__ import annotations

import dataclasses

import numpy as np

RN_NO_PARENT = 0xFFFFFFFF


@dataclasses.dataclass
class ShardColumns:
    """What capi.SceneHandle.upload needs (same attribute names as capi.RankedScene) for one rank range."""
    first: int
    position: np.ndarray
    scale: np.ndarray
    rotation: np.ndarray
    parent: np.ndarray
    flags: np.ndarray
    group: np.ndarray
    sphere: np.ndarray

    @property
    def n(self) -> int:
        return int(self.parent.shape[0])


def valid_cuts(parent: np.ndarray) -> np.ndarray:
    """Boolean array c[0..n]: c[r] is True iff no parent/child edge crosses the boundary between ranks r-1 and r."""
    parent = np.asarray(parent)
    n = parent.shape[0]
    child = np.nonzero(parent != RN_NO_PARENT)[0].astype(np.int64)
    par = parent[child].astype(np.int64)
    lo = np.minimum(child, par)
    hi = np.maximum(child, par)
    # the edge (lo, hi) crosses every boundary r with lo < r <= hi
    cover = np.zeros(n + 2, dtype=np.int64)
    np.add.at(cover, lo + 1, 1)
    np.add.at(cover, hi + 1, -1)
    crossing = np.cumsum(cover)[: n + 1]
    return crossing == 0


def plan_shards(parent: np.ndarray, world: int) -> list[tuple[int, int]]:
    """Splits ranks [0, n) into ``world`` contiguous ranges of near-equal size whose boundaries cross no edge.
    Raises ValueError if the hierarchy offers fewer than ``world`` independent pieces."""
    n = int(np.asarray(parent).shape[0])
    if world < 1:
        raise ValueError("world must be >= 1")
    ok = np.nonzero(valid_cuts(parent))[0]          # sorted, always contains 0 and n
    if ok.shape[0] < world + 1:
        raise ValueError(f"the hierarchy cannot be split into {world} edge-free rank ranges "
                         f"(only {ok.shape[0] - 1} independent pieces)")
    idx = [0]
    for k in range(1, world):
        target = (n * k) // world
        lo_i = idx[-1] + 1                          # strictly after the previous cut
        hi_i = ok.shape[0] - 1 - (world - k)        # leave one cut for every later boundary
        i = int(np.searchsorted(ok, target))
        cands = [j for j in (i - 1, i) if lo_i <= j <= hi_i]
        if not cands:
            cands = [min(max(i, lo_i), hi_i)]
        idx.append(min(cands, key=lambda j: abs(int(ok[j]) - target)))
    bounds = [int(ok[j]) for j in idx] + [n]
    return list(zip(bounds[:-1], bounds[1:]))


def extract_shard(rs, first: int, last: int) -> ShardColumns:
    """Columns of ranks [first, last) of a capi.RankedScene (or anything with its attributes)."""
    p = np.asarray(rs.parent[first:last])
    has = p != RN_NO_PARENT
    if has.any():
        inside = (p[has] >= first) & (p[has] < last)
        if not inside.all():
            raise ValueError("a parent/child edge crosses the shard boundary: use plan_shards")
    local = np.where(has, p.astype(np.int64) - first, RN_NO_PARENT).astype(np.uint32)
    sl = slice(first, last)
    return ShardColumns(first, np.ascontiguousarray(rs.position[sl]), np.ascontiguousarray(rs.scale[sl]),
                        np.ascontiguousarray(rs.rotation[sl]), local, np.ascontiguousarray(rs.flags[sl]),
                        np.ascontiguousarray(rs.group[sl]), np.ascontiguousarray(rs.sphere[sl]))


class ListExchange:
    """All-gather of per-camera visibility lists across the ranks of a torch.distributed process group."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist = dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def gather_counts(self, counts):
        """counts: int32 tensor [ncams] on the collective's device -> int32 tensor [world, ncams]."""
        import torch
        out = torch.empty(self.world * counts.numel(), dtype=counts.dtype, device=counts.device)
        self.dist.all_gather_into_tensor(out, counts.contiguous(), group=self.group)
        return out.view(self.world, counts.numel())

    def gather_lists(self, lists, all_counts_host):
        """lists[c]: 1-D int32 tensor with at least max-over-ranks(count[c]) readable entries (global ranks, only the
        first count[c] are meaningful).  all_counts_host: [world, ncams] on the host.
        Returns per camera a list of ``world`` views (rank r's visible global ranks), i.e. the concatenation order."""
        import torch
        result = []
        for c, src in enumerate(lists):
            m = int(all_counts_host[:, c].max())
            if m == 0:
                result.append([src[:0] for _ in range(self.world)])
                continue
            dst = torch.empty(self.world * m, dtype=src.dtype, device=src.device)
            self.dist.all_gather_into_tensor(dst, src[:m].contiguous(), group=self.group)
            result.append([dst[r * m: r * m + int(all_counts_host[r, c])] for r in range(self.world)])
        return result

    def exchange(self, counts, lists):
        """Full exchange.  Returns (all_counts_host [world, ncams], per-camera list of per-rank views)."""
        all_counts = self.gather_counts(counts).cpu()
        return all_counts, self.gather_lists(lists, all_counts)


class PipelinedListExchange(ListExchange):
    """The same exchange, one frame deep: frame i's lists are snapshotted right behind its compaction and gathered on
    a side stream while frame i+1 is transformed and culled on the main stream.

    No host synchronisation per frame: the gather sizes of frame i come from the newest counts table the host has
    seen (frame i-depth; identical on every rank, because every rank holds the same table) times ``margin``; the
    table of a frame is read on the host ``depth`` frames later, when it has long arrived, so the host never waits
    and keeps the device queue full.  A camera whose visible count grew beyond the margin within one frame is
    reported through ``overflows`` (its snapshot was cut) - the caller falls back to ``exchange`` for that frame.
    ``flush`` makes the current stream wait for the outstanding gathers and returns the last result.
    With CPU tensors (gloo, the CPU tests) the same protocol runs without streams."""

    def __init__(self, group=None, margin: float = 1.125, slack: int = 4096, cuda: bool = True, depth: int = 2):
        super().__init__(group)
        import torch
        self.torch = torch
        self.cuda = cuda
        self.side = torch.cuda.Stream() if cuda else None
        self.margin, self.slack = margin, slack
        self.sizes = None        # per camera: entries gathered per rank
        self.capacities = None
        self.depth = max(1, depth)   # frames in flight before the host looks at the oldest one's counts table
        self.pending = []
        self.result = None       # (counts table [world, ncams] on the host, gathered buffer, sizes used)
        self.overflows = 0
        self._host = [None] * (self.depth + 1)    # pinned count tables, one per frame in flight
        self._frame = 0

    def _set_sizes(self, table):
        mx = table.max(dim=0).values.tolist()
        self.sizes = [int(min(cap, int(m * self.margin) + self.slack)) for m, cap in zip(mx, self.capacities)]

    def _retire(self, keep: int):
        """Host side of the oldest frames in flight (all but the newest `keep`): their counts tables have arrived."""
        while len(self.pending) > keep:
            self._retire_one(self.pending.pop(0))

    def _retire_one(self, item):
        ev, host_counts, outs, used_sizes, _keep = item
        if ev is not None:
            ev.synchronize()
        table = host_counts.view(self.world, -1).clone().to(self.torch.int64)
        if any(int(table[:, c].max()) > used_sizes[c] for c in range(table.shape[1])):
            self.overflows += 1
        self.result = (table, outs, used_sizes)
        self._set_sizes(table)

    def submit(self, counts, lists):
        """counts: int32 tensor [ncams]; lists[c]: int32 tensor (full capacity) of camera c.
        ONE collective per frame: the first sizes[c] entries of every list and the counts are packed into one buffer
        (one copy kernel) and all-gathered as a whole; the counts columns of the result go to the host asynchronously."""
        torch = self.torch
        self.capacities = [int(l.numel()) for l in lists]
        if self.sizes is None:                       # first frame: one synchronous exchange seeds the size table
            table, _ = self.exchange(counts, lists)
            self._set_sizes(table)
        self._retire(self.depth - 1)
        sizes = list(self.sizes)
        ncam = counts.numel()
        packed = torch.cat([l[:m] for l, m in zip(lists, sizes)] + [counts.to(lists[0].dtype)])
        width = packed.numel()

        def gather():
            dst = torch.empty(self.world * width, dtype=packed.dtype, device=packed.device)
            self.dist.all_gather_into_tensor(dst, packed, group=self.group)
            tail = dst.view(self.world, width)[:, width - ncam:].contiguous()
            if self.cuda:
                self._frame += 1
                slot = self._frame % len(self._host)
                host_counts = self._host[slot]
                if host_counts is None or host_counts.numel() != tail.numel():
                    host_counts = torch.empty(tail.numel(), dtype=tail.dtype).pin_memory()
                    self._host[slot] = host_counts
                host_counts.copy_(tail.view(-1), non_blocking=True)
                ev_counts = torch.cuda.Event()
                ev_counts.record()
            else:
                host_counts, ev_counts = tail.view(-1), None
            return ev_counts, host_counts, dst

        if self.cuda:
            ev = torch.cuda.Event()
            ev.record()
            packed.record_stream(self.side)
            with torch.cuda.stream(self.side):
                self.side.wait_event(ev)
                ev_counts, host_counts, dst = gather()
        else:
            ev_counts, host_counts, dst = gather()
        self.pending.append((ev_counts, host_counts, dst, sizes, packed))

    def flush(self):
        """Current stream waits for every outstanding gather; returns (counts table, gathered buffer, sizes used)."""
        if self.cuda:
            self.torch.cuda.current_stream().wait_stream(self.side)
        self._retire(0)
        return self.result

    @staticmethod
    def views(result, world: int):
        """Per camera the list of per-rank views (rank r's visible global ranks) of a flushed result."""
        table, dst, sizes = result
        ncam = len(sizes)
        width = sum(sizes) + ncam
        offs = [sum(sizes[:c]) for c in range(ncam)]
        return [[dst[r * width + offs[c]: r * width + offs[c] + min(int(table[r, c]), sizes[c])] for r in range(world)]
                for c in range(ncam)]


class PipelinedMaskExchange:
    """The exchange in its compact form: instead of the lists (4 B per visible node and camera, 61 MB per GPU for the
    benchmark frame) every rank ships the rank-order visibility bitmasks of its range (rn_cull_build_masks: one bit
    per node and camera, 12.6 MB) - the same information, since a list of the cull loop is the ascending sequence of
    the ranks that passed.  Sizes are static (fixed by the shard plan), so there is no counts table, no size
    prediction and no overflow: ONE fixed-size all_gather_into_tensor per frame on a side stream, ``depth`` frames in
    flight, no host synchronisation.  The consumer expands the cameras it needs with rn_masks_expand
    (``expand``), rank range after rank range, which reproduces the single-scene list bit for bit.

    words_per_rank[r] = ceil(nodes of rank r / 32); every rank contributes ncams rows padded to the largest."""

    def __init__(self, words_per_rank, ncams: int, group=None, cuda: bool = True, depth: int = 2):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        assert len(words_per_rank) == self.world
        self.words_per_rank = [int(w) for w in words_per_rank]
        self.row = max(self.words_per_rank)            # padded words per camera
        self.ncams = int(ncams)
        self.cuda = cuda
        self.side = torch.cuda.Stream() if cuda else None
        self.depth = max(1, depth)
        self.pending = []
        self.result = None
        self._bufs = [None] * (self.depth + 1)        # gathered buffers, recycled round robin
        self._send = [None] * (self.depth + 1)
        self._frame = 0

    def begin(self):
        """Send buffer of the coming frame, [ncams, row] int32 on the device: the producer may write its masks straight
        into it (rn_cull_build_masks_to with row stride ``row``) and then call ``submit()`` without an argument."""
        torch = self.torch
        k = self._frame % len(self._bufs)
        while len(self.pending) > self.depth - 1:      # the buffers of frame i-depth are free again
            self._retire_one(self.pending.pop(0))
        if self._send[k] is None:
            dev = "cuda" if self.cuda else "cpu"
            self._send[k] = torch.zeros(self.ncams * self.row, dtype=torch.int32, device=dev)
            self._bufs[k] = torch.empty(self.world * self.ncams * self.row, dtype=torch.int32, device=dev)
        self._begun = True
        return self._send[k].view(self.ncams, self.row)

    def submit(self, masks=None):
        """masks: int32 tensor [ncams * words_local] (camera rows one behind the other, e.g. the library's buffer) -
        copied into the send buffer - or None if the producer already filled the buffer ``begin()`` returned."""
        torch = self.torch
        wl = self.words_per_rank[self.rank]
        if not getattr(self, "_begun", False):
            self.begin()
        self._begun = False
        k = self._frame % len(self._bufs)
        self._frame += 1
        send, dst = self._send[k], self._bufs[k]
        if masks is not None:
            # snapshot on the producing stream (the next cull overwrites the library's buffer), padded rows
            send.view(self.ncams, self.row)[:, :wl].copy_(masks.view(self.ncams, wl))
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record()
            with torch.cuda.stream(self.side):
                self.side.wait_event(ev)
                self.dist.all_gather_into_tensor(dst, send, group=self.group)
                done = torch.cuda.Event()
                done.record()
        else:
            self.dist.all_gather_into_tensor(dst, send, group=self.group)
            done = None
        self.pending.append((done, dst))

    def _retire_one(self, item):
        done, dst = item
        if done is not None:
            self.torch.cuda.current_stream().wait_event(done)   # device-side order only; the host does not block
        self.result = dst

    def flush(self):
        """Current stream waits for every outstanding gather; returns the newest gathered buffer
        [world, ncams, row] (rank r's words of camera c: result[r, c, :words_per_rank[r]])."""
        while self.pending:
            self._retire_one(self.pending.pop(0))
        return None if self.result is None else self.result.view(self.world, self.ncams, self.row)

    def expand(self, handle, gathered, camera: int, starts, out_list, out_count, stream: int = 0):
        """rn_masks_expand over the rank ranges in order: out_list (int32 device tensor) receives camera's global list,
        out_count (int32 device tensor [1], zeroed here) its length.  starts[r] = first global rank of rank r."""
        out_count.zero_()
        segs = [(gathered[r, camera].data_ptr(), self.words_per_rank[r], int(starts[r])) for r in range(self.world)]
        handle.masks_expand(segs, out_list.data_ptr(), out_list.numel(), out_count.data_ptr(), stream)


class MulticastMaskExchange:
    """The mask exchange with NO collective kernel: the gathered buffer lives in symmetric memory
    (torch.distributed._symmetric_memory: the same allocation on every GPU of the group, mapped behind one NVLS
    multicast address), and rn_cull_build_masks_multicast stores every mask word through that address - the NVSwitch
    replicates it into all GPUs' copies while the kernel is still computing.  What is left of the "all-gather" is a
    barrier between the GPUs (signal pads of the symmetric memory, one tiny kernel on a side stream), two frames deep
    like the NCCL variants, so the transform kernel of the next frame never shares the SMs with a collective.

    Buffer: [slots = depth + 1][world][ncams][row] int32; frame i uses slot i % slots.  Ordering contract (same as
    PipelinedMaskExchange): a consumer of frame j's gathered masks is enqueued on the producing stream before the
    masks of frame j + 1 are built; then slot reuse is safe, because building frame i waits for the barrier of frame
    i - depth, which every GPU only reaches after everything it enqueued before building frame i - depth + 1."""

    def __init__(self, words_per_rank, ncams: int, group=None, depth: int = 2):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.torch, self.dist = torch, dist
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.words_per_rank = [int(w) for w in words_per_rank]
        self.row = max(self.words_per_rank)
        self.ncams = int(ncams)
        self.depth = max(1, depth)
        self.slots = self.depth + 1
        self.block = self.ncams * self.row                        # words per rank and frame
        dev = torch.device("cuda", torch.cuda.current_device())
        self.buf = symm.empty(self.slots * self.world * self.block, dtype=torch.int32, device=dev)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, self.group)
        self.multicast_ptr = int(self.hdl.multicast_ptr or 0)
        if not self.multicast_ptr:
            raise RuntimeError("no NVLS multicast mapping for the symmetric buffer on this system")
        self.side = torch.cuda.Stream()
        self.pending = []
        self.result = None
        self._frame = 0
        torch.cuda.synchronize()
        self.hdl.barrier(channel=0)                                # every copy is zeroed before the first multicast store
        torch.cuda.synchronize()

    def begin(self) -> int:
        """Multicast address of this rank's block of the coming frame (row stride = ``row`` words); the current stream
        first waits for the barrier of frame i - depth (device side only)."""
        while len(self.pending) > self.depth - 1:
            self._retire_one(self.pending.pop(0))
        k = self._frame % self.slots
        return self.multicast_ptr + 4 * ((k * self.world + self.rank) * self.block)

    def submit(self):
        """After rn_cull_build_masks_multicast(begin(), row) on the current stream: the barrier that makes the frame's
        gathered masks readable, on the side stream."""
        torch = self.torch
        k = self._frame % self.slots
        self._frame += 1
        ev = torch.cuda.Event()
        ev.record()
        with torch.cuda.stream(self.side):
            self.side.wait_event(ev)
            self.hdl.barrier(channel=1 + k)
            done = torch.cuda.Event()
            done.record()
        self.pending.append((done, k))

    def _retire_one(self, item):
        done, k = item
        self.torch.cuda.current_stream().wait_event(done)
        self.result = k

    def flush(self):
        """Current stream waits for every outstanding barrier; returns the newest frame's gathered masks
        [world, ncams, row] (a view into the symmetric buffer)."""
        while self.pending:
            self._retire_one(self.pending.pop(0))
        if self.result is None:
            return None
        k = self.result
        return self.buf.view(self.slots, self.world, self.ncams, self.row)[k]

    expand = PipelinedMaskExchange.expand
