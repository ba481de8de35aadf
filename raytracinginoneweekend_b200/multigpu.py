"""One process per GPU: tile sharding with dynamic stealing + framebuffer gather to rank 0.

The reference has no multi-GPU path; this is new work that BASELINE.json's
north_star asks for.  Pixels are independent, so the only exchange is the final framebuffer:

* the image is cut into small tiles; the tile list is interleaved (rank r's range holds every N-th tile, so
  every range sees the image's cost mix) and split into one contiguous range per rank;
* every rank owns a cursor (one cache line in POSIX shared memory, ``rt_queue_*``); it takes guided
  chunks from its own range first (a fifth of what is left, so a rank needs only O(log n) persistent
  launches) and, when that is dry, steals chunks from the other ranks' cursors -- the path-length
  tail and the sky/glass cost gap make per-tile cost vary by >10x, static splits do not balance;
* each chunk is ONE ``rt_render_tiles`` launch (a persistent kernel over the tile list); at most
  two chunks are in flight per GPU, so stealing reacts to real progress;
* every rank resolves its tiles to RGB8 (pixels it did not render stay 0) and the frame is gathered on
  rank 0 with one NCCL ``reduce(SUM)`` over NVLink -- tiles are disjoint, so the sum is the union.

No data-path collective other than that gather.  The scheduler is CUDA-free and is exercised on CPU
with the gloo backend (tests/test_multigpu_cpu.py).
"""
from __future__ import annotations

import numpy as np

FLAG_OVERLAP = 1 << 6   # RT_FLAG_OVERLAP (include/rt_api.h): a chunk may run beside the chunks queued before it


def make_tiles(W: int, H: int, tw: int, th: int) -> np.ndarray:
    """Row-major grid of {x0,y0,w,h} tiles covering W x H; edge tiles are smaller."""
    xs = np.arange(0, W, tw, dtype=np.int64)
    ys = np.arange(0, H, th, dtype=np.int64)
    out = np.empty((len(ys) * len(xs), 4), np.uint32)
    k = 0
    for y in ys:
        for x in xs:
            out[k] = (x, y, min(tw, W - x), min(th, H - y))
            k += 1
    return out


def interleave_tiles(tiles: np.ndarray, world: int) -> np.ndarray:
    """Re-order the tile list so that rank r's contiguous range [r*T/N, (r+1)*T/N) holds every N-th tile of the
    image: each rank's own range then sees the image's cost mix (sky is ~10x cheaper than glass), and stealing
    only has to fix the residual."""
    if world <= 1:
        return tiles
    return np.concatenate([tiles[r::world] for r in range(world)], axis=0)


def rank_ranges(ntiles: int, world: int):
    """Contiguous [lo, hi) range of tile indices per rank."""
    return [(r * ntiles // world, (r + 1) * ntiles // world) for r in range(world)]


class StealingScheduler:
    """Guided self-scheduling on per-rank cursors with stealing.

    ``queue`` needs ``reset(slot, value)`` and ``next(slot, n) -> previous`` (fetch_add; n=0 reads):
    ``backend.TileQueue`` in production, any in-process fake in tests.
    """

    # a rank takes 1/OWN_FRACTION of what is left of its own range per pull.  With two chunks in flight it has
    # committed 1 - (1 - 1/5)^2 = 36 % of its range before any other rank can help: a GPU that runs at half speed (seen
    # once on a 4-GPU box: 6.3 instead of 7.6 Gsamples/s with thirds, where 56 % is committed) then costs a few per cent
    # instead of 15 %; the price is ~25 instead of ~14 persistent launches per rank and frame (< 1 % in kernel tails).
    OWN_FRACTION = 5

    def __init__(self, queue, rank: int, world: int, ntiles: int, min_chunk: int = 1):
        self.q, self.rank, self.world, self.ntiles = queue, rank, world, ntiles
        self.ranges = rank_ranges(ntiles, world)
        self.min_chunk = max(1, int(min_chunk))
        self.own_taken = 0
        self.stolen = 0

    def reset_own(self):
        """Every rank rewinds its own cursor; the caller must barrier before anyone pulls."""
        self.q.reset(self.rank, self.ranges[self.rank][0])
        self.own_taken = 0
        self.stolen = 0

    def _take(self, victim: int):
        lo, hi = self.ranges[victim]
        cur = self.q.next(victim, 0)
        remaining = hi - cur
        if remaining <= 0:
            return None
        if victim == self.rank:
            k = max(self.min_chunk, (remaining + self.OWN_FRACTION - 1) // self.OWN_FRACTION)
        else:
            k = max(self.min_chunk, remaining // (2 * self.world))
        start = self.q.next(victim, k)
        if start >= hi:
            return None
        return start, min(k, hi - start)

    def next_batch(self):
        """(start, count, victim_rank) of the next chunk of tile indices, or None when all is taken."""
        order = [self.rank] + [(self.rank + d) % self.world for d in range(1, self.world)]
        for victim in order:
            got = self._take(victim)
            if got is not None:
                if victim == self.rank:
                    self.own_taken += got[1]
                else:
                    self.stolen += got[1]
                return got[0], got[1], victim
        return None


def render_distributed(renderer, sched: StealingScheduler, tiles: np.ndarray, barrier, max_in_flight: int = 2):
    """Drive one frame on this rank.  ``renderer.render_tiles(tile_array)`` submits one asynchronous
    launch; ``renderer.throttle(n)`` blocks until at most n launches are in flight.
    Returns the number of tiles this rank rendered."""
    sched.reset_own()
    barrier()                       # all cursors are rewound before anyone pulls
    done = 0
    while True:
        b = sched.next_batch()
        if b is None:
            break
        start, count, _ = b
        renderer.render_tiles(tiles[start:start + count])
        done += count
        renderer.throttle(max_in_flight - 1)
    renderer.throttle(0)
    barrier()                       # nobody rewinds a cursor while another rank may still be stealing from it
    return done


class CudaTileRenderer:
    """Adapter: backend.Backend -> the renderer protocol of render_distributed."""

    def __init__(self, backend, tile_w, tile_h, spp, max_depth, seed, flags=0):
        self.be, self.tw, self.th = backend, tile_w, tile_h
        self.spp, self.depth, self.seed, self.flags = spp, max_depth, seed, flags

    def render_tiles(self, tiles):
        # chunks of one frame are disjoint: let each launch start beside the tail of the previous one (RT_FLAG_OVERLAP)
        self.be.render_tiles(tiles, self.tw, self.th, self.spp, self.depth, self.seed, 0, self.flags | FLAG_OVERLAP)

    def throttle(self, n):
        self.be.throttle(n)


class DeviceQueueFrame:
    """One process per GPU on the DEVICE-SIDE tile queue (include/rt_api.h): rank 0 owns the range cursors and the RGB8
    frame; the other ranks map both through CUDA IPC (handles travel once through torch.distributed).  A frame is then

        rank 0: rt_queue_dev_reset                      (stream-ordered)
        all   : all_reduce(token)                       cursors are rewound before anyone pulls; nobody is in the previous frame
        all   : rt_render_queue                         ONE persistent launch: the kernel takes tiles itself, steals when dry
        all   : rt_resolve_tiles_to(rank 0's frame)     the gather: texels go straight into rank 0's memory over NVLink
        all   : all_reduce(token)                       NCCL acknowledgement: the frame is complete on rank 0's stream

    all of it queued on one CUDA stream per rank -- no host synchronisation inside a frame, no per-chunk launches."""

    def __init__(self, backend, rank, world, dist, torch, stream):
        self.be, self.rank, self.world, self.dist, self.torch, self.stream = backend, rank, world, dist, torch, stream
        self.frame = None
        self._frame_mapped = False
        if rank == 0:
            self.cursors = backend.queue_dev_create()
            box = [backend.ipc_export(self.cursors)] if world > 1 else [None]
        else:
            box = [None]
        if world > 1:
            dist.broadcast_object_list(box, src=0)
        if rank != 0:
            self.cursors = backend.ipc_import(box[0])
        self.token = torch.zeros(1, dtype=torch.int32, device=f"cuda:{backend.device}")

    def map_frame(self):
        """(Re)map rank 0's RGB8 frame; call after every rt_framebuffer on rank 0 (a resize reallocates it)."""
        be = self.be
        if self.rank != 0 and self._frame_mapped:
            be.ipc_close(self.frame)
            self._frame_mapped = False
        if self.rank == 0:
            ptr, _ = be.device_rgb8()
            box = [be.ipc_export(ptr)] if self.world > 1 else [None]
            self.frame = ptr
        else:
            box = [None]
        if self.world > 1:
            self.dist.broadcast_object_list(box, src=0)
        if self.rank != 0:
            self.frame = be.ipc_import(box[0])
            self._frame_mapped = True

    def _ack(self):
        if self.world > 1:
            with self.torch.cuda.stream(self.stream):
                self.dist.all_reduce(self.token)

    def render_frame(self, tiles, tile_w, tile_h, spp, max_depth, seed, flags=0, trace=None):
        """trace: optional list; gets five CUDA events (start, cursors acknowledged, rendered, resolved, frame acknowledged)
        recorded on the stream, for per-phase timing of one rank (bench.py, `ranks[*].phases_ms`)."""
        be = self.be

        def mark():
            if trace is not None:
                ev = self.torch.cuda.Event(enable_timing=True)
                ev.record(self.stream)
                trace.append(ev)
        mark()
        be.clear()
        if self.rank == 0:
            be.queue_dev_reset(self.cursors, len(tiles), self.world)
        self._ack()
        mark()
        be.render_queue(tiles, tile_w, tile_h, self.cursors, self.rank, self.world, spp, max_depth, seed, 0, flags)
        mark()
        be.resolve_tiles_to(self.frame)
        mark()
        self._ack()
        mark()

    def close(self):
        be = self.be
        be.sync()
        if self.rank != 0:
            if self._frame_mapped:
                be.ipc_close(self.frame)
            be.ipc_close(self.cursors)
        else:
            be.queue_dev_free(self.cursors)


def device_rgb8_tensor(backend):
    """Zero-copy torch view (uint8, H*W*3) of the backend's device RGB8 buffer."""
    import torch

    ptr, nbytes = backend.device_rgb8()

    class _Holder:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}

    return torch.as_tensor(_Holder(), device=f"cuda:{backend.device}")


def gather_frame(backend, dist=None, dst: int = 0):
    """Resolve locally, then reduce(SUM) the RGB8 frame to rank `dst` (NCCL over NVLink when world > 1).
    Returns the torch tensor view (valid on `dst`)."""
    backend.resolve_device()
    backend.sync()                   # the resolve ran on the ctx stream; NCCL uses torch's stream
    t = device_rgb8_tensor(backend)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(t, dst=dst, op=dist.ReduceOp.SUM)
    return t
