"""Host-side multi-GPU logic on CPU: tile grid, guided stealing scheduler, and the N>1 frame gather
with torch.distributed (gloo, world_size 2).  The renderer is a fake that paints tiles with a pattern."""
import os
import socket

import numpy as np
import pytest

from raytracinginoneweekend_b200 import multigpu


class LocalQueue:
    def __init__(self):
        self.c = [0] * 64

    def reset(self, slot, value):
        self.c[slot] = value

    def next(self, slot, n):
        prev = self.c[slot]
        self.c[slot] += n
        return prev


def test_make_tiles_covers_image_once():
    for W, H, tw, th in ((1280, 720, 64, 64), (3840, 2160, 128, 64), (33, 7, 8, 4), (5, 5, 64, 64)):
        t = multigpu.make_tiles(W, H, tw, th)
        cover = np.zeros((H, W), np.int32)
        for x0, y0, w, h in t:
            assert 0 < w <= tw and 0 < h <= th
            cover[y0:y0 + h, x0:x0 + w] += 1
        assert (cover == 1).all()


def test_scheduler_hands_out_every_tile_once_with_stealing():
    q = LocalQueue()
    world, ntiles = 4, 203
    scheds = [multigpu.StealingScheduler(q, r, world, ntiles) for r in range(world)]
    for s in scheds:
        s.reset_own()
    seen = np.zeros(ntiles, np.int32)
    # rank 0 is "slow": it pulls once, the others drain everything and must steal from it
    start, count, victim = scheds[0].next_batch()
    seen[start:start + count] += 1
    assert victim == 0
    active = [1, 2, 3]
    while active:
        for r in list(active):
            b = scheds[r].next_batch()
            if b is None:
                active.remove(r)
                continue
            seen[b[0]:b[0] + b[1]] += 1
    assert scheds[0].next_batch() is None
    assert (seen == 1).all()
    assert sum(s.stolen for s in scheds[1:]) > 0
    # guided chunks: a rank needs O(log n) launches for its own range
    q2 = LocalQueue()
    s = multigpu.StealingScheduler(q2, 0, 1, 1024)
    s.reset_own()
    n = 0
    while s.next_batch() is not None:
        n += 1
    assert n <= 40


class FakeRenderer:
    """Paints pixel (x, y) with a rank-independent pattern, so the gathered frame has a known answer."""

    def __init__(self, W, H):
        self.img = np.zeros((H, W, 3), np.uint8)
        self.launches = 0

    @staticmethod
    def pattern(W, H):
        y, x = np.mgrid[0:H, 0:W]
        return np.stack([(x * 7 + y) % 251, (x + y * 13) % 241, (x ^ y) % 239], -1).astype(np.uint8)

    def render_tiles(self, tiles):
        H, W = self.img.shape[:2]
        pat = self.pattern(W, H)
        for x0, y0, w, h in tiles:
            assert (self.img[y0:y0 + h, x0:x0 + w] == 0).all(), "tile rendered twice on this rank"
            self.img[y0:y0 + h, x0:x0 + w] = pat[y0:y0 + h, x0:x0 + w]
        self.launches += 1

    def throttle(self, n):
        pass


def _worker(rank, world, port, qname, W, H, result):
    import torch
    import torch.distributed as dist
    from raytracinginoneweekend_b200.backend import TileQueue
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q = TileQueue(qname, create=True) if rank == 0 else None
        dist.barrier()                                # the shared-memory object exists before anyone opens it
        if rank != 0:
            q = TileQueue(qname, create=False)
        tiles = multigpu.make_tiles(W, H, 16, 8)
        sched = multigpu.StealingScheduler(q, rank, world, len(tiles))
        r = FakeRenderer(W, H)
        for _frame in range(2):                       # two frames: the cursors are rewound correctly
            r.img[:] = 0
            done = multigpu.render_distributed(r, sched, tiles, dist.barrier)
            frame = torch.from_numpy(r.img.copy())
            dist.reduce(frame, dst=0, op=dist.ReduceOp.SUM)
            counts = torch.tensor([done])
            dist.all_reduce(counts)
            if rank == 0:
                ok = np.array_equal(frame.numpy(), FakeRenderer.pattern(W, H))
                result.put((bool(ok), int(counts.item()), len(tiles)))
            dist.barrier()
        q.close()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_tile_stealing_and_gather():
    import torch.multiprocessing as tmp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = tmp.get_context("spawn")
    result = ctx.Queue()
    W, H = 100, 37
    qname = f"/rtiow_gloo_{os.getpid()}"
    procs = [ctx.Process(target=_worker, args=(r, 2, port, qname, W, H, result)) for r in range(2)]
    for p in procs:
        p.start()
    got = [result.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for ok, total, ntiles in got:
        assert ok, "gathered frame differs from the expected pattern"
        assert total == ntiles
