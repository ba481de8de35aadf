"""One GPU plays rank 0 of an 8-rank job: its share of the tiles rendered as ONE launch, and as the chunk schedules of
multigpu.StealingScheduler (own range only, nothing to steal), at 2..3 launches in flight.  Shows what chunking costs.
usage: python tools/chunk_overhead.py [scene] [W] [H] [spp] [world]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from raytracinginoneweekend_b200 import scenes, multigpu
from raytracinginoneweekend_b200.backend import Backend, FLAG_OVERLAP

name = sys.argv[1] if len(sys.argv) > 1 else "synthetic1m"
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 256
world = int(sys.argv[5]) if len(sys.argv) > 5 else 8
TW, TH = 128, 64
be = Backend(0)
be.upload_scene(scenes.get_scene(name, W, H)); be.framebuffer(W, H)
tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, TW, TH), world)
lo, hi = multigpu.rank_ranges(len(tiles), world)[0]
mine = tiles[lo:hi]


last = {}


def run(chunks, in_flight, flags=0):
    be.clear(); be.sync(); be.reset_stats()
    t0 = time.perf_counter()
    s = 0
    host = []
    for k in chunks:
        h0 = time.perf_counter()
        be.render_tiles(mine[s:s + k], TW, TH, spp, 50, 1, 0, flags)
        host.append((time.perf_counter() - h0) * 1e3)
        s += k
        be.throttle(in_flight - 1)
    be.throttle(0); be.sync()
    wall = (time.perf_counter() - t0) * 1e3
    st = be.stats()
    last.update(sum_launch_ms=st["render_ms"], host_ms=sum(host), host_max=max(host))
    return wall


def schedule(n, frac):
    out = []
    while n > 0:
        k = max(1, (n + frac - 1) // frac)
        out.append(k); n -= k
    return out


n = len(mine)
print(f"{name} {W}x{H} {spp} spp: rank 0 of {world} owns {n} tiles")
for rep in range(2):
    print('--- rep', rep)
    print(f"one launch: {run([n], 1):.1f} ms   {last}")
    for k in (1, 2, 4, 8):
        print(f"  a single launch of {k} tiles: {run([k], 1):.2f} ms  ({last['sum_launch_ms']:.2f} ms on the GPU)")
    for frac in (2, 3, 5):
        ch = schedule(n, frac)
        for fl in (2, 3):
            a = run(ch, fl); la = dict(last)
            b = run(ch, fl, FLAG_OVERLAP)
            print(f"1/{frac} of the rest ({len(ch)} launches), {fl} in flight: {a:.1f} ms (GPU sum {la['sum_launch_ms']:.1f}, host {la['host_ms']:.1f}, max {la['host_max']:.2f}), "
                  f"overlapped {b:.1f} ms (GPU sum {last['sum_launch_ms']:.1f}, host {last['host_ms']:.1f})")
be.close()
