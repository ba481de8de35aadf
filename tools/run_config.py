"""Run one workload through the C ABI and print timing + path statistics (GPU box)."""
import os, sys, time, json, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import Backend, FLAG_EXACT, FLAG_LBVH, FLAG_BRUTE, FLAG_FILTER_FMA, write_ppm
ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="huge"); ap.add_argument("--width", type=int, default=1280); ap.add_argument("--height", type=int, default=720)
ap.add_argument("--spp", type=int, default=16); ap.add_argument("--depth", type=int, default=50); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--exact", action="store_true"); ap.add_argument("--lbvh", action="store_true"); ap.add_argument("--brute", action="store_true")
ap.add_argument("--fma", action="store_true"); ap.add_argument("--out", default="")
ap.add_argument("--queue", action="store_true", help="render through rt_render_queue (device-side tile queue, this GPU as rank 0 of 1)")
a = ap.parse_args()
flags = (FLAG_EXACT if a.exact else 0) | (FLAG_LBVH if a.lbvh else 0) | (FLAG_BRUTE if a.brute else 0) | (FLAG_FILTER_FMA if a.fma else 0)
t0 = time.time(); sc = scenes.get_scene(a.scene, a.width, a.height); t1 = time.time()
be = Backend(0); be.upload_scene(sc); be.framebuffer(a.width, a.height)
if a.queue:
    from raytracinginoneweekend_b200 import multigpu
    tiles = multigpu.make_tiles(a.width, a.height, 128, 64)
    cur = be.queue_dev_create()
for i in range(a.reps):
    be.clear(); be.reset_stats()
    if a.queue:
        be.queue_dev_reset(cur, len(tiles), 1)
        be.render_queue(tiles, 128, 64, cur, 0, 1, a.spp, a.depth, seed=1, flags=flags); be.sync()
    else:
        be.render(a.spp, a.depth, seed=1, flags=flags); be.sync()
    st = be.stats()
    n = a.width * a.height * a.spp
    print(json.dumps({"rep": i, "scene": a.scene, "n_spheres": sc.n, "ms": st["render_ms"], "Msamples_per_s": n / st["render_ms"] / 1e3,
                      "seg_per_sample": st["segments"] / st["samples"], "nodes_per_seg": st["node_visits"] / max(1, st["segments"]),
                      "leaf_tests_per_seg": st["sphere_tests"] / max(1, st["segments"]), "bvh_build_ms": st["bvh_build_ms"],
                      "capped": st["capped"] / n, "sky": st["sky"] / n, "scene_gen_s": t1 - t0,
                      "node_visits": st["node_visits"], "filter_misses": st["filter_misses"], "segments": st["segments"]}))
if a.out:
    write_ppm(a.out, be.readback_rgb8())
