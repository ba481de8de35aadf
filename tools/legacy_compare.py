"""Legacy GPU comparator, GPU box: the reference's own CUDA path (oracle/_ref/ref_cuda_legacy, built
by oracle/build_ref_cuda.sh from /root/reference/src/CUDA/cuda_impl.cu) next to this backend on the same workload
shape: the 5-sphere scene, 1280x720 (or --width/--height), 48 spp, depth 32.  Informational: the two differ in RNG,
lambert/sky conventions and lens, so images are compared only by mean level.
usage: python tools/legacy_compare.py [--width W --height H --reps R --out profiles/x.json]"""
import os, sys, json, time, argparse, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=1280); ap.add_argument("--height", type=int, default=720)
ap.add_argument("--reps", type=int, default=5); ap.add_argument("--out", default="")
a = ap.parse_args()
exe = os.path.join(ROOT, "oracle", "_ref", "ref_cuda_legacy")
res = {"workload": {"scene": "simple (5 spheres)", "width": a.width, "height": a.height, "spp": 48, "max_depth": 32}}
if os.path.exists(exe):
    p = subprocess.run([exe, "--width", str(a.width), "--height", str(a.height), "--reps", str(a.reps)],
                       capture_output=True, text=True, timeout=600)
    line = [l for l in p.stdout.splitlines() if l.startswith("{")]
    res["reference_cuda"] = json.loads(line[-1]) if line else {"error": (p.stderr or p.stdout)[-400:]}
else:
    res["reference_cuda"] = {"unavailable": "oracle/_ref/ref_cuda_legacy not built (needs /root/reference at build time)"}

import numpy as np
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import Backend
sc = scenes.get_scene("simple", a.width, a.height)
be = Backend(0)
packed = None
best_e2e = 1e30; best_dev = 1e30
for r in range(a.reps + 1):                       # first call = warm-up (context, allocations)
    be.reset_stats()
    t0 = time.perf_counter()
    img = be.render_frame(sc, a.width, a.height, 48, max_depth=32, seed=1)   # upload + render + resolve + D2H, like cuda_impl()
    t1 = time.perf_counter()
    if r:
        best_e2e = min(best_e2e, (t1 - t0) * 1e3); best_dev = min(best_dev, be.stats()["last_render_ms"])
n = a.width * a.height * 48
res["this_backend"] = {"impl": "raytracinginoneweekend_b200", "call": "rt_render_frame (host scene in, host RGB8 out)",
                       "best_ms_end_to_end": best_e2e, "best_ms_render_kernel": best_dev,
                       "Gsamples_per_s_end_to_end": n / best_e2e / 1e6, "Gsamples_per_s_kernel": n / best_dev / 1e6,
                       "mean_texel": float(np.asarray(img, dtype=np.float64).mean())}
if "best_ms" in res["reference_cuda"]:
    res["speedup_end_to_end"] = res["reference_cuda"]["best_ms"] / best_e2e
print(json.dumps(res))
if a.out:
    json.dump(res, open(a.out, "w"), indent=1)
