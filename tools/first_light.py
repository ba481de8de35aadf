"""First-light check on a GPU box: parity spot checks + timings.  Not a test, not the bench."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import Backend, FLAG_EXACT
from oracle.loader import PortOracle, RNG_PHILOX

P = PortOracle()
be = Backend(0)
print("probe", be.fp32_peak_probe())
W, H = 320, 180
sc = scenes.huge_scene(W, H)
be.upload_scene(sc.with_camera(aperture=0.0)); be.framebuffer(W, H)
gi, gt, gn = be.primary_hits(flags=FLAG_EXACT)
oi, ot, on = P.primary_hits(sc.with_camera(aperture=0.0), W, H, nthreads=8)
print("primary idx equal", np.array_equal(gi, oi), "t equal", np.array_equal(gt.view(np.uint32), ot.view(np.uint32)),
      "n equal", np.array_equal(gn.view(np.uint32), on.view(np.uint32)))
gi2, gt2, gn2 = be.primary_hits(flags=0)
print("fast primary idx mismatches", int((gi2 != oi).sum()), "max rel t", float(np.max(np.abs(gt2-ot)/np.maximum(np.abs(ot),1e-30))))
for name, s in (("simple", scenes.simple_scene(64, 36)), ("huge", scenes.huge_scene(64, 36))):
    be.upload_scene(s); be.framebuffer(64, 36); be.reset_stats()
    be.render(4, 50, seed=99, flags=FLAG_EXACT)
    acc = be.readback_accum()
    o = P.render(s, 64, 36, 4, 50, 99, nthreads=8, rng=RNG_PHILOX, want_fixed=True)
    st = be.stats()
    print(name, "accum equal", np.array_equal(acc[..., :3], o["fixed"]), "count", acc[..., 3].min(), acc[..., 3].max(),
          "segments", st["segments"], int(o["stats"][0]), "sky", st["sky"], int(o["stats"][1]))
    if not np.array_equal(acc[..., :3], o["fixed"]):
        d = (acc[..., :3] != o["fixed"]).any(-1)
        print("  mismatching pixels", int(d.sum()), "of", d.size)
    mean, u8 = P.resolve_fixed(o["fixed"], 4)
    print("  rgb8 equal", np.array_equal(be.readback_rgb8(), u8), "f32 equal", np.array_equal(be.readback_f32(), mean))
W, H = 1280, 720
sc = scenes.huge_scene(W, H)
be.upload_scene(sc); be.framebuffer(W, H)
for mode, flags in (("fast", 0), ("exact", FLAG_EXACT)):
    for spp in (4, 16, 64):
        be.clear(); be.reset_stats()
        t0 = time.time(); be.render(spp, 50, seed=1, flags=flags); be.sync(); t1 = time.time()
        st = be.stats()
        n = W * H * spp
        flop = st["segments"] * (18 * sc.n + 64) + st["samples"] * 32
        print(mode, "spp", spp, "wall %.3fs dev %.1f ms" % (t1 - t0, st["render_ms"]), "Msamples/s %.1f" % (n / st["render_ms"] / 1e3),
              "seg/sample %.2f" % (st["segments"] / st["samples"]), "TFLOP/s %.2f" % (flop / st["render_ms"] / 1e9), st["sky"] / n, st["capped"] / n)
img = be.readback_rgb8()
from raytracinginoneweekend_b200.backend import write_ppm
os.makedirs("gpurun_out", exist_ok=True)
write_ppm("gpurun_out/huge_first_light.ppm", img)
print("mean rgb", img.reshape(-1, 3).mean(0))
