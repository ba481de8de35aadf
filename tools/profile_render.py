"""Small driver for ncu: a few renders of the bench scene at reduced spp (same kernel, same scene)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import Backend, FLAG_EXACT, FLAG_FILTER_FMA
name = sys.argv[1] if len(sys.argv) > 1 else "huge"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
mode = sys.argv[3] if len(sys.argv) > 3 else ""
flags = {"exact": FLAG_EXACT, "fma": FLAG_FILTER_FMA}.get(mode, 0)
W, H = 1280, 720
be = Backend(0)
be.upload_scene(scenes.get_scene(name, W, H)); be.framebuffer(W, H)
for i in range(4):
    be.clear(); be.reset_stats(); be.render(spp, 50, seed=1, flags=flags); be.sync()
    st = be.stats()
    print(i, "ms", st["render_ms"], "Msamples/s", W * H * spp / st["render_ms"] / 1e3, "seg/sample", st["segments"] / st["samples"])
