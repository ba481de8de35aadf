"""GPU box: tensor-filter render vs FMA-filter render of the same scene -- accumulators must be bit-identical
(both filters are conservative supersets, the hit is decided by the same candidate test)."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import Backend, FLAG_FILTER_FMA, FLAG_FILTER_CHECK
be = Backend(0)
for name, W, H, spp in (("huge", 320, 180, 8), ("simple", 320, 180, 8), ("huge", 1280, 720, 4)):
    sc = scenes.get_scene(name, W, H)
    be.upload_scene(sc); be.framebuffer(W, H)
    out = {}
    for tag, fl in (("tensor", 0), ("fma", FLAG_FILTER_FMA)):
        be.clear(); be.reset_stats(); be.render(spp, 50, seed=3, flags=fl); be.sync()
        st = be.stats()
        out[tag] = (be.readback_accum().copy(), st["render_ms"], st["segments"], st["samples"])
    miss = {}
    for tag, fl in (("tensor", FLAG_FILTER_CHECK), ("fma", FLAG_FILTER_FMA | FLAG_FILTER_CHECK)):
        be.clear(); be.reset_stats(); be.render(spp, 50, seed=5, flags=fl); be.sync()
        st = be.stats(); miss[tag] = [st["filter_misses"], st["segments"]]
    print(json.dumps({"scene": name, "filter_check_misses_of_segments": miss}), flush=True)
    same = np.array_equal(out["tensor"][0], out["fma"][0])
    nd = int((out["tensor"][0] != out["fma"][0]).any(axis=-1).sum())
    print(json.dumps({"scene": name, "W": W, "H": H, "spp": spp, "identical": bool(same), "pixels_differing": nd,
                      "tensor_ms": out["tensor"][1], "fma_ms": out["fma"][1], "segments": [out["tensor"][2], out["fma"][2]],
                      "samples": [out["tensor"][3], out["fma"][3]]}), flush=True)
be.close()
