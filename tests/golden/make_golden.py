"""Generate the committed golden fixtures from the COMPILED REFERENCE (oracle/_ref/libref_oracle.so).

The reference ships no tests, golden vectors or images, so parity is pinned
on outputs of the reference's own code run in the build container.  Run (in the container that has
/root/reference):

    bash oracle/build_ref.sh && python tests/golden/make_golden.py

Writes tests/golden/*.npz (a few hundred KB each).  The GPU box never needs /root/reference: the
tests read these files.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle.loader import RefOracle  # noqa: E402
from raytracinginoneweekend_b200 import scenes  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
NTHREADS = int(os.environ.get("GOLDEN_THREADS", os.cpu_count() or 1))


def kat(R: RefOracle):
    rng = np.random.default_rng(20260101)
    n = 4000
    o = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    sph = np.concatenate([rng.uniform(-3, 3, (n, 3)), rng.uniform(.1, 2.5, (n, 1))], 1).astype(np.float32)
    sph[::7, 3] *= -1                      # negative radius: flipped normal (hollow glass)
    inter = np.full((n, 8), np.nan, np.float32)
    for i in range(n):
        h = R.intersect(o[i], d[i], sph[i])
        if h is not None:
            inter[i, 0] = 1
            inter[i, 1] = h[0]; inter[i, 2:5] = h[1]; inter[i, 5:8] = h[2]
        else:
            inter[i, 0] = 0
    I = np.stack([R.normalize(v) for v in d])
    N = np.stack([R.normalize(v) for v in rng.normal(size=(n, 3)).astype(np.float32)])
    eta = np.where(rng.random(n) < .5, np.float32(1.5), np.float32(1 / 1.5)).astype(np.float32)
    refl = np.stack([R.reflect(I[i], N[i]) for i in range(n)])
    refr = np.stack([R.refract(I[i], N[i], eta[i]) for i in range(n)])
    cosv = rng.uniform(0, 1, n).astype(np.float32)
    sch = np.array([R.schlick(eta[i], cosv[i]) for i in range(n)], np.float32)
    tb = rng.uniform(.5, 1.5, 256).astype(np.float32)
    bg = np.stack([R.background(t) for t in tb])
    rgb = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    rgb[:4] = [[.25, .5, 1], [0, 0, 0], [1, 1, 1], [1e-6, .003, .999999]]
    gam, u8 = R.gamma8(rgb)
    # apply_material: a hit on a unit sphere at the origin for every kind, many seeds
    mats = np.array([[.1, .2, .5, 0.], [.8, .6, .2, .3], [.9, .95, 1., 1.5], [.7, .7, .7, 0.]], np.float32)
    kinds = np.array([0, 1, 2, 1], np.uint32)
    am_in, am_out = [], []
    for i in range(1200):
        k = i % 4
        nrm = N[i]
        p = nrm.copy()
        inside = (i // 4) % 2 == 1 and kinds[k] == 2
        dd = (-nrm + rng.normal(scale=.6, size=3)).astype(np.float32)
        if inside:
            dd = -dd
        oo = (p - dd).astype(np.float32)
        seed = 1000 + i
        res = R.apply_material(oo, dd, p, nrm, 1.0, mats[k], kinds[k], seed)
        am_in.append(np.concatenate([oo, dd, p, nrm, mats[k], [kinds[k], seed]]).astype(np.float64))
        am_out.append(np.concatenate([[1.], res[0], res[1], res[2]]).astype(np.float32) if res is not None
                      else np.concatenate([[0.], np.zeros(9)]).astype(np.float32))
    cams = [scenes.simple_scene().camera, scenes.huge_scene().camera,
            np.array([0, 0, 0, 0, 0, -1, 0, 1, 0, 1280 / 720, 88, 0, 1], np.float32)]
    cam_state = np.stack([R.camera_state(c) for c in cams])
    uv = rng.uniform(0, 1, (64, 2)).astype(np.float32)
    uv[:3] = [[0, 0], [.5, .5], [1279 / 1280, 719 / 720]]
    cam_rays = np.stack([np.concatenate(R.camera_ray(cams[ci], 7 + j, uv[j, 0], uv[j, 1]))
                         for ci in range(3) for j in range(64)]).reshape(3, 64, 6)
    np.savez_compressed(os.path.join(OUT, "kat.npz"), sizes=R.sizes(), o=o, d=d, sph=sph, inter=inter, I=I, N=N, eta=eta,
                        refl=refl, refr=refr, cosv=cosv, sch=sch, tb=tb, bg=bg, rgb=rgb, gam=gam, u8=u8,
                        am_in=np.stack(am_in), am_out=np.stack(am_out), cams=np.stack(cams), cam_state=cam_state,
                        uv=uv, cam_rays=cam_rays)


def primary(R: RefOracle):
    out = {}
    for name, W, H in (("huge", 256, 144), ("simple", 128, 72)):
        sc = scenes.get_scene(name, W, H).with_camera(aperture=0.0)
        idx, t, nrm = R.primary_hits(sc, W, H, nthreads=NTHREADS)
        out[f"{name}_idx"] = idx.astype(np.int16)
        out[f"{name}_t"] = t
        out[f"{name}_n"] = nrm
        out[f"{name}_wh"] = np.array([W, H])
    np.savez_compressed(os.path.join(OUT, "primary.npz"), **out)


def small_renders(R: RefOracle):
    """Bit-exact targets: reference renders with fixed mt19937 streams (seeded per row by the harness)."""
    out = {}
    for name, W, H, spp in (("simple", 48, 27, 4), ("huge", 48, 27, 2)):
        sc = scenes.get_scene(name, W, H)
        r = R.render(sc, W, H, spp, 50, 4242, nthreads=NTHREADS)
        out[f"{name}_sum"] = r["sum"]
        out[f"{name}_stats"] = r["stats"]
        out[f"{name}_cfg"] = np.array([W, H, spp, 50, 4242])
        out[f"{name}_rgb8"] = R.resolve(r["sum"], spp)
    np.savez_compressed(os.path.join(OUT, "render_small.npz"), **out)


def reference_renders(R: RefOracle):
    """High-spp reference CPU renders (mean and per-pixel sample variance, linear space, depth 50)
    for the Monte Carlo acceptance test of SURVEY section 8(c)(iii)."""
    out = {}
    for name, W, H, spp in (("simple", 160, 90, 1024), ("huge", 160, 90, 256)):
        sc = scenes.get_scene(name, W, H)
        r = R.render(sc, W, H, spp, 50, 777, nthreads=NTHREADS, want_sumsq=True)
        mean = r["sum"].astype(np.float64) / spp
        var = np.maximum(r["sumsq"] / spp - mean * mean, 0.0) * spp / (spp - 1)
        out[f"{name}_mean"] = mean.astype(np.float32)
        out[f"{name}_var"] = var.astype(np.float32)
        out[f"{name}_cfg"] = np.array([W, H, spp, 50, 777])
        out[f"{name}_stats"] = r["stats"]
        print(name, "seconds", r["seconds"], "stats", r["stats"], "segments/sample", r["stats"][0] / (W * H * spp))
    np.savez_compressed(os.path.join(OUT, "reference_render.npz"), **out)


def reference_renders_full(R: RefOracle):
    """The same at the size SURVEY section 8(c)(iii) asks for: 640x360 (a quarter of BASELINE's 1280x720 frame, same
    aspect and camera), huge scene 256 spp, simple scene 1024 spp, depth 50.  About 2 700 core-seconds.  The variance
    is stored as float16 (three digits are plenty for a noise bound) to keep the fixture small."""
    out = {}
    for name, W, H, spp in (("simple", 640, 360, 1024), ("huge", 640, 360, 256)):
        sc = scenes.get_scene(name, W, H)
        r = R.render(sc, W, H, spp, 50, 777, nthreads=NTHREADS, want_sumsq=True)
        mean = r["sum"].astype(np.float64) / spp
        var = np.maximum(r["sumsq"] / spp - mean * mean, 0.0) * spp / (spp - 1)
        out[f"{name}_mean"] = mean.astype(np.float32)
        out[f"{name}_var"] = np.minimum(var, 6.0e4).astype(np.float16)
        out[f"{name}_cfg"] = np.array([W, H, spp, 50, 777])
        out[f"{name}_stats"] = r["stats"]
        print(name, "seconds", r["seconds"], "stats", r["stats"], "segments/sample", r["stats"][0] / (W * H * spp), flush=True)
    np.savez_compressed(os.path.join(OUT, "reference_render_640.npz"), **out)


if __name__ == "__main__":
    if not RefOracle.available():
        raise SystemExit("oracle/_ref/libref_oracle.so missing: run `bash oracle/build_ref.sh` where /root/reference exists")
    R = RefOracle()
    if "--full-only" in sys.argv:          # only the big statistical fixture (minutes of CPU)
        reference_renders_full(R)
        raise SystemExit(0)
    kat(R)
    primary(R)
    small_renders(R)
    reference_renders(R)
    if "--full" in sys.argv:
        reference_renders_full(R)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
