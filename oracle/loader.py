"""oracle/loader.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings for the two CPU checkers:

* ``RefOracle``  -> oracle/_ref/libref_oracle.so : the reference's own headers compiled by
  oracle/build_ref.sh (exists in the build container and, prebuilt, on the GPU box).
* ``PortOracle`` -> oracle/_build/librt_oracle.so: the plain-C restatement oracle/rt_oracle.c.

Both expose the same numpy-level API so tests can swap one for the other.  Only tests/,
``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference`` legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libref_oracle.so")
PORT_SO = os.path.join(HERE, "_build", "librt_oracle.so")

RNG_MT, RNG_PHILOX = 0, 1
OPT_BOOK = 0x100   # OR into `rng`: textbook optics (the backend's RT_FLAG_BOOK_OPTICS); not the reference's picture

_f = np.float32
fp = C.POINTER(C.c_float)
dp = C.POINTER(C.c_double)
u32p = C.POINTER(C.c_uint32)
i32p = C.POINTER(C.c_int32)
u64p = C.POINTER(C.c_uint64)
u8p = C.POINTER(C.c_uint8)


def build_port() -> str:
    subprocess.check_call(["make", "-s", "-C", HERE, "port"])
    return PORT_SO


def build_ref() -> str | None:
    subprocess.check_call(["bash", os.path.join(HERE, "build_ref.sh")])
    return REF_SO if os.path.exists(REF_SO) else None


def _p(a, t):
    return None if a is None else a.ctypes.data_as(t)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


class _Base:
    prefix = ""
    kind = ""

    def __init__(self, path: str):
        self.lib = C.CDLL(path)
        self.path = path
        L, p = self.lib, self.prefix
        g = lambda n: getattr(L, p + n)  # noqa: E731
        g("camera_state").argtypes = [fp, fp]
        g("camera_ray").argtypes = [fp, C.c_uint32, C.c_float, C.c_float, fp, fp]
        g("intersect").argtypes = [fp, fp, fp, C.c_float, C.c_float, fp]
        g("intersect").restype = C.c_int
        g("reflect").argtypes = [fp, fp, fp]
        g("refract").argtypes = [fp, fp, C.c_float, fp]
        g("normalize").argtypes = [fp, fp]
        g("schlick").argtypes = [C.c_float, C.c_float]
        g("schlick").restype = C.c_float
        g("background").argtypes = [C.c_float, fp]
        g("gamma8").argtypes = [fp, C.c_uint64, fp, u8p]
        g("apply_material").argtypes = [fp, fp, fp, fp, C.c_float, fp, C.c_uint32, C.c_uint32, fp]
        g("apply_material").restype = C.c_int
        g("closest_hits").argtypes = [fp, u32p, C.c_uint32, fp, C.c_uint64, i32p, fp, fp, fp]
        g("resolve").argtypes = [fp, C.c_uint64, C.c_uint32, u8p]
        self._g = g

    # ---- scalar KAT helpers -------------------------------------------------
    def camera_state(self, cam):
        cam = _f32(cam)
        out = np.zeros(13, _f)
        self._g("camera_state")(_p(cam, fp), _p(out, fp))
        return out

    def camera_ray(self, cam, seed, u, v):
        cam = _f32(cam)
        o, d = np.zeros(3, _f), np.zeros(3, _f)
        self._g("camera_ray")(_p(cam, fp), int(seed), float(u), float(v), _p(o, fp), _p(d, fp))
        return o, d

    def intersect(self, o, d, sph, tmin=0.008, tmax=np.finfo(np.float32).max):
        o, d, sph = _f32(o), _f32(d), _f32(sph)
        out = np.zeros(7, _f)
        hit = self._g("intersect")(_p(o, fp), _p(d, fp), _p(sph, fp), float(np.float32(tmin)), float(tmax), _p(out, fp))
        return (out[0], out[1:4].copy(), out[4:7].copy()) if hit else None

    def _v3(self, name, *args):
        out = np.zeros(3, _f)
        self._g(name)(*args, _p(out, fp))
        return out

    def reflect(self, I, N):
        I, N = _f32(I), _f32(N)
        return self._v3("reflect", _p(I, fp), _p(N, fp))

    def refract(self, I, N, eta):
        I, N = _f32(I), _f32(N)
        return self._v3("refract", _p(I, fp), _p(N, fp), float(np.float32(eta)))

    def normalize(self, v):
        v = _f32(v)
        return self._v3("normalize", _p(v, fp))

    def schlick(self, ri, cos):
        return np.float32(self._g("schlick")(float(np.float32(ri)), float(np.float32(cos))))

    def background(self, t):
        return self._v3("background", float(np.float32(t)))

    def gamma8(self, rgb):
        rgb = _f32(rgb).reshape(-1, 3)
        g = np.zeros_like(rgb)
        u8 = np.zeros(rgb.shape, np.uint8)
        self._g("gamma8")(_p(rgb, fp), rgb.shape[0], _p(g, fp), _p(u8, u8p))
        return g, u8

    def apply_material(self, o, d, p, n, t, mat, kind, seed):
        o, d, p, n, mat = _f32(o), _f32(d), _f32(p), _f32(n), _f32(mat)
        out = np.zeros(9, _f)
        ok = self._g("apply_material")(_p(o, fp), _p(d, fp), _p(p, fp), _p(n, fp), float(np.float32(t)),
                                       _p(mat, fp), int(kind), int(seed), _p(out, fp))
        return (out[0:3].copy(), out[3:6].copy(), out[6:9].copy()) if ok else None

    # ---- batch passes -------------------------------------------------------
    def closest_hits(self, scene, rays):
        sph, sm = _f32(scene.spheres), _u32(scene.sphere_mat)
        rays = _f32(rays).reshape(-1, 6)
        n = rays.shape[0]
        idx = np.zeros(n, np.int32)
        t = np.zeros(n, _f)
        nrm = np.zeros((n, 3), _f)
        pos = np.zeros((n, 3), _f)
        self._g("closest_hits")(_p(sph, fp), _p(sm, u32p), sph.shape[0], _p(rays, fp), n,
                                _p(idx, i32p), _p(t, fp), _p(nrm, fp), _p(pos, fp))
        return idx, t, nrm, pos

    def resolve(self, total, spp):
        total = _f32(total)
        npix = total.size // 3
        out = np.zeros(total.shape, np.uint8)
        self._g("resolve")(_p(total, fp), npix, int(spp), _p(out, u8p))
        return out


class RefOracle(_Base):
    """The compiled reference (kind == "reference")."""
    prefix = "ref_"
    kind = "reference"

    def __init__(self, path: str = REF_SO):
        super().__init__(path)
        L = self.lib
        L.ref_primary_hits.argtypes = [fp, u32p, C.c_uint32, fp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                       i32p, fp, fp, C.c_int]
        L.ref_render.argtypes = [fp, u32p, C.c_uint32, fp, u32p, C.c_uint32, fp, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64,
                                 fp, dp, u64p, C.c_int, C.c_uint32]
        L.ref_render.restype = C.c_double
        L.ref_sizes.argtypes = [u32p]
        L.ref_sizes.restype = C.c_int
        L.ref_save_ppm.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, u8p]
        L.ref_save_ppm.restype = C.c_int
        L.ref_hardware_concurrency.restype = C.c_int

    @staticmethod
    def available() -> bool:
        return os.path.exists(REF_SO)

    def sizes(self):
        out = np.zeros(10, np.uint32)
        self.lib.ref_sizes(_p(out, u32p))
        return out

    def primary_hits(self, scene, W, H, y0=0, rows=None, nthreads=1):
        rows = H - y0 if rows is None else rows
        sph, sm, cam = _f32(scene.spheres), _u32(scene.sphere_mat), _f32(scene.camera)
        idx = np.zeros((rows, W), np.int32)
        t = np.zeros((rows, W), _f)
        nrm = np.zeros((rows, W, 3), _f)
        self.lib.ref_primary_hits(_p(sph, fp), _p(sm, u32p), sph.shape[0], _p(cam, fp), W, H, y0, rows,
                                  _p(idx, i32p), _p(t, fp), _p(nrm, fp), int(nthreads))
        return idx, t, nrm

    def render(self, scene, W, H, spp, depth, seed, y0=0, rows=None, nthreads=1, want_sumsq=False, ystep=1):
        """rows y0, y0 + ystep, ... (`rows` of them; default: all rows from y0 with that stride)"""
        rows = (H - y0 + ystep - 1) // ystep if rows is None else rows
        sph, sm, cam = _f32(scene.spheres), _u32(scene.sphere_mat), _f32(scene.camera)
        mats, mk = _f32(scene.materials), _u32(scene.mat_kind)
        total = np.zeros((rows, W, 3), _f)
        sumsq = np.zeros((rows, W, 3), np.float64) if want_sumsq else None
        stats = np.zeros(4, np.uint64)
        secs = self.lib.ref_render(_p(sph, fp), _p(sm, u32p), sph.shape[0], _p(mats, fp), _p(mk, u32p), mats.shape[0],
                                   _p(cam, fp), W, H, y0, rows, spp, depth, int(seed),
                                   _p(total, fp), _p(sumsq, dp), _p(stats, u64p), int(nthreads), int(ystep))
        return dict(sum=total, sumsq=sumsq, stats=stats, seconds=secs)

    def save_ppm(self, path, rgb8):
        rgb8 = np.ascontiguousarray(rgb8, np.uint8)
        H, W = rgb8.shape[:2]
        return self.lib.ref_save_ppm(os.fsencode(path), W, H, _p(rgb8, u8p))


class PortOracle(_Base):
    """The plain-C restatement (kind == "port")."""
    prefix = "orc_"
    kind = "port"

    def __init__(self, path: str = PORT_SO):
        if not os.path.exists(path):
            build_port()
        super().__init__(path)
        L = self.lib
        L.orc_primary_hits.argtypes = [fp, u32p, C.c_uint32, fp, C.c_uint32, C.c_uint32,
                                       C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, i32p, fp, fp, C.c_int]
        L.orc_render.argtypes = [fp, u32p, C.c_uint32, fp, u32p, C.c_uint32, fp, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                 C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, C.c_int,
                                 fp, dp, u64p, u64p, C.c_int, C.c_uint32]
        L.orc_render.restype = C.c_double
        L.orc_resolve_fixed.argtypes = [u64p, C.c_uint64, C.c_uint32, fp, u8p]
        L.orc_philox.argtypes = [u32p, u32p, u32p]

    def philox(self, ctr, key):
        ctr, key = _u32(ctr), _u32(key)
        out = np.zeros(4, np.uint32)
        self.lib.orc_philox(_p(ctr, u32p), _p(key, u32p), _p(out, u32p))
        return out

    def primary_hits(self, scene, W, H, y0=0, rows=None, nthreads=1, x0=0, w=None):
        rows = H - y0 if rows is None else rows
        w = W - x0 if w is None else w
        sph, sm, cam = _f32(scene.spheres), _u32(scene.sphere_mat), _f32(scene.camera)
        idx = np.zeros((rows, w), np.int32)
        t = np.zeros((rows, w), _f)
        nrm = np.zeros((rows, w, 3), _f)
        self.lib.orc_primary_hits(_p(sph, fp), _p(sm, u32p), sph.shape[0], _p(cam, fp), W, H, x0, y0, w, rows,
                                  _p(idx, i32p), _p(t, fp), _p(nrm, fp), int(nthreads))
        return idx, t, nrm

    def render(self, scene, W, H, spp, depth, seed, y0=0, rows=None, nthreads=1, want_sumsq=False,
               x0=0, w=None, sample_base=0, rng=RNG_MT, want_fixed=False, ystep=1):
        rows = (H - y0 + ystep - 1) // ystep if rows is None else rows
        w = W - x0 if w is None else w
        sph, sm, cam = _f32(scene.spheres), _u32(scene.sphere_mat), _f32(scene.camera)
        mats, mk = _f32(scene.materials), _u32(scene.mat_kind)
        total = np.zeros((rows, w, 3), _f)
        sumsq = np.zeros((rows, w, 3), np.float64) if want_sumsq else None
        fixed = np.zeros((rows, w, 3), np.uint64) if want_fixed else None
        stats = np.zeros(4, np.uint64)
        secs = self.lib.orc_render(_p(sph, fp), _p(sm, u32p), sph.shape[0], _p(mats, fp), _p(mk, u32p), mats.shape[0],
                                   _p(cam, fp), W, H, x0, y0, w, rows, spp, depth, int(seed), int(sample_base), int(rng),
                                   _p(total, fp), _p(sumsq, dp), _p(fixed, u64p), _p(stats, u64p), int(nthreads), int(ystep))
        return dict(sum=total, sumsq=sumsq, fixed=fixed, stats=stats, seconds=secs)

    def resolve_fixed(self, fixed, spp):
        fixed = np.ascontiguousarray(fixed, np.uint64)
        npix = fixed.size // 3
        mean = np.zeros(fixed.shape, _f)
        u8 = np.zeros(fixed.shape, np.uint8)
        self.lib.orc_resolve_fixed(_p(fixed, u64p), npix, int(spp), _p(mean, fp), _p(u8, u8p))
        return mean, u8


def best_oracle():
    """The compiled reference if present, else the port."""
    return RefOracle() if RefOracle.available() else PortOracle()
