"""Scene and camera definitions, numpy float32, camera always at the origin.

The reference hard-codes its scenes inside ``main()`` / ``cuda_impl`` (src/main.cxx:120-183,
src/CUDA/cuda_impl.cu:348-352,404-417); its "huge-scene" is an ``#if 0`` block that does not
compile.  This module is the deterministic restatement every consumer shares
(tests, bench, oracle drivers).  ``host/scenes.hxx`` holds the same generator in C++ for the
``cuda_impl`` drop-in; tests check both produce identical bits.

Because ``camera::ray`` omits ``- origin`` (src/camera.hxx:56) every scene is
translated so the eye sits at (0,0,0); the reference formula is then exact.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

LAMBERT, METAL, DIELECTRIC = 0, 1, 2
F = np.float32


@dataclass
class Scene:
    name: str
    spheres: np.ndarray      # (n,4) float32  cx,cy,cz,r   (already translated by -eye)
    sphere_mat: np.ndarray   # (n,)  uint32   material index
    materials: np.ndarray    # (m,4) float32  r,g,b,param  (roughness | refraction index | 0)
    mat_kind: np.ndarray     # (m,)  uint32   0 lambert, 1 metal, 2 dielectric
    camera: np.ndarray       # (13,) float32  pos[3], lookat[3], up[3], aspect, vfov, aperture, focus

    @property
    def n(self) -> int:
        return int(self.spheres.shape[0])

    @property
    def m(self) -> int:
        return int(self.materials.shape[0])

    def with_camera(self, *, aspect: float | None = None, aperture: float | None = None) -> "Scene":
        cam = self.camera.copy()
        if aspect is not None:
            cam[9] = F(aspect)
        if aperture is not None:
            cam[11] = F(aperture)
        return Scene(self.name, self.spheres, self.sphere_mat, self.materials, self.mat_kind, cam)


class _MT:
    """std::mt19937(seed) raw 32-bit stream; U() = (g() >> 8) * 2^-24 (App. D)."""

    def __init__(self, seed: int):
        self.bg = np.random.MT19937()
        self.bg._legacy_seeding(int(seed))

    def u(self) -> np.float32:
        raw = int(self.bg.random_raw())
        return F(raw >> 8) * F(2.0 ** -24)


def _camera(eye, lookat, up, aspect, vfov, aperture, focus) -> np.ndarray:
    eye = np.asarray(eye, F)
    lookat = np.asarray(lookat, F)
    cam = np.zeros(13, F)
    cam[0:3] = 0
    cam[3:6] = lookat - eye
    cam[6:9] = np.asarray(up, F)
    cam[9] = F(aspect)
    cam[10] = F(vfov)
    cam[11] = F(aperture)
    cam[12] = F(focus)
    return cam


def _base_materials():
    mats = [(.1, .2, .5, 0.), (.8, .6, .2, 0.), (1., 1., 1., 1.5), (.64, .8, .0, 0.)]
    kinds = [LAMBERT, METAL, DIELECTRIC, LAMBERT]
    return mats, kinds


def simple_scene(width: int = 1280, height: int = 720) -> Scene:
    """Config 1: ground + 3 spheres (hollow glass) -- src/CUDA/cuda_impl.cu:404-417, camera :348-352."""
    mats, kinds = _base_materials()
    sph = [(0, 0, -1, .5), (0, -100.5, -1, 100), (1, 0, -1, .5), (-1, 0, -1, .5), (-1, 0, -1, -.499)]
    smat = [0, 3, 1, 2, 2]
    cam = _camera((0, 0, 0), (0, 0, -1), (0, 1, 0), F(width) / F(height), 88., .0625, 1.)
    return Scene("simple", np.asarray(sph, F), np.asarray(smat, np.uint32),
                 np.asarray(mats, F), np.asarray(kinds, np.uint32), cam)


def huge_scene(width: int = 1280, height: int = 720, seed: int = 20260101) -> Scene:
    """Configs 2-4: the book's final scene as src/main.cxx:120-183 intends it (App. D)."""
    mats, kinds = _base_materials()
    sph = [(0, 1, 0, 1.), (0, -1000.125, 0, 1000.), (2, 1, 0, 1.), (-2, 1, 0, 1.), (-2, 1, 0, -.99)]
    smat = [0, 3, 1, 2, 2]
    g = _MT(seed)
    for a in range(-11, 11):
        for b in range(-11, 11):
            k = min(2, int(g.u() * F(3)))
            cx = F(.9) * g.u() + F(a)
            cz = F(.9) * g.u() + F(b)
            r0, r1, r2, r3 = g.u(), g.u(), g.u(), g.u()
            dx, dy, dz = cx - F(0), F(.2) - F(1), cz - F(0)
            dist = np.sqrt(F(dx * dx + dy * dy + dz * dz), dtype=F)
            if dist < F(1):
                continue
            sph.append((cx, F(.2), cz, F(.2)))
            smat.append(len(mats))
            if k == 0:
                mats.append((r0, r1, r2, 0.)); kinds.append(LAMBERT)
            elif k == 1:
                mats.append((r0, r1, r2, F(.5) * r3)); kinds.append(METAL)
            else:
                mats.append((r0, r1, r2, 1.5)); kinds.append(DIELECTRIC)
    eye = np.asarray((-4, 3.2, 5), F)
    lookat = np.asarray((0, 1, 0), F)
    d = eye - lookat
    focus = np.sqrt(F(F(d[0] * d[0]) + F(d[1] * d[1])) + F(d[2] * d[2]), dtype=F)
    spheres = np.asarray(sph, F)
    spheres[:, 0:3] -= eye
    cam = _camera(eye, lookat, (0, 1, 0), F(width) / F(height), 42., .0625, focus)
    return Scene("huge", spheres, np.asarray(smat, np.uint32),
                 np.asarray(mats, F), np.asarray(kinds, np.uint32), cam)


def synthetic_scene(n_side: int = 1000, width: int = 1920, height: int = 1080, seed: int = 7) -> Scene:
    """Config 5: one sphere per cell of a jittered n_side x n_side grid + ground (App. D); no
    reference counterpart.  Draw order per cell: r, x-jitter, z-jitter, kind, r0..r3."""
    bg = np.random.MT19937()
    bg._legacy_seeding(int(seed))
    ncell = n_side * n_side
    raw = bg.random_raw(ncell * 8).astype(np.uint64)
    u = ((raw >> np.uint64(8)).astype(F) * F(2.0 ** -24)).reshape(ncell, 8)
    half = n_side // 2
    i = (np.arange(ncell) // n_side).astype(F)
    j = (np.arange(ncell) % n_side).astype(F)
    r = F(.1) + F(.3) * u[:, 0]
    cx = (i - F(half)) + F(.8) * u[:, 1]
    cz = (j - F(half)) + F(.8) * u[:, 2]
    kind = np.minimum(2, (u[:, 3] * F(3)).astype(np.int64)).astype(np.uint32)
    spheres = np.empty((ncell + 1, 4), F)
    spheres[0] = (0, -10000, 0, 10000)
    spheres[1:, 0] = cx
    spheres[1:, 1] = r
    spheres[1:, 2] = cz
    spheres[1:, 3] = r
    materials = np.empty((ncell + 1, 4), F)
    materials[0] = (.64, .8, .0, 0.)
    materials[1:, 0:3] = u[:, 4:7]
    materials[1:, 3] = np.where(kind == METAL, F(.5) * u[:, 7], np.where(kind == DIELECTRIC, F(1.5), F(0)))
    mat_kind = np.concatenate([[np.uint32(LAMBERT)], kind]).astype(np.uint32)
    sphere_mat = np.arange(ncell + 1, dtype=np.uint32)
    scale = F(n_side) / F(1000)
    eye = np.asarray((0, 30 * scale, 120 * scale), F)
    spheres[:, 0:3] -= eye
    cam = _camera(eye, (0, 0, 0), (0, 1, 0), F(width) / F(height), 40., 0., 1.)
    return Scene(f"synthetic-{ncell}", spheres, sphere_mat, materials, mat_kind, cam)


def get_scene(name: str, width: int, height: int) -> Scene:
    if name == "simple":
        return simple_scene(width, height)
    if name == "huge":
        return huge_scene(width, height)
    if name.startswith("synthetic"):
        n_side = 1000
        if ":" in name:
            n_side = int(name.split(":")[1])
        return synthetic_scene(n_side, width, height)
    raise ValueError(f"unknown scene {name!r}")
