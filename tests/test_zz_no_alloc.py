"""rt_closest_hits_static (include/rt_api.h, csrc/rt_static.cu): the closest-hit pass with no context and no allocation
-- its scratch is a fixed device array inside the library.  No counterpart in the reference,
so the checker is the EXISTING entry point rt_closest_hits on the same spheres and rays, at the bar the
closest-hit tests of test_gpu_parity.py use: index, t and normal bit for bit.

Collected last on purpose.  Every call shares one scratch array that keeps the last
call's contents, so the calls below follow each other with DIFFERENT sizes (large, small, large again) and each is
checked: stale records, rays or results of the previous call must not show.
"""
import numpy as np
import pytest

from conftest import assert_bits_equal
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import (FLAG_BRUTE, FLAG_EXACT, STATIC_MAX_RAYS, STATIC_MAX_SPHERES, RtError,
                                                 closest_hits_static)

pytestmark = pytest.mark.gpu


def bounce_like_rays(sc, n, seed):
    """Origins on / inside / outside spheres, directions of mixed length (as test_closest_hits_arbitrary_rays)."""
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, sc.n, n)
    c, r = sc.spheres[pick, :3], sc.spheres[pick, 3:4]
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    o = (c + dirs * np.abs(r) * rng.choice(np.float32([0.0, 0.5, 1.0, 1.0, 1.5, 8.0]), (n, 1))).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32) * rng.choice(np.float32([1e-3, 1.0, 1.0, 50.0]), (n, 1))
    return np.concatenate([o, d], 1).astype(np.float32)


def check_against_existing_entry_point(be, sc, rays, what, mixed=True):
    be.upload_scene(sc)
    want_i, want_t, want_n = be.closest_hits(rays, flags=FLAG_EXACT | FLAG_BRUTE)      # the existing entry point
    got_i, got_t, got_n = closest_hits_static(sc.spheres, rays)                         # the one that allocates nothing
    if mixed:
        assert (want_i >= 0).mean() > .2 and (want_i < 0).any(), what                   # the case means something: hits and misses
    assert_bits_equal(got_i, want_i, f"index ({what})")
    assert_bits_equal(got_t, want_t, f"t ({what})")
    assert_bits_equal(got_n, want_n, f"normal ({what})")
    # ... which is also what the throughput build's closest hit returns
    fast_i, fast_t, _ = be.closest_hits(rays, flags=FLAG_BRUTE)
    assert_bits_equal(got_i, fast_i, f"index vs throughput build ({what})")
    assert_bits_equal(got_t, fast_t, f"t vs throughput build ({what})")
    return got_i, got_t, got_n


def test_static_closest_hits_equal_existing_entry_point_on_calls_in_a_row(be):
    big = scenes.synthetic_scene(n_side=30, width=64, height=36)        # 901 spheres: several 512-sphere chunks
    small = scenes.synthetic_scene(n_side=5, width=64, height=36)       # 26 spheres: not a multiple of 32
    assert big.n <= STATIC_MAX_SPHERES and big.n % 32 != 0 and small.n % 32 != 0
    rays_big, rays_small = bounce_like_rays(big, 20000, 1), bounce_like_rays(small, 777, 2)
    first = check_against_existing_entry_point(be, big, rays_big, "call 1: 901 spheres, 20000 rays")
    check_against_existing_entry_point(be, small, rays_small, "call 2: 26 spheres, 777 rays, after the larger call")
    again = check_against_existing_entry_point(be, big, rays_big, "call 3: the first inputs again")
    for a, b, name in zip(first, again, ("index", "t", "normal")):
        assert_bits_equal(a, b, f"{name}: call 3 vs call 1")


def test_static_closest_hits_at_the_stated_maximum(be):
    """RT_STATIC_MAX_SPHERES spheres and RT_STATIC_MAX_RAYS rays in one call -- the last byte of the scratch is in use."""
    sc = scenes.synthetic_scene(n_side=64, width=64, height=36)          # 4097 spheres
    keep = STATIC_MAX_SPHERES
    assert sc.n >= keep
    sc.spheres, sc.sphere_mat = np.ascontiguousarray(sc.spheres[:keep]), np.ascontiguousarray(sc.sphere_mat[:keep])
    rays = bounce_like_rays(sc, STATIC_MAX_RAYS, 3)
    check_against_existing_entry_point(be, sc, rays, "maximum size")
    check_against_existing_entry_point(be, sc, rays[:1], "one ray after the maximum", mixed=False)


def test_static_closest_hits_rejects_what_does_not_fit():
    sph = np.tile(np.float32([0, 0, -1, .5]), (STATIC_MAX_SPHERES + 1, 1))
    rays = np.tile(np.float32([0, 0, 0, 0, 0, -1]), (STATIC_MAX_RAYS + 1, 1))
    with pytest.raises(RtError, match="RT_STATIC_MAX_SPHERES"):
        closest_hits_static(sph, rays[:4])
    with pytest.raises(RtError, match="RT_STATIC_MAX_RAYS"):
        closest_hits_static(sph[:4], rays)
    with pytest.raises(RtError, match="empty scene"):
        closest_hits_static(sph[:0], rays[:4])
    with pytest.raises(RtError, match="unknown flag"):
        closest_hits_static(sph[:4], rays[:4], flags=2)
    idx, t, nrm = closest_hits_static(sph[:4], rays[:0])                 # no rays: nothing to do, no error
    assert idx.shape == (0,) and t.shape == (0,) and nrm.shape == (0, 3)
    # and a call that fits still works afterwards: one sphere straight ahead
    idx, t, nrm = closest_hits_static(sph[:1], rays[:1])
    assert idx[0] == 0 and t[0] == np.float32(.5) and (nrm[0] == np.float32([0, 0, 1])).all()
