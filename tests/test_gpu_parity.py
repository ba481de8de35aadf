"""GPU parity tests (run on the B200 box with `-m gpu`).  Everything goes through the C ABI
(librtiow_b200.so via ctypes); the CPU oracle is only the checker.

Bars (BASELINE.json north_star):
  * primary rays, 1 spp, no jitter: hit sphere index bit-exact, t / normal within 1e-5 relative -- asserted at 0 ulp
    for EVERY closest-hit path of the library, the timed ones first: the tensor-core kernel's closest hit (flags = 0 on
    a 48..1024-sphere scene: k_hits_tc calls the very function k_render_tc calls), the FMA-chain filter, the
    all-spheres loop of the small-scene kernel, the LBVH, and the straightforward parity kernels;
  * full renders: per-pixel RMSE and mean luminance within the Monte Carlo noise bound against the
    reference CPU render; with RT_FLAG_EXACT (parity shading: the reference's rejection sampler and IEEE normalisations)
    additionally BIT-EXACT accumulators against the oracle port driven by the same counter-based RNG -- on the
    tensor-core kernel too (rtx::k_render_tc).
"""
import numpy as np
import pytest

from conftest import assert_bits_equal
from raytracinginoneweekend_b200 import scenes
from raytracinginoneweekend_b200.backend import (FLAG_BOOK_OPTICS, FLAG_BRUTE, FLAG_EXACT, FLAG_FILTER_CHECK, FLAG_FILTER_FMA, FLAG_LBVH, FLAG_OVERLAP,
                                                 KERNEL_NAMES, RtError)

# every closest-hit path: (flags, id)
HIT_PATHS = [(0, "throughput-default"), (FLAG_FILTER_FMA, "throughput-fma-filter"), (FLAG_LBVH, "throughput-lbvh"),
             (FLAG_EXACT, "parity"), (FLAG_EXACT | FLAG_LBVH, "parity-lbvh")]

pytestmark = pytest.mark.gpu

RNG_PHILOX = 1


def setup(be, sc, W, H):
    be.upload_scene(sc)
    be.framebuffer(W, H)
    be.reset_stats()


# ------------------------------------------------------------------ config 2: primary rays
def test_primary_hits_match_reference_fixture(be, golden):
    """Against tests/golden/primary.npz, produced by the compiled reference's hit_world/intersect."""
    g = golden("primary")
    for name in ("huge", "simple"):
        W, H = (int(x) for x in g[f"{name}_wh"])
        setup(be, scenes.get_scene(name, W, H).with_camera(aperture=0.0), W, H)
        for flags, what in HIT_PATHS:
            idx, t, nrm = be.primary_hits(flags=flags)
            assert_bits_equal(idx.astype(np.int16), g[f"{name}_idx"], f"{name} hit index ({what})")
            assert_bits_equal(t, g[f"{name}_t"], f"{name} t ({what})")
            assert_bits_equal(nrm, g[f"{name}_n"], f"{name} normal ({what})")


def test_camera_state_matches_reference(be, golden):
    """Row a1: rt_camera_set derives {origin, lower-left corner, horizontal, vertical, lens radius} exactly as the
    reference constructor does (src/camera.hxx:24-44); fixture from the compiled reference."""
    g = golden("kat")
    for ci, cam in enumerate(g["cams"]):
        be.set_camera(cam)
        assert_bits_equal(be.camera_state(), g["cam_state"][ci], f"camera {ci}")


def test_primary_hits_config2_full_size(be, port):
    """BASELINE config 2: huge-scene 1280x720, 1 spp, no jitter -- hit index bit-exact, t and normal
    within 1e-5 relative (asserted at 0 ulp) against the oracle."""
    W, H = 1280, 720
    sc = scenes.huge_scene(W, H).with_camera(aperture=0.0)
    setup(be, sc, W, H)
    oi, ot, on = port.primary_hits(sc, W, H, nthreads=64)
    assert len(np.unique(oi)) > 100 and (oi < 0).any()
    for flags, what in HIT_PATHS:            # flags = 0 first: the closest hit of the timed kernel (k_render_tc)
        idx, t, nrm = be.primary_hits(flags=flags)
        assert_bits_equal(idx, oi, f"hit index ({what})")
        rel_t = np.abs(t - ot) / np.maximum(np.abs(ot), 1e-30)
        rel_n = np.abs(nrm - on) / np.maximum(np.abs(on), 1e-30)
        assert rel_t.max() <= 1e-5 and rel_n.max() <= 1e-5, what     # the stated tolerance
        assert_bits_equal(t, ot, f"t ({what})")                      # ... and in fact 0 ulp
        assert_bits_equal(nrm, on, f"normal ({what})")
        # tiles of the same pass
        ti, tt, tn = be.primary_hits(tile=(37, 11, 501, 97), flags=flags)
        assert_bits_equal(ti, oi[11:108, 37:538], f"tile hit index ({what})")
        assert_bits_equal(tt, ot[11:108, 37:538], f"tile t ({what})")


@pytest.mark.parametrize("n_side", [1, 5, 30, 46])
def test_closest_hits_arbitrary_rays(be, port, n_side):
    """Bounce-like rays (origins on/inside/outside spheres) over scenes of 2..2117 spheres: covers
    n not a multiple of 32, one chunk (<=512) and several chunks, ties, the second root and kMIN."""
    sc = scenes.synthetic_scene(n_side=n_side, width=64, height=36)
    rng = np.random.default_rng(n_side)
    n = 20000
    o = np.empty((n, 3), np.float32)
    pick = rng.integers(0, sc.n, n)
    c, r = sc.spheres[pick, :3], sc.spheres[pick, 3:4]
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    o[:] = c + dirs * np.abs(r) * rng.choice(np.float32([0.0, 0.5, 1.0, 1.0, 1.5, 8.0]), (n, 1))
    d = rng.normal(size=(n, 3)).astype(np.float32) * rng.choice(np.float32([1e-3, 1.0, 1.0, 50.0]), (n, 1))
    rays = np.concatenate([o, d], 1).astype(np.float32)
    be.upload_scene(sc)
    oi, ot, on, _ = port.closest_hits(sc, rays)
    assert (oi >= 0).mean() > .2
    # parity kernels, and the throughput kernels' closest hit: all spheres (2), tensor filter (126, 901), FMA filter (2117)
    for flags in (FLAG_EXACT | FLAG_BRUTE, FLAG_BRUTE, FLAG_BRUTE | FLAG_FILTER_FMA):
        gi, gt, gn = be.closest_hits(rays, flags=flags)
        assert_bits_equal(gi, oi, f"index (flags {flags})")
        assert_bits_equal(gt, ot, f"t (flags {flags})")
        assert_bits_equal(gn, on, f"normal (flags {flags})")


# ------------------------------------------------------------------ full renders, exact mode: bit-exact
@pytest.mark.parametrize("filt", [0, FLAG_FILTER_FMA], ids=["tensor-kernel", "parity-kernel"])
@pytest.mark.parametrize("name,W,H,spp,depth", [("simple", 96, 54, 4, 50), ("huge", 96, 54, 3, 50),
                                                 ("huge", 33, 7, 2, 5), ("simple", 31, 17, 5, 1)])
def test_exact_render_is_bit_identical_to_oracle(be, port, name, W, H, spp, depth, filt):
    """RT_FLAG_EXACT: on the huge scene the dispatch takes the tensor-core kernel with parity shading
    (rtx::k_render_tc) -- the timed kernel's closest hit and loop structure -- or, with RT_FLAG_FILTER_FMA, the
    straightforward parity kernel; both must reproduce the oracle's accumulators, statistics and 8-bit image exactly."""
    sc = scenes.get_scene(name, W, H)
    setup(be, sc, W, H)
    be.render(spp, depth, seed=0xC0FFEE12345, flags=FLAG_EXACT | filt)
    acc = be.readback_accum()
    o = port.render(sc, W, H, spp, depth, 0xC0FFEE12345, nthreads=64, rng=RNG_PHILOX, want_fixed=True)
    assert_bits_equal(acc[..., :3], o["fixed"], "32.32 radiance sums")
    assert (acc[..., 3] == spp).all()
    st = be.stats()
    want = "rtx::k_render_tc" if (name == "huge" and not filt) else "rtx::k_render"
    assert KERNEL_NAMES[st["last_kernel"]] == want
    assert st["samples"] == W * H * spp
    assert st["segments"] == int(o["stats"][0])
    assert st["sky"] == int(o["stats"][1])
    assert st["absorbed"] + st["capped"] == int(o["stats"][2] + o["stats"][3])
    assert st["sphere_tests"] == st["segments"] * sc.n
    mean, u8 = port.resolve_fixed(o["fixed"], spp)
    assert_bits_equal(be.readback_f32(), mean, "mean radiance")
    assert_bits_equal(be.readback_rgb8(), u8, "8-bit image")


def test_exact_render_tiles_and_sample_ranges_compose(be, port):
    """Integer accumulation + (pixel, sample)-keyed RNG: any split into tiles, tile lists or sample
    ranges gives exactly the accumulators of one full call."""
    W, H, spp = 70, 41, 4
    sc = scenes.huge_scene(W, H)
    setup(be, sc, W, H)
    be.render(spp, 50, seed=5, flags=FLAG_EXACT)
    full = be.readback_accum()
    be.clear()
    be.render(spp, 50, seed=5, tile=(0, 0, 30, 41), flags=FLAG_EXACT)
    be.render(spp, 50, seed=5, tile=(30, 0, 40, 20), flags=FLAG_EXACT)
    from raytracinginoneweekend_b200.multigpu import make_tiles
    rest = make_tiles(40, 21, 16, 8)
    rest[:, 0] += 30
    rest[:, 1] += 20
    be.render_tiles(rest, 16, 8, spp, 50, seed=5, flags=FLAG_EXACT)
    assert_bits_equal(be.readback_accum(), full, "tiles")
    be.clear()
    be.render(1, 50, seed=5, flags=FLAG_EXACT)
    be.render(3, 50, seed=5, sample_base=1, flags=FLAG_EXACT)
    assert_bits_equal(be.readback_accum(), full, "sample ranges")
    # and the oracle agrees with the whole
    o = port.render(sc, W, H, spp, 50, 5, nthreads=64, rng=RNG_PHILOX, want_fixed=True)
    assert_bits_equal(full[..., :3], o["fixed"], "oracle")


@pytest.mark.parametrize("flags", [FLAG_EXACT, 0, FLAG_LBVH], ids=["exact", "tensor", "lbvh"])
def test_overlapped_chunks_equal_one_launch(be, flags):
    """RT_FLAG_OVERLAP: chunk launches alternate between two internal streams and run beside each other; clear, readback
    and non-overlapped renders on the context stay ordered around them without any call from the user.  Many small
    chunks (more than the two streams, several frames) must give the accumulators of one launch, bit for bit."""
    from raytracinginoneweekend_b200.multigpu import make_tiles
    W, H, spp = 160, 96, 6
    sc = scenes.huge_scene(W, H)
    setup(be, sc, W, H)
    tiles = make_tiles(W, H, 32, 16)
    be.render_tiles(tiles, 32, 16, spp, 50, seed=9, flags=flags)
    full = be.readback_accum()
    for frame in range(3):
        be.clear()                                            # ordered behind the previous frame's chunks
        for s in range(0, len(tiles), 4):
            be.render_tiles(tiles[s:s + 4], 32, 16, spp, 50, seed=9, flags=flags | FLAG_OVERLAP)
        assert_bits_equal(be.readback_accum(), full, f"overlapped chunks, frame {frame}")   # no throttle / sync by the caller
    # mixed: half the samples overlapped, the other half as a plain render of the whole frame
    be.clear()
    for s in range(0, len(tiles), 7):
        be.render_tiles(tiles[s:s + 7], 32, 16, 3, 50, seed=9, flags=flags | FLAG_OVERLAP)
    be.render(3, 50, seed=9, sample_base=3, flags=flags)
    assert_bits_equal(be.readback_accum(), full, "overlapped + plain")
    st = be.stats()
    assert st["samples"] >= W * H * spp


def test_fast_render_full_size_composition(be):
    """BASELINE size (1280x720): the throughput build is deterministic and split-invariant too."""
    W, H = 1280, 720
    sc = scenes.huge_scene(W, H)
    setup(be, sc, W, H)
    be.render(2, 50, seed=1)
    a = be.readback_accum()
    be.clear()
    be.render(2, 50, seed=1)
    assert_bits_equal(be.readback_accum(), a, "run-to-run determinism")
    be.clear()
    from raytracinginoneweekend_b200.multigpu import make_tiles
    tiles = make_tiles(W, H, 128, 64)
    k = len(tiles) // 3
    be.render_tiles(tiles[:k], 128, 64, 2, 50, seed=1)
    be.render_tiles(tiles[k:], 128, 64, 1, 50, seed=1)
    be.render_tiles(tiles[k:], 128, 64, 1, 50, seed=1, sample_base=1)
    assert_bits_equal(be.readback_accum(), a, "tile lists + sample ranges")
    assert (a[..., 3] == 2).all()
    # checksum-of-checksums style sanity: radiance is within [0, 1] per sample
    assert a[..., :3].max() <= 2 * (1 << 32)


# ------------------------------------------------------------------ full renders vs the reference CPU render
def mc_check(mean_gpu, n_gpu, g, name):
    """RMSE <= 1.15 sqrt(mean_px(var) (1/Nref + 1/Ngpu)); |d luminance| <= 4 sigma;
    16x16 block means within 5 sigma of their own noise."""
    ref_mean, var = g[f"{name}_mean"].astype(np.float64), g[f"{name}_var"].astype(np.float64)
    n_ref = int(g[f"{name}_cfg"][2])
    k = 1.0 / n_ref + 1.0 / n_gpu
    diff = mean_gpu.astype(np.float64) - ref_mean
    rmse = np.sqrt((diff ** 2).mean())
    bound = 1.15 * np.sqrt(var.mean() * k)
    assert rmse <= bound, f"{name}: RMSE {rmse:.5f} exceeds the noise bound {bound:.5f}"
    assert rmse >= 0.5 * np.sqrt(var.mean() * k), f"{name}: RMSE {rmse:.5f} implausibly small for independent renders"
    wy = np.array([.2126, .7152, .0722])
    dlum = abs((diff @ wy).mean())
    sig = np.sqrt(((var * wy ** 2).sum(-1) * k).sum()) / diff.shape[0] / diff.shape[1]
    assert dlum <= 4 * sig, f"{name}: mean luminance differs by {dlum:.6f} > 4 sigma = {4 * sig:.6f}"
    H, W = diff.shape[:2]
    hb, wb = H // 16 * 16, W // 16 * 16
    bd = diff[:hb, :wb].reshape(hb // 16, 16, wb // 16, 16, 3).mean((1, 3))
    bs = np.sqrt(var[:hb, :wb].reshape(hb // 16, 16, wb // 16, 16, 3).mean((1, 3)) * k / 256)
    z = np.abs(bd) / np.maximum(bs, 1e-6)
    assert np.percentile(z, 99) < 5.0, f"{name}: low-frequency bias, block z-score p99 = {np.percentile(z, 99):.2f}"
    return rmse, bound, dlum, sig


@pytest.mark.parametrize("flags", [0, FLAG_EXACT], ids=["fast", "exact"])
@pytest.mark.parametrize("name", ["simple", "huge"])
def test_render_matches_reference_cpu_render_statistically(be, golden, name, flags):
    g = golden("reference_render")
    W, H, _, depth, _ = (int(x) for x in g[f"{name}_cfg"])
    setup(be, scenes.get_scene(name, W, H), W, H)
    n_gpu = 2048
    be.render(n_gpu, depth, seed=31337, flags=flags)
    mean = be.readback_f32()
    rmse, bound, dlum, sig = mc_check(mean, n_gpu, g, name)
    st = be.stats()
    ref_seg = float(g[f"{name}_stats"][0]) / (W * H * int(g[f"{name}_cfg"][2]))
    assert abs(st["segments"] / st["samples"] - ref_seg) < 0.02 * ref_seg     # path-length statistics agree
    ref_cap = float(g[f"{name}_stats"][3] + g[f"{name}_stats"][2]) / (W * H * int(g[f"{name}_cfg"][2]))
    assert abs((st["capped"] + st["absorbed"]) / st["samples"] - ref_cap) < 0.05 * ref_cap + 1e-4


def mc_check_masked(mean_gpu, n_gpu, g, name, mask, what):
    """The RMSE and luminance bounds of mc_check restricted to the pixels of `mask` (one material class)."""
    ref_mean, var = g[f"{name}_mean"].astype(np.float64)[mask], g[f"{name}_var"].astype(np.float64)[mask]
    n_ref = int(g[f"{name}_cfg"][2])
    k = 1.0 / n_ref + 1.0 / n_gpu
    diff = mean_gpu.astype(np.float64)[mask] - ref_mean                       # (npx, 3)
    rmse, bound = np.sqrt((diff ** 2).mean()), 1.15 * np.sqrt(var.mean() * k)
    assert rmse <= bound, f"{name}/{what}: RMSE {rmse:.5f} exceeds the noise bound {bound:.5f} on {mask.sum()} pixels"
    wy = np.array([.2126, .7152, .0722])
    dlum = abs((diff @ wy).mean())
    sig = np.sqrt(((var * wy ** 2).sum(-1) * k).sum()) / diff.shape[0]
    assert dlum <= 4 * sig, f"{name}/{what}: mean luminance differs by {dlum:.6f} > 4 sigma = {4 * sig:.6f}"
    return rmse, bound, dlum, sig


@pytest.mark.parametrize("flags", [0, FLAG_EXACT], ids=["throughput-shading", "parity-shading"])
@pytest.mark.parametrize("name", ["simple", "huge"])
def test_render_matches_reference_cpu_render_at_baseline_geometry(be, golden, port, name, flags):
    """SURVEY 8(c)(iii) at the size it asks for: the compiled reference rendered 640x360 (BASELINE's 1280x720 camera and
    aspect, a quarter of the pixels) at 256 spp (huge) / 1024 spp (simple), depth 50 -- tests/golden/reference_render_640.npz,
    ~2 700 core-seconds.  Whole-image bounds, then the same bounds PER MATERIAL CLASS of the primary hit (sky, lambert,
    metal, glass), so that a bias confined to one material cannot hide in another's variance."""
    g = golden("reference_render_640")
    W, H, _, depth, _ = (int(x) for x in g[f"{name}_cfg"])
    sc = scenes.get_scene(name, W, H)
    setup(be, sc, W, H)
    n_gpu = 2048
    be.render(n_gpu, depth, seed=2718, flags=flags)
    mean = be.readback_f32()
    mc_check(mean, n_gpu, g, name)
    st = be.stats()
    ref_seg = float(g[f"{name}_stats"][0]) / (W * H * int(g[f"{name}_cfg"][2]))
    assert abs(st["segments"] / st["samples"] - ref_seg) < 0.01 * ref_seg       # path-length statistics agree
    idx, _, _ = port.primary_hits(sc.with_camera(aperture=0.0), W, H, nthreads=64)
    kind = np.where(idx < 0, 3, sc.mat_kind[sc.sphere_mat[np.maximum(idx, 0)]])
    checked = 0
    for k, what in enumerate(("lambert", "metal", "dielectric", "sky")):
        mask = kind == k
        if mask.sum() >= 400:
            mc_check_masked(mean, n_gpu, g, name, mask, what)
            checked += 1
    assert checked >= 3


def test_fast_and_exact_builds_agree_statistically(be):
    W, H = 160, 90
    setup(be, scenes.huge_scene(W, H), W, H)
    be.render(512, 50, seed=3, flags=FLAG_EXACT)
    a = be.readback_f32().astype(np.float64)
    be.clear()
    be.render(512, 50, seed=3, flags=0)
    b = be.readback_f32().astype(np.float64)
    # independent sample sets (the throughput shading draws the unit ball by inversion): agreement to Monte Carlo precision
    assert np.sqrt(((a - b) ** 2).mean()) < 0.01
    assert abs(a.mean() - b.mean()) < 1e-3


# ------------------------------------------------------------------ tensor-core candidate filter (throughput build)
def reference_ball_numpy(n, seed):
    """random_in_unit_sphere restated in numpy (src/raytracer.hxx:32-43): 3 x U(-1,1) until length <= 1."""
    rng = np.random.default_rng(seed)
    out = np.empty((0, 3), np.float32)
    while len(out) < n:
        v = rng.uniform(-1.0, 1.0, size=(2 * n, 3)).astype(np.float32)
        out = np.concatenate([out, v[np.sqrt((v * v).sum(1)) <= 1.0]])
    return out[:n]


@pytest.mark.parametrize("lens", [False, True], ids=["ball", "lens"])
def test_inversion_sampler_has_the_law_of_the_rejection_sampler(be, lens):
    """The throughput build draws random_in_unit_sphere by inversion (one Philox block) instead of the reference's
    rejection loop.  Same law: checked against the closed-form CDFs and, two-sample, against the rejection loop."""
    from scipy import stats
    n = 1 << 20
    p = be.sample_ball(n, seed=12345, lens=lens).astype(np.float64)
    ref = reference_ball_numpy(n, 7).astype(np.float64)
    if lens:
        assert np.all(p[:, 2] == 0.0)
        ref = ref.copy(); ref[:, 2] = 0.0
        rho = np.hypot(p[:, 0], p[:, 1])
        assert rho.max() <= 1.0 + 1e-6
        assert stats.kstest(rho, lambda t: 1.0 - (1.0 - np.clip(t, 0, 1) ** 2) ** 1.5).pvalue > 1e-4
    else:
        r = np.sqrt((p * p).sum(1))
        assert r.max() <= 1.0 + 1e-6
        assert stats.kstest(r, lambda t: np.clip(t, 0, 1) ** 3).pvalue > 1e-4           # volume grows as r^3
        assert stats.kstest(p[:, 2] / np.maximum(r, 1e-30), stats.uniform(-1, 2).cdf).pvalue > 1e-4
        assert abs((p * p).sum(1).mean() - 0.6) < 2e-3                                   # E r^2 = 3/5
    phi = np.arctan2(p[:, 1], p[:, 0])
    assert stats.kstest(phi, stats.uniform(-np.pi, 2 * np.pi).cdf).pvalue > 1e-4
    for k in range(2 if lens else 3):                                                    # marginals against the rejection loop
        assert stats.ks_2samp(p[:, k], ref[:, k]).pvalue > 1e-4, k
        assert abs(p[:, k].mean()) < 3e-3 and abs((p[:, k] ** 2).mean() - 0.2) < 2e-3
    c = np.corrcoef(p[:, :2].T)[0, 1]
    assert abs(c) < 5e-3


def random_scene(n, seed, W, H):
    """n spheres of mixed sizes and materials around the camera's view, plus a huge ground sphere when n > 3."""
    rng = np.random.default_rng(seed)
    sc = scenes.huge_scene(W, H)
    sph = np.zeros((n, 4), dtype=np.float32)
    sph[:, 0] = rng.uniform(-12, 12, n); sph[:, 1] = rng.uniform(-4, 3, n); sph[:, 2] = rng.uniform(-14, 6, n)
    sph[:, 3] = rng.choice([0.2, 0.2, 0.2, 0.7, 1.5], n) * rng.choice([1.0, 1.0, 1.0, -0.95], n)
    if n > 3:
        sph[0] = (3.0, -1003.2, -4.0, 1000.0)
    mat = rng.integers(0, sc.m, n).astype(np.uint32)
    return scenes.Scene(f"random{n}", sph, mat, sc.materials, sc.mat_kind, sc.camera)


@pytest.mark.parametrize("n", [48, 63, 64, 65, 128, 129, 488, 777, 1024])
def test_tensor_filter_never_drops_a_hit(be, n):
    """RT_FLAG_FILTER_CHECK re-runs the candidate test over ALL spheres for every segment and counts the
    segments whose closest hit differs from the filtered one: the tcgen05 filter is a superset filter, so 0.
    Sphere counts straddle the 64-column accumulator tiles and the 1024-sphere limit of the tensor path."""
    W, H = 192, 108
    sc = scenes.huge_scene(W, H) if n == 488 else random_scene(n, 1000 + n, W, H)
    setup(be, sc, W, H)
    be.render(6, 50, seed=11 + n, flags=FLAG_FILTER_CHECK | FLAG_BRUTE)
    st = be.stats()
    assert st["segments"] > 2 * W * H and st["node_visits"] == 0
    assert st["filter_misses"] == 0, f"{st['filter_misses']} of {st['segments']} segments lost their closest hit"
    be.clear(); be.reset_stats()
    be.render(6, 50, seed=11 + n, flags=FLAG_FILTER_CHECK | FLAG_FILTER_FMA | FLAG_BRUTE)
    st = be.stats()
    assert st["filter_misses"] == 0, "FMA-pipe filter dropped a hit"


def test_tensor_fma_and_lbvh_kernels_render_the_same_image(be):
    """The three throughput kernels trace the same counter-based paths with the same closest hit (the IEEE reference
    test decides in all of them: identical hits for identical rays).  Their shading is the same source compiled into three
    different kernels -- two with FMA contraction, the LBVH kernel without (its unit, rt_kernels_bvh.cu, says why) -- so the
    low bits of a radiance sample -- and, at a grazing hit, a whole path -- can differ: most accumulator words of the two
    contracted kernels are bit-equal, about half of the LBVH kernel's, the rest agree to rounding, statistics to a few 1e-4."""
    W, H = 160, 90
    setup(be, scenes.huge_scene(W, H), W, H)
    out = {}
    for flags, want in ((0, "rtf::k_render_tc"), (FLAG_FILTER_FMA, "rtf::k_render"), (FLAG_LBVH, "rtb::k_render_bvh")):
        be.clear(); be.reset_stats()
        be.render(64, 50, seed=3, flags=flags)
        out[flags] = (be.readback_accum().astype(np.float64), be.stats())
        assert KERNEL_NAMES[out[flags][1]["last_kernel"]] == want
        assert out[flags][1]["samples"] == W * H * 64
    a, sa = out[0]
    for flags in (FLAG_FILTER_FMA, FLAG_LBVH):
        b, sb = out[flags]
        assert (a == b).mean() > (0.30 if flags == FLAG_LBVH else 0.80), flags
        # rounding differences of up to 50 bounces of float shading: 1e-6 of a sum between the two contracted kernels, 1e-4 against
        # the uncontracted one
        close = np.abs(a - b) <= (1e-4 if flags == FLAG_LBVH else 1e-6) * np.maximum(a, 1.0)
        # the rest: a path that took another turn (a coin or a total reflection decided by the last bit): < 0.5 % of the words
        # between the contracted kernels, 1.5 % against the uncontracted one (measured)
        assert close.mean() > (0.97 if flags == FLAG_LBVH else 0.995), flags
        assert abs(a[..., :3].mean() - b[..., :3].mean()) < 1e-3 * a[..., :3].mean()
        for k in ("segments", "sky", "absorbed", "capped"):
            assert abs(sa[k] - sb[k]) <= 2e-3 * max(sa[k], 1) + 3 * max(sa[k], 1) ** .5, k      # (+ counting noise of a rare event)


# ------------------------------------------------------------------ resolve + e2e call
def test_resolve_full_size_matches_oracle(be, port):
    W, H = 1280, 720
    setup(be, scenes.simple_scene(W, H), W, H)
    be.render(3, 50, seed=8)
    acc = be.readback_accum()
    mean, u8 = port.resolve_fixed(acc[..., :3].copy(), 3)
    assert_bits_equal(be.readback_f32(), mean, "mean")
    got = be.readback_rgb8()
    assert (got != u8).mean() < 1e-6                     # double pow vs glibc powf: no visible difference
    assert_bits_equal(got, u8, "8-bit")


def test_render_frame_equals_staged_calls(be):
    """rt_render_frame (the cuda_impl-shaped call, host in / host out) == upload + render + readback."""
    W, H = 200, 120
    sc = scenes.huge_scene(W, H)
    img = be.render_frame(sc, W, H, 8, 50, seed=77)
    setup(be, sc, W, H)
    be.render(8, 50, seed=77)
    assert_bits_equal(be.readback_rgb8(), img, "frame")
    assert img.shape == (H, W, 3) and img[0].mean() > img[-1].mean() * 0.5   # row 0 is the TOP (sky) row
    sky = img[0].astype(int)
    assert (sky[:, 2] >= sky[:, 0]).all()                # bluish sky at the top, not the ground


# ------------------------------------------------------------------ edge cases and error behaviour
def test_progressive_render_equals_one_shot_and_previews_are_exact(be):
    """rt_render_progressive: the accumulators after all passes are bit-identical to one rt_render of the total spp
    (integer accumulation is order independent), every preview is exactly the image of the samples so far, previews
    arrive while later passes render, and a callback can stop the render."""
    W, H = 160, 90
    sc = scenes.huge_scene(W, H)
    for flags in (0, FLAG_EXACT):
        setup(be, sc, W, H)
        be.clear()
        be.render(12, 50, seed=3, flags=flags)
        one_shot_acc = be.readback_accum().copy()
        one_shot_img = be.readback_rgb8().copy()
        partial = {}
        for spp_done in (4, 8):
            be.clear(); be.render(spp_done, 50, seed=3, flags=flags)
            partial[spp_done] = be.readback_rgb8().copy()
        be.clear()
        seen = []
        img, done = be.render_progressive(12, 4, preview_every=1, seed=3, flags=flags,
                                          on_preview=lambda k, n, im: seen.append((k, n, im.copy())) and False)
        assert done == 12
        assert_bits_equal(be.readback_accum(), one_shot_acc, "progressive accumulators")
        assert np.array_equal(img, one_shot_img)
        assert [(k, n) for k, n, _ in seen] == [(1, 4), (2, 8), (3, 12)]
        assert np.array_equal(seen[0][2], partial[4]) and np.array_equal(seen[1][2], partial[8])
        assert np.array_equal(seen[2][2], one_shot_img)
    # ragged last pass, sparse previews
    be.clear()
    seen = []
    img, done = be.render_progressive(10, 4, preview_every=2, seed=3, on_preview=lambda k, n, im: seen.append((k, n)) and False)
    assert done == 10 and seen == [(2, 8), (3, 10)]
    # cancellation: stop at the first preview; at most the pass already in flight is added
    be.clear()
    img, done = be.render_progressive(40, 4, preview_every=1, seed=3, on_preview=lambda k, n, im: True)
    assert done in (4, 8)
    acc = be.readback_accum()
    assert int(acc[..., 3].max()) == done and int(acc[..., 3].min()) == done
    # no callback: plain multi-pass render
    be.clear()
    img, done = be.render_progressive(12, 5, preview_every=0, seed=3)
    assert done == 12
    assert int(be.readback_accum()[..., 3].min()) == 12
    with pytest.raises(RtError):
        be.render_progressive(4, 0)


# ------------------------------------------------------------------ opt-in textbook optics
def moved_camera_scene(name, W, H):
    """The scene with the camera NOT at the origin (all centres and the camera shifted): the reference's camera::ray
    (quirk F3) breaks there, the book-optics ray does not."""
    sc = scenes.get_scene(name, W, H)
    shift = np.array([3.0, -2.0, 5.0], np.float32)
    sph = sc.spheres.copy(); sph[:, :3] += shift
    cam = sc.camera.copy(); cam[0:3] += shift; cam[3:6] += shift
    return scenes.Scene(sc.name + "-moved", sph, sc.sphere_mat.copy(), sc.materials.copy(), sc.mat_kind.copy(), cam)


@pytest.mark.parametrize("name,W,H,spp,depth", [("simple", 96, 54, 4, 50), ("huge", 64, 36, 3, 50)])
def test_book_optics_exact_render_is_bit_identical_to_oracle(be, port, name, W, H, spp, depth):
    """RT_FLAG_BOOK_OPTICS in the parity build == the oracle restatement with the same three substitutions
    (Snell refraction, eye-relative camera ray with a disk lens, [0,1] sky blend), bit for bit."""
    from oracle.loader import OPT_BOOK
    sc = moved_camera_scene(name, W, H)
    setup(be, sc, W, H)
    be.clear()
    be.render(spp, depth, seed=77, flags=FLAG_EXACT | FLAG_BOOK_OPTICS)
    acc = be.readback_accum()
    o = port.render(sc, W, H, spp, depth, 77, nthreads=64, rng=RNG_PHILOX | OPT_BOOK, want_fixed=True)
    assert_bits_equal(acc[..., :3], o["fixed"], "book optics: 32.32 radiance sums")
    st = be.stats()
    assert st["segments"] == int(o["stats"][0]) and st["sky"] == int(o["stats"][1])


def test_book_optics_changes_what_it_should(be, port):
    """Property checks: (1) with the camera at the origin, zero aperture and no glass, book optics differs from the
    reference only by the sky blend; (2) glass stops trapping paths (quirk F5 makes 5-8 % of the huge scene's paths
    hit the depth cap); (3) the throughput build agrees with the parity build statistically; (4) default renders are
    untouched by the flag's existence."""
    W, H = 160, 90
    sc = scenes.huge_scene(W, H)
    setup(be, sc, W, H)
    be.clear(); be.reset_stats()
    be.render(64, 50, seed=5, flags=0)
    ref_stats = be.stats(); ref_img = be.readback_f32().copy()
    be.clear(); be.reset_stats()
    be.render(64, 50, seed=5, flags=FLAG_BOOK_OPTICS)
    book_stats = be.stats(); book_img = be.readback_f32().copy()
    assert ref_stats["capped"] / ref_stats["samples"] > 0.03
    assert book_stats["capped"] / book_stats["samples"] < 0.2 * ref_stats["capped"] / ref_stats["samples"]
    # sky: the reference's .5*y + 1.0 extrapolates past pure blue (red channel down to .25); the [0,1] blend stays whiter
    assert book_img[0, :, 0].mean() > ref_img[0, :, 0].mean() + 0.15 and book_img[0, :, 2].mean() == 1.0
    be.clear()
    be.render(64, 50, seed=5, flags=FLAG_BOOK_OPTICS | FLAG_EXACT)
    exact_img = be.readback_f32().copy()
    assert abs(float(exact_img.mean()) - float(book_img.mean())) < 0.01
    blk = lambda im: im.reshape(H // 10, 10, W // 10, 10, 3).mean((1, 3))
    assert np.abs(blk(exact_img) - blk(book_img)).max() < 0.06
    # lambert-only scene, pinhole camera at the origin: only the sky term differs -> identical path statistics
    lam = scenes.Scene("lambert", sc.spheres[:40].copy(), np.zeros(40, np.uint32), np.array([[.5, .5, .5, 0]], np.float32),
                       np.zeros(1, np.uint32), sc.with_camera(aperture=0.0).camera)
    setup(be, lam, W, H)
    out = {}
    for f in (0, FLAG_BOOK_OPTICS):
        be.clear(); be.reset_stats(); be.render(8, 50, seed=9, flags=FLAG_EXACT | f)
        out[f] = be.stats()
    for k in ("segments", "sky", "capped", "absorbed"):
        assert out[0][k] == out[FLAG_BOOK_OPTICS][k], k


def test_edge_cases_and_errors(be):
    from raytracinginoneweekend_b200.backend import Backend
    sc = scenes.simple_scene(16, 9)
    fresh = Backend(0)
    try:
        with pytest.raises(RtError, match="no scene"):
            fresh.framebuffer(16, 9) or fresh.render(1)
        fresh.upload_scene(sc)
        fresh.framebuffer(16, 9)
        fresh.render(0)                                  # zero samples: no-op
        fresh.render(1, tile=(3, 3, 0, 0))               # empty tile: no-op
        assert fresh.readback_accum().sum() == 0
        with pytest.raises(RtError, match="outside"):
            fresh.render(1, tile=(10, 0, 16, 9))
        with pytest.raises(RtError, match="max_depth"):
            fresh.render(1, max_depth=0)
        with pytest.raises(RtError, match="both"):
            fresh.render(1, flags=FLAG_BRUTE | FLAG_LBVH)
        fresh.render(1, tile=(15, 8, 1, 1))              # a single pixel
        acc = fresh.readback_accum()
        assert acc[8, 15, 3] == 1 and acc[..., 3].sum() == 1
    finally:
        fresh.close()


# ------------------------------------------------------------------ LBVH (row a10): must equal brute force
def bounce_like_rays(sc, n, seed):
    rng = np.random.default_rng(seed)
    pick = rng.integers(0, sc.n, n)
    c, r = sc.spheres[pick, :3], sc.spheres[pick, 3:4]
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    o = (c + dirs * np.abs(r) * rng.choice(np.float32([0.0, 0.5, 1.0, 1.0, 1.5, 8.0]), (n, 1))).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32) * rng.choice(np.float32([1e-3, 1.0, 1.0, 50.0]), (n, 1))
    o[: n // 8] = 0                                             # camera-like rays from the origin
    return np.concatenate([o, d], 1).astype(np.float32)


@pytest.mark.parametrize("n_side", [1, 5, 46, 80])
def test_lbvh_equals_brute_force(be, n_side):
    """Same closest hit, bit for bit (index, t, normal), as the shared-memory brute-force loop."""
    sc = scenes.synthetic_scene(n_side=n_side, width=64, height=36)
    rays = bounce_like_rays(sc, 40000, n_side)
    be.upload_scene(sc)
    for flags in (FLAG_EXACT, 0):
        bi, bt, bn = be.closest_hits(rays, flags=flags | FLAG_BRUTE)
        li, lt, ln = be.closest_hits(rays, flags=flags | FLAG_LBVH)
        assert_bits_equal(li, bi, f"index (flags {flags})")
        assert_bits_equal(lt, bt, f"t (flags {flags})")
        assert_bits_equal(ln, bn, f"normal (flags {flags})")
    st = be.stats()
    assert st["node_visits"] > 0


def test_lbvh_20k_spheres_equals_oracle(be, port):
    """SURVEY App. D gate for config 5: LBVH == the oracle's brute-force hit_world on a ~20k-sphere scene."""
    sc = scenes.synthetic_scene(n_side=141, width=64, height=36)
    assert sc.n == 19882
    rays = bounce_like_rays(sc, 6000, 141)
    be.upload_scene(sc)
    li, lt, ln = be.closest_hits(rays, flags=FLAG_EXACT)       # auto -> LBVH above RT_BRUTE_MAX_SPHERES
    oi, ot, on, _ = port.closest_hits(sc, rays)
    assert_bits_equal(li, oi, "index")
    assert_bits_equal(lt, ot, "t")
    assert_bits_equal(ln, on, "normal")
    be.framebuffer(64, 36)
    pi, pt, pn = be.primary_hits(flags=FLAG_EXACT)
    qi, qt, qn = port.primary_hits(sc, 64, 36, nthreads=64)
    assert_bits_equal(pi, qi, "primary index")
    assert_bits_equal(pt, qt, "primary t")


def test_lbvh_render_equals_brute_render(be):
    """Identical hits => identical paths => identical accumulators, exact build."""
    W, H = 96, 54
    sc = scenes.synthetic_scene(n_side=40, width=W, height=H)
    setup(be, sc, W, H)
    be.render(4, 50, seed=11, flags=FLAG_EXACT | FLAG_BRUTE)
    a = be.readback_accum()
    be.clear()
    be.render(4, 50, seed=11, flags=FLAG_EXACT | FLAG_LBVH)
    assert_bits_equal(be.readback_accum(), a, "accumulators")
    sc2 = scenes.huge_scene(W, H)
    setup(be, sc2, W, H)
    be.render(3, 50, seed=2, flags=FLAG_EXACT | FLAG_BRUTE)
    a = be.readback_accum()
    be.clear()
    be.render(3, 50, seed=2, flags=FLAG_EXACT | FLAG_LBVH)
    assert_bits_equal(be.readback_accum(), a, "huge-scene accumulators")


@pytest.mark.parametrize("exact", [FLAG_EXACT, 0], ids=["parity-build", "throughput-build"])
def test_lbvh_straggler_mode_does_not_change_a_bit(be, port, exact):
    """k_render_bvh finishes a warp's last closest-hit queries with the whole warp (one ray on 32 lanes, a shared work list;
    DESIGN section 6).  On small frames the mode is rarely entered, so RT_FLAG_TEST_STRAGGLERS forces it from the moment a
    warp's work pool is dry -- most of a small frame -- and gives the odd warps of every CTA a work list of 8 entries, so that
    their queries overflow into the fallback.  The accumulators must not change by a bit: the closest hit is the (t, index)
    minimum whatever the order in which candidates are met, and the throughput LBVH kernel is compiled without FMA
    contraction so that both of its shading sites are the same arithmetic.  Parity build: equal to the oracle as well."""
    from raytracinginoneweekend_b200.backend import FLAG_TEST_STRAGGLERS
    for sc, W, H, spp, seed in ((scenes.synthetic_scene(n_side=40, width=96, height=54), 96, 54, 4, 11),
                                (scenes.huge_scene(160, 90), 160, 90, 5, 3),
                                (scenes.synthetic_scene(n_side=141, width=128, height=72), 128, 72, 2, 7)):
        setup(be, sc, W, H)
        be.render(spp, 50, seed=seed, flags=exact | FLAG_LBVH)
        plain, st_plain = be.readback_accum().copy(), be.stats()
        be.clear(); be.reset_stats()
        be.render(spp, 50, seed=seed, flags=exact | FLAG_LBVH | FLAG_TEST_STRAGGLERS)
        forced, st_forced = be.readback_accum(), be.stats()
        assert_bits_equal(forced, plain, f"accumulators, straggler mode forced ({sc.name})")
        for k in ("samples", "segments", "sky", "absorbed", "capped"):
            assert st_forced[k] == st_plain[k], k
        assert st_forced["node_visits"] > 0
        if exact and sc.n < 5000:
            o = port.render(sc, W, H, spp, 50, seed, nthreads=64, rng=RNG_PHILOX, want_fixed=True)
            assert_bits_equal(forced[..., :3], o["fixed"], f"32.32 radiance sums against the oracle ({sc.name})")
    # the same through the tile lists of the multi-GPU path (queue instantiation is covered by test_multigpu_gpu.py)
    from raytracinginoneweekend_b200.multigpu import make_tiles
    sc, W, H = scenes.synthetic_scene(n_side=40, width=96, height=54), 96, 54
    setup(be, sc, W, H)
    be.render(3, 50, seed=5, flags=exact | FLAG_LBVH)
    plain = be.readback_accum().copy()
    be.clear()
    be.render_tiles(make_tiles(W, H, 32, 16), 32, 16, 3, 50, seed=5, flags=exact | FLAG_LBVH | FLAG_TEST_STRAGGLERS)
    assert_bits_equal(be.readback_accum(), plain, "tile list, straggler mode forced")


def test_lbvh_thin_deep_tree_equals_brute_force(be):
    """Sphere centres in geometric progression (a long, thin Karras tree) plus a random cloud: the resumable traversal
    of the render kernel and the one-shot traversal of rt_closest_hits must agree with brute force bit for bit.
    (Run once with a build that has RT_BVH_SMEM_DEPTH=3, i.e. nearly every push in the local overflow stack: same
    results -- tools/build_variant.sh ovf -DRT_BVH_SMEM_DEPTH=3; RTIOW_LIB=... pytest -k lbvh.)"""
    W, H = 64, 36
    k = np.arange(62, dtype=np.float64)
    xs = np.concatenate([-(2.0 ** -k), 2.0 ** -k]) * 40.0                 # 124 centres crowding towards x = 0
    rng = np.random.default_rng(3)
    sph = np.zeros((len(xs) + 200, 4), np.float32)
    sph[:len(xs), 0] = xs; sph[:len(xs), 1] = 0.5 * np.sin(np.arange(len(xs))); sph[:len(xs), 2] = -6.0
    sph[:len(xs), 3] = np.maximum(np.abs(xs) * 0.3, 1e-3)
    sph[len(xs):, :3] = rng.uniform([-20, -3, -25], [20, 3, -3], size=(200, 3)); sph[len(xs):, 3] = rng.uniform(0.05, 0.6, 200)
    base = scenes.huge_scene(W, H)
    m = base.materials.shape[0]
    sc = scenes.Scene("deep", sph, rng.integers(0, m, len(sph)).astype(np.uint32), base.materials, base.mat_kind,
                      base.with_camera(aperture=0.0).camera)
    setup(be, sc, W, H)
    be.clear(); be.render(6, 50, seed=4, flags=FLAG_EXACT | FLAG_BRUTE)
    a = be.readback_accum().copy()
    be.clear(); be.reset_stats(); be.render(6, 50, seed=4, flags=FLAG_EXACT | FLAG_LBVH)
    assert_bits_equal(be.readback_accum(), a, "deep-tree accumulators")
    assert be.stats()["node_visits"] > 0
    rays = bounce_like_rays(sc, 4096, 9)
    ib, tb, nb = be.closest_hits(rays, flags=FLAG_EXACT | FLAG_BRUTE)
    il, tl, nl = be.closest_hits(rays, flags=FLAG_EXACT | FLAG_LBVH)
    assert np.array_equal(ib, il)
    assert_bits_equal(tb, tl, "t")
    # throughput kernels: the LBVH leaves and the tensor kernel's candidates run the same IEEE test
    ib2, tb2, nb2 = be.closest_hits(rays, flags=FLAG_BRUTE)
    il2, tl2, nl2 = be.closest_hits(rays, flags=FLAG_LBVH)
    assert np.array_equal(ib2, ib) and np.array_equal(il2, ib)
    assert_bits_equal(tb2, tb, "t, tensor kernel"); assert_bits_equal(tl2, tb, "t, throughput LBVH")
    be.clear(); be.reset_stats(); be.render(6, 50, seed=4, flags=FLAG_LBVH)
    st = be.stats()
    assert st["samples"] == W * H * 6 and st["node_visits"] > 0


def test_lbvh_million_spheres_builds_and_renders(be):
    """BASELINE config 5 shape: 1M-sphere synthetic scene, GPU-built LBVH, 1920x1080 (1 spp here)."""
    W, H = 1920, 1080
    sc = scenes.synthetic_scene(n_side=1000, width=W, height=H)
    assert sc.n == 1000001
    setup(be, sc, W, H)
    be.render(1, 50, seed=1)
    img = be.readback_f32()
    st = be.stats()
    assert st["samples"] == W * H and st["bvh_build_ms"] > 0
    assert np.isfinite(img).all() and 0.2 < img.mean() < 0.9
    assert st["node_visits"] / st["segments"] < 400           # a hierarchy, not a list
    # primary rays: sampled pixels agree with a brute-force reference kernel run on the same rays in chunks
    idx, t, nrm = be.primary_hits(tile=(0, 540, 1920, 2), flags=FLAG_EXACT)
    assert (idx >= 0).all()                                     # below the horizon everything hits ground or spheres
