"""Pin the oracle (CPU, no GPU): the plain-C port against (a) the committed golden fixtures generated
from the compiled reference and (b) the compiled reference itself when oracle/_ref is present.
Everything is compared BIT FOR BIT."""
import numpy as np
import pytest

from conftest import assert_bits_equal
from raytracinginoneweekend_b200 import scenes

KMIN = np.float32(.008)


# ---------------------------------------------------------------- App. B known answers of the survey
def test_survey_known_answers(port):
    t, p, n = port.intersect([0, 0, 0], [0, 0, -1], [0, 0, -1, .5])
    assert t.view(np.uint32) == 0x3f000000 and tuple(n) == (0, 0, 1)
    t, p, n = port.intersect([0, 0, 0], [0.3, -0.2, -1.5], [0.2, -0.1, -1, .5])
    assert t.view(np.uint32) == 0x3eae526a
    I = port.normalize([1, -1, 0])
    np.testing.assert_array_equal(port.reflect(I, [0, 1, 0]), np.float32([0.707106769, 0.707106769, 0]))
    np.testing.assert_allclose(port.refract(I, [0, 1, 0], 1 / 1.5), [0.942809045, -0.881917119, 0.471404523], rtol=1e-7)
    assert np.isnan(port.refract(I, [0, -1, 0], 1.5)).all()          # quirk F5: TIR gives NaN, not 0
    assert port.schlick(1 / 1.5, .5) == np.float32(0.0699999928)
    assert port.schlick(1.5, .1) == np.float32(0.606870353)
    np.testing.assert_array_equal(port.background(1.15), np.float32([0.425000012, 0.654999971, 1]))
    g, u8 = port.gamma8([.25, .5, 1])
    assert tuple(u8[0]) == (135, 186, 255)
    cam = np.float32([0, 0, 0, 0, 0, -1, 0, 1, 0, 1280 / 720, 88, 0, 1])
    o, d = port.camera_ray(cam, 1, 0, 0)
    assert [int(x) for x in d.view(np.uint32)] == [0xbfdbbf73, 0x3f773761, 0xbf800000]
    o, d = port.camera_ray(cam, 1, 1279 / 1280, 719 / 720)
    assert [int(x) for x in d.view(np.uint32)] == [0x3fdb678d, 0xbf768794, 0xbf800000]


def test_reference_struct_sizes(golden):
    # vec3, u8vec3, ray, sphere, hit, lambert, metal, dielectric, material::types, camera (App. B)
    assert list(golden("kat")["sizes"]) == [12, 3, 24, 24, 40, 12, 16, 16, 20, 10064]


# ---------------------------------------------------------------- port vs golden KATs (from the compiled reference)
def test_port_intersect_matches_golden(port, golden):
    g = golden("kat")
    for i in range(len(g["o"])):
        h = port.intersect(g["o"][i], g["d"][i], g["sph"][i])
        exp = g["inter"][i]
        if exp[0] == 0:
            assert h is None
        else:
            got = np.concatenate([[h[0]], h[1], h[2]]).astype(np.float32)
            assert_bits_equal(got, exp[1:], f"intersect[{i}]")
    assert 200 < int(g["inter"][:, 0].sum()) < 3800     # the fixture exercises both outcomes


def test_port_vector_math_matches_golden(port, golden):
    g = golden("kat")
    n = len(g["I"])
    assert_bits_equal(np.stack([port.reflect(g["I"][i], g["N"][i]) for i in range(n)]), g["refl"], "reflect")
    assert_bits_equal(np.stack([port.refract(g["I"][i], g["N"][i], g["eta"][i]) for i in range(n)]), g["refr"], "refract")
    assert np.isnan(g["refr"]).any() and not np.isnan(g["refr"]).all()
    assert_bits_equal(np.float32([port.schlick(g["eta"][i], g["cosv"][i]) for i in range(n)]), g["sch"], "schlick")
    assert_bits_equal(np.stack([port.background(t) for t in g["tb"]]), g["bg"], "background")
    gam, u8 = port.gamma8(g["rgb"])
    assert_bits_equal(gam, g["gam"], "gamma")
    assert_bits_equal(u8, g["u8"], "8-bit")


def test_port_apply_material_matches_golden(port, golden):
    g = golden("kat")
    scattered = 0
    for row, exp in zip(g["am_in"], g["am_out"]):
        row32 = row.astype(np.float32)
        o, d, p, n, mat = row32[0:3], row32[3:6], row32[6:9], row32[9:12], row32[12:16]
        kind, seed = int(row[16]), int(row[17])
        res = port.apply_material(o, d, p, n, 1.0, mat, kind, seed)
        if exp[0] == 0:
            assert res is None
        else:
            scattered += 1
            assert_bits_equal(np.concatenate(res), exp[1:], "apply_material")
    assert 0 < scattered < len(g["am_in"])              # metal absorption is covered


def test_port_camera_matches_golden(port, golden):
    g = golden("kat")
    for ci, cam in enumerate(g["cams"]):
        assert_bits_equal(port.camera_state(cam), g["cam_state"][ci], "camera state")
        for j, (u, v) in enumerate(g["uv"]):
            o, d = port.camera_ray(cam, 7 + j, u, v)
            assert_bits_equal(np.concatenate([o, d]), g["cam_rays"][ci, j], "camera ray")


def test_port_primary_hits_match_golden(port, golden):
    g = golden("primary")
    for name in ("huge", "simple"):
        W, H = (int(x) for x in g[f"{name}_wh"])
        sc = scenes.get_scene(name, W, H).with_camera(aperture=0.0)
        idx, t, nrm = port.primary_hits(sc, W, H, nthreads=8)
        assert_bits_equal(idx.astype(np.int16), g[f"{name}_idx"], "hit index")
        assert_bits_equal(t, g[f"{name}_t"], "t")
        assert_bits_equal(nrm, g[f"{name}_n"], "normal")
        assert (idx >= 0).mean() > .5 and (idx < 0).any()


def test_port_render_matches_golden_bit_exact(port, golden):
    """Whole renders with the reference's std::mt19937 streams: sums, path statistics and 8-bit image."""
    from oracle.loader import RNG_MT
    g = golden("render_small")
    for name in ("simple", "huge"):
        W, H, spp, depth, seed = (int(x) for x in g[f"{name}_cfg"])
        sc = scenes.get_scene(name, W, H)
        r = port.render(sc, W, H, spp, depth, seed, nthreads=8, rng=RNG_MT)
        assert_bits_equal(r["sum"], g[f"{name}_sum"], f"{name} radiance sums")
        assert list(r["stats"]) == list(g[f"{name}_stats"])
        assert_bits_equal(port.resolve(r["sum"], spp), g[f"{name}_rgb8"], "8-bit image")


# ---------------------------------------------------------------- port vs the live compiled reference
def test_port_vs_compiled_reference_render(port, ref):
    from oracle.loader import RNG_MT
    sc = scenes.huge_scene(40, 24)
    a = ref.render(sc, 40, 24, 3, 50, 2024, nthreads=8)
    b = port.render(sc, 40, 24, 3, 50, 2024, nthreads=8, rng=RNG_MT)
    assert_bits_equal(a["sum"], b["sum"], "radiance sums")
    assert list(a["stats"]) == list(b["stats"])
    # a different depth cap changes the image and both sides agree on how
    a2 = ref.render(sc, 40, 24, 3, 5, 2024, nthreads=8)
    b2 = port.render(sc, 40, 24, 3, 5, 2024, nthreads=8, rng=RNG_MT)
    assert_bits_equal(a2["sum"], b2["sum"], "depth-5 sums")
    assert a2["stats"][3] > a["stats"][3]


def test_port_vs_compiled_reference_closest_hits(port, ref):
    rng = np.random.default_rng(5)
    sc = scenes.huge_scene(64, 36)
    o = rng.uniform(-8, 8, (3000, 3)).astype(np.float32)
    o[:, 1] = np.abs(o[:, 1]) - np.float32(3.0)
    d = rng.normal(size=(3000, 3)).astype(np.float32)
    rays = np.concatenate([o, d], 1)
    a, b = ref.closest_hits(sc, rays), port.closest_hits(sc, rays)
    for x, y, what in zip(a, b, ("index", "t", "normal", "position")):
        assert_bits_equal(x, y, what)


def test_schlick_multiplication_chain_equals_libm_pow(port, ref):
    """The port (and the CUDA exact mode) restate std::pow(x, 2|5) as multiplications; after the
    conversion to float the results must equal the reference's libm pow."""
    rng = np.random.default_rng(11)
    for ri in (np.float32(1.5), np.float32(1 / 1.5), np.float32(2.4), np.float32(1.33)):
        for c in rng.uniform(0, 1, 4000).astype(np.float32):
            assert port.schlick(ri, c) == ref.schlick(ri, c)


# ---------------------------------------------------------------- RNG building blocks
def test_philox_known_answer_vectors(port):
    """Random123 kat_vectors for Philox4x32-10."""
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, exp in kat:
        assert tuple(int(x) for x in port.philox(ctr, key)) == exp


def test_scene_generators():
    h = scenes.huge_scene()
    assert (h.n, h.m) == (488, 487)                      # SURVEY F10 / App. D, seed 20260101
    assert set(np.unique(h.mat_kind)) == {0, 1, 2}
    assert h.sphere_mat.max() == h.m - 1
    np.testing.assert_array_equal(h.camera[0:3], 0)     # camera at the origin
    s = scenes.simple_scene()
    assert (s.n, s.m) == (5, 4) and s.spheres[4, 3] < 0  # hollow glass shell
    y = scenes.synthetic_scene(n_side=20)
    assert y.n == 401 and y.m == 401


def test_philox_render_is_tile_and_thread_invariant(port):
    """The counter-based stream is keyed by (pixel, sample): any tiling / threading gives the same image."""
    from oracle.loader import RNG_PHILOX
    sc = scenes.simple_scene(32, 18)
    full = port.render(sc, 32, 18, 3, 50, 9, nthreads=1, rng=RNG_PHILOX, want_fixed=True)
    part = port.render(sc, 32, 18, 3, 50, 9, nthreads=5, rng=RNG_PHILOX, want_fixed=True, x0=8, w=16, y0=4, rows=10)
    np.testing.assert_array_equal(full["fixed"][4:14, 8:24], part["fixed"])
    a = port.render(sc, 32, 18, 2, 50, 9, nthreads=3, rng=RNG_PHILOX, want_fixed=True)
    b = port.render(sc, 32, 18, 1, 50, 9, nthreads=3, rng=RNG_PHILOX, want_fixed=True, sample_base=2)
    np.testing.assert_array_equal(full["fixed"], a["fixed"] + b["fixed"])
