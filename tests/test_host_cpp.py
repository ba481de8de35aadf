"""The C++ host layer: the reference's scene API (camera / sphere / lambert / metal / dielectric /
raytracer::data), the scene generators, the cuda_impl shim and the reference-main drop-in."""
import os
import subprocess

import numpy as np
import pytest

from conftest import assert_bits_equal
from raytracinginoneweekend_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "raytracinginoneweekend_b200")
DEMO = os.path.join(PKG, "rtiow_demo")
DROPIN = os.path.join(PKG, "_build", "reference_main_dropin")


def read_scene_dump(path):
    raw = open(path, "rb").read()
    n, m = (int(x) for x in np.frombuffer(raw[:8], np.uint32))
    sp = np.frombuffer(raw[8:8 + 20 * n], np.dtype([("v", "<f4", 4), ("m", "<u4")]))
    mt = np.frombuffer(raw[8 + 20 * n:8 + 20 * n + 20 * m], np.dtype([("v", "<f4", 4), ("k", "<u4")]))
    cam = np.frombuffer(raw[8 + 20 * n + 20 * m:], np.float32)
    return sp, mt, cam


def read_ppm(path):
    raw = open(path, "rb").read()
    assert raw[:3] == b"P6\n"
    header_end = raw.index(b"\n255\n") + 5
    w, h = (int(x) for x in raw[3:header_end - 5].split())
    return np.frombuffer(raw[header_end:], np.uint8).reshape(h, w, 3)


@pytest.mark.parametrize("name", ["simple", "huge", "synthetic:30"])
def test_cpp_scene_generators_match_python(tmp_path, name):
    """host/scenes.hxx (std::mt19937) and scenes.py (numpy) define the same scenes, bit for bit."""
    assert os.path.exists(DEMO), "rtiow_demo not built (run __graft_entry__.build())"
    out = str(tmp_path / "scene.bin")
    subprocess.check_call([DEMO, "--scene", name, "--width", "640", "--height", "360", "--dump-scene", out],
                          stdout=subprocess.DEVNULL)
    sp, mt, cam = read_scene_dump(out)
    s = scenes.get_scene(name, 640, 360)
    assert_bits_equal(np.ascontiguousarray(sp["v"]), s.spheres, "spheres")
    assert_bits_equal(sp["m"], s.sphere_mat, "material index")
    assert_bits_equal(np.ascontiguousarray(mt["v"]), s.materials, "materials")
    assert_bits_equal(mt["k"], s.mat_kind, "material kind")
    assert_bits_equal(cam, s.camera, "camera")


def test_shim_exports_reference_signature():
    out = subprocess.check_output(["nm", "-DC", os.path.join(PKG, "librtiow_shim.so")], text=True)
    assert "cuda_impl(unsigned int, unsigned int, std::vector<math::vec<3ul, unsigned char>" in out


@pytest.mark.gpu
def test_demo_main_equals_c_abi_render(tmp_path, be):
    """rtiow_demo (C++ host: scene API -> cuda_impl overload -> C ABI -> PPM) == the same frame via ctypes."""
    out = str(tmp_path / "demo.ppm")
    subprocess.check_call([DEMO, "--scene", "huge", "--width", "320", "--height", "180", "--spp", "8", "--depth", "50",
                           "--seed", "5", "--out", out], stdout=subprocess.DEVNULL)
    img = read_ppm(out)
    ref = be.render_frame(scenes.huge_scene(320, 180), 320, 180, 8, 50, seed=5)
    assert_bits_equal(img, ref, "demo frame")


@pytest.mark.gpu
def test_demo_progressive_passes_give_the_one_shot_image(tmp_path, be):
    """rtiow_demo --spp-per-pass: the C++ cuda_impl overload drives rt_render_progressive; previews are written as PPMs
    and the final frame equals the one-shot frame bit for bit."""
    out = str(tmp_path / "prog.ppm")
    pre = str(tmp_path / "prev_")
    res = subprocess.run([DEMO, "--scene", "huge", "--width", "320", "--height", "180", "--spp", "8", "--depth", "50",
                          "--seed", "5", "--spp-per-pass", "3", "--preview-prefix", pre, "--out", out],
                         capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    assert [l for l in res.stdout.splitlines() if l.startswith("preview:")] == \
        ["preview: pass 1, 3 spp", "preview: pass 2, 6 spp", "preview: pass 3, 8 spp"]
    ref = be.render_frame(scenes.huge_scene(320, 180), 320, 180, 8, 50, seed=5)
    assert_bits_equal(read_ppm(out), ref, "progressive demo frame")
    assert_bits_equal(read_ppm(pre + "8.ppm"), ref, "last preview")
    part = be.render_frame(scenes.huge_scene(320, 180), 320, 180, 3, 50, seed=5)
    assert_bits_equal(read_ppm(pre + "3.ppm"), part, "first preview")


@pytest.mark.gpu
def test_reference_main_runs_on_this_backend(tmp_path, be):
    """The reference's OWN main.cxx, linked against our cuda_impl, writes image_cuda.ppm (src/main.cxx:104-118):
    1920x1080, the reference GPU path's scene at 48 spp / 32 bounces."""
    if not os.path.exists(DROPIN):
        pytest.skip("reference_main_dropin not built (needs /root/reference at build time)")
    res = subprocess.run([DROPIN], cwd=str(tmp_path), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr
    assert "started..." in res.stdout
    img = read_ppm(str(tmp_path / "image_cuda.ppm"))
    assert img.shape == (1080, 1920, 3)
    ref = be.render_frame(scenes.simple_scene(1920, 1080), 1920, 1080, 48, 32, seed=1)
    assert_bits_equal(img, ref, "image_cuda.ppm")
