"""Multi-GPU paths on real GPUs.

On any box (1 GPU is enough): the device-side tile queue (rt_render_queue: a persistent launch that pulls tiles from range
cursors and steals) renders exactly the accumulators of a plain render; rt_multi (several contexts behind one call, one
host thread each) works with the same device listed twice.  With >= 2 GPUs: the N-GPU frame -- host-scheduled chunks +
NCCL reduce, device-side queue + peer-store gather between processes (CUDA IPC), rt_multi in one process, and
`rtiow_demo --gpus 2` -- is bit-identical to the single-GPU frame."""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import assert_bits_equal

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def test_device_queue_on_one_gpu_equals_plain_render(be):
    """Three 'ranks' played one after the other on one GPU: the first launch renders its own range and then steals every
    other range through the cursors, the later ones find the queue dry.  Union = the plain render, bit for bit; the
    sample count lands on every pixel once; resolving each launch's tiles into one frame gives the plain 8-bit image."""
    from raytracinginoneweekend_b200 import multigpu, scenes
    from raytracinginoneweekend_b200.backend import FLAG_EXACT, FLAG_LBVH, RtError
    W, H, spp = 200, 90, 5                                   # ragged: edge tiles narrower and shorter than 32 x 8
    sc = scenes.huge_scene(W, H)
    be.upload_scene(sc); be.framebuffer(W, H)
    tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, 32, 8), 3)
    cur = be.queue_dev_create()
    try:
        for flags in (0, FLAG_EXACT, FLAG_LBVH):
            be.clear(); be.render(spp, 50, seed=21, flags=flags)
            want = be.readback_accum().copy()
            want_img = be.readback_rgb8().copy()
            be.clear()
            be.queue_dev_reset(cur, len(tiles), 3)
            took = []
            for rank in (1, 0, 2):
                be.render_queue(tiles, 32, 8, cur, rank, 3, spp, 50, seed=21, flags=flags)
                be.resolve_tiles_to(None)
                took.append(be.tiles_rendered())
            assert took == [len(tiles), 0, 0]
            assert_bits_equal(be.readback_accum(), want, f"accumulators through the device queue, flags {flags}")
            assert np.array_equal(be.copy_rgb8(), want_img)
        with pytest.raises(RtError, match="multiple of 256"):
            be.render_queue(tiles, 32, 12, cur, 0, 1, spp)
    finally:
        be.queue_dev_free(cur)


def test_rt_multi_one_process_equals_single_context(be):
    """rt_multi with the same device listed twice: two contexts, two host threads, one queue -- the frame and both gather
    modes' plumbing (NCCL needs distinct devices, so peer stores here) against rt_render_frame."""
    from raytracinginoneweekend_b200 import scenes
    from raytracinginoneweekend_b200.backend import MultiBackend
    W, H, spp = 320, 192, 6
    sc = scenes.huge_scene(W, H)
    want = be.render_frame(sc, W, H, spp, 50, seed=4)
    for devices in ([0], [0, 0], [0, 0, 0]):
        mb = MultiBackend(devices)
        try:
            img = mb.render_frame(sc, W, H, spp, 50, seed=4, tile_w=64, tile_h=32)
            assert np.array_equal(img, want), devices
            info = mb.last_frame()
            assert sum(info["tiles_per_device"]) == 5 * 6
            again = mb.render_again(spp, 50, seed=4)
            assert np.array_equal(again, want)
            # (with one device listed several times the contexts share the GPU: any of them may end up rendering everything)
            assert sum(mb.ctx(i).stats()["samples"] for i in range(len(devices))) >= W * H * spp
        finally:
            mb.close()


def _torchrun(script_args, port):
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", str(port),
                           os.path.join(ROOT, "tools", "multigpu_check.py")] + script_args,
                          capture_output=True, text=True, timeout=900)


@pytest.mark.parametrize("mode", ["host", "device"])
def test_two_gpu_frame_is_identical_to_single_gpu(mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = _torchrun(["640", "360", "8", mode], 29533 if mode == "host" else 29534)
    assert out.returncode == 0, out.stderr[-2000:]
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    res = json.loads(line)
    assert res["identical_to_single_gpu"]
    assert sum(res["tiles_per_rank"]) == 30


def test_two_gpu_rt_multi_and_demo(tmp_path, be):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from raytracinginoneweekend_b200 import scenes
    from raytracinginoneweekend_b200.backend import MultiBackend
    W, H, spp = 640, 360, 8
    sc = scenes.huge_scene(W, H)
    want = be.render_frame(sc, W, H, spp, 50, seed=4)
    mb = MultiBackend([0, 1])
    try:
        for gather in (0, MultiBackend.GATHER_NCCL):
            img = mb.render_frame(sc, W, H, spp, 50, seed=4, gather=gather)
            assert np.array_equal(img, want), f"gather {gather}"
        info = mb.last_frame()
        assert info["nccl"] and all(t > 0 for t in info["tiles_per_device"])
    finally:
        mb.close()
    demo = os.path.join(ROOT, "raytracinginoneweekend_b200", "rtiow_demo")
    outs = []
    for extra in ([], ["--gpus", "2"]):
        ppm = tmp_path / f"demo{len(extra)}.ppm"
        r = subprocess.run([demo, "--scene", "huge", "--width", "640", "--height", "360", "--spp", "8", "--depth", "50",
                            "--out", str(ppm)] + extra, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr
        outs.append(hashlib.sha256(ppm.read_bytes()).hexdigest())
    assert outs[0] == outs[1]
