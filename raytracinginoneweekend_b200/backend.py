"""ctypes binding of librtiow_b200.so (include/rt_api.h) -- the Python face of the C ABI.

This is plumbing for tests, bench.py and the torch.distributed multi-GPU driver; the product is the
CUDA library.  There is NO fallback: if the shared library is missing or no sm_100 GPU is usable,
construction raises.  Nothing here imports anything under ``oracle/``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTIOW_LIB") or os.path.join(HERE, "librtiow_b200.so")   # RTIOW_LIB: experimental variant builds (tools/build_variant.sh)

FLAG_EXACT = 1 << 0
FLAG_BRUTE = 1 << 1
FLAG_LBVH = 1 << 2
FLAG_FILTER_CHECK = 1 << 4  # count segments whose filtered closest hit differs from the unfiltered one (stats: filter_misses)
FLAG_BOOK_OPTICS = 1 << 5   # opt-in textbook optics (Snell refraction, eye-relative camera rays, disk lens, [0,1] sky blend)
FLAG_OVERLAP = 1 << 6       # rt_render_tiles: this chunk may run beside the chunks queued before it (disjoint tiles; two internal streams)
FLAG_TEST_STRAGGLERS = 1 << 7  # test hook: LBVH kernels enter the straggler mode as soon as a warp's pool is dry; odd warps overflow their work list
FLAG_FILTER_FMA = 1 << 3   # candidate filter on the FP32 pipe instead of the tensor cores
# rt_stats.last_kernel (rt_kernel_id, include/rt_api.h)
KERNEL_NAMES = {0: "none", 1: "rtx::k_render", 2: "rtf::k_render", 3: "rtf::k_render_small", 4: "rtf::k_render_tc",
                5: "rtx::k_render_tc", 6: "rtx::k_render_bvh", 7: "rtb::k_render_bvh"}

EXPORTS = [
    "rt_last_error", "rt_version", "rt_create", "rt_destroy", "rt_set_stream", "rt_sync", "rt_throttle",
    "rt_scene_upload", "rt_camera_set", "rt_camera_get", "rt_framebuffer", "rt_clear", "rt_render",
    "rt_render_tiles",
    "rt_primary_hits", "rt_closest_hits", "rt_readback_rgb8", "rt_readback_f32", "rt_readback_accum",
    "rt_device_accum", "rt_device_rgb8", "rt_resolve_device", "rt_render_frame", "rt_stats_get",
    "rt_stats_reset", "rt_write_ppm", "rt_fp32_peak_probe", "rt_tf32_peak_probe", "rt_filter_info", "rt_debug_sample_ball", "rt_render_progressive",
    "rt_queue_open", "rt_queue_reset", "rt_queue_next", "rt_queue_close",
    "rt_get_stream", "rt_queue_dev_create", "rt_queue_dev_reset", "rt_queue_dev_free", "rt_render_queue",
    "rt_resolve_tiles_to", "rt_queue_tiles_rendered", "rt_queue_trace", "rt_copy_rgb8", "rt_ipc_export", "rt_ipc_import", "rt_ipc_close",
    "rt_multi_create", "rt_multi_destroy", "rt_multi_device_count", "rt_multi_ctx", "rt_multi_render_frame",
    "rt_multi_render_again", "rt_multi_last_frame",
    "rt_closest_hits_static",
]


class RtSphere(C.Structure):
    _fields_ = [("cx", C.c_float), ("cy", C.c_float), ("cz", C.c_float), ("r", C.c_float), ("mat", C.c_uint32)]


class RtMaterial(C.Structure):
    _fields_ = [("r", C.c_float), ("g", C.c_float), ("b", C.c_float), ("param", C.c_float), ("kind", C.c_uint32)]


class RtTile(C.Structure):
    _fields_ = [("x0", C.c_uint32), ("y0", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


class RtRenderParams(C.Structure):
    _fields_ = [("spp", C.c_uint32), ("max_depth", C.c_uint32), ("seed", C.c_uint64),
                ("sample_base", C.c_uint32), ("flags", C.c_uint32)]


class RtStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("segments", C.c_uint64), ("sphere_tests", C.c_uint64),
                ("node_visits", C.c_uint64), ("sky", C.c_uint64), ("absorbed", C.c_uint64),
                ("capped", C.c_uint64), ("launches", C.c_uint64), ("render_ms", C.c_double),
                ("last_render_ms", C.c_double), ("bvh_build_ms", C.c_double), ("filter_misses", C.c_uint64),
                ("last_kernel", C.c_uint32), ("reserved", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


SPHERE_DTYPE = np.dtype([("cx", "<f4"), ("cy", "<f4"), ("cz", "<f4"), ("r", "<f4"), ("mat", "<u4")])
MATERIAL_DTYPE = np.dtype([("r", "<f4"), ("g", "<f4"), ("b", "<f4"), ("param", "<f4"), ("kind", "<u4")])


PREVIEW_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint8))


class RtError(RuntimeError):
    pass


_lib = None


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """Load the C-ABI library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise RtError(f"{path} is missing: build it with `make -C {HERE}` (or __graft_entry__.build()); "
                      "there is no CPU fallback")
    lib = C.CDLL(path)
    if os.environ.get("RTIOW_LIB_OLD_ABI") == "1":      # A/B timing against a library built from an older tree (tools/ab_bench.sh)
        class _Tolerant:
            def __init__(self, real): object.__setattr__(self, "_real", real)
            def __getattr__(self, name):
                try:
                    return getattr(self._real, name)
                except AttributeError:
                    def missing(*a, **k): raise RtError(f"{name} is not exported by {path}")
                    return missing
        lib = _Tolerant(lib)
    vp, u32, u64, i32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    fp = C.POINTER(C.c_float)
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_version.restype = C.c_char_p
    lib.rt_create.argtypes = [C.POINTER(vp), i32]
    lib.rt_destroy.argtypes = [vp]
    lib.rt_set_stream.argtypes = [vp, vp]
    lib.rt_sync.argtypes = [vp]
    lib.rt_throttle.argtypes = [vp, u32]
    lib.rt_scene_upload.argtypes = [vp, vp, u32, vp, u32]
    lib.rt_camera_set.argtypes = [vp, fp, fp, fp, C.c_float, C.c_float, C.c_float, C.c_float]
    lib.rt_camera_get.argtypes = [vp, fp]
    lib.rt_framebuffer.argtypes = [vp, u32, u32]
    lib.rt_clear.argtypes = [vp]
    lib.rt_render.argtypes = [vp, C.POINTER(RtTile), C.POINTER(RtRenderParams)]
    lib.rt_render_tiles.argtypes = [vp, vp, u32, u32, u32, C.POINTER(RtRenderParams)]
    lib.rt_primary_hits.argtypes = [vp, C.POINTER(RtTile), u32, vp, vp, vp]
    lib.rt_closest_hits.argtypes = [vp, vp, u64, u32, vp, vp, vp]
    lib.rt_readback_rgb8.argtypes = [vp, vp]
    lib.rt_readback_f32.argtypes = [vp, vp]
    lib.rt_readback_accum.argtypes = [vp, vp]
    lib.rt_device_accum.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]
    lib.rt_device_rgb8.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_size_t)]
    lib.rt_resolve_device.argtypes = [vp]
    lib.rt_render_frame.argtypes = [vp, vp, u32, vp, u32, fp, u32, u32, C.POINTER(RtRenderParams), vp]
    lib.rt_stats_get.argtypes = [vp, C.POINTER(RtStats)]
    lib.rt_stats_reset.argtypes = [vp]
    lib.rt_write_ppm.argtypes = [C.c_char_p, u32, u32, vp]
    lib.rt_fp32_peak_probe.argtypes = [vp, C.POINTER(C.c_double)]
    lib.rt_tf32_peak_probe.argtypes = [vp, C.POINTER(C.c_double)]
    lib.rt_filter_info.argtypes = [vp, C.POINTER(C.c_uint32)]
    lib.rt_render_progressive.argtypes = [vp, C.POINTER(RtRenderParams), u32, u32, PREVIEW_FN, vp, vp, C.POINTER(u32)]
    lib.rt_debug_sample_ball.argtypes = [vp, C.c_uint64, C.c_uint64, C.c_int, C.POINTER(C.c_float)]
    lib.rt_queue_open.argtypes = [C.POINTER(vp), C.c_char_p, i32]
    lib.rt_queue_reset.argtypes = [vp, u32, u64]
    lib.rt_queue_next.argtypes = [vp, u32, u64, C.POINTER(u64)]
    lib.rt_queue_close.argtypes = [vp, i32]
    lib.rt_get_stream.argtypes = [vp, C.POINTER(vp)]
    lib.rt_queue_dev_create.argtypes = [vp, C.POINTER(vp)]
    lib.rt_queue_dev_reset.argtypes = [vp, vp, u32, u32]
    lib.rt_queue_dev_free.argtypes = [vp, vp]
    lib.rt_render_queue.argtypes = [vp, vp, u32, u32, u32, C.POINTER(RtRenderParams), vp, u32, u32]
    lib.rt_resolve_tiles_to.argtypes = [vp, vp]
    lib.rt_queue_tiles_rendered.argtypes = [vp, C.POINTER(u32)]
    lib.rt_queue_trace.argtypes = [vp, vp, vp, u32, C.POINTER(u32)]
    lib.rt_copy_rgb8.argtypes = [vp, vp]
    lib.rt_ipc_export.argtypes = [vp, vp, vp]
    lib.rt_ipc_import.argtypes = [vp, vp, C.POINTER(vp)]
    lib.rt_ipc_close.argtypes = [vp, vp]
    lib.rt_multi_create.argtypes = [C.POINTER(vp), C.POINTER(i32), i32]
    lib.rt_multi_destroy.argtypes = [vp]
    lib.rt_multi_device_count.argtypes = [vp]
    lib.rt_multi_ctx.argtypes = [vp, i32]
    lib.rt_multi_render_frame.argtypes = [vp, vp, u32, vp, u32, fp, u32, u32, C.POINTER(RtRenderParams), u32, u32, u32, vp]
    lib.rt_multi_render_again.argtypes = [vp, C.POINTER(RtRenderParams), u32, vp]
    lib.rt_multi_last_frame.argtypes = [vp, C.POINTER(u32), C.POINTER(i32), C.POINTER(C.c_double)]
    lib.rt_closest_hits_static.argtypes = [vp, u32, vp, u32, vp, vp, vp, vp, u32]
    for name in EXPORTS:
        fn = getattr(lib, name)
        if name not in ("rt_last_error", "rt_version", "rt_multi_ctx"):
            fn.restype = C.c_int
    lib.rt_multi_ctx.restype = vp
    _lib = lib
    return lib


def _check(lib, rc):
    if rc != 0:
        raise RtError(f"rt error {rc}: {lib.rt_last_error().decode()}")


def pack_scene(scene):
    """scenes.Scene -> (rt_sphere[n], rt_material[m]) structured arrays."""
    sph = np.zeros(scene.n, SPHERE_DTYPE)
    sph["cx"], sph["cy"], sph["cz"], sph["r"] = scene.spheres.T
    sph["mat"] = scene.sphere_mat
    mat = np.zeros(scene.m, MATERIAL_DTYPE)
    mat["r"], mat["g"], mat["b"], mat["param"] = scene.materials.T
    mat["kind"] = scene.mat_kind
    return sph, mat


class Backend:
    """One rt_ctx on one GPU."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.ctx = C.c_void_p()
        _check(self.lib, self.lib.rt_create(C.byref(self.ctx), int(device)))
        self.device = device
        self.W = self.H = 0

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.rt_destroy(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- scene / camera / framebuffer ---------------------------------------
    def upload_scene(self, scene):
        sph, mat = pack_scene(scene)
        _check(self.lib, self.lib.rt_scene_upload(self.ctx, sph.ctypes.data, sph.shape[0], mat.ctypes.data, mat.shape[0]))
        self.set_camera(scene.camera)

    def set_camera(self, cam13):
        cam = np.ascontiguousarray(cam13, np.float32)
        fp = C.POINTER(C.c_float)
        pos, lookat, up = cam[0:3].copy(), cam[3:6].copy(), cam[6:9].copy()
        _check(self.lib, self.lib.rt_camera_set(self.ctx, pos.ctypes.data_as(fp), lookat.ctypes.data_as(fp),
                                                 up.ctypes.data_as(fp),
                                                 float(cam[9]), float(cam[10]), float(cam[11]), float(cam[12])))

    def camera_state(self):
        out = np.zeros(13, np.float32)
        _check(self.lib, self.lib.rt_camera_get(self.ctx, out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def framebuffer(self, W, H):
        _check(self.lib, self.lib.rt_framebuffer(self.ctx, int(W), int(H)))
        self.W, self.H = int(W), int(H)

    def clear(self):
        _check(self.lib, self.lib.rt_clear(self.ctx))

    def set_stream(self, cuda_stream_ptr):
        _check(self.lib, self.lib.rt_set_stream(self.ctx, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        _check(self.lib, self.lib.rt_sync(self.ctx))

    def throttle(self, max_in_flight):
        _check(self.lib, self.lib.rt_throttle(self.ctx, int(max_in_flight)))

    # ---- rendering ----------------------------------------------------------
    def render(self, spp, max_depth=50, seed=1, tile=None, sample_base=0, flags=0):
        x0, y0, w, h = tile if tile is not None else (0, 0, self.W, self.H)
        t = RtTile(x0, y0, w, h)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), int(sample_base), int(flags))
        _check(self.lib, self.lib.rt_render(self.ctx, C.byref(t), C.byref(p)))

    def render_tiles(self, tiles, tile_w, tile_h, spp, max_depth=50, seed=1, sample_base=0, flags=0):
        """tiles: (n,4) uint32 array of {x0,y0,w,h}; one persistent launch for the whole list."""
        tiles = np.ascontiguousarray(tiles, np.uint32).reshape(-1, 4)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), int(sample_base), int(flags))
        _check(self.lib, self.lib.rt_render_tiles(self.ctx, tiles.ctypes.data, tiles.shape[0], int(tile_w), int(tile_h), C.byref(p)))

    # ---- device-side tile queue (multi-GPU) ---------------------------------
    def get_stream(self):
        p = C.c_void_p()
        _check(self.lib, self.lib.rt_get_stream(self.ctx, C.byref(p)))
        return p.value

    def queue_dev_create(self):
        """Range cursors of the device-side tile queue, allocated on THIS context's device (rank 0); -> device pointer."""
        p = C.c_void_p()
        _check(self.lib, self.lib.rt_queue_dev_create(self.ctx, C.byref(p)))
        return p.value

    def queue_dev_reset(self, cursors, ntiles, nranks):
        _check(self.lib, self.lib.rt_queue_dev_reset(self.ctx, C.c_void_p(cursors), int(ntiles), int(nranks)))

    def queue_dev_free(self, cursors):
        _check(self.lib, self.lib.rt_queue_dev_free(self.ctx, C.c_void_p(cursors)))

    def render_queue(self, tiles, tile_w, tile_h, cursors, rank, nranks, spp, max_depth=50, seed=1, sample_base=0, flags=0):
        """ONE persistent launch that pulls its tiles from the device-side queue (own range first, then stealing)."""
        tiles = np.ascontiguousarray(tiles, np.uint32).reshape(-1, 4)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), int(sample_base), int(flags))
        _check(self.lib, self.lib.rt_render_queue(self.ctx, tiles.ctypes.data, tiles.shape[0], int(tile_w), int(tile_h), C.byref(p),
                                                   C.c_void_p(cursors), int(rank), int(nranks)))

    def resolve_tiles_to(self, rgb8_frame_dev=None):
        """Resolve the tiles the last render_queue launch rendered into a device RGB8 frame (None: this context's own)."""
        _check(self.lib, self.lib.rt_resolve_tiles_to(self.ctx, C.c_void_p(rgb8_frame_dev) if rgb8_frame_dev else None))

    def tiles_rendered(self):
        n = C.c_uint32(0)
        _check(self.lib, self.lib.rt_queue_tiles_rendered(self.ctx, C.byref(n)))
        return int(n.value)

    def queue_trace(self, cap=65536):
        """(tile indices, acquisition times in ~us of this GPU's globaltimer) of the last render_queue launch, in order."""
        tiles = np.zeros(cap, np.uint32); when = np.zeros(cap, np.uint32); n = C.c_uint32(0)
        _check(self.lib, self.lib.rt_queue_trace(self.ctx, tiles.ctypes.data, when.ctypes.data, cap, C.byref(n)))
        k = min(cap, int(n.value))
        return tiles[:k], when[:k]

    def copy_rgb8(self):
        out = np.zeros((self.H, self.W, 3), np.uint8)
        _check(self.lib, self.lib.rt_copy_rgb8(self.ctx, out.ctypes.data))
        return out

    def ipc_export(self, devptr) -> bytes:
        h = (C.c_ubyte * 64)()
        _check(self.lib, self.lib.rt_ipc_export(self.ctx, C.c_void_p(devptr), h))
        return bytes(h)

    def ipc_import(self, handle: bytes):
        h = (C.c_ubyte * 64).from_buffer_copy(handle)
        p = C.c_void_p()
        _check(self.lib, self.lib.rt_ipc_import(self.ctx, h, C.byref(p)))
        return p.value

    def ipc_close(self, devptr):
        _check(self.lib, self.lib.rt_ipc_close(self.ctx, C.c_void_p(devptr)))

    def primary_hits(self, tile=None, flags=FLAG_EXACT):
        x0, y0, w, h = tile if tile is not None else (0, 0, self.W, self.H)
        t = RtTile(x0, y0, w, h)
        idx = np.zeros((h, w), np.int32)
        tt = np.zeros((h, w), np.float32)
        nrm = np.zeros((h, w, 3), np.float32)
        _check(self.lib, self.lib.rt_primary_hits(self.ctx, C.byref(t), int(flags), idx.ctypes.data, tt.ctypes.data, nrm.ctypes.data))
        return idx, tt, nrm

    def closest_hits(self, rays, flags=FLAG_EXACT):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        idx = np.zeros(n, np.int32)
        tt = np.zeros(n, np.float32)
        nrm = np.zeros((n, 3), np.float32)
        _check(self.lib, self.lib.rt_closest_hits(self.ctx, rays.ctypes.data, n, int(flags), idx.ctypes.data, tt.ctypes.data, nrm.ctypes.data))
        return idx, tt, nrm

    def readback_rgb8(self):
        out = np.zeros((self.H, self.W, 3), np.uint8)
        _check(self.lib, self.lib.rt_readback_rgb8(self.ctx, out.ctypes.data))
        return out

    def readback_f32(self):
        out = np.zeros((self.H, self.W, 3), np.float32)
        _check(self.lib, self.lib.rt_readback_f32(self.ctx, out.ctypes.data))
        return out

    def readback_accum(self):
        out = np.zeros((self.H, self.W, 4), np.uint64)
        _check(self.lib, self.lib.rt_readback_accum(self.ctx, out.ctypes.data))
        return out

    def resolve_device(self):
        _check(self.lib, self.lib.rt_resolve_device(self.ctx))

    def device_accum(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(self.lib, self.lib.rt_device_accum(self.ctx, C.byref(p), C.byref(n)))
        return p.value, n.value

    def device_rgb8(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(self.lib, self.lib.rt_device_rgb8(self.ctx, C.byref(p), C.byref(n)))
        return p.value, n.value

    def render_frame(self, scene, W, H, spp, max_depth=50, seed=1, flags=0, out=None, packed=None):
        """The cuda_impl()-shaped call: host scene in, host RGB8 out, copies inside."""
        sph, mat = packed if packed is not None else pack_scene(scene)
        cam = np.ascontiguousarray(scene.camera, np.float32)
        if out is None:
            out = np.zeros((H, W, 3), np.uint8)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), 0, int(flags))
        _check(self.lib, self.lib.rt_render_frame(self.ctx, sph.ctypes.data, sph.shape[0], mat.ctypes.data, mat.shape[0],
                                                   cam.ctypes.data_as(C.POINTER(C.c_float)), int(W), int(H), C.byref(p),
                                                   out.ctypes.data))
        self.W, self.H = int(W), int(H)
        return out

    # ---- statistics / probes ------------------------------------------------
    def render_progressive(self, spp, spp_per_pass, preview_every=1, on_preview=None, max_depth=50, seed=1,
                           sample_base=0, flags=0):
        """rt_render_progressive: on_preview(pass_index, spp_done, image[H,W,3] uint8) -> truthy to stop.
        Returns (final image, spp_done)."""
        W, H = self.W, self.H
        out = np.empty((H, W, 3), np.uint8)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), int(sample_base), int(flags))
        err = []

        def tramp(_user, npass, done, ptr):
            try:
                img = np.ctypeslib.as_array(ptr, shape=(H, W, 3))
                return 1 if on_preview(int(npass), int(done), img) else 0
            except BaseException as ex:   # never unwind through the C frame
                err.append(ex)
                return 1
        cb = PREVIEW_FN(tramp) if on_preview else PREVIEW_FN(0)
        done = C.c_uint32(0)
        _check(self.lib, self.lib.rt_render_progressive(self.ctx, C.byref(p), int(spp_per_pass), int(preview_every), cb, None,
                                                        out.ctypes.data_as(C.c_void_p), C.byref(done)))
        if err:
            raise err[0]
        return out, int(done.value)

    def stats(self):
        s = RtStats()
        _check(self.lib, self.lib.rt_stats_get(self.ctx, C.byref(s)))
        return s.as_dict()

    def reset_stats(self):
        _check(self.lib, self.lib.rt_stats_reset(self.ctx))

    def tf32_peak_probe(self):
        out = (C.c_double * 3)()
        _check(self.lib, self.lib.rt_tf32_peak_probe(self.ctx, out))
        return dict(tflops=out[0], ms=out[1], sm_count=int(out[2]))

    def filter_info(self):
        out = (C.c_uint32 * 4)()
        _check(self.lib, self.lib.rt_filter_info(self.ctx, out))
        return dict(tensor=bool(out[0]), columns=int(out[1]), k=int(out[2]), rays_per_cta=int(out[3]))

    def sample_ball(self, n, seed=1, lens=False):
        """n points of the throughput build's ball (or lens) sampler, (n, 3) float32."""
        out = np.empty((int(n), 3), np.float32)
        _check(self.lib, self.lib.rt_debug_sample_ball(self.ctx, int(n), int(seed), 1 if lens else 0,
                                                       out.ctypes.data_as(C.POINTER(C.c_float))))
        return out

    def fp32_peak_probe(self):
        out = (C.c_double * 3)()
        _check(self.lib, self.lib.rt_fp32_peak_probe(self.ctx, out))
        return dict(tflops=out[0], ms=out[1], sm_count=int(out[2]))


class _CtxView(Backend):
    """A Backend face on a context owned by an rt_multi (never destroyed from here)."""

    def __init__(self, lib, ctx, device, W, H):
        self.lib, self.ctx, self.device, self.W, self.H = lib, ctx, device, W, H

    def close(self):
        self.ctx = None


def _preload_bundled_nccl():
    """rt_multi dlopen()s "libnccl.so.2".  In a Python process that also imports torch, the library of that SONAME must be the
    one torch was built against (its wheel bundles a newer NCCL than the system's: loading the system one first makes a
    later `import torch` fail on a missing symbol).  So load the bundled library first, globally, when there is one."""
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia.nccl")
    except (ImportError, ValueError):
        spec = None
    if spec is None or not spec.submodule_search_locations:
        return None
    for base in spec.submodule_search_locations:
        cand = os.path.join(base, "lib", "libnccl.so.2")
        if os.path.exists(cand):
            try:
                C.CDLL(cand, mode=C.RTLD_GLOBAL)
                return cand
            except OSError:
                return None
    return None


class MultiBackend:
    """rt_multi: several GPUs of ONE process behind one call (device-side tile queue, peer-store gather, NCCL ack)."""
    GATHER_NCCL = 1

    def __init__(self, devices):
        _preload_bundled_nccl()
        self.lib = load_library()
        self.devices = [int(d) for d in devices]
        arr = (C.c_int * len(self.devices))(*self.devices)
        self.m = C.c_void_p()
        _check(self.lib, self.lib.rt_multi_create(C.byref(self.m), arr, len(self.devices)))
        self.W = self.H = 0

    def close(self):
        if getattr(self, "m", None):
            self.lib.rt_multi_destroy(self.m)
            self.m = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def ctx(self, i):
        return _CtxView(self.lib, C.c_void_p(self.lib.rt_multi_ctx(self.m, int(i))), self.devices[i], self.W, self.H)

    def render_frame(self, scene, W, H, spp, max_depth=50, seed=1, flags=0, tile_w=128, tile_h=64, gather=0, out=None, packed=None):
        sph, mat = packed if packed is not None else pack_scene(scene)
        cam = np.ascontiguousarray(scene.camera, np.float32)
        if out is None:
            out = np.zeros((H, W, 3), np.uint8)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), 0, int(flags))
        _check(self.lib, self.lib.rt_multi_render_frame(self.m, sph.ctypes.data, sph.shape[0], mat.ctypes.data, mat.shape[0],
                                                         cam.ctypes.data_as(C.POINTER(C.c_float)), int(W), int(H), C.byref(p),
                                                         int(tile_w), int(tile_h), int(gather), out.ctypes.data))
        self.W, self.H = int(W), int(H)
        return out

    def render_again(self, spp, max_depth=50, seed=1, flags=0, gather=0, out=None, want_image=True):
        if out is None and want_image:
            out = np.zeros((self.H, self.W, 3), np.uint8)
        p = RtRenderParams(int(spp), int(max_depth), int(seed), 0, int(flags))
        _check(self.lib, self.lib.rt_multi_render_again(self.m, C.byref(p), int(gather), out.ctypes.data if out is not None else None))
        return out

    def last_frame(self):
        n = len(self.devices)
        tiles = (C.c_uint32 * n)()
        nccl, ms = C.c_int(0), C.c_double(0)
        _check(self.lib, self.lib.rt_multi_last_frame(self.m, tiles, C.byref(nccl), C.byref(ms)))
        return {"tiles_per_device": list(tiles), "nccl": bool(nccl.value), "frame_ms": ms.value}


STATIC_MAX_SPHERES, STATIC_MAX_RAYS, STATIC_NO_SYNC = 4096, 65536, 1


def closest_hits_static(spheres_xyzr, rays, stream=None, flags=0):
    """rt_closest_hits_static (include/rt_api.h): closest hit with no context and no allocation -- the scratch is a
    fixed device array inside the library, so at most STATIC_MAX_SPHERES spheres and STATIC_MAX_RAYS rays per call.
    Host arrays in, host arrays out; runs on the current CUDA device.  No counterpart in the reference."""
    lib = load_library()
    sph = np.ascontiguousarray(spheres_xyzr, np.float32).reshape(-1, 4)
    rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
    n = rays.shape[0]
    idx = np.zeros(n, np.int32)
    tt = np.zeros(n, np.float32)
    nrm = np.zeros((n, 3), np.float32)
    _check(lib, lib.rt_closest_hits_static(sph.ctypes.data, sph.shape[0], rays.ctypes.data, n, idx.ctypes.data,
                                           tt.ctypes.data, nrm.ctypes.data, stream, int(flags)))
    return idx, tt, nrm


def write_ppm(path, rgb8):
    lib = load_library()
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    H, W = rgb8.shape[:2]
    _check(lib, lib.rt_write_ppm(os.fsencode(path), W, H, rgb8.ctypes.data))


class TileQueue:
    """Node-level work counter shared by the ranks of one box (rt_queue_*)."""

    def __init__(self, name: str, create: bool):
        self.lib = load_library()
        self.q = C.c_void_p()
        self.create = create
        _check(self.lib, self.lib.rt_queue_open(C.byref(self.q), name.encode(), 1 if create else 0))

    def reset(self, slot=0, value=0):
        _check(self.lib, self.lib.rt_queue_reset(self.q, int(slot), int(value)))

    def next(self, slot=0, n=1) -> int:
        """fetch_add(n) on counter `slot`; returns the previous value (n = 0 just reads)."""
        prev = C.c_uint64()
        _check(self.lib, self.lib.rt_queue_next(self.q, int(slot), int(n), C.byref(prev)))
        return int(prev.value)

    def close(self):
        if self.q:
            self.lib.rt_queue_close(self.q, 1 if self.create else 0)
            self.q = None
