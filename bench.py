#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 path-tracing backend.

Metric (BASELINE.json): Gsamples/s, a sample = one camera path with max depth 50, huge-scene.
  N = 1 : configs[2]  huge-scene 1280x720, 500 spp, depth 50 on one B200          ("huge-720p-500spp")
  N > 1 : configs[3]  huge-scene 3840x2160, 1024 spp, tile-sharded over N B200 with dynamic tile
                      stealing and an NCCL reduce of the RGB8 frame to rank 0        ("huge-4k-1024spp")
A step = one full pass of the hot path over the frame: clear accumulators, render every sample,
resolve (mean -> gamma -> 8 bit) and, for N > 1, gather the frame on rank 0.

  value  : W*H*spp / device time, scene already resident in HBM (CUDA events on the launching stream)
  e2e    : the same through the reference-facing call (rt_render_frame == cuda_impl with the scene as
           an argument): host scene in, host RGB8 out, host<->device copies inside the timed region
  --impl reference : the reference's own CPU code (oracle/_ref, else the C port) on the host cores.

One JSON line on stdout (rank 0).  Under torchrun every rank runs this file.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from raytracinginoneweekend_b200 import scenes  # noqa: E402

METRIC = "Gsamples/s (camera paths, max depth 50) huge-scene"
UNIT = "Gsamples/s"
DEPTH = 50
SEED = 1
WORKLOADS = {
    # name: (scene, W, H, spp)
    "huge-720p-500spp": ("huge", 1280, 720, 500),
    "huge-4k-1024spp": ("huge", 3840, 2160, 1024),
    "simple-720p-64spp": ("simple", 1280, 720, 64),
    "synthetic1m-1080p-256spp": ("synthetic1m", 1920, 1080, 256),      # BASELINE configs[4]: GPU-built LBVH
}
TILE_W, TILE_H = 128, 64


def workload_for(n_gpus: int, name: str | None):
    if name in (None, "auto"):
        name = "huge-720p-500spp" if n_gpus == 1 else "huge-4k-1024spp"
    return name, WORKLOADS[name]


def algorithmic_flop(segments, samples, pixels, n_spheres):
    """FLOP = segments*(18N + 64) + samples*32 + pixels*8  (FMA = 2, `a` hoisted)."""
    return segments * (18 * n_spheres + 64) + samples * 32 + pixels * 8


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc = gpu_index, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=10)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nme, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples in the upper half of the observed power range
        thr = (max(pw) + min(pw)) / 2
        load = [s for s, p in zip(sm, pw) if p >= thr] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------- CPU arms
def cpu_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def load_cpu_oracle():
    """The compiled reference if it was built (kind 'reference'), else the C port (kind 'port')."""
    from oracle.loader import PortOracle, RefOracle
    if RefOracle.available():
        return RefOracle()
    return PortOracle()


def spread_rows(H, n_strips, rows_per_strip):
    """Evenly spaced strips so a bounded sample sees sky, spheres and ground in the frame's proportions."""
    n_strips = max(1, min(n_strips, H // rows_per_strip))
    return [(int(i * H / n_strips), rows_per_strip) for i in range(n_strips)]


def cpu_sample(orc, sc, W, H, strips, spp, threads):
    """Render the given row strips at `spp`; returns (samples, seconds)."""
    samples, secs = 0, 0.0
    for y0, rows in strips:
        r = orc.render(sc, W, H, spp, DEPTH, SEED, y0=y0, rows=rows, nthreads=threads)
        samples += W * rows * spp
        secs += r["seconds"]
    return samples, secs


def cpu_plan(orc, sc, W, H, threads, target_s):
    """Choose a bounded sample (strips x spp) that takes about target_s on this box."""
    probe = spread_rows(H, 24, 1)
    n, s = cpu_sample(orc, sc, W, H, probe, 1, threads)
    rate = n / max(s, 1e-9)
    budget = rate * target_s                        # samples affordable
    if budget >= W * H:
        spp = int(max(1, min(64, budget // (W * H))))
        return [(0, H)], spp, f"full {W}x{H} frame at {spp} spp, depth {DEPTH}"
    rows = int(max(24, budget // W))
    n_strips = max(1, rows // 2)
    strips = spread_rows(H, n_strips, 2)
    return strips, 1, f"{len(strips)} evenly spaced 2-row strips of the {W}x{H} frame ({len(strips) * 2} rows) at 1 spp, depth {DEPTH}"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wname, (scene_name, W, H, spp) = workload_for(args.gpus, args.workload)
    sc = scenes.get_scene(scene_name, W, H)
    orc = load_cpu_oracle()
    threads = cpu_threads()
    per_step = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    strips, s_spp, desc = cpu_plan(orc, sc, W, H, threads, per_step)
    for _ in range(args.warmup):
        cpu_sample(orc, sc, W, H, strips, s_spp, threads)
    tot_n, tot_s = 0, 0.0
    for _ in range(args.steps):
        n, s = cpu_sample(orc, sc, W, H, strips, s_spp, threads)
        tot_n += n; tot_s += s
    value = tot_n / tot_s / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wname, "scene": scene_name, "width": W, "height": H, "spp": spp, "max_depth": DEPTH,
                   "note": "reference CPU path (hit_world/apply_material/camera::ray) on the host cores; each step is a bounded sample of the workload"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": orc.kind, "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------- GPU arm
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from raytracinginoneweekend_b200 import multigpu
    from raytracinginoneweekend_b200.backend import Backend, TileQueue, pack_scene

    os.environ["NCCL_DEBUG"] = "WARN"       # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run (one rank per GPU)")
        raise SystemExit(f"--gpus {args.gpus} does not match WORLD_SIZE {world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the backend has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))

    wname, (scene_name, W, H, spp) = workload_for(world, args.workload)
    if args.spp:
        spp = args.spp
    sc = scenes.get_scene(scene_name, W, H)
    packed = pack_scene(sc)
    be = Backend(local)
    stream = torch.cuda.Stream(device=local)
    be.set_stream(stream.cuda_stream)
    be.upload_scene(sc)
    be.framebuffer(W, H)
    flags = 0

    def barrier():
        if world > 1:
            dist.barrier()

    tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, TILE_W, TILE_H), world)
    queue = sched = renderer = None
    if world > 1:
        qname = f"/rtiow_bench_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            queue = TileQueue(qname, create=True)
        barrier()
        if rank != 0:
            queue = TileQueue(qname, create=False)
        sched = multigpu.StealingScheduler(queue, rank, world, len(tiles))
        renderer = multigpu.CudaTileRenderer(be, TILE_W, TILE_H, spp, DEPTH, SEED, flags)

    def step():
        """clear -> render (all samples) -> resolve -> (N>1) gather on rank 0; all on `stream`."""
        be.clear()
        if world == 1:
            be.render(spp, DEPTH, seed=SEED, flags=flags)
            be.resolve_device()
        else:
            multigpu.render_distributed(renderer, sched, tiles, barrier)
            be.resolve_device()
            with torch.cuda.stream(stream):
                dist.reduce(multigpu.device_rgb8_tensor(be), dst=0, op=dist.ReduceOp.SUM)

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")   # > 126 MB L2

    for _ in range(args.warmup):
        step()
    be.sync()
    torch.cuda.synchronize()

    probe = be.fp32_peak_probe()
    finfo = be.filter_info()
    tprobe = be.tf32_peak_probe() if finfo["tensor"] else None
    be.reset_stats()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    for i in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()                         # L2 flush between timed iterations (outside the event pair)
            starts[i].record(stream)
        step()
        with torch.cuda.stream(stream):
            ends[i].record(stream)
    be.sync()
    torch.cuda.synchronize()
    barrier()
    t_wall1 = time.perf_counter()
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    st = be.stats()
    clk = clocks.stop() if rank == 0 else None

    # ---- e2e through the reference-facing call, host buffers, copies inside the timed region
    out = np.zeros((H, W, 3), np.uint8)
    h2d = packed[0].nbytes + packed[1].nbytes + sc.camera.nbytes
    e2e_steps = max(1, min(args.steps, 3)) if world > 1 else args.steps

    def e2e_step():
        if world == 1:
            be.render_frame(sc, W, H, spp, DEPTH, seed=SEED, flags=flags, out=out, packed=packed)
        else:
            be.upload_scene(sc)                   # H2D of the scene on every rank
            step()
            if rank == 0:
                p, nbytes = be.device_rgb8()
                t = multigpu.device_rgb8_tensor(be)
                with torch.cuda.stream(stream):
                    host = t.cpu()                # D2H of the gathered frame
                stream.synchronize()
                out[...] = host.numpy().reshape(H, W, 3)
            be.sync()

    e2e_step()
    barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    be.sync(); torch.cuda.synchronize(); barrier()
    e2e_s = time.perf_counter() - t0

    # ---- reduce over ranks: max time, summed work
    t = torch.tensor([dev_ms, e2e_s * 1e3, t_wall1 - t_wall0], dtype=torch.float64, device=f"cuda:{local}")
    work = torch.tensor([st["samples"], st["segments"], st["launches"]], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(work, op=dist.ReduceOp.SUM)
    # per-rank view (diagnosis of stragglers): device time of the timed steps, time inside the render kernels, tiles taken
    mine = torch.tensor([dev_ms, st["render_ms"], float(sched.own_taken if sched else 0), float(sched.stolen if sched else 0)],
                        dtype=torch.float64, device=f"cuda:{local}")
    per_rank = [torch.zeros_like(mine) for _ in range(world)] if world > 1 else [mine]
    if world > 1:
        dist.all_gather(per_rank, mine)
    ranks_info = [{"rank": i, "step_ms": float(v[0]) / max(1, args.steps), "render_kernel_ms": float(v[1]) / max(1, args.steps),
                   "own_tiles_last_step": int(v[2]), "stolen_tiles_last_step": int(v[3])} for i, v in enumerate(per_rank)]
    dev_ms, e2e_ms, wall_s = (float(x) for x in t.tolist())
    samples, segments, launches = (float(x) for x in work.tolist())

    if rank == 0:
        total = float(W) * H * spp * args.steps
        value = total / (dev_ms * 1e-3) / 1e9
        e2e_value = float(W) * H * spp * e2e_steps / (e2e_ms * 1e-3) / 1e9
        # roofline of the dominant kernel (one launch per step at N=1), live: device counters / CUDA-event launch time.
        #  * brute force <= 1024 spheres: k_render_tc.  The sphere tests run on the tensor pipe as a TF32x3 product
        #    (64 flop per ray x padded sphere column: K = 32 slots), so the bound is "tensor": executed tensor FLOP/s
        #    against the TF32 peak measured in this run.  The SURVEY 8(d) algorithmic FP32 count is reported next to it.
        #  * otherwise k_render on the FP32 FMA pipe: SURVEY 8(d) algorithmic FLOP/s against the measured FMA peak.
        render_ms_per_launch = st["render_ms"] / max(1, args.steps)
        flop = algorithmic_flop(st["segments"], st["samples"], W * H * args.steps, sc.n) / max(1, args.steps)
        achieved = flop / (render_ms_per_launch * 1e-3) / 1e12
        traffic = None
        prof = os.path.join(ROOT, "profiles", "ncu_summary.json")
        kname = "k_render_tc" if finfo["tensor"] else "k_render"
        if os.path.exists(prof) and world == 1:      # the committed capture is of the N=1 workload's launch
            try:
                traffic = json.load(open(prof)).get(kname, {}).get("dram_bytes_per_launch")
            except Exception:
                traffic = None
        fp32_model = {"achieved": achieved, "peak": probe["tflops"], "unit": "TFLOP/s", "frac": achieved / probe["tflops"],
                      "flop_model": "SURVEY 8(d): segments*(18*N+64) + samples*32 + pixels*8, segments/samples from device counters",
                      "peak_source": "FMA-chain probe measured in this run (rt_fp32_peak_probe); nominal 148 SM x 128 x 2 x 1.965 GHz = 74.45"}
        if finfo["tensor"]:
            tflop = st["segments"] * finfo["columns"] * 2.0 * finfo["k"] / max(1, args.steps)
            tach = tflop / (render_ms_per_launch * 1e-3) / 1e12
            roofline = {"bound": "tensor", "achieved": tach, "peak": tprobe["tflops"], "unit": "TFLOP/s",
                        "frac": tach / tprobe["tflops"], "traffic": traffic,
                        "kernel": "k_render_tc (persistent bounce loop; candidate filter = TF32x3 tcgen05 product, accumulators in TMEM)",
                        "peak_source": "tcgen05.mma kind::tf32 stream measured in this run (rt_tf32_peak_probe); nominal dense TF32 1.1 PFLOP/s",
                        "flop_model": f"segments * {finfo['columns']} sphere columns * K={finfo['k']} * 2 (split-TF32 product of the ray x sphere discriminant form)",
                        "note": "the kernel is bound by issue slots of the sign-mask epilogue and the shading code, not by the tensor pipe; "
                                "fp32_algorithmic is SURVEY 8(d)'s FP32 model (a fraction above 1 = more sphere tests per second than the FMA pipe could execute)",
                        "fp32_algorithmic": fp32_model, "ms_per_launch": render_ms_per_launch, "rank": 0}
            accel = "brute force: tensor-core candidate filter (sphere matrix in shared memory via TMA, accumulators in TMEM) + reference test on candidates"
        elif st["node_visits"] > 0:
            # LBVH: bytes = node visits x 64 B (one node = the boxes of both children) + leaf
            # tests x 16 B, against the measured HBM copy bandwidth.  The tree's hot levels are L2/L1 resident, so this
            # is traffic the SMs request, not DRAM traffic (ncu: profiles/r01_ncu_k_render_bvh.json).
            peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
            hbm = None
            if os.path.exists(peaks_path):
                try:
                    hbm = float(json.load(open(peaks_path)).get("hbm_gbs"))
                except Exception:
                    hbm = None
            peak_src = "MEASURED_PEAKS.json hbm_gbs" if hbm else "fallback 6550 GB/s (B200_PROFILING.md)"
            hbm = hbm or 6550.0
            nbytes = (st["node_visits"] * 64.0 + st["sphere_tests"] * 16.0) / max(1, args.steps)
            ach = nbytes / (render_ms_per_launch * 1e-3) / 1e9
            try:
                traffic = json.load(open(prof)).get("k_render_bvh", {}).get("dram_bytes_per_launch") if world == 1 else None
            except Exception:
                traffic = None
            roofline = {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": traffic,
                        "kernel": "k_render_bvh (persistent bounce loop, resumable LBVH traversal, stacks in shared memory)",
                        "peak_source": peak_src,
                        "byte_model": "node_visits * 64 B + leaf_tests * 16 B (device counters)",
                        "note": "requested bytes are served mostly by L1/L2 (the capture's DRAM traffic is for 8 spp); the kernel is issue bound (two conservative box tests per node)",
                        "node_visits_per_segment": st["node_visits"] / max(1, st["segments"]),
                        "leaf_tests_per_segment": st["sphere_tests"] / max(1, st["segments"]),
                        "ms_per_launch": render_ms_per_launch, "rank": 0}
            accel = "GPU-built LBVH (Morton + radix sort + Karras + refit), conservative traversal, reference test at the leaves"
        else:
            roofline = dict(fp32_model, bound="fp32", traffic=traffic, kernel="k_render / k_render_small (persistent bounce loop: sphere tests + shading)",
                            ms_per_launch=render_ms_per_launch, rank=0)
            accel = "brute-force (shared-memory SoA spheres)"
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": wname, "scene": scene_name, "n_spheres": sc.n, "width": W, "height": H, "spp": spp,
                       "max_depth": DEPTH, "seed": SEED, "accel": accel,
                       "parallelism": f"tile-sharded x{world}, stealing, NCCL reduce to rank 0" if world > 1 else "1 GPU",
                       "l2": "flushed between timed steps (256 MiB write)",
                       "segments_per_sample": segments / max(1.0, samples)},
            "clocks": clk,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d * world),
                    "d2h_bytes_per_step": int(W * H * 3), "steps": e2e_steps,
                    "call": "rt_render_frame (cuda_impl-shaped: host scene in, host RGB8 out)" if world == 1
                            else "rt_scene_upload + tile-stolen rt_render_tiles + NCCL reduce + D2H on rank 0"},
            "gpu_launches": int(launches),
            "roofline": roofline,
            "wall_s": wall_s,
        }
        if world > 1:
            line["ranks"] = ranks_info
        if world == 1 and not args.no_cpu_baseline:
            try:
                orc = load_cpu_oracle()
                threads = cpu_threads()
                strips, s_spp, desc = cpu_plan(orc, sc, W, H, threads, 12.0)
                n, s = cpu_sample(orc, sc, W, H, strips, s_spp, threads)
                line["cpu_baseline"] = {"value": n / s / 1e9, "unit": UNIT, "cores": threads, "kind": orc.kind, "sample": desc}
            except Exception as exc:  # the checker is optional for the number, never for the product
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": cpu_threads(), "kind": "port", "sample": f"failed: {exc}"}
        print(json.dumps(line), flush=True)

    if queue is not None:
        barrier()
        queue.close()
    be.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", default="auto", choices=["auto"] + list(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override spp (debug only; invalidates the headline number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        print(f"bench.py: warning: --warmup {args.warmup} < 3 breaks the timing rules", file=sys.stderr)
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
