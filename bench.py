#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 path-tracing backend.

Metric (BASELINE.json): Gsamples/s, a sample = one camera path with max depth 50, huge-scene.
  N = 1 : configs[2]  huge-scene 1280x720, 500 spp, depth 50 on one B200          ("huge-720p-500spp")
  N > 1 : configs[3]  huge-scene 3840x2160, 1024 spp, tile-sharded over N B200 with dynamic tile
                      stealing and the framebuffer gathered on rank 0 over NVLink      ("huge-4k-1024spp")
A step = one full pass of the hot path over the frame: clear accumulators, render every sample,
resolve (mean -> gamma -> 8 bit) and, for N > 1, gather the frame on rank 0.

  value  : W*H*spp / device time, scene already resident in HBM (CUDA events on the launching stream)
  e2e    : the same through the reference-facing call (rt_render_frame == cuda_impl with the scene as
           an argument): host scene in, host RGB8 out, host<->device copies inside the timed region
  --impl reference : the reference's own CPU code (oracle/_ref, else the C port) on the host cores.

Everything else BASELINE.json names is measured in the same run at reduced steps and reported under
`secondary` (configs[0] simple scene, configs[4] 1M-sphere LBVH, the FMA-pipe filter kernel that answers the
">= 50 % of FP32 peak" target, the reference's own CUDA path), `parity_build` (the same tensor-core kernel with
the parity shading, bit-identical to the oracle) and, at N = 1, `scaling_anchor` (the N > 1 workload on one GPU,
so strong-scaling efficiency can be recomputed on ONE workload).  `frame_sha256` is the hash of rank 0's RGB8
frame: the N-GPU frame of huge-4k-1024spp must carry the hash the 1-GPU anchor prints.

One JSON line on stdout (rank 0), printed last.  Under torchrun every rank runs this file.
"""
from __future__ import annotations

import argparse
import hashlib
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
    "huge-720p-500spp": ("huge", 1280, 720, 500),                       # BASELINE configs[2]
    "huge-4k-1024spp": ("huge", 3840, 2160, 1024),                      # BASELINE configs[3]
    "simple-720p-64spp": ("simple", 1280, 720, 64),                     # BASELINE configs[0]
    "synthetic1m-1080p-256spp": ("synthetic1m", 1920, 1080, 256),      # BASELINE configs[4]: GPU-built LBVH
}
TILE_W, TILE_H = 128, 64


def workload_for(n_gpus: int, name: str | None):
    if name in (None, "auto"):
        name = "huge-720p-500spp" if n_gpus == 1 else "huge-4k-1024spp"
    return name, WORKLOADS[name]


def workload_config(wname, scene_name, n_spheres, W, H, spp):
    """The workload definition, identical for both arms (implementation details live outside `config`)."""
    return {"workload": wname, "scene": scene_name, "n_spheres": int(n_spheres), "width": W, "height": H, "spp": spp,
            "max_depth": DEPTH, "seed": SEED}


def algorithmic_flop(segments, samples, pixels, n_spheres):
    """FLOP = segments*(18N + 64) + samples*32 + pixels*8  (FMA = 2, `a` hoisted)."""
    return segments * (18 * n_spheres + 64) + samples * 32 + pixels * 8


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


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


class CpuSample:
    """A bounded, unbiased sample of one workload for the CPU arms: every `ystep`-th row of the frame at `spp`
    samples per pixel, rendered by ONE call -- the harness deals rows to its worker threads one at a time, so a call
    must hold many more rows than threads for all of them to stay busy (round 1 issued 2-row calls and kept 2 of
    16-32 threads busy).  `threads_busy` is measured: process CPU time / wall time of the call."""

    def __init__(self, orc, sc, W, H, threads):
        self.orc, self.sc, self.W, self.H, self.threads = orc, sc, W, H, threads
        self.ystep, self.spp = 1, 1

    def rows(self):
        return (self.H + self.ystep - 1) // self.ystep

    def run(self, threads=None):
        """-> (samples, wall seconds, threads busy on average)"""
        threads = threads or self.threads
        c0, t0 = time.process_time(), time.perf_counter()
        r = self.orc.render(self.sc, self.W, self.H, self.spp, DEPTH, SEED, y0=0, nthreads=threads, ystep=self.ystep)
        c1, t1 = time.process_time(), time.perf_counter()
        secs = r["seconds"]
        return self.W * self.rows() * self.spp, secs, (c1 - c0) / max(t1 - t0, 1e-9)

    def plan(self, target_s):
        """Choose (ystep, spp) so that one call takes about target_s on this box; at least 8 rows per thread."""
        max_step = max(1, self.H // (8 * self.threads))
        self.ystep, self.spp = max_step, 1
        n, s, _ = self.run()                                   # probe with the same shape (all threads busy)
        rate = n / max(s, 1e-9)
        budget = rate * target_s                               # samples affordable per step
        full = self.W * self.H
        if budget >= full:
            self.ystep, self.spp = 1, int(max(1, min(64, budget // full)))
        else:
            self.ystep, self.spp = int(min(max_step, max(1, -(-full // int(max(budget, 1))))) ), 1
        return self.describe()

    def describe(self):
        what = f"full {self.W}x{self.H} frame" if self.ystep == 1 else \
            f"every {self.ystep}th row of the {self.W}x{self.H} frame ({self.rows()} rows, one call)"
        return f"{what} at {self.spp} spp, depth {DEPTH}"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wname, (scene_name, W, H, spp) = workload_for(args.gpus, args.workload)
    sc = scenes.get_scene(scene_name, W, H)
    orc = load_cpu_oracle()
    threads = cpu_threads()
    per_step = max(2.0, min(20.0, 150.0 / max(1, args.steps + args.warmup)))
    smp = CpuSample(orc, sc, W, H, threads)
    desc = smp.plan(per_step)
    for _ in range(args.warmup):
        smp.run()
    tot_n, tot_s, busy = 0, 0.0, []
    for _ in range(args.steps):
        n, s, b = smp.run()
        tot_n += n; tot_s += s; busy.append(b)
    value = tot_n / tot_s / 1e9
    threads_busy = float(np.mean(busy)) if busy else 0.0
    # what the UNMODIFIED reference does: its pixel loop is single-threaded (src/main.cxx:197-203; the row threading is
    # the harness's) -- one thread on a thinner sample of the same frame
    one = CpuSample(orc, sc, W, H, 1)
    one.ystep, one.spp = max(1, H // 24), 1
    n1, s1, _ = one.run(threads=1)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(1, args.steps),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(wname, scene_name, sc.n, W, H, spp),
        "impl_detail": {"note": "reference CPU path (hit_world/apply_material/camera::ray) on the host cores; each step is a "
                                "bounded sample of the workload (strided rows, one call, all threads busy)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "threads_busy": round(threads_busy, 2),
                         "kind": orc.kind, "sample": desc,
                         "single_thread": {"value": n1 / s1 / 1e9, "unit": UNIT,
                                           "note": "the reference's own loop is single-threaded (src/main.cxx:197-203); "
                                                   f"{one.describe()}"}},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    if threads > 2 and threads_busy < 0.5 * threads:
        print(f"bench.py: reference arm kept only {threads_busy:.1f} of {threads} threads busy -- the number is not an "
              f"all-cores baseline", file=sys.stderr)
        return 3
    return 0


# ----------------------------------------------------------------------------- GPU arm
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    from raytracinginoneweekend_b200 import multigpu
    from raytracinginoneweekend_b200.backend import (FLAG_EXACT, FLAG_FILTER_FMA, KERNEL_NAMES, Backend, TileQueue,
                                                     pack_scene)

    # NCCL's INFO lines (if the caller asked for them) go wherever NCCL_DEBUG_FILE says, stdout by default; the JSON
    # line is printed LAST, after the process group is destroyed, so it stays the final line of rank 0's stdout.
    os.environ.setdefault("NCCL_DEBUG", "WARN")
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
    dev = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(dev))

    def barrier():
        if world > 1:
            dist.barrier()

    be = Backend(local)
    stream = torch.cuda.Stream(device=local)
    be.set_stream(stream.cuda_stream)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)   # > 126 MB L2
    queue = dq = None
    if world > 1 and args.mgpu == "device":
        # device-side tile queue: cursors + frame in rank 0's memory, mapped by the other ranks through CUDA IPC
        dq = multigpu.DeviceQueueFrame(be, rank, world, dist, torch, stream)
    elif world > 1:
        qname = f"/rtiow_bench_{os.environ.get('MASTER_PORT', '0')}"
        if rank == 0:
            queue = TileQueue(qname, create=True)
        barrier()
        if rank != 0:
            queue = TileQueue(qname, create=False)

    def measure(wname, steps, warmup, flags=0, want_e2e=False, want_hash=False, e2e_steps=None):
        """One workload on all ranks: -> dict (rank 0) with value / ms_per_step / stats / optional e2e and frame hash."""
        scene_name, W, H, spp = WORKLOADS[wname]
        if args.spp:
            spp = args.spp
        sc = scenes.get_scene(scene_name, W, H)
        packed = pack_scene(sc)
        be.upload_scene(sc)
        be.framebuffer(W, H)
        tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, TILE_W, TILE_H), world)
        sched = renderer = None
        if dq is not None:
            dq.map_frame()
        elif world > 1:
            sched = multigpu.StealingScheduler(queue, rank, world, len(tiles))
            renderer = multigpu.CudaTileRenderer(be, TILE_W, TILE_H, spp, DEPTH, SEED, flags)
        phase_trace = []
        r_start = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        r_end = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]

        def step(i=None):
            """clear -> render (all samples) -> resolve -> (N>1) gather on rank 0; all ordered on `stream`.
            r_start/r_end bracket the render phase as ONE interval (overlapped chunk launches are not summed)."""
            if dq is not None:                # ONE persistent launch per GPU: tiles pulled from the device-side queue,
                if i is not None:             # texels stored into rank 0's frame, NCCL all-reduce as the acknowledgement
                    r_start[i].record(stream)
                dq.render_frame(tiles, TILE_W, TILE_H, spp, DEPTH, SEED, flags, trace=(phase_trace if i == steps - 1 else None))
                if i is not None:
                    r_end[i].record(stream)
                return
            be.clear()
            if i is not None:
                r_start[i].record(stream)
            if world == 1:
                be.render(spp, DEPTH, seed=SEED, flags=flags)
                be.resolve_device()
                if i is not None:
                    r_end[i].record(stream)
            else:
                multigpu.render_distributed(renderer, sched, tiles, barrier)
                be.resolve_device()               # joins the overlapped chunk launches, then resolves this rank's tiles
                if i is not None:
                    r_end[i].record(stream)
                with torch.cuda.stream(stream):
                    dist.reduce(multigpu.device_rgb8_tensor(be), dst=0, op=dist.ReduceOp.SUM)

        for _ in range(warmup):
            step()
        be.sync()
        torch.cuda.synchronize()
        be.reset_stats()
        starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
        barrier()
        torch.cuda.synchronize()
        t_wall0 = time.perf_counter()
        for i in range(steps):
            with torch.cuda.stream(stream):
                flush.zero_()                     # L2 flush between timed iterations (outside the event pair)
                starts[i].record(stream)
            step(i)
            with torch.cuda.stream(stream):
                ends[i].record(stream)
        be.sync()
        torch.cuda.synchronize()
        barrier()
        wall = time.perf_counter() - t_wall0
        dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
        st = be.stats()
        # render phase of this rank per step: the kernel's own event pair when a step is one launch; the interval around
        # the chunk sequence when the host schedules overlapped chunks (their event pairs overlap and must not be summed)
        render_ms = st["render_ms"] if (world == 1 or dq is not None) else sum(s.elapsed_time(e) for s, e in zip(r_start, r_end))

        frame_hash = None
        if want_hash and rank == 0:               # rank 0's device frame after the last step (gathered when N > 1)
            with torch.cuda.stream(stream):
                host = multigpu.device_rgb8_tensor(be).cpu()
            stream.synchronize()
            frame_hash = hashlib.sha256(host.numpy().tobytes()).hexdigest()

        e2e = None
        if want_e2e:
            out = np.zeros((H, W, 3), np.uint8)
            h2d = packed[0].nbytes + packed[1].nbytes + sc.camera.nbytes
            n_e2e = e2e_steps or steps

            def e2e_step():
                if world == 1:
                    be.render_frame(sc, W, H, spp, DEPTH, seed=SEED, flags=flags, out=out, packed=packed)
                else:
                    be.upload_scene(sc)           # H2D of the scene on every rank
                    step()
                    if rank == 0:
                        with torch.cuda.stream(stream):
                            host = multigpu.device_rgb8_tensor(be).cpu()      # D2H of the gathered frame
                        stream.synchronize()
                        out[...] = host.numpy().reshape(H, W, 3)
                    be.sync()

            e2e_step()
            barrier(); torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(n_e2e):
                e2e_step()
            be.sync(); torch.cuda.synchronize(); barrier()
            e2e_ms = (time.perf_counter() - t0) * 1e3
            e2e = {"ms": e2e_ms, "steps": n_e2e, "h2d": int(h2d * world), "d2h": int(W * H * 3)}

        # ---- reduce over ranks: max time, summed work
        t = torch.tensor([dev_ms, e2e["ms"] if e2e else 0.0, wall, render_ms], dtype=torch.float64, device=dev)
        work = torch.tensor([st["samples"], st["segments"], st["launches"], st["node_visits"], st["sphere_tests"]],
                            dtype=torch.float64, device=dev)
        own_tiles = be.tiles_rendered() if dq is not None else (sched.own_taken if sched else 0)
        ph = [phase_trace[k].elapsed_time(phase_trace[k + 1]) for k in range(4)] if len(phase_trace) == 5 else [0.0] * 4
        mine = torch.tensor([dev_ms, render_ms, float(own_tiles), float(sched.stolen if sched else 0)] + ph,
                            dtype=torch.float64, device=dev)
        per_rank = [torch.zeros_like(mine) for _ in range(world)] if world > 1 else [mine]
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(work, op=dist.ReduceOp.SUM)
            dist.all_gather(per_rank, mine)
        dev_ms, e2e_ms, wall, _ = (float(x) for x in t.tolist())
        samples, segments, launches, node_visits, sphere_tests = (float(x) for x in work.tolist())
        total = float(W) * H * spp * steps
        res = {
            "workload": wname, "scene": sc, "W": W, "H": H, "spp": spp, "steps": steps,
            "value": total / (dev_ms * 1e-3) / 1e9, "ms_per_step": dev_ms / steps, "wall_s": wall,
            "render_ms_rank0": render_ms / steps,            # this rank's render phase per step, one interval
            "samples": samples, "segments": segments, "launches": launches, "node_visits": node_visits,
            "sphere_tests": sphere_tests, "stats_rank0": st, "kernel": KERNEL_NAMES.get(st["last_kernel"], "?"),
            "frame_sha256": frame_hash,
            "ranks": [{"rank": i, "step_ms": float(v[0]) / steps, "render_phase_ms": float(v[1]) / steps,
                       "tiles_last_step": int(v[2]), "stolen_tiles_last_step": int(v[3]),
                       "phases_ms_last_step": {"clear+ack": float(v[4]), "render": float(v[5]), "resolve_to_rank0": float(v[6]),
                                               "ack": float(v[7])}} for i, v in enumerate(per_rank)],
        }
        if e2e:
            res["e2e"] = {"value": float(W) * H * spp * e2e["steps"] / (e2e_ms * 1e-3) / 1e9, "unit": UNIT,
                          "h2d_bytes_per_step": e2e["h2d"], "d2h_bytes_per_step": e2e["d2h"], "steps": e2e["steps"],
                          "call": "rt_render_frame (cuda_impl-shaped: host scene in, host RGB8 out)" if world == 1
                                  else "rt_scene_upload + rt_render_queue + rt_resolve_tiles_to(rank 0) + NCCL ack + D2H on rank 0" if dq is not None
                                  else "rt_scene_upload + tile-stolen rt_render_tiles + NCCL reduce + D2H on rank 0"}
        return res

    def roofline_for(res, probe, tprobe):
        """Roofline of the dominant kernel of `res` (rank 0's launches), live: device counters / CUDA-event time."""
        sc, W, H, steps = res["scene"], res["W"], res["H"], res["steps"]
        st = res["stats_rank0"]
        ms = res["render_ms_rank0"]
        kernel = res["kernel"]
        flop = algorithmic_flop(st["segments"], st["samples"], W * H * steps, sc.n) / steps
        fp32_model = {"achieved": flop / (ms * 1e-3) / 1e12, "peak": probe["tflops"], "unit": "TFLOP/s",
                      "frac": flop / (ms * 1e-3) / 1e12 / probe["tflops"],
                      "flop_model": "SURVEY 8(d): segments*(18*N+64) + samples*32 + pixels*8, segments/samples from device counters",
                      "peak_source": "FMA-chain probe measured in this run (rt_fp32_peak_probe); nominal 148 SM x 128 x 2 x 1.965 GHz = 74.45"}
        traffic = None
        try:
            prof = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json")))
            key = "k_render_tc" if "k_render_tc" in kernel else "k_render_bvh" if "bvh" in kernel else "k_render"
            traffic = prof.get(key, {}).get("dram_bytes_per_launch") if world == 1 else None
        except Exception:
            pass
        if "k_render_tc" in kernel:
            # the sphere tests run on the tensor pipe as a TF32x3 product: 64 flop per ray x padded sphere column (K = 32)
            cols = ((sc.n + 63) // 64) * 64
            tflop = st["segments"] * cols * 2.0 * 32 / steps
            tach = tflop / (ms * 1e-3) / 1e12
            peaks = measured_peaks()
            half_bf16 = 0.5 * float(peaks.get("bf16_tflops_sustained", 0) or 0) or None
            return {"bound": "tensor", "achieved": tach, "peak": tprobe["tflops"], "unit": "TFLOP/s",
                    "frac": tach / tprobe["tflops"], "traffic": traffic,
                    "kernel": f"{kernel} (persistent bounce loop; candidate filter = TF32x3 tcgen05 product, accumulators in TMEM; "
                              "IEEE reference test on the candidates)",
                    "peak_source": "tcgen05.mma kind::tf32 stream (M128 N256 K8, one CTA per SM) measured in this run "
                                   "(rt_tf32_peak_probe); MEASURED_PEAKS.json has no TF32 entry -- half of its sustained bf16 figure "
                                   f"is {half_bf16}",
                    "frac_of_half_bf16_sustained": (tach / half_bf16) if half_bf16 else None,
                    "flop_model": f"segments * {cols} sphere columns * K=32 * 2 (EXECUTED split-TF32 flop; about a third of it is "
                                  "algorithmic: 488/512 columns x 11 of 32 K slots)",
                    "note": "the kernel is bound by issue slots (sign-mask epilogue, shading, candidate tests), not by the tensor pipe; "
                            "fp32_algorithmic is SURVEY 8(d)'s FP32 model (a fraction above 1 = more sphere tests per second than "
                            "the FMA pipe could execute)",
                    "fp32_algorithmic": fp32_model, "ms_per_launch": ms, "rank": 0}
        if st["node_visits"] > 0:
            # LBVH.  HBM is NOT what bounds this kernel, and the line says so with the honest figure: the bytes that MUST come
            # from DRAM once per launch (the tree: (n-1) nodes x 64 B, the sphere records: n x 28 B, the accumulators: W*H*32 B)
            # against the measured HBM copy bandwidth -- a fraction of a per cent.  The SMs REQUEST node_visits x 64 B + leaf
            # tests x 16 B (SURVEY 8(d)'s model, 7-8 TB/s here), served by L1/L2 (ncu: L2 hit rate 87 %, DRAM ~30 GB/s); the limiter
            # is issue slots (ncu: 71 % issue-active, 16 of 32 lanes per instruction, L2 hit rate 93 %; profiles/r02_ncu_k_render_bvh.json).
            hbm = float(measured_peaks().get("hbm_gbs", 0) or 0)
            peak_src = "MEASURED_PEAKS.json hbm_gbs" if hbm else "fallback 6550 GB/s (B200_PROFILING.md)"
            hbm = hbm or 6550.0
            compulsory = (sc.n - 1) * 64.0 + sc.n * 28.0 + W * H * 32.0
            ach = compulsory / (ms * 1e-3) / 1e9
            requested = (st["node_visits"] * 64.0 + st["sphere_tests"] * 16.0) / steps
            cap = {}
            try:
                cap = json.load(open(os.path.join(ROOT, "profiles", "ncu_summary.json"))).get("k_render_bvh", {})
            except Exception:
                pass
            return {"bound": "hbm", "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm,
                    "traffic": None,       # the committed ncu capture is of a smaller launch (see ncu_capture): not this launch's bytes
                    "ncu_capture": {"dram_bytes_of_the_captured_launch": cap.get("dram_bytes_per_launch"), "what": cap.get("source"),
                                    "issue_active_pct": cap.get("issue_active_pct"), "lanes_per_instruction": cap.get("lanes_per_instruction")},
                    "issue_slots": {"frac": (cap.get("issue_active_pct") or 0) / 100.0,
                                    "note": "the bound that matters: share of cycles in which a scheduler issues (ncu); 16 of 32 lanes per instruction"},
                    "kernel": f"{kernel} (persistent bounce loop, resumable LBVH traversal, stacks in shared memory; a warp whose work pool is dry "
                              "finishes its last queries cooperatively, one ray on 32 lanes)",
                    "peak_source": peak_src,
                    "byte_model": "compulsory DRAM bytes per launch: (n-1) nodes * 64 B + n sphere records * 28 B + W*H*32 B of accumulators",
                    "note": "not HBM bound (frac << 1 on purpose): the tree is L2-resident; the kernel is issue bound -- 71 % issue-active at "
                            "16 of 32 lanes (ncu, profiles/r02_ncu_k_render_bvh.json)",
                    "l2_request_model": {"bytes_per_launch": requested, "GBps": requested / (ms * 1e-3) / 1e9,
                                         "model": "node_visits * 64 B + leaf_tests * 16 B (device counters): what the SMs request, served by L1/L2"},
                    "node_visits_per_segment": st["node_visits"] / max(1, st["segments"]),
                    "leaf_tests_per_segment": st["sphere_tests"] / max(1, st["segments"]),
                    "ms_per_launch": ms, "rank": 0}
        return dict(fp32_model, bound="fp32", traffic=traffic, kernel=f"{kernel} (persistent bounce loop: sphere tests + shading)",
                    ms_per_launch=ms, rank=0)

    def brief(res, probe, tprobe, extra=None):
        d = {"workload": res["workload"], "n_gpus": world, "steps": res["steps"], "value": res["value"], "unit": UNIT,
             "ms_per_step": res["ms_per_step"], "kernel": res["kernel"],
             "segments_per_sample": res["segments"] / max(1.0, res["samples"])}
        if "e2e" in res:
            d["e2e"] = res["e2e"]
        if res.get("frame_sha256"):
            d["frame_sha256"] = res["frame_sha256"]
        if rank == 0:
            d["roofline"] = roofline_for(res, probe, tprobe)
        if extra:
            d.update(extra)
        return d

    # ------------------------------------------------------------------ headline
    wname, _ = workload_for(world, args.workload)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    head = measure(wname, args.steps, args.warmup, want_e2e=True, want_hash=True,
                   e2e_steps=(max(1, min(args.steps, 3)) if world > 1 else args.steps))
    clk = clocks.stop() if rank == 0 else None
    probe = be.fp32_peak_probe()
    tprobe = be.tf32_peak_probe()

    line = None
    if rank == 0:
        sc = head["scene"]
        accel = {"rtf::k_render_tc": "brute force: tensor-core candidate filter (sphere matrix in shared memory via TMA, "
                                            "accumulators in TMEM) + IEEE reference test on the candidates",
                 "rtb::k_render_bvh": "GPU-built LBVH (Morton + radix sort + Karras + refit), conservative traversal, IEEE reference "
                                      "test at the leaves; stragglers finished by whole warps (DESIGN 6)"}.get(head["kernel"], head["kernel"])
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(wname, WORKLOADS[wname][0], sc.n, head["W"], head["H"], head["spp"]),
            "impl_detail": {"accel": accel, "kernel": head["kernel"],
                            "parallelism": ("1 GPU" if world == 1 else
                                            f"x{world}: device-side tile queue (cursors in rank 0's memory over NVLink, in-kernel stealing), "
                                            "one persistent launch per GPU and frame, resolve stores into rank 0's frame (CUDA IPC), NCCL all-reduce ack"
                                            if dq is not None else f"tile-sharded x{world}, host-scheduled stealing, NCCL reduce to rank 0"),
                            "l2": "flushed between timed steps (256 MiB write)",
                            "segments_per_sample": head["segments"] / max(1.0, head["samples"]),
                            "closest_hit": "bit-identical to the oracle (index, t, normal) in every kernel; shading of the timed "
                                           "kernel samples the unit ball by inversion (statistical parity); parity_build = same "
                                           "kernel, reference sampler, bit-identical renders"},
            "clocks": clk,
            "e2e": head["e2e"],
            "gpu_launches": int(head["launches"]),
            "roofline": roofline_for(head, probe, tprobe),
            "frame_sha256": head["frame_sha256"],
            "wall_s": head["wall_s"],
        }
        if world > 1:
            line["ranks"] = head["ranks"]

    # ------------------------------------------------------------------ everything else BASELINE names, same run
    if not args.no_secondary:
        sec = []
        few = max(1, min(2, args.steps))
        # the same tensor-core kernel with parity shading: the price of full bit-identity with the oracle
        pb = measure(wname, few, 1, flags=FLAG_EXACT)
        if rank == 0:
            line["parity_build"] = {"value": pb["value"], "unit": UNIT, "ms_per_step": pb["ms_per_step"], "steps": few,
                                    "kernel": pb["kernel"], "workload": wname,
                                    "note": "RT_FLAG_EXACT: rejection sampler + IEEE normalisations; accumulators bit-identical to the oracle"}
        if world == 1:
            sec.append(brief(measure("simple-720p-64spp", max(3, few), 2, want_e2e=True), probe, tprobe))
            sec.append(brief(measure("huge-720p-500spp", few, 1, flags=FLAG_FILTER_FMA), probe, tprobe,
                             {"flags": "RT_FLAG_FILTER_FMA", "note": "the FP32-pipe kernel: the number behind north_star's '>= 50 % of FP32 peak'"}))
        sec.append(brief(measure("synthetic1m-1080p-256spp", few, 1, want_e2e=(world == 1), want_hash=True), probe, tprobe))
        if world == 1:
            anchor = measure("huge-4k-1024spp", few, 1, want_hash=True)
            if rank == 0:
                line["scaling_anchor"] = brief(anchor, probe, tprobe, {"note": "the N > 1 workload on ONE GPU: strong-scaling "
                                               "efficiency = value(N) / (N * this); frame_sha256 must match the N-GPU lines"})
            if rank == 0:
                sec.append(legacy_comparator())
        if rank == 0:
            line["secondary"] = sec

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            orc = load_cpu_oracle()
            threads = cpu_threads()
            smp = CpuSample(orc, head["scene"], head["W"], head["H"], threads)
            desc = smp.plan(12.0)
            n, s, busy = smp.run()
            line["cpu_baseline"] = {"value": n / s / 1e9, "unit": UNIT, "cores": threads, "threads_busy": round(busy, 2),
                                    "kind": orc.kind, "sample": desc}
        except Exception as exc:  # the checker is optional for the number, never for the product
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": cpu_threads(), "kind": "port", "sample": f"failed: {exc}"}

    if dq is not None:
        barrier()
        dq.close()
    if queue is not None:
        barrier()
        queue.close()
    be.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


def legacy_comparator():
    """the reference's own CUDA path (src/CUDA/cuda_impl.cu compiled for sm_100a, oracle/_ref/ref_cuda_legacy)
    next to this backend on ITS workload shape: 5 spheres, 48 spp, depth 32, 1280x720.  Informational (different RNG, sky
    and lens conventions: SURVEY App. E): only speed is compared."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_cuda_legacy")
    out = {"workload": "legacy-comparator simple 1280x720 48spp depth32", "unit": UNIT}
    if not os.path.exists(exe):
        out["unavailable"] = "oracle/_ref/ref_cuda_legacy not built (needs /root/reference at build time)"
        return out
    try:
        p = subprocess.run([exe, "--width", "1280", "--height", "720", "--reps", "5"], capture_output=True, text=True, timeout=300)
        rec = [l for l in p.stdout.splitlines() if l.startswith("{")]
        ref = json.loads(rec[-1])
        from raytracinginoneweekend_b200.backend import Backend
        sc = scenes.get_scene("simple", 1280, 720)
        be = Backend(0)
        best = 1e30
        for r in range(6):
            t0 = time.perf_counter()
            be.render_frame(sc, 1280, 720, 48, max_depth=32, seed=1)
            if r:
                best = min(best, (time.perf_counter() - t0) * 1e3)
        be.close()
        n = 1280 * 720 * 48
        out.update({"reference_cuda": {"best_ms": ref.get("best_ms"), "value": n / ref["best_ms"] / 1e6 if ref.get("best_ms") else None},
                    "this_backend_e2e": {"best_ms": best, "value": n / best / 1e6},
                    "speedup_end_to_end": ref["best_ms"] / best if ref.get("best_ms") else None})
    except Exception as exc:
        out["unavailable"] = f"failed: {exc}"
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["b200", "reference"], default="b200")
    ap.add_argument("--workload", default="auto", choices=["auto"] + list(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override spp (debug only; invalidates the headline number)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="headline workload only (profiling runs)")
    ap.add_argument("--mgpu", choices=["device", "host"], default="device",
                    help="N > 1: device-side tile queue + peer-store gather (default) | host-scheduled chunks + NCCL reduce (round 1)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        print(f"bench.py: warning: --warmup {args.warmup} < 3 breaks the timing rules", file=sys.stderr)
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
