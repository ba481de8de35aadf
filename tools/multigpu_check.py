"""Launched with torchrun: renders a frame on WORLD_SIZE GPUs and checks it is bit-identical to the same frame rendered
by rank 0 alone.  Prints one JSON line on rank 0.
usage: multigpu_check.py W H spp [host|device]   host   = host-scheduled chunks + NCCL reduce (round 1)
                                                 device = device-side tile queue + peer-store gather over CUDA IPC"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from raytracinginoneweekend_b200 import multigpu, scenes
from raytracinginoneweekend_b200.backend import Backend, TileQueue

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
W, H, spp = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (1280, 720, 16)))
mode = sys.argv[4] if len(sys.argv) > 4 else "host"
sc = scenes.huge_scene(W, H)
be = Backend(local)
stream = torch.cuda.Stream(device=local)
be.set_stream(stream.cuda_stream)
be.upload_scene(sc); be.framebuffer(W, H)
qname = f"/rtiow_check_{os.environ.get('MASTER_PORT', '0')}"
q = TileQueue(qname, create=True) if rank == 0 else None
dist.barrier()
if rank != 0:
    q = TileQueue(qname, create=False)
tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, 128, 64), world)
sched = multigpu.StealingScheduler(q, rank, world, len(tiles))
rend = multigpu.CudaTileRenderer(be, 128, 64, spp, 50, 1)
times = []
dq = None
if mode == "device":
    dq = multigpu.DeviceQueueFrame(be, rank, world, dist, torch, stream)
    dq.map_frame()
for it in range(3 if mode == "host" else 0):
    be.clear()
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    done = multigpu.render_distributed(rend, sched, tiles, dist.barrier)
    be.resolve_device()
    frame = multigpu.device_rgb8_tensor(be)
    with torch.cuda.stream(stream):
        dist.reduce(frame, dst=0, op=dist.ReduceOp.SUM)
    stream.synchronize(); torch.cuda.synchronize(); dist.barrier()
    times.append(time.perf_counter() - t0)
for it in range(3 if mode == "device" else 0):
    torch.cuda.synchronize(); dist.barrier(); t0 = time.perf_counter()
    dq.render_frame(tiles, 128, 64, spp, 50, 1)
    stream.synchronize(); torch.cuda.synchronize(); dist.barrier()
    times.append(time.perf_counter() - t0)
    frame = multigpu.device_rgb8_tensor(be)
    done = be.tiles_rendered()
counts = torch.tensor([done, sched.stolen if mode == "host" else 0], device=f"cuda:{local}")
all_counts = [torch.zeros_like(counts) for _ in range(world)]
dist.all_gather(all_counts, counts)
if rank == 0:
    multi = frame.cpu().numpy().reshape(H, W, 3).copy()
    # the single-GPU frame through the same kernel instantiation (the shading is compiled with FMA contraction, so the
    # tile-list and the tile-queue instantiations may differ in the last bit of a radiance sample)
    be.clear()
    if mode == "device":
        be.render(spp, 50, seed=1)
    else:
        be.render_tiles(tiles, 128, 64, spp, 50, seed=1)
    single = be.readback_rgb8()
    print(json.dumps({"mode": mode, "world": world, "W": W, "H": H, "spp": spp, "identical_to_single_gpu": bool(np.array_equal(multi, single)),
                      "tiles_per_rank": [int(c[0]) for c in all_counts], "stolen_per_rank": [int(c[1]) for c in all_counts],
                      "seconds": times, "Gsamples_per_s": W * H * spp / min(times) / 1e9}), flush=True)
dist.barrier()
if dq is not None:
    dq.close()
q.close(); be.close()
dist.destroy_process_group()
