"""torchrun diagnostic: one frame of a workload through the device-side tile queue; prints per rank when (relative to its
first acquisition) it acquired how many own / stolen tiles, and its kernel time.
usage: torchrun --nproc-per-node N tools/queue_trace.py [workload]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from raytracinginoneweekend_b200 import multigpu, scenes
from raytracinginoneweekend_b200.backend import Backend
import bench
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
wname = sys.argv[1] if len(sys.argv) > 1 else "synthetic1m-1080p-256spp"
scene_name, W, H, spp = bench.WORKLOADS[wname]
sc = scenes.get_scene(scene_name, W, H)
be = Backend(local); stream = torch.cuda.Stream(device=local); be.set_stream(stream.cuda_stream)
be.upload_scene(sc); be.framebuffer(W, H)
tiles = multigpu.interleave_tiles(multigpu.make_tiles(W, H, 128, 64), world)
dq = multigpu.DeviceQueueFrame(be, rank, world, dist, torch, stream); dq.map_frame()
for it in range(3):
    be.reset_stats(); dq.render_frame(tiles, 128, 64, spp, 50, 1); be.sync()
st = be.stats(); t, w = be.queue_trace()
lo, hi = rank * len(tiles) // world, (rank + 1) * len(tiles) // world
w = (w.astype(np.int64) - int(w[0])) * 1.024e-3 if len(w) else w
own = (t >= lo) & (t < hi)
rec = {"rank": rank, "kernel_ms": st["last_render_ms"], "tiles": int(len(t)), "own": int(own.sum()), "stolen": int((~own).sum()),
       "last_own_acq_ms": float(w[own][-1]) if own.any() else None, "first_steal_ms": float(w[~own][0]) if (~own).any() else None,
       "last_acq_ms": float(w[-1]) if len(w) else None, "acq_ms_every_8th": [round(float(x), 1) for x in w[::8]]}
out = [None] * world
dist.all_gather_object(out, rec)
if rank == 0:
    for r in out: print(json.dumps(r))
dist.barrier(); dq.close(); be.close(); dist.destroy_process_group()
