"""tools/pathstat.py: run the RT_BVH_PATHSTAT measurement build (tools/build_variant.sh pathstat -DRT_BVH_PATHSTAT=20000) on the
1M-sphere frame and summarise the heavy tail of path cost: how many paths visit more than the threshold of nodes, how
long the longest ones run, and when they end relative to the kernel (GPU box).
usage: RTIOW_LIB=build/variants/pathstat/librtiow_b200.so python tools/pathstat.py [spp] > log"""
import ctypes, os, re, subprocess, sys
HERE = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, HERE)
    rt = ctypes.CDLL("libcudart.so.12")
    assert rt.cudaDeviceSetLimit(1, ctypes.c_size_t(256 << 20)) == 0      # cudaLimitPrintfFifoSize
    from raytracinginoneweekend_b200 import scenes
    from raytracinginoneweekend_b200.backend import Backend
    spp = int(sys.argv[2])
    W, H = 1920, 1080
    sc = scenes.get_scene("synthetic1m", W, H)
    be = Backend(0); be.upload_scene(sc); be.framebuffer(W, H)
    be.clear(); be.reset_stats(); be.render(spp, 50, seed=1); be.sync()
    st = be.stats()
    print("KERNEL ms %.3f samples %d node_visits %d segments %d" % (st["render_ms"], st["samples"], st["node_visits"], st["segments"]), flush=True)
    sys.exit(0)
spp = sys.argv[1] if len(sys.argv) > 1 else "256"
out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", spp], capture_output=True, text=True)
start, paths, kern = None, [], None
for l in out.stdout.splitlines():
    if l.startswith("PATHSTAT start"):
        start = int(l.split()[2])
    elif l.startswith("PATHSTAT path"):
        f = l.split()
        paths.append(dict(pix=int(f[3]), smp=int(f[5]), nodes=int(f[7]), bounces=int(f[9]), us=int(f[11]), end=int(f[13])))
    elif l.startswith("KERNEL"):
        kern = l
print(kern); print("stderr tail:", out.stderr[-300:])
print("paths above the threshold:", len(paths), " kernel start (us):", start)
if paths:
    import statistics
    tot = sum(p["nodes"] for p in paths)
    print("node visits in those paths:", tot)
    for key in ("us", "nodes"):
        print("top 15 by", key)
        for p in sorted(paths, key=lambda p: -p[key])[:15]:
            print("   pix (%4d,%4d) smp %3d nodes %8d bounces %2d  %8.2f ms  ends at %8.2f ms" % (p["pix"] % 1920, p["pix"] // 1920, p["smp"], p["nodes"], p["bounces"], p["us"] / 1e3, (p["end"] - start) / 1e3))
    edges = [1, 2, 5, 10, 20, 50, 100, 200, 500]
    print("duration histogram (ms): ", {e: sum(1 for p in paths if p["us"] / 1e3 >= e) for e in edges})
    rows = {}
    for p in paths:
        rows.setdefault(p["pix"] // 1920 // 64, []).append(p["us"])
    print("per 64-row band of the image: count, max ms:", {k: (len(v), max(v) / 1e3) for k, v in sorted(rows.items())})
