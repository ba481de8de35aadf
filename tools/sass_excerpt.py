"""profiles/rNN_k_render_tc.sass: what the shipped tensor kernel is made of, from the built library (no GPU needed).
usage: python tools/sass_excerpt.py [lib.so] > profiles/r02_k_render_tc.sass
Mnemonic histogram of rtf::k_render_tc<true> (the timed instantiation) and the instruction lines that prove the
tcgen05 / TMEM / TMA-bulk path: UTCHMMA (tcgen05.mma), UTCBAR (tcgen05.commit), LDTM / STTM (tcgen05.ld / st),
UBLKCP (cp.async.bulk), SYNCS (mbarrier), I2IP (cvt.pack.sat sign masks)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "raytracinginoneweekend_b200", "librtiow_b200.so")
fun = "_ZN3rtf11k_render_tcILb1EEEv12RtRenderArgs"   # the tile-queue instantiation: what rt_render and the multi-GPU frames run
out = subprocess.run(["cuobjdump", "-sass", "-fun", fun, lib], capture_output=True, text=True).stdout
lines = [l for l in out.splitlines() if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l)]
ins = []
for l in lines:
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+(.*?);", l)
    if m:
        ins.append((m.group(1), re.sub(r"\s+", " ", m.group(2)).strip()))
hist = collections.Counter()
for _, t in ins:
    t = re.sub(r"^@!?U?P\d+\s+", "", t)
    hist[t.split()[0].split(".")[0] + ("." + t.split()[0].split(".")[1] if t.split()[0].startswith(("LDTM", "STTM", "I2IP", "MUFU", "UTC", "SYNCS", "UBLKCP")) and "." in t.split()[0] else "")] += 1
print(f"# {fun}  ({os.path.relpath(lib, ROOT)})")
print(f"# {len(ins)} SASS instructions, {len(ins) * 16} bytes of code")
print("# mnemonic histogram (top 40):")
for k, v in hist.most_common(40):
    print(f"#   {k:16s} {v}")
keys = ("UTCHMMA", "UTCBAR", "LDTM", "STTM", "UBLKCP", "SYNCS", "I2IP", "UTCATOM", "ELECT", "USETMAXREG", "REDUX", "RED.E")
print("# tcgen05 / TMEM / TMA / mbarrier instructions:")
seen = collections.Counter()
for a, t in ins:
    for k in keys:
        if k in t:
            seen[k] += 1
            if seen[k] <= 6:
                print(f"/*{a}*/  {t}")
print("# counts:", dict(seen))
