"""Per-source-line instruction / stall-sample shares of one kernel from an ncu report (needs -lineinfo + --import-source on).
usage: ncu_lines.py <cubin> <mangled kernel name> <report.ncu-rep> [top]"""
import re, csv, subprocess, collections, sys
cub, fun, rep = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
dis = subprocess.run(['nvdisasm', '-g', '-c', cub], capture_output=True, text=True).stdout.splitlines()
start = [i for i, l in enumerate(dis) if l.startswith('\t.section\t.text.' + fun)][0]
cur = None; off2line = {}
for l in dis[start + 1:]:
    if l.startswith('//---------------------'): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: off2line[int(m.group(1), 16)] = cur
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'Address'][0]
hdr = rows[hi]; data = rows[hi + 1:]
ia = hdr.index('Address'); iex = hdr.index('Instructions Executed'); ith = hdr.index('Thread Instructions Executed'); ismp = hdr.index('# Samples')
first = int(data[0][ia], 16)
agg = collections.defaultdict(lambda: [0, 0, 0]); tots = 0; tote = 0
for r in data:
    off = int(r[ia], 16) - first; ln = off2line.get(off)
    key = (ln[0], ln[1]) if ln else ('?', 0)
    a = agg[key]; a[0] += int(r[iex]); a[1] += int(r[ith]); a[2] += int(r[ismp]); tots += int(r[ismp]); tote += int(r[iex])
byfile = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    b = byfile[f]; b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
for f, b in byfile.items(): print(f, f"inst {b[0]/tote*100:.1f}% samples {b[2]/tots*100:.1f}% lanes {b[1]/max(b[0],1):.1f}")
print('--- top lines by stall samples')
for (f, l), a in sorted(agg.items(), key=lambda x: -x[1][2])[:top]:
    print(f"{f}:{l:4d} inst {a[0]/tote*100:5.2f}% samples {a[2]/tots*100:5.2f}% lanes {a[1]/max(a[0],1):5.1f}")
print('--- instruction share by 20-line block')
blk = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    b = blk[(f, l // 20 * 20)]; b[0] += a[0]; b[1] += a[1]; b[2] += a[2]
for (f, l), a in sorted(blk.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{f}:{l:4d}-{l+19:4d} inst {a[0]/tote*100:5.2f}% samples {a[2]/tots*100:5.2f}% lanes {a[1]/max(a[0],1):5.1f}")
