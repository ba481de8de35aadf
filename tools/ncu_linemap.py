import re, csv, sys, subprocess, collections
cub = sys.argv[1]; fun = sys.argv[2]; rep = sys.argv[3]
dis = subprocess.run(['nvdisasm','-g','-c',cub],capture_output=True,text=True).stdout.splitlines()
# find function section
start = [i for i,l in enumerate(dis) if l.startswith('\t.section\t.text.'+fun)][0]
cur_line = None; inl = None; off2line = {}
for l in dis[start+1:]:
    if l.startswith('//---------------------') : break
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
    if m:
        cur_line = (m.group(1).split('/')[-1], int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        off2line[int(m.group(1),16)] = cur_line
# ncu sass csv
raw = subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hi = [i for i,r in enumerate(rows) if r and r[0]=='Address'][0]
hdr = rows[hi]; data = rows[hi+1:]
ia=hdr.index('Address'); isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); ith=hdr.index('Thread Instructions Executed'); ismp=hdr.index('# Samples')
first = int(data[0][ia],16)
def cost(op):
    base = op.split('.')[0]
    if base in ('FFMA','FMUL','FADD','IMAD','HFMA2','FFMA2'): return 1
    if base in ('LDS',): return 4 if '128' in op else (2 if '64' in op else 1)
    if base in ('MUFU','F2I','I2F','I2FP','F2F','FRND','POPC','FLO','BREV'): return 4
    if base in ('DFMA','DMUL','DADD'): return 2
    return 2
agg = collections.defaultdict(lambda:[0,0,0,0.0])
tot=0; totc=0
for r in data:
    off = int(r[ia],16)-first
    ln = off2line.get(off)
    src = r[isrc].strip()
    op = src.split()[1] if src.startswith('@') else src.split()[0]
    ex = int(r[iex]); th=int(r[ith]); sm=int(r[ismp])
    a = agg[ln]; a[0]+=ex; a[1]+=th; a[2]+=sm; a[3]+=ex*cost(op)
    tot+=ex; totc+=ex*cost(op)
print("total warp-instr", tot, "est dispatch cycles", totc)
srcs = {}
for ln,a in sorted(agg.items(), key=lambda x:-x[1][3])[:45]:
    f,l = ln if ln else ('?',0)
    if f not in srcs:
        try: srcs[f]=open('/root/repo/raytracinginoneweekend_b200/csrc/'+f).read().splitlines()
        except Exception: srcs[f]=[]
    text = srcs[f][l-1].strip()[:90] if 0<l<=len(srcs[f]) else ''
    print(f"{f}:{l:4d} inst={a[0]/tot*100:5.2f}% cyc={a[3]/totc*100:5.2f}% thr/inst={a[1]/max(a[0],1):5.1f} samples={a[2]:7d} | {text}")
