import re, csv, sys, subprocess, collections
cub, fun, rep, src = sys.argv[1:5]
dis = subprocess.run(['nvdisasm','-g','-c',cub],capture_output=True,text=True).stdout.splitlines()
start = [i for i,l in enumerate(dis) if l.startswith('\t.section\t.text.'+fun)][0]
cur=None; off2line={}
for l in dis[start+1:]:
    if l.startswith('//---------------------'): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur=(m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m: off2line[int(m.group(1),16)] = cur
raw = subprocess.run(['ncu','-i',rep,'--page','source','--csv'],capture_output=True,text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hi = [i for i,r in enumerate(rows) if r and r[0]=='Address'][0]
hdr = rows[hi]; data = rows[hi+1:]
ia=hdr.index('Address'); isrc=hdr.index('Source'); iex=hdr.index('Instructions Executed'); ith=hdr.index('Thread Instructions Executed'); ismp=hdr.index('# Samples')
first=int(data[0][ia],16)
lines=open(src).read().splitlines()
# phase boundaries by function names in source
def find(pat):
    for i,l in enumerate(lines):
        if pat in l: return i+1
    return None
marks = sorted([(find('struct FastRay'),'fast_setup'),(find('void sphere_candidate_fast'),'cand_test'),(find('void closest_hit_brute_fast'),'loop_setup'),
 (find('for (uint32_t chunk = 0; chunk < nfilt'),'loop'),(find('// one candidate per lane per iteration: the warp iterates max-over-lanes of the candidate COUNT\n') or 0,'x'),
 (find('struct WarpPool'),'pool'),(find('void shade_slot'),'shade_A'),(find('// ---- regeneration'),'regen'),(find('// ---- one Philox loop'),'rng_loop'),(find('// ---- next ray'),'next_ray'),
 (find('void __launch_bounds__(RT_THREADS, 3) k_render'),'kernel_main'),(find('uint4 philox4x32_10'),'philox'),(find('float u01('),'u01_s11'),(find('float dot3('),'vecmath'),
 (find('float sphere_disc('),'sphere_disc'),(find('void sphere_candidate('),'cand_exact'),(1,'helpers')])
marks=[m for m in marks if m[0]]
# locate fast post-phase start (second occurrence of the comment)
occ=[i+1 for i,l in enumerate(lines) if 'one candidate per lane per iteration' in l]
if len(occ)>=2: marks.append((occ[1],'post'))
marks=sorted(marks)
def phase(ln):
    p='?'
    for m,name in marks:
        if ln>=m: p=name
    return p
def cost(op):
    base=op.split('.')[0]
    if base in ('FFMA','FMUL','FADD','IMAD','HFMA2','FFMA2'): return 1
    if base=='LDS': return 4 if '128' in op else (2 if '64' in op else 1)
    if base in ('MUFU','F2I','I2F','I2FP','F2F','FRND','POPC','FLO','BREV'): return 4
    return 2
agg=collections.defaultdict(lambda:[0,0,0,0.0]); tot=0; totc=0
for r in data:
    off=int(r[ia],16)-first; ln=off2line.get(off); s=r[isrc].strip()
    op=s.split()[1] if s.startswith('@') else s.split()[0]
    ex=int(r[iex]); th=int(r[ith]); sm=int(r[ismp])
    p=phase(ln[1]) if ln and ln[0].startswith('rt_device') else 'other'
    a=agg[p]; a[0]+=ex; a[1]+=th; a[2]+=sm; a[3]+=ex*cost(op); tot+=ex; totc+=ex*cost(op)
tots=sum(a[2] for a in agg.values())
for p,a in sorted(agg.items(), key=lambda x:-x[1][3]):
    print(f"{p:12s} inst={a[0]/tot*100:6.2f}% est_cyc={a[3]/totc*100:6.2f}% lanes/inst={a[1]/max(a[0],1):5.1f} stall_samples={a[2]/tots*100:6.2f}%")
