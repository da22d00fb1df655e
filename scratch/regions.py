#!/usr/bin/env python
"""usage: regions.py src.csv dis.txt kernel-substring name:lo-hi ...  -> instruction / stall-sample share by source region"""
import csv,re,collections,sys
src,dis,K=sys.argv[1:4]
regs=[]
for spec in sys.argv[4:]:
    n,r=spec.split(':'); lo,hi=map(int,r.split('-')); regs.append((n,lo,hi))
addr2line={}; cur=None; inside=False
for ln in open(dis):
    if ln.startswith('//---') and '.text.' in ln: inside=K in ln
    if not inside: continue
    m=re.search(r'//## File "([^"]+)", line (\d+)',ln)
    if m: cur=(m.group(1).split('/')[-1],int(m.group(2))); continue
    m=re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);',ln)
    if m: addr2line[int(m.group(1),16)]=(cur,m.group(2))
rows=list(csv.reader(open(src)))
h=next(i for i,r in enumerate(rows) if 'Instructions Executed' in r); hdr=rows[h]
ia=hdr.index('Address'); ie=hdr.index('Instructions Executed'); isamp=hdr.index('# Samples'); ith=hdr.index('Thread Instructions Executed')
names=['stall_no_inst','stall_long_sb','stall_wait','stall_short_sb','stall_branch_resolving','stall_math','stall_not_selected','stall_lg','stall_mio','stall_dispatch']
sc=[hdr.index(n) for n in names]
def region(k):
    if not k: return '?'
    f,l=k
    if f!='pair_kernel.cuh' and f!='ring_kernels.cuh': return f
    for name,lo,hi in regs:
        if lo<=l<=hi: return name
    return 'other'
agg=collections.defaultdict(lambda:[0]*(3+len(sc))); base=None
for r in rows[h+1:]:
    try: a=int(r[ia],16)
    except: continue
    if base is None: base=a
    k=addr2line.get(a-base,(None,''))[0]
    v=agg[region(k)]
    for j,c in enumerate([ie,isamp,ith]+sc): v[j]+=int(r[c].replace(',',''))
T=[sum(v[j] for v in agg.values()) for j in range(3+len(sc))]
print("total warp instr %.1fM, samples %d"%(T[0]/1e6,T[1]))
print("region                 inst%  Minst  act  smp% | "+" ".join(n[6:14].rjust(8) for n in names))
for k,v in sorted(agg.items(), key=lambda kv:-kv[1][0]):
    print(f"{k:22s} {100*v[0]/T[0]:5.1f} {v[0]/1e6:6.2f} {v[2]/max(v[0],1):4.1f} {100*v[1]/T[1]:5.1f} | "+" ".join(f"{100*v[3+j]/T[1]:8.1f}" for j in range(len(sc))))
print("all".ljust(22), " "*24, "| "+" ".join(f"{100*T[3+j]/T[1]:8.1f}" for j in range(len(sc))))
